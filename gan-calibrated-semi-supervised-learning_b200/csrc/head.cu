// PatchGAN head: Conv2d(512,1,k4,s1,p1,bias=False) (cgan/models.py:252).  N=1 cannot fill an MMA tile, so these are
// bandwidth kernels: a GEMV per output pixel (fprop), a 16-tap scatter of a scalar map (dgrad) and a channel-wise
// reduction (wgrad).  a4 / g_a4 are NHWC bf16 (N,h,h,C); the score map is fp32 (N,1,h-1,h-1).
#include "common.cuh"

using namespace b2c;

// Stage w[c][16 taps] as w_s[tap][C] with the 8-float groups of a row XOR-swizzled by the tap (the plain transpose is a
// 16-way bank conflict); readers use hw_off(t, c) for an aligned group of 4 or 8 channels.
__device__ __forceinline__ int hw_off(int t, int c, int C) { return t * C + ((((c >> 3) ^ t) << 3) | (c & 7)); }
__device__ __forceinline__ void head_stage_w(const float* __restrict__ w, float* w_s, int C) {
  for (int i = threadIdx.x; i < 4 * C; i += blockDim.x) {  // one float4 = 4 consecutive taps of a channel
    const int c = i / 4, t0 = (i % 4) * 4;
    const float4 v = __ldg(reinterpret_cast<const float4*>(w) + i);
    w_s[hw_off(t0, c, C)] = v.x;
    w_s[hw_off(t0 + 1, c, C)] = v.y;
    w_s[hw_off(t0 + 2, c, C)] = v.z;
    w_s[hw_off(t0 + 3, c, C)] = v.w;
  }
  __syncthreads();
}

// Each input pixel is read ONCE: a warp owns HP consecutive pixels, computes their 16 per-tap dot products
// P[t] = a[n,iy,ix,:] . w[:,t] (one pass over the shared-memory weights serves all HP pixels) and scatters them to the
// (up to 16) score pixels they belong to, out[n,iy-ky+1,ix-kx+1] += P[ky*4+kx].  `out` must be zeroed by the caller
// (the C entry point does it).
constexpr int HP = 4;
__global__ void __launch_bounds__(256) head_fprop_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ w,
                                                         float* __restrict__ out, int N, int h, int C) {
  extern __shared__ float w_s[];  // [16 taps][C], swizzled
  head_stage_w(w, w_s, C);
  const int ho = h - 1;
  const long long npix = (long long)N * h * h;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  for (long long p0 = ((long long)blockIdx.x * wpb + warp) * HP; p0 < npix; p0 += (long long)gridDim.x * wpb * HP) {
    float acc[HP][16];
#pragma unroll
    for (int q = 0; q < HP; ++q)
#pragma unroll
      for (int t = 0; t < 16; ++t) acc[q][t] = 0.f;
    for (int c = lane * 8; c < C; c += 256) {
      float f[HP][8];
#pragma unroll
      for (int q = 0; q < HP; ++q) {
        const uint4 u = p0 + q < npix ? __ldg(reinterpret_cast<const uint4*>(a + (p0 + q) * C + c)) : make_uint4(0, 0, 0, 0);
        const float2 f0 = unpack_bf16(u.x), f1 = unpack_bf16(u.y), f2 = unpack_bf16(u.z), f3 = unpack_bf16(u.w);
        f[q][0] = f0.x; f[q][1] = f0.y; f[q][2] = f1.x; f[q][3] = f1.y;
        f[q][4] = f2.x; f[q][5] = f2.y; f[q][6] = f3.x; f[q][7] = f3.y;
      }
#pragma unroll
      for (int t = 0; t < 16; ++t) {
        const float* wp = w_s + hw_off(t, c, C);
        const float4 w0 = *reinterpret_cast<const float4*>(wp);
        const float4 w1 = *reinterpret_cast<const float4*>(wp + 4);
#pragma unroll
        for (int q = 0; q < HP; ++q)
          acc[q][t] += f[q][0] * w0.x + f[q][1] * w0.y + f[q][2] * w0.z + f[q][3] * w0.w + f[q][4] * w1.x + f[q][5] * w1.y +
                       f[q][6] * w1.z + f[q][7] * w1.w;
      }
    }
#pragma unroll
    for (int q = 0; q < HP; ++q) {
      // 16 values x 32 lanes -> lane t (< 16) ends up with the full sum of tap t (halving butterfly)
#pragma unroll
      for (int half = 8, off = 16; half >= 1; half >>= 1, off >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int j = 0; j < half; ++j) {
          const float send = upper ? acc[q][j] : acc[q][j + half];
          const float keep = upper ? acc[q][j + half] : acc[q][j];
          acc[q][j] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
      }
      const float v = acc[q][0] + __shfl_xor_sync(0xffffffffu, acc[q][0], 1);
      // after the butterfly lane L holds tap index built from its bits: bit4->+8, bit3->+4, bit2->+2, bit1->+1
      const int t = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
      const long long pix = p0 + q;
      if ((lane & 1) == 0 && pix < npix) {
        const int ix = (int)(pix % h), iy = (int)((pix / h) % h), n = (int)(pix / ((long long)h * h));
        const int oy = iy - (t >> 2) + 1, ox = ix - (t & 3) + 1;
        if (oy >= 0 && oy < ho && ox >= 0 && ox < ho) atomicAdd(out + ((long long)n * ho + oy) * ho + ox, v);
      }
    }
  }
}

// ga[n,iy,ix,c] = sum_{ky,kx} gy[n,iy-ky+1,ix-kx+1] * w[c][ky][kx].  A thread owns 8 channels of DP horizontally adjacent
// pixels, so every 32-byte weight read from shared memory feeds DP outputs; h must be a multiple of DP.
constexpr int DP = 2;
__global__ void __launch_bounds__(256) head_dgrad_kernel(const float* __restrict__ gy, const float* __restrict__ w,
                                                         __nv_bfloat16* __restrict__ ga, int N, int h, int C) {
  extern __shared__ float w_s[];  // [16 taps][C], swizzled
  head_stage_w(w, w_s, C);
  const int ho = h - 1;
  const int cvecs = C / 8;
  const long long total = (long long)N * h * (h / DP) * cvecs;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(i % cvecs);
    const long long pp = i / cvecs;  // pixel pair
    const int ix0 = (int)(pp % (h / DP)) * DP, iy = (int)((pp / (h / DP)) % h), n = (int)(pp / ((long long)(h / DP) * h));
    float acc[DP][8];
#pragma unroll
    for (int q = 0; q < DP; ++q)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[q][j] = 0.f;
#pragma unroll
    for (int t = 0; t < 16; ++t) {
      const int oy = iy - (t >> 2) + 1;
      if (oy < 0 || oy >= ho) continue;
      float g[DP];
      bool any = false;
#pragma unroll
      for (int q = 0; q < DP; ++q) {
        const int ox = ix0 + q - (t & 3) + 1;
        g[q] = (ox >= 0 && ox < ho) ? __ldg(gy + ((long long)n * ho + oy) * ho + ox) : 0.f;
        any |= (ox >= 0 && ox < ho);
      }
      if (!any) continue;
      const float* wp = w_s + hw_off(t, cv * 8, C);
      const float4 w0 = *reinterpret_cast<const float4*>(wp);
      const float4 w1 = *reinterpret_cast<const float4*>(wp + 4);
#pragma unroll
      for (int q = 0; q < DP; ++q) {
        acc[q][0] += g[q] * w0.x; acc[q][1] += g[q] * w0.y; acc[q][2] += g[q] * w0.z; acc[q][3] += g[q] * w0.w;
        acc[q][4] += g[q] * w1.x; acc[q][5] += g[q] * w1.y; acc[q][6] += g[q] * w1.z; acc[q][7] += g[q] * w1.w;
      }
    }
#pragma unroll
    for (int q = 0; q < DP; ++q) {
      const long long pix = ((long long)n * h + iy) * h + ix0 + q;
      *reinterpret_cast<uint4*>(ga + pix * C + cv * 8) = make_uint4(pack_bf16(acc[q][0], acc[q][1]), pack_bf16(acc[q][2], acc[q][3]),
                                                                  pack_bf16(acc[q][4], acc[q][5]), pack_bf16(acc[q][6], acc[q][7]));
    }
  }
}

// dw[c][ky][kx] += sum_{n,iy,ix} a[n,iy,ix,c] * gy[n,iy-ky+1,ix-kx+1].  Each input pixel is read once: a thread owns
// two channels and keeps all 16 taps in registers over its block's pixel range (8 pixels per batch so that the loads
// of a batch are in flight together); one atomic flush per block.  Requires C == 2 * blockDim.x.
__global__ void __launch_bounds__(256) head_wgrad_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ gy,
                                                         float* __restrict__ dw, int N, int h, int C) {
  constexpr int PB = 8;
  const int ho = h - 1;
  const long long npix = (long long)N * h * h;
  const long long p_begin = npix * blockIdx.x / gridDim.x, p_end = npix * (blockIdx.x + 1) / gridDim.x;
  __shared__ float g_s[2][PB][16];
  const int c = threadIdx.x * 2;
  float acc0[16], acc1[16];
#pragma unroll
  for (int t = 0; t < 16; ++t) acc0[t] = acc1[t] = 0.f;
  int buf = 0;
  for (long long p0 = p_begin; p0 < p_end; p0 += PB, buf ^= 1) {
    if (threadIdx.x < PB * 16) {
      const long long pix = p0 + (threadIdx.x >> 4);
      float g = 0.f;
      if (pix < p_end) {
        const int ix = (int)(pix % h), iy = (int)((pix / h) % h), n = (int)(pix / ((long long)h * h));
        const int t = threadIdx.x & 15, oy = iy - (t >> 2) + 1, ox = ix - (t & 3) + 1;
        if (oy >= 0 && oy < ho && ox >= 0 && ox < ho) g = __ldg(gy + ((long long)n * ho + oy) * ho + ox);
      }
      g_s[buf][threadIdx.x >> 4][threadIdx.x & 15] = g;
    }
    uint32_t u[PB];
#pragma unroll
    for (int j = 0; j < PB; ++j)
      u[j] = (p0 + j < p_end) ? __ldg(reinterpret_cast<const uint32_t*>(a + (p0 + j) * C + c)) : 0u;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < PB; ++j) {
      const float2 f = unpack_bf16(u[j]);
#pragma unroll
      for (int t = 0; t < 16; ++t) {
        const float g = g_s[buf][j][t];
        acc0[t] += g * f.x;
        acc1[t] += g * f.y;
      }
    }
  }
  // the 2 x 16 taps of this thread are 128 contiguous bytes of dw: 8 vector reductions instead of 32 scalar atomics
  float4* d4 = reinterpret_cast<float4*>(dw + (long long)c * 16);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    atomicAdd(d4 + j, make_float4(acc0[4 * j], acc0[4 * j + 1], acc0[4 * j + 2], acc0[4 * j + 3]));
    atomicAdd(d4 + 4 + j, make_float4(acc1[4 * j], acc1[4 * j + 1], acc1[4 * j + 2], acc1[4 * j + 3]));
  }
}


// ================================================================================================ round-2 rewrites
// ncu on the kernels above at (64, 16, 512) (profiles/r02_head_ncu.txt): fprop 23.6 us - 10.2 M instructions for 4.2 M
// FMAs, half its shared-memory wavefronts bank conflicts of the transposing weight staging; dgrad 29.5 us - 18.5 M
// instructions, ALU-bound on per-tap index arithmetic and predicated single loads of gy; wgrad 25.6 us - seven dependent
// load -> __syncthreads -> FMA rounds per block.  The rewrites keep the arithmetic (fp32 accumulation of bf16 activations
// times fp32 weights) and remove the overhead:
//   fprop : (experimental, B2C_HEAD_V2=2; not faster yet) the weights stay in the tensor's own [c][16 taps] order (a plain
//           32 KB copy, 16-byte skew per channel quarter); four adjacent lanes own FP pixels x a quarter of the channels,
//           and the 16 partial taps are transpose-reduced over the lane quad with 12 shuffles
//   dgrad : a thread owns 4 channels - its 64 weights live in registers for the whole kernel - and DP adjacent pixels;
//           the 4 x (DP+3) window of gy it needs is loaded up front (warp-uniform addresses), then 64*DP FMAs
//   wgrad : one 1024-thread block per SM: four pixel sub-ranges x 256 channel pairs, all gathered gy taps of the block's
//           range staged once, activation loads software-pipelined 8 pixels ahead, the four partial sums combined in
//           shared memory, one vector-atomic flush per block (148 x 2048 instead of 296 x 2048)

// A lane owns FP pixels x a quarter of the channels: every LDS.128 of four weights feeds 4*FP FMAs (a shared-memory load
// returns 32 x 4 bytes per cycle however many lanes share an address, so with one pixel per lane the kernel ran at the
// LDS rate: 28.8 us).
constexpr int FP = 4;
template <int C>
__global__ void __launch_bounds__(128) head_fprop2_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ w,
                                                          float* __restrict__ out, int N, int h) {
  constexpr int CQ = C / 4;                      // channels per lane of a quad
  __shared__ __align__(16) float w_s[C * 16 + 16];  // (c, t) at c*16 + t + (c / CQ)*4: the quad's four rows hit disjoint banks
  for (int i = threadIdx.x; i < C * 4; i += blockDim.x) {
    const int c = i >> 2, t0 = (i & 3) * 4;
    *reinterpret_cast<float4*>(w_s + c * 16 + t0 + (c / CQ) * 4) = __ldg(reinterpret_cast<const float4*>(w) + i);
  }
  __syncthreads();
  const int ho = h - 1;
  const long long npix = (long long)N * h * h;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const int sub = lane & 3;
  const float* wq = w_s + sub * CQ * 16 + sub * 4;
  for (long long p0 = ((long long)blockIdx.x * wpb + warp) * (8 * FP); p0 < npix; p0 += (long long)gridDim.x * wpb * (8 * FP)) {
    const long long pixb = p0 + (lane >> 2) * FP;
    const __nv_bfloat16* ap = a + sub * CQ;
    float acc[FP][16];
#pragma unroll
    for (int q = 0; q < FP; ++q)
#pragma unroll
      for (int t = 0; t < 16; ++t) acc[q][t] = 0.f;
#pragma unroll 1
    for (int i0 = 0; i0 < CQ; i0 += 32) {
      uint4 u[FP][4];
#pragma unroll
      for (int q = 0; q < FP; ++q)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          u[q][j] = pixb + q < npix ? __ldg(reinterpret_cast<const uint4*>(ap + (pixb + q) * C + i0 + j * 8)) : make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          float f[FP];
#pragma unroll
          for (int q = 0; q < FP; ++q) {
            const uint32_t word = k < 2 ? u[q][j].x : k < 4 ? u[q][j].y : k < 6 ? u[q][j].z : u[q][j].w;
            f[q] = (k & 1) ? __uint_as_float(word & 0xffff0000u) : __uint_as_float(word << 16);
          }
          const float* wp = wq + (i0 + j * 8 + k) * 16;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const float4 wv = *reinterpret_cast<const float4*>(wp + g * 4);
#pragma unroll
            for (int q = 0; q < FP; ++q) {
              acc[q][4 * g] = fmaf(f[q], wv.x, acc[q][4 * g]);
              acc[q][4 * g + 1] = fmaf(f[q], wv.y, acc[q][4 * g + 1]);
              acc[q][4 * g + 2] = fmaf(f[q], wv.z, acc[q][4 * g + 2]);
              acc[q][4 * g + 3] = fmaf(f[q], wv.w, acc[q][4 * g + 3]);
            }
          }
        }
      }
    }
    // transpose-reduce over the quad: after xor 1 a lane keeps 8 taps, after xor 2 four: taps of one ky
    const bool up1 = sub & 1, up2 = sub & 2;
    const int ky = (up1 ? 2 : 0) + (up2 ? 1 : 0);
#pragma unroll
    for (int q = 0; q < FP; ++q) {
      float r8[8], r4[4];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float send = up1 ? acc[q][j] : acc[q][j + 8], keep = up1 ? acc[q][j + 8] : acc[q][j];
        r8[j] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float send = up2 ? r8[j] : r8[j + 4], keep = up2 ? r8[j + 4] : r8[j];
        r4[j] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
      }
      const long long pix = pixb + q;
      if (pix < npix) {
        const int ix = (int)(pix % h), iy = (int)((pix / h) % h), n = (int)(pix / ((long long)h * h));
        const int oy = iy - ky + 1;
        if (oy >= 0 && oy < ho) {
#pragma unroll
          for (int kx = 0; kx < 4; ++kx) {
            const int ox = ix - kx + 1;
            if (ox >= 0 && ox < ho) atomicAdd(out + ((long long)n * ho + oy) * ho + ox, r4[kx]);
          }
        }
      }
    }
  }
}

template <int DPX>
__global__ void __launch_bounds__(256) head_dgrad2_kernel(const float* __restrict__ gy, const float* __restrict__ w,
                                                          __nv_bfloat16* __restrict__ ga, int N, int h, int C) {
  const int ho = h - 1;
  const int cqs = C / 4;                       // 256 % cqs == 0 (host-checked): a thread keeps its channel quad
  const int cq = threadIdx.x % cqs;
  float wr[4][16];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(w + (long long)(cq * 4 + j) * 16 + g * 4));
      wr[j][4 * g] = v.x; wr[j][4 * g + 1] = v.y; wr[j][4 * g + 2] = v.z; wr[j][4 * g + 3] = v.w;
    }
  const int gpr = h / DPX;                     // pixel groups per row
  const long long ngroups = (long long)N * h * gpr;
  const int gpb = blockDim.x / cqs;            // pixel groups per block and sweep
  for (long long g0 = (long long)blockIdx.x * gpb + threadIdx.x / cqs; g0 < ngroups; g0 += (long long)gridDim.x * gpb) {
    const int ix0 = (int)(g0 % gpr) * DPX, iy = (int)((g0 / gpr) % h), n = (int)(g0 / ((long long)gpr * h));
    // window of gy: rows oy = iy+1-ky (ky = 0..3), columns ox = ix0-2 .. ix0+DPX
    float G[4][DPX + 3];
#pragma unroll
    for (int ky = 0; ky < 4; ++ky) {
      const int oy = iy + 1 - ky;
      const bool rok = oy >= 0 && oy < ho;
      const float* gr = gy + ((long long)n * ho + (rok ? oy : 0)) * ho;
#pragma unroll
      for (int cidx = 0; cidx < DPX + 3; ++cidx) {
        const int ox = ix0 - 2 + cidx;
        G[ky][cidx] = (rok && ox >= 0 && ox < ho) ? __ldg(gr + ox) : 0.f;
      }
    }
    float acc[DPX][4];
#pragma unroll
    for (int q = 0; q < DPX; ++q)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[q][j] = 0.f;
#pragma unroll
    for (int ky = 0; ky < 4; ++ky)
#pragma unroll
      for (int kx = 0; kx < 4; ++kx)
#pragma unroll
        for (int q = 0; q < DPX; ++q) {
          const float g = G[ky][q + 3 - kx];   // ox = ix0 + q + 1 - kx
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[q][j] = fmaf(g, wr[j][ky * 4 + kx], acc[q][j]);
        }
#pragma unroll
    for (int q = 0; q < DPX; ++q) {
      const long long pix = ((long long)n * h + iy) * h + ix0 + q;
      *reinterpret_cast<uint2*>(ga + pix * C + cq * 4) = make_uint2(pack_bf16(acc[q][0], acc[q][1]), pack_bf16(acc[q][2], acc[q][3]));
    }
  }
}

// blockDim = 4 * (C/2): sub-range sr = tid / (C/2), channel pair cp = tid % (C/2).
constexpr int HW_MAXPIX = 160;  // pixels per block (host sizes the grid accordingly)
__global__ void __launch_bounds__(1024) head_wgrad2_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ gy,
                                                           float* __restrict__ dw, int N, int h, int C) {
  extern __shared__ __align__(16) float hw_smem[];
  float* g_s = hw_smem;                 // [HW_MAXPIX][16] gathered taps of the block's pixel range
  float* red = g_s + HW_MAXPIX * 16;    // [C][16] block sums
  const int ho = h - 1;
  const long long npix = (long long)N * h * h;
  const long long p_begin = npix * blockIdx.x / gridDim.x, p_end = npix * (blockIdx.x + 1) / gridDim.x;
  const int len = (int)(p_end - p_begin);
  const int half_c = C / 2, sr = threadIdx.x / half_c, cp = threadIdx.x % half_c;
  for (int i = threadIdx.x; i < len * 16; i += blockDim.x) {
    const long long pix = p_begin + (i >> 4);
    const int t = i & 15;
    const int ix = (int)(pix % h), iy = (int)((pix / h) % h), n = (int)(pix / ((long long)h * h));
    const int oy = iy - (t >> 2) + 1, ox = ix - (t & 3) + 1;
    g_s[i] = (oy >= 0 && oy < ho && ox >= 0 && ox < ho) ? __ldg(gy + ((long long)n * ho + oy) * ho + ox) : 0.f;
  }
  __syncthreads();
  const int per = (len + 3) / 4;
  const int q0 = sr * per, q1 = min(len, q0 + per);
  float acc0[16], acc1[16];
#pragma unroll
  for (int t = 0; t < 16; ++t) acc0[t] = acc1[t] = 0.f;
  constexpr int PB = 8;
  const __nv_bfloat16* ap = a + p_begin * C + cp * 2;
  uint32_t u[PB], un[PB];
#pragma unroll
  for (int j = 0; j < PB; ++j) u[j] = (q0 + j < q1) ? __ldg(reinterpret_cast<const uint32_t*>(ap + (long long)(q0 + j) * C)) : 0u;
  for (int q = q0; q < q1; q += PB) {
#pragma unroll
    for (int j = 0; j < PB; ++j)
      un[j] = (q + PB + j < q1) ? __ldg(reinterpret_cast<const uint32_t*>(ap + (long long)(q + PB + j) * C)) : 0u;
#pragma unroll
    for (int j = 0; j < PB; ++j) {
      if (q + j < q1) {
        const float f0 = __uint_as_float(u[j] << 16), f1 = __uint_as_float(u[j] & 0xffff0000u);
        const float4* gp = reinterpret_cast<const float4*>(g_s + (q + j) * 16);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const float4 gv = gp[g];
          acc0[4 * g] = fmaf(gv.x, f0, acc0[4 * g]); acc1[4 * g] = fmaf(gv.x, f1, acc1[4 * g]);
          acc0[4 * g + 1] = fmaf(gv.y, f0, acc0[4 * g + 1]); acc1[4 * g + 1] = fmaf(gv.y, f1, acc1[4 * g + 1]);
          acc0[4 * g + 2] = fmaf(gv.z, f0, acc0[4 * g + 2]); acc1[4 * g + 2] = fmaf(gv.z, f1, acc1[4 * g + 2]);
          acc0[4 * g + 3] = fmaf(gv.w, f0, acc0[4 * g + 3]); acc1[4 * g + 3] = fmaf(gv.w, f1, acc1[4 * g + 3]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < PB; ++j) u[j] = un[j];
  }
  // combine the four sub-ranges in turn: (tap, channel) at red[t*C + ch], so a warp's float2 accesses are contiguous
  for (int r = 0; r < 4; ++r) {
    if (sr == r) {
#pragma unroll
      for (int t = 0; t < 16; ++t) {
        float2* p2 = reinterpret_cast<float2*>(red + t * C + cp * 2);
        float2 v = make_float2(acc0[t], acc1[t]);
        if (r > 0) { const float2 o = *p2; v.x += o.x; v.y += o.y; }
        *p2 = v;
      }
    }
    __syncthreads();
  }
  float4* d4 = reinterpret_cast<float4*>(dw);
  for (int i = threadIdx.x; i < C * 4; i += blockDim.x) {
    const int ch = i >> 2, g = i & 3;
    atomicAdd(d4 + i, make_float4(red[(4 * g) * C + ch], red[(4 * g + 1) * C + ch], red[(4 * g + 2) * C + ch], red[(4 * g + 3) * C + ch]));
  }
}

// ------------------------------------------------------------------------------------------------ fprop on mma.sync
// The per-tap dot products of head_fprop_kernel are a GEMM with N = 16: P[pixel][tap] = a[pixel][:] . w[:][tap].  On the
// CUDA cores it is FFMA- and LDS-bound (22 us for 17 MB at (64, 16, 512)); here a warp owns 16 pixels x 16 taps as two
// m16n8k16 bf16 MMAs per 16 channels, so what is left is the activation stream.  The contraction index may be permuted
// freely as long as A and B agree, which lets every thread feed its A fragments from ONE 16-byte load per pixel row (8
// consecutive channels: 4 k-pairs = two MMAs) and its B fragments from one LDS.128 of the [tap][channel] weight rows.
// The fp32 weights enter as hi + lo bf16 pairs (two MMAs, 16 mantissa bits): results agree with the fp32-weight kernel to
// ~1e-5.  The channel range is split over KS warps per pixel tile (the taps are scattered with atomics anyway).
constexpr int H3_U = 4;        // 16-byte loads in flight per pixel row and thread
constexpr int H3_THREADS = 512;
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                               uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(H3_THREADS) head_fprop3_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ w,
                                                                 float* __restrict__ out, int N, int h, int C, int KS) {
  extern __shared__ __align__(16) uint8_t h3_smem[];
  __nv_bfloat16* ws = reinterpret_cast<__nv_bfloat16*>(h3_smem);  // [hi, lo][16 taps][C + 32]: rows 64 B apart mod 128 B
  const int ldw = C + 32;
  for (int i = threadIdx.x; i < 4 * C; i += blockDim.x) {  // one float4 = 4 consecutive taps of a channel
    const int c = i >> 2, t0 = (i & 3) * 4;
    const float4 v = __ldg(reinterpret_cast<const float4*>(w) + i);
    const float f[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __nv_bfloat16 hi = __float2bfloat16_rn(f[j]);
      ws[(t0 + j) * ldw + c] = hi;
      ws[(16 + t0 + j) * ldw + c] = __float2bfloat16_rn(f[j] - __bfloat162float(hi));
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int ho = h - 1, CK = C / KS;
  const long long npix = (long long)N * h * h;
  const long long items = ((npix + 15) / 16) * KS;
  for (long long item = (long long)blockIdx.x * (H3_THREADS / 32) + warp; item < items; item += (long long)gridDim.x * (H3_THREADS / 32)) {
    const long long p0 = (item / KS) * 16;
    const int cbase = (int)(item % KS) * CK + 8 * t;
    const long long r0 = p0 + g, r1 = p0 + g + 8;
    const bool v0 = r0 < npix, v1 = r1 < npix;
    const uint4* pa = reinterpret_cast<const uint4*>(a + (v0 ? r0 : 0) * C + cbase);
    const uint4* pb = reinterpret_cast<const uint4*>(a + (v1 ? r1 : 0) * C + cbase);
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    for (int c = 0; c < CK; c += 32 * H3_U) {
      uint4 ra[H3_U], rb[H3_U];
#pragma unroll
      for (int u = 0; u < H3_U; ++u) {
        ra[u] = v0 ? __ldg(pa + (c + 32 * u) / 8) : make_uint4(0, 0, 0, 0);
        rb[u] = v1 ? __ldg(pb + (c + 32 * u) / 8) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < H3_U; ++u) {
        const int cc = cbase + c + 32 * u;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const uint4 bh = *reinterpret_cast<const uint4*>(ws + (nt * 8 + g) * ldw + cc);
          const uint4 bl = *reinterpret_cast<const uint4*>(ws + (16 + nt * 8 + g) * ldw + cc);
          mma_bf16_16816(acc[nt], ra[u].x, rb[u].x, ra[u].y, rb[u].y, bh.x, bh.y);
          mma_bf16_16816(acc[nt], ra[u].z, rb[u].z, ra[u].w, rb[u].w, bh.z, bh.w);
          mma_bf16_16816(acc[nt], ra[u].x, rb[u].x, ra[u].y, rb[u].y, bl.x, bl.y);
          mma_bf16_16816(acc[nt], ra[u].z, rb[u].z, ra[u].w, rb[u].w, bl.z, bl.w);
        }
      }
    }
    // C fragment: acc[nt][j] = P[row g + 8 (j >> 1)][tap nt * 8 + 2 t + (j & 1)]
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const long long pix = half ? r1 : r0;
      if (pix >= npix) continue;
      const int ix = (int)(pix % h), iy = (int)((pix / h) % h);
      float* on = out + (pix / ((long long)h * h)) * ho * ho;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int tap = nt * 8 + 2 * t + j;
          const int oy = iy - (tap >> 2) + 1, ox = ix - (tap & 3) + 1;
          if (oy >= 0 && oy < ho && ox >= 0 && ox < ho) atomicAdd(on + oy * ho + ox, acc[nt][half * 2 + j]);
        }
    }
  }
}

// ------------------------------------------------------------------------------------------------ dgrad on mma.sync
// ga[pixel][c] = sum_tap G[pixel][tap] w[c][tap] with G[pixel][tap] = gy[n, iy - ky + 1, ix - kx + 1] (0 outside): a GEMM with
// K = 16 taps = ONE k16 step.  A warp owns 16 pixels: it gathers its A fragments (G, split hi + lo) once and then walks the
// channels in groups of 4 n-tiles whose columns are PERMUTED so that a thread's 8 accumulators of a pixel row are 8
// consecutive channels (logical column 2t'+e of n-tile j = channel base + 8t' + 2j + e): one 16-byte store per row and group.
// Weights: bf16x2 tap pairs, hi and lo, as [pair][C + 2] words in shared memory (conflict-free for the B fragment reads);
// three MMAs per n-tile (hi*hi + lo*hi + hi*lo) keep 16 mantissa bits of both operands.
__device__ __forceinline__ void split_bf16x2(float x, float y, uint32_t& hi, uint32_t& lo) {
  const __nv_bfloat16 xh = __float2bfloat16_rn(x), yh = __float2bfloat16_rn(y);
  hi = (uint32_t)__bfloat16_as_ushort(xh) | ((uint32_t)__bfloat16_as_ushort(yh) << 16);
  lo = pack_bf16(x - __bfloat162float(xh), y - __bfloat162float(yh));
}

__global__ void __launch_bounds__(H3_THREADS) head_dgrad3_kernel(const float* __restrict__ gy, const float* __restrict__ w,
                                                                 __nv_bfloat16* __restrict__ ga, int N, int h, int C, int NS) {
  extern __shared__ __align__(16) uint8_t h3_smem[];
  uint32_t* wsd = reinterpret_cast<uint32_t*>(h3_smem);  // [hi, lo][8 tap pairs][C + 2]
  const int ldw = C + 2;
  for (int i = threadIdx.x; i < 8 * C; i += blockDim.x) {
    const int c = i >> 3, tp = i & 7;
    const float2 v = __ldg(reinterpret_cast<const float2*>(w) + i);
    uint32_t hi, lo;
    split_bf16x2(v.x, v.y, hi, lo);
    wsd[tp * ldw + c] = hi;
    wsd[(8 + tp) * ldw + c] = lo;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int ho = h - 1, CN = C / NS;
  const long long npix = (long long)N * h * h;
  const long long items = ((npix + 15) / 16) * NS;
  for (long long item = (long long)blockIdx.x * (H3_THREADS / 32) + warp; item < items; item += (long long)gridDim.x * (H3_THREADS / 32)) {
    const long long p0 = (item / NS) * 16;
    const int cbeg = (int)(item % NS) * CN;
    uint32_t ah[4], al[4];  // a0: row g, taps 2t,2t+1; a1: row g+8, same taps; a2 / a3: taps 2t+8, 2t+9
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const long long pix = p0 + g + 8 * half;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (pix < npix) {
        const int ix = (int)(pix % h), iy = (int)((pix / h) % h);
        const float* gn = gy + (pix / ((long long)h * h)) * ho * ho;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int tap = 2 * t + (q & 1) + 8 * (q >> 1);
          const int oy = iy - (tap >> 2) + 1, ox = ix - (tap & 3) + 1;
          if (oy >= 0 && oy < ho && ox >= 0 && ox < ho) v[q] = __ldg(gn + oy * ho + ox);
        }
      }
      split_bf16x2(v[0], v[1], ah[half], al[half]);
      split_bf16x2(v[2], v[3], ah[2 + half], al[2 + half]);
    }
    const long long pr0 = p0 + g, pr1 = p0 + g + 8;
    for (int base = cbeg; base < cbeg + CN; base += 32) {
      float acc[4][4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
        const int ch = base + 8 * (g >> 1) + 2 * j + (g & 1);
        const uint32_t b0h = wsd[t * ldw + ch], b1h = wsd[(t + 4) * ldw + ch];
        const uint32_t b0l = wsd[(8 + t) * ldw + ch], b1l = wsd[(12 + t) * ldw + ch];
        mma_bf16_16816(acc[j], ah[0], ah[1], ah[2], ah[3], b0h, b1h);
        mma_bf16_16816(acc[j], al[0], al[1], al[2], al[3], b0h, b1h);
        mma_bf16_16816(acc[j], ah[0], ah[1], ah[2], ah[3], b0l, b1l);
      }
      if (pr0 < npix)
        *reinterpret_cast<uint4*>(ga + pr0 * C + base + 8 * t) =
            make_uint4(pack_bf16(acc[0][0], acc[0][1]), pack_bf16(acc[1][0], acc[1][1]), pack_bf16(acc[2][0], acc[2][1]),
                       pack_bf16(acc[3][0], acc[3][1]));
      if (pr1 < npix)
        *reinterpret_cast<uint4*>(ga + pr1 * C + base + 8 * t) =
            make_uint4(pack_bf16(acc[0][2], acc[0][3]), pack_bf16(acc[1][2], acc[1][3]), pack_bf16(acc[2][2], acc[2][3]),
                       pack_bf16(acc[3][2], acc[3][3]));
    }
  }
}

// ------------------------------------------------------------------------------------------------ wgrad on mma.sync
// dw[c][tap] += sum_pixel a[pixel][c] G[pixel][tap]: D[16 taps][8 channels] += A[16 taps][16 pixels] B[16 pixels][8 channels].
// B wants pixel PAIRS of one channel per register, memory has channel pairs of one pixel: a thread loads 8 consecutive
// channels of pixel row g (one 16-byte load = the row-major fragments of four 8x8 tiles under the channel permutation of
// the dgrad kernel above) and `movmatrix.trans` turns each into the B fragment.  A warp owns a 128-channel slice (64
// accumulators) and a quarter of the CTA's pixel tiles; the CTA's warps meet in a shared [16][C + 1] array, then one global
// atomic per element and CTA (dw is accumulated into, as before).  G is split hi + lo (two MMAs), the activations are exact.
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t x) {
  uint32_t y;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(y) : "r"(x));
  return y;
}

constexpr int W3_CW = 128;  // channels per warp
__global__ void __launch_bounds__(H3_THREADS) head_wgrad3_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ gy,
                                                                 float* __restrict__ dw, int N, int h, int C) {
  extern __shared__ __align__(16) uint8_t h3_smem[];
  float* sacc = reinterpret_cast<float*>(h3_smem);  // [16 taps][C + 1]
  const int lds = C + 1;
  for (int i = threadIdx.x; i < 16 * lds; i += blockDim.x) sacc[i] = 0.f;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int nslices = C / W3_CW, PG = (H3_THREADS / 32) / nslices;  // pixel groups per CTA
  const int slice = warp % nslices, pg = warp / nslices;
  const int ho = h - 1, cbeg = slice * W3_CW;
  const long long npix = (long long)N * h * h, ntiles = (npix + 15) / 16;
  float acc[W3_CW / 32][4][4];
#pragma unroll
  for (int q = 0; q < W3_CW / 32; ++q)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[q][j][0] = acc[q][j][1] = acc[q][j][2] = acc[q][j][3] = 0.f;
  if (pg < PG)
    for (long long tile = (long long)blockIdx.x * PG + pg; tile < ntiles; tile += (long long)gridDim.x * PG) {
      const long long p0 = tile * 16;
      const long long r0 = p0 + g, r1 = p0 + g + 8;
      const bool v0 = r0 < npix, v1 = r1 < npix;
      const uint4* pa = reinterpret_cast<const uint4*>(a + (v0 ? r0 : 0) * C + cbeg + 8 * t);
      const uint4* pb = reinterpret_cast<const uint4*>(a + (v1 ? r1 : 0) * C + cbeg + 8 * t);
      uint4 ra[W3_CW / 32], rb[W3_CW / 32];
#pragma unroll
      for (int q = 0; q < W3_CW / 32; ++q) {
        ra[q] = v0 ? __ldg(pa + 4 * q) : make_uint4(0, 0, 0, 0);
        rb[q] = v1 ? __ldg(pb + 4 * q) : make_uint4(0, 0, 0, 0);
      }
      // A fragments: G[pixel][tap] for pixels 2t, 2t+1 (a0, a1) and 2t+8, 2t+9 (a2, a3), taps g (a0, a2) and g+8 (a1, a3)
      uint32_t ah[4], al[4];
#pragma unroll
      for (int kh = 0; kh < 2; ++kh) {
        float v[2][2] = {{0.f, 0.f}, {0.f, 0.f}};  // [tap half][pixel of the pair]
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const long long pix = p0 + 2 * t + e + 8 * kh;
          if (pix < npix) {
            const int ix = (int)(pix % h), iy = (int)((pix / h) % h);
            const float* gn = gy + (pix / ((long long)h * h)) * ho * ho;
#pragma unroll
            for (int th = 0; th < 2; ++th) {
              const int tap = g + 8 * th;
              const int oy = iy - (tap >> 2) + 1, ox = ix - (tap & 3) + 1;
              if (oy >= 0 && oy < ho && ox >= 0 && ox < ho) v[th][e] = __ldg(gn + oy * ho + ox);
            }
          }
        }
        split_bf16x2(v[0][0], v[0][1], ah[2 * kh], al[2 * kh]);
        split_bf16x2(v[1][0], v[1][1], ah[2 * kh + 1], al[2 * kh + 1]);
      }
#pragma unroll
      for (int q = 0; q < W3_CW / 32; ++q) {
        const uint32_t x0[4] = {ra[q].x, ra[q].y, ra[q].z, ra[q].w}, x1[4] = {rb[q].x, rb[q].y, rb[q].z, rb[q].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t b0 = movmatrix_trans(x0[j]), b1 = movmatrix_trans(x1[j]);
          mma_bf16_16816(acc[q][j], ah[0], ah[1], ah[2], ah[3], b0, b1);
          mma_bf16_16816(acc[q][j], al[0], al[1], al[2], al[3], b0, b1);
        }
      }
    }
  // acc[q][j][e] = D[tap g][channel cbeg + 32 q + 8 t + 2 j + e], acc[q][j][2 + e]: tap g + 8
  if (pg < PG) {
#pragma unroll
    for (int q = 0; q < W3_CW / 32; ++q)
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int c = cbeg + 32 * q + 8 * t + 2 * j + e;
          atomicAdd(sacc + g * lds + c, acc[q][j][e]);
          atomicAdd(sacc + (g + 8) * lds + c, acc[q][j][2 + e]);
        }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 16 * C; i += blockDim.x) atomicAdd(dw + i, sacc[(i & 15) * lds + (i >> 4)]);
}

static bool head_mma() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B2C_HEAD_MMA");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

static bool head_v2() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B2C_HEAD_V2");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

static bool head_v2_fprop() {
  const char* e = getenv("B2C_HEAD_V2");
  return e && e[0] == '2';
}

extern "C" int b2c_head_fprop(const void* a, const float* w, float* out, int N, int h, int C, void* stream) {
  B2C_REQUIRE(C % 256 == 0 && h >= 2, "head_fprop: C=%d must be a multiple of 256, h=%d >= 2", C, h);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(out, 0, sizeof(float) * (size_t)N * (h - 1) * (h - 1), st);
  if (e != cudaSuccess) { b2c_set_error("head_fprop memset: %s", cudaGetErrorString(e)); return B2C_ERR_CUDA; }
  const long long npix = (long long)N * h * h;
  // measured at (64, 16, 512): 31.7 us against 29.7 us for the warp-per-4-pixels kernel below - opt-in only (B2C_HEAD_V2=2)
  if (head_v2_fprop() && C == 512 && ((uintptr_t)a & 15) == 0) {
    long long blocks = (npix + 4 * 8 * FP - 1) / (4 * 8 * FP);  // 4 warps x 8 quads x FP pixels
    if (blocks > 148) blocks = 148;
    head_fprop2_kernel<512><<<(unsigned)blocks, 128, 0, st>>>((const __nv_bfloat16*)a, w, out, N, h);
    return b2c_check_launch("head_fprop2");
  }
  if (head_mma() && C % (32 * H3_U) == 0 && C <= 1024 && ((uintptr_t)a & 15) == 0) {
    const int KS = C % (64 * H3_U) == 0 ? 2 : 1;
    const size_t smem3 = (size_t)32 * (C + 32) * sizeof(__nv_bfloat16);
    static bool attr3 = false;
    if (!attr3) {
      cudaFuncSetAttribute(head_fprop3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
      attr3 = true;
    }
    const long long items = ((npix + 15) / 16) * KS, wpb = H3_THREADS / 32;
    long long blocks = (items + wpb - 1) / wpb;
    if (blocks > 148) blocks = 148;  // one CTA per SM: the weight staging is paid per CTA
    head_fprop3_kernel<<<(unsigned)blocks, H3_THREADS, smem3, st>>>((const __nv_bfloat16*)a, w, out, N, h, C, KS);
    return b2c_check_launch("head_fprop3");
  }
  const size_t smem = (size_t)16 * C * sizeof(float);
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(head_fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    attr_done = true;
  }
  B2C_REQUIRE(smem <= 64 * 1024, "head_fprop: C=%d too large", C);
  long long blocks = (npix + 8 * HP - 1) / (8 * HP);
  if (blocks > 148 * 2) blocks = 148 * 2;  // the 32 KB weight staging is paid per block
  head_fprop_kernel<<<(unsigned)blocks, 256, smem, st>>>((const __nv_bfloat16*)a, w, out, N, h, C);
  return b2c_check_launch("head_fprop");
}

extern "C" int b2c_head_dgrad(const float* gy, const float* w, void* ga, int N, int h, int C, void* stream) {
  B2C_REQUIRE(C % 128 == 0 && (size_t)16 * C * sizeof(float) <= 64 * 1024, "head_dgrad: C=%d must be a multiple of 128, <= 1024", C);
  if (head_mma() && C % 64 == 0 && ((uintptr_t)ga & 15) == 0) {
    const int NS = 2;
    const size_t smem3 = (size_t)16 * (C + 2) * sizeof(uint32_t);
    static bool attr3 = false;
    if (!attr3) {
      cudaFuncSetAttribute(head_dgrad3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
      attr3 = true;
    }
    const long long npix = (long long)N * h * h;
    const long long items = ((npix + 15) / 16) * NS, wpb = H3_THREADS / 32;
    long long blocks = (items + wpb - 1) / wpb;
    if (blocks > 148) blocks = 148;
    head_dgrad3_kernel<<<(unsigned)blocks, H3_THREADS, smem3, (cudaStream_t)stream>>>(gy, w, (__nv_bfloat16*)ga, N, h, C, NS);
    return b2c_check_launch("head_dgrad3");
  }
  if (head_v2() && 256 % (C / 4) == 0) {
    const int dpx = h % 4 == 0 ? 4 : (h % 2 == 0 ? 2 : 1);  // odd maps (img_size = 16 * odd): one pixel per thread
    const long long ngroups = (long long)N * h * (h / dpx);
    const int gpb = 256 / (C / 4);
    long long blocks = (ngroups + gpb - 1) / gpb;
    if (blocks > 148 * 2) blocks = 148 * 2;
    if (dpx == 4)
      head_dgrad2_kernel<4><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(gy, w, (__nv_bfloat16*)ga, N, h, C);
    else if (dpx == 2)
      head_dgrad2_kernel<2><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(gy, w, (__nv_bfloat16*)ga, N, h, C);
    else
      head_dgrad2_kernel<1><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(gy, w, (__nv_bfloat16*)ga, N, h, C);
    return b2c_check_launch("head_dgrad2");
  }
  B2C_REQUIRE(h % DP == 0, "head_dgrad: h=%d must be even", h);
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(head_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    attr_done = true;
  }
  const long long total = (long long)N * h * (h / DP) * (C / 8);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 3) blocks = 148 * 3;  // the 32 KB weight staging is paid per block
  head_dgrad_kernel<<<(unsigned)blocks, 256, (size_t)16 * C * sizeof(float), (cudaStream_t)stream>>>(
      gy, w, (__nv_bfloat16*)ga, N, h, C);
  return b2c_check_launch("head_dgrad");
}

// dw fp32 [C][4][4] is ACCUMULATED into (caller zeroes).
extern "C" int b2c_head_wgrad(const void* a, const float* gy, float* dw, int N, int h, int C, void* stream) {
  B2C_REQUIRE(C == 256 || C == 512, "head_wgrad: C=%d must be 256 or 512", C);
  const long long npix = (long long)N * h * h;
  if (head_mma() && ((uintptr_t)a & 15) == 0) {
    const size_t smem3 = (size_t)16 * (C + 1) * sizeof(float);
    static bool attr3 = false;
    if (!attr3) {
      cudaFuncSetAttribute(head_wgrad3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
      attr3 = true;
    }
    const int PG = (H3_THREADS / 32) / (C / W3_CW);
    long long blocks = ((npix + 15) / 16 + PG - 1) / PG;
    if (blocks > 148) blocks = 148;  // fewer CTAs are slower (96: +3, 74: +5 us at (64, 16, 512)): the stream, not the atomics, is the cost
    if (blocks < 1) blocks = 1;
    head_wgrad3_kernel<<<(unsigned)blocks, H3_THREADS, smem3, (cudaStream_t)stream>>>((const __nv_bfloat16*)a, gy, dw, N, h, C);
    return b2c_check_launch("head_wgrad3");
  }
  if (head_v2()) {
    long long blocks = (npix + HW_MAXPIX - 1) / HW_MAXPIX;  // at most HW_MAXPIX pixels per block ...
    if (blocks < 148 && npix >= 148 * 8) blocks = 148;        // ... and every SM busy when there is enough work
    if (blocks < 1) blocks = 1;
    const size_t smem = sizeof(float) * ((size_t)HW_MAXPIX * 16 + (size_t)C * 16);
    static bool attr_done = false;
    if (!attr_done) {
      cudaFuncSetAttribute(head_wgrad2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
      attr_done = true;
    }
    head_wgrad2_kernel<<<(unsigned)blocks, 4 * (C / 2), smem, (cudaStream_t)stream>>>((const __nv_bfloat16*)a, gy, dw, N, h, C);
    return b2c_check_launch("head_wgrad2");
  }
  long long blocks = npix / 32;
  if (blocks > 148 * 2) blocks = 148 * 2;
  if (blocks < 1) blocks = 1;
  head_wgrad_kernel<<<(unsigned)blocks, C / 2, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)a, gy, dw, N, h, C);
  return b2c_check_launch("head_wgrad");
}
