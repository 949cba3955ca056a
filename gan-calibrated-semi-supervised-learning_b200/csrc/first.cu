// First-layer convolutions (Cin = 3 for G.down1, 6 for D.model.0; cgan/models.py:103,245,257).  K = 16*Cin is 48/96:
// these layers are bandwidth-bound and read the caller's NCHW fp32 patches directly, so they run on
// CUDA cores with the channel concat (models.py:257), the GP interpolation (losses.py:203-204), the spectral-norm
// 1/sigma, bias and LeakyReLU(0.2) fused in.  Output / incoming gradients are NHWC bf16 like every other activation.
#include "common.cuh"

using namespace b2c;

struct FirstIn {
  const float* a[2];   // up to two (N,3,H,W) fp32 tensors = channels 0-2 and 3-5
  const float* b[2];   // optional interpolation partners
  const float* alpha;  // optional (N) : x = alpha*a + (1-alpha)*b
  int nparts;
};

__device__ __forceinline__ float first_in_load(const FirstIn& in, int n, int c, int h, int w, int H, int W) {
  const int part = c >= 3 ? 1 : 0, cc = c - 3 * part;
  const long long off = (((long long)n * 3 + cc) * H + h) * W + w;
  float v = __ldg(in.a[part] + off);
  if (in.alpha) {
    const float al = __ldg(in.alpha + n);
    v = al * v + (1.f - al) * __ldg(in.b[part] + off);
  }
  return v;
}

// y[n,oy,ox,co] = act( s * sum x[n,c,2oy+ky-1,2ox+kx-1] w[co][c][ky][kx] + bias[co] ), 64 output channels.
template <int CIN>
__global__ void __launch_bounds__(256) first_fprop_kernel(FirstIn in, const float* __restrict__ w,
                                                          const float* __restrict__ inv_sigma,
                                                          const float* __restrict__ bias, __nv_bfloat16* __restrict__ y,
                                                          int N, int H, int W, float slope) {
  constexpr int K = CIN * 16;
  __shared__ float w_s[K][64];
  const float s = inv_sigma ? *inv_sigma : 1.f;
  for (int i = threadIdx.x; i < K * 64; i += 256) {
    const int co = i / K, k = i % K;  // w is [co][k] contiguous in k
    w_s[k][co] = w[i] * s;
  }
  __syncthreads();
  const int Ho = H / 2, Wo = W / 2;
  const long long npix = (long long)N * Ho * Wo;
  const int cg = threadIdx.x >> 6;  // 4 groups of 16 output channels
  for (long long pix = (long long)blockIdx.x * 64 + (threadIdx.x & 63); pix < npix; pix += (long long)gridDim.x * 64) {
    const int ox = (int)(pix % Wo), oy = (int)((pix / Wo) % Ho), n = (int)(pix / ((long long)Wo * Ho));
    float acc[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = bias ? __ldg(bias + cg * 16 + j) : 0.f;
#pragma unroll
    for (int c = 0; c < CIN; ++c) {
#pragma unroll
      for (int ky = 0; ky < 4; ++ky) {
        const int iy = 2 * oy + ky - 1;
#pragma unroll
        for (int kx = 0; kx < 4; ++kx) {
          const int ix = 2 * ox + kx - 1;
          float xv = 0.f;
          if (iy >= 0 && iy < H && ix >= 0 && ix < W) xv = first_in_load(in, n, c, iy, ix, H, W);
          const float4* wr = reinterpret_cast<const float4*>(&w_s[c * 16 + ky * 4 + kx][cg * 16]);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 wv = wr[j];
            acc[4 * j] += xv * wv.x; acc[4 * j + 1] += xv * wv.y; acc[4 * j + 2] += xv * wv.z; acc[4 * j + 3] += xv * wv.w;
          }
        }
      }
    }
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float a = acc[2 * j], b = acc[2 * j + 1];
      a = a > 0.f ? a : a * slope;
      b = b > 0.f ? b : b * slope;
      o[j] = pack_bf16(a, b);
    }
    uint4* d = reinterpret_cast<uint4*>(y + pix * 64 + cg * 16);
    d[0] = make_uint4(o[0], o[1], o[2], o[3]);
    d[1] = make_uint4(o[4], o[5], o[6], o[7]);
  }
}

// gx[n,c,h,w] = s * sum_{co,ky,kx} gz[n,oy,ox,co] w[co][c][ky][kx],  h = 2oy+ky-1, w = 2ox+kx-1.  fp32 NCHW out.
template <int CIN>
__global__ void __launch_bounds__(256) first_dgrad_kernel(const __nv_bfloat16* __restrict__ gz, const float* __restrict__ w,
                                                          const float* __restrict__ inv_sigma, float* __restrict__ gx0,
                                                          float* __restrict__ gx1, int N, int H, int W) {
  constexpr int K = CIN * 16;
  __shared__ float w_s[16][CIN][64];  // [tap][c][co]
  const float s = inv_sigma ? *inv_sigma : 1.f;
  for (int i = threadIdx.x; i < K * 64; i += 256) {
    const int co = i / K, k = i % K, c = k / 16, tap = k % 16;
    w_s[tap][c][co] = w[i] * s;
  }
  __syncthreads();
  const int Ho = H / 2, Wo = W / 2;
  const long long npix = (long long)N * H * W;
  for (long long pix = (long long)blockIdx.x * 256 + threadIdx.x; pix < npix; pix += (long long)gridDim.x * 256) {
    const int x = (int)(pix % W), yy = (int)((pix / W) % H), n = (int)(pix / ((long long)W * H));
    float acc[CIN];
#pragma unroll
    for (int c = 0; c < CIN; ++c) acc[c] = 0.f;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int ky = ((yy + 1) & 1) + 2 * (t >> 1), kx = ((x + 1) & 1) + 2 * (t & 1);
      const int oy = (yy + 1 - ky) >> 1, ox = (x + 1 - kx) >> 1;
      if (oy < 0 || oy >= Ho || ox < 0 || ox >= Wo) continue;
      const uint4* g4 = reinterpret_cast<const uint4*>(gz + (((long long)n * Ho + oy) * Wo + ox) * 64);
      const int tap = ky * 4 + kx;
#pragma unroll
      for (int v = 0; v < 8; ++v) {
        const uint4 u = __ldg(g4 + v);
        const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 g = unpack_bf16(uu[e]);
          const int co = v * 8 + e * 2;
#pragma unroll
          for (int c = 0; c < CIN; ++c) acc[c] += g.x * w_s[tap][c][co] + g.y * w_s[tap][c][co + 1];
        }
      }
    }
#pragma unroll
    for (int c = 0; c < CIN; ++c) {
      float* dst = c >= 3 ? gx1 : gx0;
      dst[(((long long)n * 3 + (c % 3)) * H + yy) * W + x] = acc[c];
    }
  }
}

// dw[co][c][ky][kx] += sum_pix gz[pix,co] * x[n,c,2oy+ky-1,2ox+kx-1];  db[co] += sum_pix gz[pix,co].
// Block: 256 threads = 64 co x 4 k-groups; pixels staged 32 at a time in shared memory; fp32 atomics at the end.
template <int CIN>
__global__ void __launch_bounds__(256) first_wgrad_kernel(FirstIn in, const __nv_bfloat16* __restrict__ gz,
                                                          float* __restrict__ dw, float* __restrict__ db, int N, int H,
                                                          int W, long long pix_per_block) {
  constexpr int K = CIN * 16, KG = K / 4, P = 32;
  __shared__ float x_s[P][K];
  __shared__ float g_s[P][65];
  const int Ho = H / 2, Wo = W / 2;
  const long long npix = (long long)N * Ho * Wo;
  const long long p_begin = (long long)blockIdx.x * pix_per_block;
  long long p_end = p_begin + pix_per_block;
  if (p_end > npix) p_end = npix;
  const int co = threadIdx.x & 63, kg = threadIdx.x >> 6;
  float acc[KG];
#pragma unroll
  for (int j = 0; j < KG; ++j) acc[j] = 0.f;
  float bacc = 0.f;
  for (long long p0 = p_begin; p0 < p_end; p0 += P) {
    const int np = (int)((p_end - p0) < P ? (p_end - p0) : P);
    for (int i = threadIdx.x; i < P * K; i += 256) {
      const int pp = i / K, k = i % K;
      float v = 0.f;
      if (pp < np) {
        const long long pix = p0 + pp;
        const int ox = (int)(pix % Wo), oy = (int)((pix / Wo) % Ho), n = (int)(pix / ((long long)Wo * Ho));
        const int c = k / 16, ky = (k / 4) & 3, kx = k & 3;
        const int iy = 2 * oy + ky - 1, ix = 2 * ox + kx - 1;
        if (iy >= 0 && iy < H && ix >= 0 && ix < W) v = first_in_load(in, n, c, iy, ix, H, W);
      }
      x_s[pp][k] = v;
    }
    for (int i = threadIdx.x; i < P * 64; i += 256) {
      const int pp = i >> 6, c = i & 63;
      g_s[pp][c] = pp < np ? __bfloat162float(gz[(p0 + pp) * 64 + c]) : 0.f;
    }
    __syncthreads();
#pragma unroll 4
    for (int pp = 0; pp < P; ++pp) {
      const float g = g_s[pp][co];
      if (kg == 0) bacc += g;
#pragma unroll
      for (int j = 0; j < KG; ++j) acc[j] += g * x_s[pp][kg * KG + j];
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < KG; ++j) atomicAdd(dw + (long long)co * K + kg * KG + j, acc[j]);
  if (kg == 0 && db) atomicAdd(db + co, bacc);
}

static FirstIn make_in(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha) {
  FirstIn in;
  in.a[0] = a0; in.a[1] = a1; in.b[0] = b0; in.b[1] = b1; in.alpha = alpha; in.nparts = a1 ? 2 : 1;
  return in;
}

// a0/a1: (N,3,H,W) fp32 NCHW halves (a1 null => Cin=3).  b0/b1/alpha: optional interpolation (all or none).
extern "C" int b2c_first_fprop(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                               const float* w, const float* inv_sigma, const float* bias, void* y, int N, int H, int W,
                               float slope, void* stream) {
  B2C_REQUIRE(H % 2 == 0 && W % 2 == 0, "first_fprop: odd size");
  FirstIn in = make_in(a0, a1, b0, b1, alpha);
  const long long npix = (long long)N * (H / 2) * (W / 2);
  long long blocks = (npix + 63) / 64;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (a1)
    first_fprop_kernel<6><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(in, w, inv_sigma, bias, (__nv_bfloat16*)y, N, H, W, slope);
  else
    first_fprop_kernel<3><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(in, w, inv_sigma, bias, (__nv_bfloat16*)y, N, H, W, slope);
  return b2c_check_launch("first_fprop");
}

extern "C" int b2c_first_dgrad(const void* gz, const float* w, const float* inv_sigma, float* gx0, float* gx1, int N,
                               int H, int W, void* stream) {
  const long long npix = (long long)N * H * W;
  long long blocks = (npix + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (gx1)
    first_dgrad_kernel<6><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)gz, w, inv_sigma, gx0, gx1, N, H, W);
  else
    first_dgrad_kernel<3><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)gz, w, inv_sigma, gx0, gx1, N, H, W);
  return b2c_check_launch("first_dgrad");
}

// dw: fp32 [64][Cin][4][4], db: fp32 [64] or null; both are ACCUMULATED into (caller zeroes).
extern "C" int b2c_first_wgrad(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                               const void* gz, float* dw, float* db, int N, int H, int W, void* stream) {
  FirstIn in = make_in(a0, a1, b0, b1, alpha);
  const long long npix = (long long)N * (H / 2) * (W / 2);
  long long blocks = 148 * 4;
  long long ppb = (npix + blocks - 1) / blocks;
  ppb = (ppb + 31) / 32 * 32;
  blocks = (npix + ppb - 1) / ppb;
  if (a1)
    first_wgrad_kernel<6><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(in, (const __nv_bfloat16*)gz, dw, db, N, H, W, ppb);
  else
    first_wgrad_kernel<3><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(in, (const __nv_bfloat16*)gz, dw, db, N, H, W, ppb);
  return b2c_check_launch("first_wgrad");
}
