// PatchGAN head: Conv2d(512,1,k4,s1,p1,bias=False) (cgan/models.py:252).  N=1 cannot fill an MMA tile, so these are
// bandwidth kernels: a GEMV per output pixel (fprop), a 16-tap scatter of a scalar map (dgrad) and a channel-wise
// reduction (wgrad).  a4 / g_a4 are NHWC bf16 (N,h,h,C); the score map is fp32 (N,1,h-1,h-1).
#include "common.cuh"

using namespace b2c;

// out[n,oy,ox] = sum_{ky,kx,c} a[n,oy+ky-1,ox+kx-1,c] * w[c][ky][kx]
__global__ void __launch_bounds__(256) head_fprop_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ w,
                                                         float* __restrict__ out, int N, int h, int C) {
  extern __shared__ float w_s[];  // [16][C]
  for (int i = threadIdx.x; i < 16 * C; i += blockDim.x) {
    const int c = i / 16, t = i % 16;
    w_s[t * C + c] = w[i];
  }
  __syncthreads();
  const int ho = h - 1;
  const long long npix = (long long)N * ho * ho;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  for (long long pix = (long long)blockIdx.x * wpb + warp; pix < npix; pix += (long long)gridDim.x * wpb) {
    const int ox = (int)(pix % ho), oy = (int)((pix / ho) % ho), n = (int)(pix / ((long long)ho * ho));
    float acc = 0.f;
#pragma unroll
    for (int t = 0; t < 16; ++t) {
      const int iy = oy + (t >> 2) - 1, ix = ox + (t & 3) - 1;
      if (iy < 0 || iy >= h || ix < 0 || ix >= h) continue;
      const __nv_bfloat16* row = a + (((long long)n * h + iy) * h + ix) * C;
      for (int c = lane * 8; c < C; c += 256) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(row + c));
        const float* ws = w_s + t * C + c;
        const float2 f0 = unpack_bf16(u.x), f1 = unpack_bf16(u.y), f2 = unpack_bf16(u.z), f3 = unpack_bf16(u.w);
        acc += f0.x * ws[0] + f0.y * ws[1] + f1.x * ws[2] + f1.y * ws[3] + f2.x * ws[4] + f2.y * ws[5] + f3.x * ws[6] +
               f3.y * ws[7];
      }
    }
    acc = warp_sum(acc);
    if (lane == 0) out[pix] = acc;
  }
}

// ga[n,iy,ix,c] = sum_{ky,kx} gy[n,iy-ky+1,ix-kx+1] * w[c][ky][kx]
__global__ void __launch_bounds__(256) head_dgrad_kernel(const float* __restrict__ gy, const float* __restrict__ w,
                                                         __nv_bfloat16* __restrict__ ga, int N, int h, int C) {
  const int ho = h - 1;
  const int cvecs = C / 8;
  const long long total = (long long)N * h * h * cvecs;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(i % cvecs);
    const long long pix = i / cvecs;
    const int ix = (int)(pix % h), iy = (int)((pix / h) % h), n = (int)(pix / ((long long)h * h));
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int t = 0; t < 16; ++t) {
      const int oy = iy - (t >> 2) + 1, ox = ix - (t & 3) + 1;
      if (oy < 0 || oy >= ho || ox < 0 || ox >= ho) continue;
      const float g = __ldg(gy + ((long long)n * ho + oy) * ho + ox);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += g * __ldg(w + (long long)(cv * 8 + e) * 16 + t);
    }
    uint4 o = make_uint4(pack_bf16(acc[0], acc[1]), pack_bf16(acc[2], acc[3]), pack_bf16(acc[4], acc[5]),
                         pack_bf16(acc[6], acc[7]));
    *reinterpret_cast<uint4*>(ga + pix * C + cv * 8) = o;
  }
}

// dw[c][ky][kx] += sum_{n,oy,ox} gy[n,oy,ox] * a[n,oy+ky-1,ox+kx-1,c].  grid = (16 taps, image splits).
__global__ void __launch_bounds__(256) head_wgrad_kernel(const __nv_bfloat16* __restrict__ a, const float* __restrict__ gy,
                                                         float* __restrict__ dw, int N, int h, int C) {
  const int t = blockIdx.x, ky = t >> 2, kx = t & 3;
  const int ho = h - 1;
  const int n_begin = (int)((long long)N * blockIdx.y / gridDim.y), n_end = (int)((long long)N * (blockIdx.y + 1) / gridDim.y);
  for (int c = threadIdx.x * 2; c < C; c += 512) {
    float a0 = 0.f, a1 = 0.f;
    for (int n = n_begin; n < n_end; ++n)
      for (int oy = 0; oy < ho; ++oy) {
        const int iy = oy + ky - 1;
        if (iy < 0 || iy >= h) continue;
        for (int ox = 0; ox < ho; ++ox) {
          const int ix = ox + kx - 1;
          if (ix < 0 || ix >= h) continue;
          const float g = __ldg(gy + ((long long)n * ho + oy) * ho + ox);
          const float2 f = unpack_bf16(__ldg(reinterpret_cast<const uint32_t*>(a + (((long long)n * h + iy) * h + ix) * C + c)));
          a0 += g * f.x;
          a1 += g * f.y;
        }
      }
    atomicAdd(dw + (long long)c * 16 + t, a0);
    atomicAdd(dw + (long long)(c + 1) * 16 + t, a1);
  }
}

extern "C" int b2c_head_fprop(const void* a, const float* w, float* out, int N, int h, int C, void* stream) {
  B2C_REQUIRE(C % 256 == 0 && h >= 2, "head_fprop: C=%d must be a multiple of 256, h=%d >= 2", C, h);
  const size_t smem = (size_t)16 * C * sizeof(float);
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(head_fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    attr_done = true;
  }
  B2C_REQUIRE(smem <= 64 * 1024, "head_fprop: C=%d too large", C);
  const long long npix = (long long)N * (h - 1) * (h - 1);
  long long blocks = (npix + 7) / 8;
  if (blocks > 148 * 2) blocks = 148 * 2;
  head_fprop_kernel<<<(unsigned)blocks, 256, smem, (cudaStream_t)stream>>>((const __nv_bfloat16*)a, w, out, N, h, C);
  return b2c_check_launch("head_fprop");
}

extern "C" int b2c_head_dgrad(const float* gy, const float* w, void* ga, int N, int h, int C, void* stream) {
  B2C_REQUIRE(C % 8 == 0, "head_dgrad: C=%d", C);
  const long long total = (long long)N * h * h * (C / 8);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  head_dgrad_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(gy, w, (__nv_bfloat16*)ga, N, h, C);
  return b2c_check_launch("head_dgrad");
}

// dw fp32 [C][4][4] is ACCUMULATED into (caller zeroes).
extern "C" int b2c_head_wgrad(const void* a, const float* gy, float* dw, int N, int h, int C, void* stream) {
  B2C_REQUIRE(C % 2 == 0, "head_wgrad: C=%d", C);
  int splits = N < 16 ? N : 16;
  head_wgrad_kernel<<<dim3(16, splits), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)a, gy, dw, N, h, C);
  return b2c_check_launch("head_wgrad");
}
