// Weight packing: fp32 master weights (torch layout [D0][D1][4][4]) -> bf16 K-major GEMM operands for tapgemm.cu,
// with the spectral-norm 1/sigma (torch spectral_norm.py:111-112, attached at cgan/models.py:238) folded in.
#include "common.cuh"

using namespace b2c;

// C-form: out[d0][(ky*4+kx)*D1 + d1] = w[d0][d1][ky][kx] * s      (conv fprop: d0=Cout; convT dgrad: d0=Cin)
__device__ __forceinline__ void pack_c_one(const float* __restrict__ w, const float* __restrict__ inv_sigma,
                                           __nv_bfloat16* __restrict__ out, int D0, int D1, long long i) {  // i over D0*D1
  if (i >= (long long)D0 * D1) return;
  const int d1 = (int)(i % D1);
  const long long d0 = i / D1;
  const float s = inv_sigma ? *inv_sigma : 1.f;
  const float4* src = reinterpret_cast<const float4*>(w + i * 16);
  float v[16];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float4 f = src[j];
    v[4 * j] = f.x; v[4 * j + 1] = f.y; v[4 * j + 2] = f.z; v[4 * j + 3] = f.w;
  }
  __nv_bfloat16* o = out + d0 * 16 * D1 + d1;
#pragma unroll
  for (int t = 0; t < 16; ++t) o[(long long)t * D1] = __float2bfloat16(v[t] * s);
}
__global__ void pack_w_c_kernel(const float* __restrict__ w, const float* __restrict__ inv_sigma,
                                __nv_bfloat16* __restrict__ out, int D0, int D1) {
  pack_c_one(w, inv_sigma, out, D0, D1, (long long)blockIdx.x * blockDim.x + threadIdx.x);
}

// T-form: out[py*2+px][d1][(a*2+b)*D0 + d0] = w[d0][d1][3-py-2a][3-px-2b] * s   (convT fprop: d0=Cin; conv dgrad: d0=Cout)
__device__ __forceinline__ void pack_t_one(const float* __restrict__ w, const float* __restrict__ inv_sigma,
                                           __nv_bfloat16* __restrict__ out, int D0, int D1, long long i) {  // i over D1*D0, d0 fastest
  if (i >= (long long)D0 * D1) return;
  const int d0 = (int)(i % D0);
  const int d1 = (int)(i / D0);
  const float s = inv_sigma ? *inv_sigma : 1.f;
  const float4* src = reinterpret_cast<const float4*>(w + ((long long)d0 * D1 + d1) * 16);
  float v[16];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float4 f = src[j];
    v[4 * j] = f.x; v[4 * j + 1] = f.y; v[4 * j + 2] = f.z; v[4 * j + 3] = f.w;
  }
#pragma unroll
  for (int par = 0; par < 4; ++par) {
    const int py = par >> 1, px = par & 1;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int a = t >> 1, b = t & 1;
      const int ky = 3 - py - 2 * a, kx = 3 - px - 2 * b;
      out[(((long long)par * D1 + d1) * 4 + t) * D0 + d0] = __float2bfloat16(v[ky * 4 + kx] * s);
    }
  }
}
__global__ void pack_w_t_kernel(const float* __restrict__ w, const float* __restrict__ inv_sigma,
                                __nv_bfloat16* __restrict__ out, int D0, int D1) {
  pack_t_one(w, inv_sigma, out, D0, D1, (long long)blockIdx.x * blockDim.x + threadIdx.x);
}

// Every stale pack of a network in ONE launch (after its optimizer step): a by-value job table, blocks assigned by prefix.
constexpr int PACK_MAX_JOBS = 64;
struct PackJobs {
  const float* w[PACK_MAX_JOBS];
  __nv_bfloat16* out[PACK_MAX_JOBS];
  int D0[PACK_MAX_JOBS], D1[PACK_MAX_JOBS], kind[PACK_MAX_JOBS], blk0[PACK_MAX_JOBS + 1];
  int n;
};
__global__ void pack_w_multi_kernel(const __grid_constant__ PackJobs J) {
  int j = 0;
  while (j + 1 < J.n && (int)blockIdx.x >= J.blk0[j + 1]) ++j;
  const long long i = (long long)((int)blockIdx.x - J.blk0[j]) * blockDim.x + threadIdx.x;
  if (J.kind[j] == 0) pack_c_one(J.w[j], nullptr, J.out[j], J.D0[j], J.D1[j], i);
  else pack_t_one(J.w[j], nullptr, J.out[j], J.D0[j], J.D1[j], i);
}

extern "C" int b2c_pack_w_c(const float* w, const float* inv_sigma, void* out, int D0, int D1, void* stream) {
  const long long n = (long long)D0 * D1;
  pack_w_c_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(w, inv_sigma, (__nv_bfloat16*)out, D0, D1);
  return b2c_check_launch("pack_w_c");
}
extern "C" int b2c_pack_w_t(const float* w, const float* inv_sigma, void* out, int D0, int D1, void* stream) {
  const long long n = (long long)D0 * D1;
  pack_w_t_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(w, inv_sigma, (__nv_bfloat16*)out, D0, D1);
  return b2c_check_launch("pack_w_t");
}

// Host arrays of njobs entries: kind 0 = C-form, 1 = T-form of the un-scaled weight (what b2c_pack_w_c / _t write with
// inv_sigma = NULL).  More than 64 jobs go out as several launches.
extern "C" int b2c_pack_w_multi(const void* const* w, void* const* out, const int* D0, const int* D1, const int* kind, int njobs,
                                void* stream) {
  for (int base = 0; base < njobs; base += PACK_MAX_JOBS) {
    PackJobs J;
    J.n = njobs - base < PACK_MAX_JOBS ? njobs - base : PACK_MAX_JOBS;
    int blocks = 0;
    for (int j = 0; j < J.n; ++j) {
      B2C_REQUIRE(kind[base + j] == 0 || kind[base + j] == 1, "pack_w_multi: kind[%d]=%d", base + j, kind[base + j]);
      J.w[j] = (const float*)w[base + j];
      J.out[j] = (__nv_bfloat16*)out[base + j];
      J.D0[j] = D0[base + j]; J.D1[j] = D1[base + j]; J.kind[j] = kind[base + j];
      J.blk0[j] = blocks;
      blocks += (int)(((long long)D0[base + j] * D1[base + j] + 255) / 256);
    }
    J.blk0[J.n] = blocks;
    if (!blocks) continue;
    pack_w_multi_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(J);
    const int rc = b2c_check_launch("pack_w_multi");
    if (rc) return rc;
  }
  return B2C_OK;
}
