// Legacy spectral-norm hook (torch/nn/utils/spectral_norm.py:92-114, attached at cgan/models.py:236-238):
// one power iteration per train-mode forward on W viewed as [R][K], sigma = u.(W v), and its backward
//   dW = (dWe - <dWe, We> u v^T) / sigma      (u, v constants for autograd).
// All fp32, tiny matrices (<= 512 x 4096): three short launches, no host sync; 1/sigma stays on the device and is
// folded into the bf16 weight packing (pack.cu).
#include "common.cuh"

using namespace b2c;

__device__ __forceinline__ float block_sum(float v, float* sh) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  float r = 0.f;
  if (w == 0) {
    r = l < nw ? sh[l] : 0.f;
    r = warp_sum(r);
    if (l == 0) sh[0] = r;
  }
  __syncthreads();
  r = sh[0];
  return r;
}

// Power iteration in three atomic-free launches (deterministic, so u/v stay bit-identical on every data-parallel rank):
//   wtu   : part[rs][k] = sum_{r in row chunk rs} W[r][k] u[r]           grid (K/128, RS), 8 independent loads in flight
//   wv    : t = sum_rs part[rs]  (every block re-derives it into shared memory, K <= 4096 floats), v = t / max(|t|, eps),
//           s[r] = W[r] . v for the block's SN_ROWS rows (W streamed once, coalesced)
//   finish: u = s / max(|s|, eps), sigma = u . s
constexpr int SN_CHUNK = 32;  // rows per wtu block
constexpr int SN_ROWS = 4;    // rows per wv block
constexpr int SN_MAXK = 4096;

__global__ void __launch_bounds__(128) sn_wtu_kernel(const float* __restrict__ W, const float* __restrict__ u,
                                                      float* __restrict__ part, int R, int K) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int r0 = blockIdx.y * SN_CHUNK;
  if (k >= K) return;
  float acc = 0.f;
  const int r1 = min(r0 + SN_CHUNK, R);
  int r = r0;
  for (; r + 8 <= r1; r += 8) {
    float w[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = __ldg(W + (long long)(r + j) * K + k);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += w[j] * __ldg(u + r + j);
  }
  for (; r < r1; ++r) acc += __ldg(W + (long long)r * K + k) * __ldg(u + r);
  part[(long long)blockIdx.y * K + k] = acc;
}

__global__ void __launch_bounds__(256) sn_wv_kernel(const float* __restrict__ W, const float* __restrict__ part,
                                                     float* __restrict__ v, float* __restrict__ s, int R, int K, int RS,
                                                     int update_v) {
  __shared__ float ts[SN_MAXK];
  __shared__ float sh[32];
  __shared__ float red[SN_ROWS][8];
  float inv = 1.f;
  if (update_v) {
    float ss = 0.f;
    for (int k = threadIdx.x; k < K; k += 256) {
      float t = 0.f;
      for (int j = 0; j < RS; ++j) t += part[(long long)j * K + k];
      ts[k] = t;
      ss += t * t;
    }
    ss = block_sum(ss, sh);
    inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
    for (int k = threadIdx.x; k < K; k += 256) {
      const float vk = ts[k] * inv;
      ts[k] = vk;
      if (blockIdx.x == 0) v[k] = vk;
    }
  } else {
    for (int k = threadIdx.x; k < K; k += 256) ts[k] = v[k];
  }
  __syncthreads();
  const int r0 = blockIdx.x * SN_ROWS;
  float acc[SN_ROWS];
#pragma unroll
  for (int j = 0; j < SN_ROWS; ++j) acc[j] = 0.f;
  for (int k = threadIdx.x; k < K; k += 256) {
    const float vk = ts[k];
#pragma unroll
    for (int j = 0; j < SN_ROWS; ++j)
      if (r0 + j < R) acc[j] += __ldg(W + (long long)(r0 + j) * K + k) * vk;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < SN_ROWS; ++j) {
    const float w = warp_sum(acc[j]);
    if (lane == 0) red[j][warp] = w;
  }
  __syncthreads();
  if (threadIdx.x < SN_ROWS && r0 + threadIdx.x < R) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[threadIdx.x][w];
    s[r0 + threadIdx.x] = t;
  }
}

// train: u = s / max(||s||, eps), sigma = u.s ; eval: sigma = u_old.s.   sig[0] = sigma, sig[1] = 1/sigma
__global__ void sn_finish_kernel(const float* __restrict__ s, float* __restrict__ u, float* __restrict__ sig, int R,
                                 int update_u) {
  __shared__ float sh[32];
  float inv = 1.f;
  if (update_u) {
    float ss = 0.f;
    for (int r = threadIdx.x; r < R; r += blockDim.x) ss += s[r] * s[r];
    ss = block_sum(ss, sh);
    inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
  }
  float acc = 0.f;
  for (int r = threadIdx.x; r < R; r += blockDim.x) {
    float ur;
    if (update_u) { ur = s[r] * inv; u[r] = ur; } else ur = u[r];
    acc += ur * s[r];
  }
  const float sigma = block_sum(acc, sh);
  if (threadIdx.x == 0) { sig[0] = sigma; sig[1] = 1.f / sigma; }
}

// ws: fp32 scratch of ceil(R/32)*K + R floats.  u[R], v[K] updated in place when train != 0.  sig[2] <- (sigma, 1/sigma).
extern "C" int b2c_sn_power_iter(const float* W, float* u, float* v, float* sig, float* ws, int R, int K, int train,
                                 void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B2C_REQUIRE(K <= SN_MAXK, "sn_power_iter: K=%d exceeds %d", K, SN_MAXK);
  const int RS = (R + SN_CHUNK - 1) / SN_CHUNK;
  float* part = ws; float* s = ws + (long long)RS * K;
  if (train) sn_wtu_kernel<<<dim3((K + 127) / 128, RS), 128, 0, st>>>(W, u, part, R, K);
  sn_wv_kernel<<<(R + SN_ROWS - 1) / SN_ROWS, 256, 0, st>>>(W, part, v, s, R, K, RS, train ? 1 : 0);
  sn_finish_kernel<<<1, 256, 0, st>>>(s, u, sig, R, train ? 1 : 0);
  return b2c_check_launch("sn_power_iter");
}

__global__ void sn_dot_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, long long n) {
  __shared__ float sh[32];
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    acc += __ldg(a + i) * __ldg(b + i);
  const float r = block_sum(acc, sh);
  if (threadIdx.x == 0) atomicAdd(out, r);
}
__global__ void sn_bwd_kernel(const float* __restrict__ dWe, const float* __restrict__ u, const float* __restrict__ v,
                              const float* __restrict__ sig, const float* __restrict__ dot, float* __restrict__ dW, int R,
                              int K, int accumulate) {
  const float inv = sig[1];
  const float coef = dot[0] * inv;  // <dWe, We> = <dWe, W> / sigma
  const long long n = (long long)R * K;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / K), k = (int)(i % K);
    const float g = (dWe[i] - coef * __ldg(u + r) * __ldg(v + k)) * inv;
    dW[i] = accumulate ? dW[i] + g : g;
  }
}
// u, v, sig: the values used by THAT forward instance (saved copies).  ws: 1 float scratch.
extern "C" int b2c_sn_bwd(const float* dWe, const float* W, const float* u, const float* v, const float* sig, float* ws,
                          float* dW, int R, int K, int accumulate, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(ws, 0, sizeof(float), st);
  if (e != cudaSuccess) { b2c_set_error("sn_bwd memset: %s", cudaGetErrorString(e)); return B2C_ERR_CUDA; }
  const long long n = (long long)R * K;
  int blocks = (int)((n + 1023) / 1024);
  if (blocks > 592) blocks = 592;
  sn_dot_kernel<<<blocks, 256, 0, st>>>(dWe, W, ws, n);
  sn_bwd_kernel<<<blocks, 256, 0, st>>>(dWe, u, v, sig, ws, dW, R, K, accumulate);
  return b2c_check_launch("sn_bwd");
}
