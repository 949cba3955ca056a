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

// t[k] = sum_r W[r][k] u[r];  scal[0] += sum t^2
__global__ void sn_wtu_kernel(const float* __restrict__ W, const float* __restrict__ u, float* __restrict__ t,
                              float* __restrict__ scal, int R, int K) {
  __shared__ float sh[32];
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  float acc = 0.f;
  if (k < K)
    for (int r = 0; r < R; ++r) acc += __ldg(W + (long long)r * K + k) * __ldg(u + r);
  if (k < K) t[k] = acc;
  const float ss = block_sum(k < K ? acc * acc : 0.f, sh);
  if (threadIdx.x == 0) atomicAdd(scal, ss);
}

// v = t / max(||t||, eps) (written by block 0 when update_v);  s[r] = W[r].v ; scal[1] += sum s^2
__global__ void sn_wv_kernel(const float* __restrict__ W, const float* __restrict__ t, float* __restrict__ v,
                             float* __restrict__ s, float* __restrict__ scal, int R, int K, int update_v) {
  const float inv = update_v ? 1.f / fmaxf(sqrtf(scal[0]), 1e-12f) : 1.f;
  const float* src = update_v ? t : v;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const int r = blockIdx.x * wpb + warp;
  if (r < R) {
    float acc = 0.f;
    for (int k = lane; k < K; k += 32) acc += __ldg(W + (long long)r * K + k) * (src[k] * inv);
    acc = warp_sum(acc);
    if (lane == 0) {
      s[r] = acc;
      atomicAdd(scal + 1, acc * acc);
    }
  }
  if (update_v) {
    __syncthreads();  // all reads of t by this block are done; v != t so other blocks are unaffected
    if (blockIdx.x == 0)
      for (int k = threadIdx.x; k < K; k += blockDim.x) v[k] = t[k] * inv;
  }
}

// train: u = s / max(||s||, eps), sigma = u.s ; eval: sigma = u_old.s.   sig[0] = sigma, sig[1] = 1/sigma
__global__ void sn_finish_kernel(const float* __restrict__ s, float* __restrict__ u, const float* __restrict__ scal,
                                 float* __restrict__ sig, int R, int update_u) {
  __shared__ float sh[32];
  const float inv = update_u ? 1.f / fmaxf(sqrtf(scal[1]), 1e-12f) : 1.f;
  float acc = 0.f;
  for (int r = threadIdx.x; r < R; r += blockDim.x) {
    float ur;
    if (update_u) { ur = s[r] * inv; u[r] = ur; } else ur = u[r];
    acc += ur * s[r];
  }
  const float sigma = block_sum(acc, sh);
  if (threadIdx.x == 0) { sig[0] = sigma; sig[1] = 1.f / sigma; }
}

// ws: fp32 scratch of K + R + 2 floats.  u[R], v[K] updated in place when train != 0.  sig[2] <- (sigma, 1/sigma).
extern "C" int b2c_sn_power_iter(const float* W, float* u, float* v, float* sig, float* ws, int R, int K, int train,
                                 void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  float* t = ws; float* s = ws + K; float* scal = ws + K + R;
  cudaError_t e = cudaMemsetAsync(scal, 0, 2 * sizeof(float), st);
  if (e != cudaSuccess) { b2c_set_error("sn memset: %s", cudaGetErrorString(e)); return B2C_ERR_CUDA; }
  if (train) sn_wtu_kernel<<<(K + 127) / 128, 128, 0, st>>>(W, u, t, scal, R, K);
  sn_wv_kernel<<<(R + 7) / 8, 256, 0, st>>>(W, t, v, s, scal, R, K, train ? 1 : 0);
  sn_finish_kernel<<<1, 256, 0, st>>>(s, u, scal, sig, R, train ? 1 : 0);
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
