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
//   wtu   : one block per 32-column strip of W: t[k] = sum_r W[r][k] u[r] over ALL rows (16 warps take rows r = warp mod 16,
//           8 independent 128-byte row loads in flight per warp, fixed-order combine in shared memory) and the strip's
//           sum of t^2 - no partial-sum array for the next kernel to re-reduce
//   wv    : v = t / max(|t|, eps) (|t|^2 = fixed-order sum of the <= 128 strip partials, re-derived by every block into
//           shared memory, K <= 4096 floats), s[r] = W[r] . v for the block's SN_ROWS rows (W streamed once, coalesced)
//   finish: u = s / max(|s|, eps), sigma = u . s
constexpr int SN_STRIP = 32;  // columns per wtu block
constexpr int SN_ROWS = 4;    // rows per wv block
constexpr int SN_MAXK = 4096;

__global__ void __launch_bounds__(512) sn_wtu_kernel(const float* __restrict__ W, const float* __restrict__ u,
                                                      float* __restrict__ t, float* __restrict__ ssp, int R, int K) {
  __shared__ float red[16][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k = blockIdx.x * SN_STRIP + lane;
  float acc = 0.f;
  if (k < K) {
    int r = warp;
    for (; r + 7 * 16 < R; r += 8 * 16) {
      float w[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = __ldg(W + (long long)(r + 16 * j) * K + k);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc += w[j] * __ldg(u + r + 16 * j);
    }
    for (; r < R; r += 16) acc += __ldg(W + (long long)r * K + k) * __ldg(u + r);
  }
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0) {
    float tk = 0.f;
#pragma unroll
    for (int w = 0; w < 16; ++w) tk += red[w][lane];
    if (k < K) t[k] = tk; else tk = 0.f;
    const float ss = warp_sum(tk * tk);
    if (lane == 0) ssp[blockIdx.x] = ss;
  }
}

__global__ void __launch_bounds__(256) sn_wv_kernel(const float* __restrict__ W, const float* __restrict__ t,
                                                     const float* __restrict__ ssp, float* __restrict__ v,
                                                     float* __restrict__ s, int R, int K, int NS, int update_v) {
  __shared__ float ts[SN_MAXK];
  __shared__ float red[SN_ROWS][8];
  if (update_v) {
    float ss = 0.f;  // every warp derives the same value in the same order
    for (int i = threadIdx.x & 31; i < NS; i += 32) ss += ssp[i];
    ss = warp_sum(ss);
    const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
    for (int k = threadIdx.x; k < K; k += 256) {
      const float vk = t[k] * inv;
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
    float tt = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) tt += red[threadIdx.x][w];
    s[r0 + threadIdx.x] = tt;
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

// ws: fp32 scratch of K + ceil(K/32) + R floats (ceil(R/32)*K + R is enough whenever R > 32).  u[R], v[K] updated in place
// when train != 0.  sig[2] <- (sigma, 1/sigma).
extern "C" int b2c_sn_power_iter(const float* W, float* u, float* v, float* sig, float* ws, int R, int K, int train,
                                 void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B2C_REQUIRE(K <= SN_MAXK, "sn_power_iter: K=%d exceeds %d", K, SN_MAXK);
  const int NS = (K + SN_STRIP - 1) / SN_STRIP;
  float* t = ws; float* ssp = ws + K; float* s = ssp + NS;
  if (train) sn_wtu_kernel<<<NS, 512, 0, st>>>(W, u, t, ssp, R, K);
  sn_wv_kernel<<<(R + SN_ROWS - 1) / SN_ROWS, 256, 0, st>>>(W, t, ssp, v, s, R, K, NS, train ? 1 : 0);
  sn_finish_kernel<<<1, 256, 0, st>>>(s, u, sig, R, train ? 1 : 0);
  return b2c_check_launch("sn_power_iter");
}

// <dWe, W> in two atomic-free stages: one partial per block, summed in a fixed order by every block of sn_bwd_kernel.
__global__ void sn_dot_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ part, long long n) {
  __shared__ float sh[32];
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    acc += __ldg(a + i) * __ldg(b + i);
  const float r = block_sum(acc, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = r;
}
__global__ void sn_bwd_kernel(const float* __restrict__ dWe, const float* __restrict__ u, const float* __restrict__ v,
                              const float* __restrict__ sig, const float* __restrict__ part, int nparts,
                              float* __restrict__ dW, int R, int K, int accumulate) {
  float dot = 0.f;  // every warp derives the same value in the same order
  for (int i = threadIdx.x & 31; i < nparts; i += 32) dot += __ldg(part + i);
  dot = warp_sum(dot);
  const float inv = sig[1];
  const float coef = dot * inv;  // <dWe, We> = <dWe, W> / sigma
  const long long n = (long long)R * K;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / K), k = (int)(i % K);
    const float g = (dWe[i] - coef * __ldg(u + r) * __ldg(v + k)) * inv;
    dW[i] = accumulate ? dW[i] + g : g;
  }
}
// u, v, sig: the values used by THAT forward instance (saved copies).  ws: min(592, ceil(R*K/1024)) floats of scratch.
extern "C" int b2c_sn_bwd(const float* dWe, const float* W, const float* u, const float* v, const float* sig, float* ws,
                          float* dW, int R, int K, int accumulate, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const long long n = (long long)R * K;
  int blocks = (int)((n + 1023) / 1024);
  if (blocks > 592) blocks = 592;
  sn_dot_kernel<<<blocks, 256, 0, st>>>(dWe, W, ws, n);
  sn_bwd_kernel<<<blocks, 256, 0, st>>>(dWe, u, v, sig, ws, blocks, dW, R, K, accumulate);
  return b2c_check_launch("sn_bwd");
}

// ------------------------------------------------------------------------------------------------ batched variants
// All spectral-norm layers of one Discriminator forward in THREE launches (blockIdx.y = layer) instead of three per layer,
// and the per-forward-instance snapshots of u / v (the backward of THAT forward needs them; torch clones the buffers at
// spectral_norm.py:108-110) written by the same kernels instead of separate copies.
#define B2C_SN_MAX_LAYERS 8
struct SnLayer {
  const float* W; float* u; float* v; float* sig; float* ws; float* u_snap; float* v_snap;
  int R, K;
};
struct SnTable { SnLayer l[B2C_SN_MAX_LAYERS]; };

__global__ void __launch_bounds__(512) sn_wtu_multi_kernel(SnTable tb) {
  const SnLayer& L = tb.l[blockIdx.y];
  const int R = L.R, K = L.K;
  if ((int)blockIdx.x * SN_STRIP >= K) return;
  __shared__ float red[16][33];
  const float* __restrict__ W = L.W; const float* __restrict__ u = L.u;
  float* t = L.ws; float* ssp = L.ws + K;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k = blockIdx.x * SN_STRIP + lane;
  float acc = 0.f;
  if (k < K) {
    int r = warp;
    for (; r + 7 * 16 < R; r += 8 * 16) {
      float w[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = __ldg(W + (long long)(r + 16 * j) * K + k);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc += w[j] * __ldg(u + r + 16 * j);
    }
    for (; r < R; r += 16) acc += __ldg(W + (long long)r * K + k) * __ldg(u + r);
  }
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0) {
    float tk = 0.f;
#pragma unroll
    for (int w = 0; w < 16; ++w) tk += red[w][lane];
    if (k < K) t[k] = tk; else tk = 0.f;
    const float ss = warp_sum(tk * tk);
    if (lane == 0) ssp[blockIdx.x] = ss;
  }
}

__global__ void __launch_bounds__(256) sn_wv_multi_kernel(SnTable tb, int update_v) {
  const SnLayer& L = tb.l[blockIdx.y];
  const int R = L.R, K = L.K;
  const int r0 = blockIdx.x * SN_ROWS;
  if (r0 >= R) return;
  __shared__ float ts[SN_MAXK];
  __shared__ float red[SN_ROWS][8];
  const int NS = (K + SN_STRIP - 1) / SN_STRIP;
  const float* __restrict__ W = L.W;
  const float* t = L.ws; const float* ssp = L.ws + K; float* s = L.ws + K + NS;
  if (update_v) {
    float ss = 0.f;  // every warp derives the same value in the same order
    for (int i = threadIdx.x & 31; i < NS; i += 32) ss += ssp[i];
    ss = warp_sum(ss);
    const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
    for (int k = threadIdx.x; k < K; k += 256) {
      const float vk = t[k] * inv;
      ts[k] = vk;
      if (blockIdx.x == 0) { L.v[k] = vk; if (L.v_snap) L.v_snap[k] = vk; }
    }
  } else {
    for (int k = threadIdx.x; k < K; k += 256) {
      const float vk = L.v[k];
      ts[k] = vk;
      if (blockIdx.x == 0 && L.v_snap) L.v_snap[k] = vk;
    }
  }
  __syncthreads();
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
    float tt = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) tt += red[threadIdx.x][w];
    s[r0 + threadIdx.x] = tt;
  }
}

__global__ void __launch_bounds__(256) sn_finish_multi_kernel(SnTable tb, int update_u) {
  const SnLayer& L = tb.l[blockIdx.x];
  const int R = L.R, K = L.K;
  const float* s = L.ws + K + (K + SN_STRIP - 1) / SN_STRIP;
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
    if (update_u) { ur = s[r] * inv; L.u[r] = ur; } else ur = L.u[r];
    if (L.u_snap) L.u_snap[r] = ur;
    acc += ur * s[r];
  }
  const float sigma = block_sum(acc, sh);
  if (threadIdx.x == 0) { L.sig[0] = sigma; L.sig[1] = 1.f / sigma; }
}

// count <= 8 layers; per layer: W [R][K], u [R], v [K] (updated in place when train != 0), sig [2] <- (sigma, 1/sigma),
// ws: K + ceil(K/32) + R floats of scratch, u_snap [R] / v_snap [K] (nullable) <- the u / v this forward used.
extern "C" int b2c_sn_power_iter_multi(const float* const* W, float* const* u, float* const* v, float* const* sig,
                                       float* const* ws, float* const* u_snap, float* const* v_snap, const int* R,
                                       const int* K, int count, int train, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B2C_REQUIRE(count >= 1 && count <= B2C_SN_MAX_LAYERS, "sn_power_iter_multi: %d layers (max %d)", count, B2C_SN_MAX_LAYERS);
  SnTable tb;
  int maxK = 0, maxR = 0;
  for (int i = 0; i < count; ++i) {
    B2C_REQUIRE(K[i] <= SN_MAXK, "sn_power_iter_multi: K=%d exceeds %d", K[i], SN_MAXK);
    tb.l[i].W = W[i]; tb.l[i].u = u[i]; tb.l[i].v = v[i]; tb.l[i].sig = sig[i]; tb.l[i].ws = ws[i];
    tb.l[i].u_snap = u_snap ? u_snap[i] : nullptr; tb.l[i].v_snap = v_snap ? v_snap[i] : nullptr;
    tb.l[i].R = R[i]; tb.l[i].K = K[i];
    if (K[i] > maxK) maxK = K[i];
    if (R[i] > maxR) maxR = R[i];
  }
  if (train) sn_wtu_multi_kernel<<<dim3((maxK + SN_STRIP - 1) / SN_STRIP, count), 512, 0, st>>>(tb);
  sn_wv_multi_kernel<<<dim3((maxR + SN_ROWS - 1) / SN_ROWS, count), 256, 0, st>>>(tb, train ? 1 : 0);
  sn_finish_multi_kernel<<<count, 256, 0, st>>>(tb, train ? 1 : 0);
  return b2c_check_launch("sn_power_iter_multi");
}

// Deferred rank-1 term of the spectral-norm backward.  The weight-gradient kernels accumulate dWe / sigma_i straight into
// the gradient of weight_orig and leave block partials of <dWe, W> (b2c_wgrad4s2_bf16 / b2c_first_wgrad: `part`); once per
// backward this kernel subtracts, for every layer, sum_i (<dWe_i, W> / sigma_i^2) u_i v_i^T over the forward instances i
// (real / fake / interpolated critic passes each carry their own u, v, sigma): dW = (dWe - <dWe, We> u v^T) / sigma.
#define B2C_SN_MAX_INST 8
struct SnR1Inst { const float* u; const float* v; const float* sig; const float* part; int nparts; };
struct SnR1Layer { float* grad; int R, K, ninst; SnR1Inst inst[B2C_SN_MAX_INST]; };
struct SnR1Table { SnR1Layer l[B2C_SN_MAX_LAYERS]; };

__global__ void __launch_bounds__(256) sn_rank1_kernel(SnR1Table tb) {
  const SnR1Layer& L = tb.l[blockIdx.y];
  float coef[B2C_SN_MAX_INST];
#pragma unroll
  for (int j = 0; j < B2C_SN_MAX_INST; ++j) {
    coef[j] = 0.f;
    if (j < L.ninst) {
      float dot = 0.f;  // every warp derives the same value in the same order
      for (int i = threadIdx.x & 31; i < L.inst[j].nparts; i += 32) dot += __ldg(L.inst[j].part + i);
      dot = warp_sum(dot);
      const float inv = __ldg(L.inst[j].sig + 1);
      coef[j] = dot * inv * inv;
    }
  }
  const long long n = (long long)L.R * L.K;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / L.K), k = (int)(i % L.K);
    float corr = 0.f;
#pragma unroll
    for (int j = 0; j < B2C_SN_MAX_INST; ++j)
      if (j < L.ninst) corr += coef[j] * __ldg(L.inst[j].u + r) * __ldg(L.inst[j].v + k);
    L.grad[i] -= corr;
  }
}

// nlayers <= 8 gradient tensors grad[l] ([R][K] fp32, updated in place); their ninst[l] <= 8 instances are given as flat
// arrays in layer order: u [R], v [K], sig [2] = (sigma, 1/sigma), part [nparts] (block partials of <dWe, W>).
extern "C" int b2c_sn_rank1(float* const* grad, const int* R, const int* K, const int* ninst, int nlayers,
                            const float* const* u, const float* const* v, const float* const* sig,
                            const float* const* part, const int* nparts, void* stream) {
  B2C_REQUIRE(nlayers >= 1 && nlayers <= B2C_SN_MAX_LAYERS, "sn_rank1: %d layers (max %d)", nlayers, B2C_SN_MAX_LAYERS);
  SnR1Table tb;
  int k = 0;
  long long maxn = 0;
  for (int l = 0; l < nlayers; ++l) {
    B2C_REQUIRE(ninst[l] >= 0 && ninst[l] <= B2C_SN_MAX_INST, "sn_rank1: %d instances (max %d)", ninst[l], B2C_SN_MAX_INST);
    tb.l[l].grad = grad[l]; tb.l[l].R = R[l]; tb.l[l].K = K[l]; tb.l[l].ninst = ninst[l];
    for (int j = 0; j < ninst[l]; ++j, ++k) {
      tb.l[l].inst[j].u = u[k]; tb.l[l].inst[j].v = v[k]; tb.l[l].inst[j].sig = sig[k];
      tb.l[l].inst[j].part = part[k]; tb.l[l].inst[j].nparts = nparts[k];
    }
    const long long n = (long long)R[l] * K[l];
    if (n > maxn) maxn = n;
  }
  long long blocks = (maxn + 1023) / 1024;
  if (blocks > 592) blocks = 592;
  if (blocks < 1) blocks = 1;
  sn_rank1_kernel<<<dim3((unsigned)blocks, nlayers), 256, 0, (cudaStream_t)stream>>>(tb);
  return b2c_check_launch("sn_rank1");
}
