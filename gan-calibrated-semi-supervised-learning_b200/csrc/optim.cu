// clip_grad_norm_(max_norm) + torch-default Adam as used at cgan/cgan_train_enhanced.py:256-257,331-332,368-369,
// as two multi-tensor launches per optimizer step: (1) global sum of squares over all gradient tensors,
// (2) fused clip-scale + Adam update (reads p,g,m,v; writes p,m,v: 28 B/param, pure HBM bandwidth).
#include "common.cuh"

using namespace b2c;

#define B2C_MAX_TENSORS 16
struct TensorTable {
  float* p[B2C_MAX_TENSORS];
  float* g[B2C_MAX_TENSORS];
  float* m[B2C_MAX_TENSORS];
  float* v[B2C_MAX_TENSORS];
  long long n[B2C_MAX_TENSORS];
  int count;
};

__global__ void __launch_bounds__(256) multi_sumsq_kernel(TensorTable t, float* __restrict__ out) {
  __shared__ float sh[8];
  const float* g = t.g[blockIdx.y];
  const long long n = t.n[blockIdx.y];
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const float x = g[i];
    acc += x * x;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += sh[w];
    atomicAdd(out, s);
  }
}

__global__ void __launch_bounds__(256) multi_adam_kernel(TensorTable t, const float* __restrict__ gnorm_sq, float max_norm,
                                                         float lr, float beta1, float beta2, float eps, int step_index,
                                                         const long long* __restrict__ step_base) {
  // bias corrections from the step count; the count may live on the device so a captured CUDA graph can be replayed
  const double stepd = (double)step_index + (step_base ? (double)(*step_base) : 0.0);
  const float bc1 = (float)(1.0 - pow((double)beta1, stepd)), bc2 = (float)(1.0 - pow((double)beta2, stepd));
  float* p = t.p[blockIdx.y]; float* m = t.m[blockIdx.y]; float* v = t.v[blockIdx.y];
  const float* g = t.g[blockIdx.y];
  const long long n = t.n[blockIdx.y];
  float coef = 1.f;
  if (gnorm_sq) coef = fminf(max_norm / (sqrtf(*gnorm_sq) + 1e-6f), 1.f);
  const float step = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const float gi = g[i] * coef;
    const float mi = beta1 * m[i] + (1.f - beta1) * gi;
    const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
    m[i] = mi; v[i] = vi;
    p[i] -= step * mi / (sqrtf(vi) * inv_sqrt_bc2 + eps);
  }
}

static int fill_table(TensorTable& t, float* const* p, float* const* g, float* const* m, float* const* v,
                      const long long* n, int count) {
  B2C_REQUIRE(count >= 1 && count <= B2C_MAX_TENSORS, "optimizer: %d tensors (max %d)", count, B2C_MAX_TENSORS);
  t.count = count;
  for (int i = 0; i < count; ++i) {
    t.p[i] = p ? p[i] : nullptr; t.g[i] = g[i]; t.m[i] = m ? m[i] : nullptr; t.v[i] = v ? v[i] : nullptr; t.n[i] = n[i];
  }
  return B2C_OK;
}

// out[0] <- sum over all tensors of g^2 (out is zeroed here).
extern "C" int b2c_multi_sumsq(float* const* g, const long long* n, int count, float* out, void* stream) {
  TensorTable t;
  if (int rc = fill_table(t, nullptr, g, nullptr, nullptr, n, count)) return rc;
  cudaError_t e = cudaMemsetAsync(out, 0, sizeof(float), (cudaStream_t)stream);
  if (e != cudaSuccess) { b2c_set_error("multi_sumsq memset: %s", cudaGetErrorString(e)); return B2C_ERR_CUDA; }
  multi_sumsq_kernel<<<dim3(64, count), 256, 0, (cudaStream_t)stream>>>(t, out);
  return b2c_check_launch("multi_sumsq");
}

// One Adam step on `count` tensors; gnorm_sq (device scalar, nullable) drives the clip coefficient
// min(1, max_norm / (sqrt(gnorm_sq) + 1e-6)).  The optimizer's step count AFTER the increment is
// step_index + *step_base (step_base: device scalar, nullable) and must be >= 1.
extern "C" int b2c_multi_adam(float* const* p, float* const* g, float* const* m, float* const* v, const long long* n,
                              int count, const float* gnorm_sq, float max_norm, float lr, float beta1, float beta2,
                              float eps, int step_index, const long long* step_base, void* stream) {
  TensorTable t;
  if (int rc = fill_table(t, p, g, m, v, n, count)) return rc;
  multi_adam_kernel<<<dim3(128, count), 256, 0, (cudaStream_t)stream>>>(t, gnorm_sq, max_norm, lr, beta1, beta2, eps,
                                                                        step_index, step_base);
  return b2c_check_launch("multi_adam");
}
