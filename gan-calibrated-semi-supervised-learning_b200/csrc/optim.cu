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

// Stage 1 of the global gradient norm: one partial sum of squares per block, part[blockIdx.y * gridDim.x + blockIdx.x];
// stage 2 (the fixed-order sum of the partials) runs inside multi_adam_kernel, so the clip coefficient is bit-identical
// on every data-parallel rank and from run to run (no float atomics).
__global__ void __launch_bounds__(256) multi_sumsq_kernel(TensorTable t, float* __restrict__ part) {
  __shared__ float sh[8];
  const float* g = t.g[blockIdx.y];
  const long long n = t.n[blockIdx.y];
  float acc = 0.f;
  if ((n & 3) == 0 && (reinterpret_cast<uintptr_t>(g) & 15) == 0) {
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n / 4; i += (long long)gridDim.x * 256) {
      const float4 x = __ldg(g4 + i);
      acc += (x.x * x.x + x.y * x.y) + (x.z * x.z + x.w * x.w);
    }
  } else {
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
      const float x = g[i];
      acc += x * x;
    }
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += sh[w];
    part[blockIdx.y * gridDim.x + blockIdx.x] = s;
  }
}

__device__ __forceinline__ void adam1(float& p, float& m, float& v, float g, float beta1, float beta2, float step,
                                      float inv_sqrt_bc2, float eps) {
  m = beta1 * m + (1.f - beta1) * g;
  v = beta2 * v + (1.f - beta2) * g * g;
  p -= step * m / (sqrtf(v) * inv_sqrt_bc2 + eps);
}

// 128-bit vectorised (float4 on p, g, m, v whenever n % 4 == 0 and the pointers are 16-byte aligned - always the case for
// the flat parameter / gradient buckets of train_step.py), grid sized to the tensors: 28 B/param of pure HBM traffic.
__global__ void __launch_bounds__(256) multi_adam_kernel(TensorTable t, const float* __restrict__ gnorm_part, int nparts,
                                                         float* __restrict__ gnorm_sq_out, float grad_scale, float max_norm,
                                                         float lr, const float* __restrict__ lr_dev, float beta1, float beta2,
                                                         float eps, int step_index, const long long* __restrict__ step_base) {
  // bias corrections from the step count; the count may live on the device so a captured CUDA graph can be replayed
  const double stepd = (double)step_index + (step_base ? (double)(*step_base) : 0.0);
  const float bc1 = (float)(1.0 - pow((double)beta1, stepd)), bc2 = (float)(1.0 - pow((double)beta2, stepd));
  float coef = grad_scale;
  if (gnorm_part) {
    float ss = 0.f;  // every warp sums the partials in the same fixed order
    for (int i = threadIdx.x & 31; i < nparts; i += 32) ss += __ldg(gnorm_part + i);
    ss = warp_sum(ss) * grad_scale * grad_scale;  // squared norm of the (averaged) gradient
    if (gnorm_sq_out && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) gnorm_sq_out[0] = ss;
    coef = grad_scale * fminf(max_norm / (sqrtf(ss) + 1e-6f), 1.f);
  }
  if (lr_dev) lr = __ldg(lr_dev);  // device-side learning rate: a scheduler can change it under a captured CUDA graph
  const float step = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
  float* p = t.p[blockIdx.y]; float* m = t.m[blockIdx.y]; float* v = t.v[blockIdx.y];
  const float* g = t.g[blockIdx.y];
  const long long n = t.n[blockIdx.y];
  const bool vec = (n & 3) == 0 && ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                                     reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0;
  if (vec) {
    float4* p4 = reinterpret_cast<float4*>(p); float4* m4 = reinterpret_cast<float4*>(m); float4* v4 = reinterpret_cast<float4*>(v);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n / 4; i += (long long)gridDim.x * 256) {
      float4 pi = p4[i], mi = m4[i], vi = v4[i];
      const float4 gi = __ldg(g4 + i);
      adam1(pi.x, mi.x, vi.x, gi.x * coef, beta1, beta2, step, inv_sqrt_bc2, eps);
      adam1(pi.y, mi.y, vi.y, gi.y * coef, beta1, beta2, step, inv_sqrt_bc2, eps);
      adam1(pi.z, mi.z, vi.z, gi.z * coef, beta1, beta2, step, inv_sqrt_bc2, eps);
      adam1(pi.w, mi.w, vi.w, gi.w * coef, beta1, beta2, step, inv_sqrt_bc2, eps);
      p4[i] = pi; m4[i] = mi; v4[i] = vi;
    }
  } else {
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
      float pi = p[i], mi = m[i], vi = v[i];
      adam1(pi, mi, vi, g[i] * coef, beta1, beta2, step, inv_sqrt_bc2, eps);
      p[i] = pi; m[i] = mi; v[i] = vi;
    }
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

static unsigned blocks_for(const long long* n, int count, long long per_block, unsigned cap) {
  long long mx = 0;
  for (int i = 0; i < count; ++i) if (n[i] > mx) mx = n[i];
  long long b = (mx + per_block - 1) / per_block;
  if (b > cap) b = cap;
  return (unsigned)(b < 1 ? 1 : b);
}

// part[0..*nparts_host) <- per-block partial sums of g^2 over all `count` tensors (room for 1024 floats); every slot is
// written, none needs zeroing.  b2c_multi_adam sums them in a fixed order.
extern "C" int b2c_multi_sumsq(float* const* g, const long long* n, int count, float* part, int* nparts_host, void* stream) {
  TensorTable t;
  if (int rc = fill_table(t, nullptr, g, nullptr, nullptr, n, count)) return rc;
  const unsigned bx = blocks_for(n, count, 4096, 1024u / (unsigned)count);
  if (nparts_host) *nparts_host = (int)(bx * count);
  multi_sumsq_kernel<<<dim3(bx, count), 256, 0, (cudaStream_t)stream>>>(t, part);
  return b2c_check_launch("multi_sumsq");
}

// One Adam step on `count` tensors.  Gradients are first multiplied by grad_scale (data parallel: 1 / world_size after a
// SUM all-reduce), then clipped: with gnorm_part (nparts partials from b2c_multi_sumsq of the UNSCALED gradients, nullable)
// the coefficient is min(1, max_norm / (|grad_scale * g| + 1e-6)) and gnorm_sq_out[0] (nullable) receives |grad_scale * g|^2.
// lr_dev (device scalar, nullable) overrides lr.  The optimizer's step count AFTER the increment is step_index + *step_base
// (step_base: device scalar, nullable) and must be >= 1.
extern "C" int b2c_multi_adam(float* const* p, float* const* g, float* const* m, float* const* v, const long long* n,
                              int count, const float* gnorm_part, int nparts, float* gnorm_sq_out, float grad_scale,
                              float max_norm, float lr, const float* lr_dev, float beta1, float beta2, float eps,
                              int step_index, const long long* step_base, void* stream) {
  TensorTable t;
  if (int rc = fill_table(t, p, g, m, v, n, count)) return rc;
  const unsigned bx = blocks_for(n, count, 2048, 148u * 16u);
  multi_adam_kernel<<<dim3(bx, count), 256, 0, (cudaStream_t)stream>>>(t, gnorm_part, nparts, gnorm_sq_out, grad_scale,
                                                                       max_norm, lr, lr_dev, beta1, beta2, eps, step_index,
                                                                       step_base);
  return b2c_check_launch("multi_adam");
}
