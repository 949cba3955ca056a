// Loss-side kernels: delta->box transform and EIoU with their gradients (cgan/losses.py:19-73,99-150), IoU metric
// (:152-183), gradient-penalty reductions (:223-231), score-map means for the Wasserstein terms
// (cgan/cgan_train_enhanced.py:327-328,362), and G's pooled head AvgPool->Linear(64,4)->Tanh->*delta_scale
// (cgan/models.py:118-123,139-141).  All fp32 and tiny; gradients use forward-mode duals (4 partials), which keeps
// the chain sigmoid-clamp -> exp -> sigmoid-clamp -> EIoU exact without hand-derived formulas.
#include "common.cuh"

using namespace b2c;

struct D4 {
  float v, d[4];
};
__device__ __forceinline__ D4 dconst(float v) { D4 r; r.v = v; r.d[0] = r.d[1] = r.d[2] = r.d[3] = 0.f; return r; }
__device__ __forceinline__ D4 dvar(float v, int i) { D4 r = dconst(v); r.d[i] = 1.f; return r; }
__device__ __forceinline__ D4 operator+(D4 a, D4 b) { D4 r; r.v = a.v + b.v; for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] + b.d[i]; return r; }
__device__ __forceinline__ D4 operator-(D4 a, D4 b) { D4 r; r.v = a.v - b.v; for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] - b.d[i]; return r; }
__device__ __forceinline__ D4 operator*(D4 a, D4 b) { D4 r; r.v = a.v * b.v; for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i]; return r; }
__device__ __forceinline__ D4 operator/(D4 a, D4 b) {
  D4 r; r.v = a.v / b.v;
  for (int i = 0; i < 4; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) / b.v;
  return r;
}
__device__ __forceinline__ D4 operator*(D4 a, float s) { D4 r; r.v = a.v * s; for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] * s; return r; }
__device__ __forceinline__ D4 operator+(D4 a, float s) { a.v += s; return a; }
__device__ __forceinline__ D4 dchain(D4 a, float v, float dv) { D4 r; r.v = v; for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] * dv; return r; }
__device__ __forceinline__ D4 dsigmoid(D4 a) { const float s = 1.f / (1.f + expf(-a.v)); return dchain(a, s, s * (1.f - s)); }
__device__ __forceinline__ D4 dexp(D4 a) { const float e = expf(a.v); return dchain(a, e, e); }
__device__ __forceinline__ D4 dclamp(D4 a, float lo, float hi) {
  return dchain(a, fminf(fmaxf(a.v, lo), hi), (a.v >= lo && a.v <= hi) ? 1.f : 0.f);
}
__device__ __forceinline__ D4 dmax(D4 a, D4 b) { return a.v > b.v ? a : (a.v < b.v ? b : (a + b) * 0.5f); }
__device__ __forceinline__ D4 dmin(D4 a, D4 b) { return a.v < b.v ? a : (a.v > b.v ? b : (a + b) * 0.5f); }
__device__ __forceinline__ D4 drelu(D4 a) { return dchain(a, fmaxf(a.v, 0.f), a.v >= 0.f ? 1.f : 0.f); }
__device__ __forceinline__ D4 dsmooth_clamp(D4 x, float lo, float hi) {  // losses.py:99-106, temperature 0.5
  return dsigmoid((x + (-(lo + hi) * 0.5f)) * 2.f) * (hi - lo) + lo;
}

__device__ __forceinline__ void apply_delta_dual(const float* bb, const D4 (&dl)[4], int training, D4 (&out)[4]) {
  D4 dc[4];
  for (int i = 0; i < 4; ++i) dc[i] = training ? dsmooth_clamp(dl[i], -1.5f, 1.5f) : dclamp(dl[i], -1.5f, 1.5f);
  D4 cx = dc[0] * bb[2] + bb[0], cy = dc[1] * bb[3] + bb[1];
  D4 w = dexp(dclamp(dc[2], -1.f, 1.f)) * bb[2], h = dexp(dclamp(dc[3], -1.f, 1.f)) * bb[3];
  if (training) {
    out[0] = dsmooth_clamp(cx, 0.05f, 0.95f); out[1] = dsmooth_clamp(cy, 0.05f, 0.95f);
    out[2] = dsmooth_clamp(w, 0.02f, 0.8f); out[3] = dsmooth_clamp(h, 0.02f, 0.8f);
  } else {
    out[0] = dclamp(cx, 0.05f, 0.95f); out[1] = dclamp(cy, 0.05f, 0.95f);
    out[2] = dclamp(w, 0.02f, 0.8f); out[3] = dclamp(h, 0.02f, 0.8f);
  }
}

// out = apply_delta_to_bbox(bbox, delta, training); if gout: gdelta[b][j] = sum_i gout[b][i] d out_i / d delta_j
__global__ void apply_delta_kernel(const float* __restrict__ bbox, const float* __restrict__ delta, int training,
                                   float* __restrict__ out, const float* __restrict__ gout, float* __restrict__ gdelta, int B) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  float bb[4];
  D4 dl[4], o[4];
  for (int i = 0; i < 4; ++i) { bb[i] = bbox[b * 4 + i]; dl[i] = dvar(delta[b * 4 + i], i); }
  apply_delta_dual(bb, dl, training, o);
  if (out) for (int i = 0; i < 4; ++i) out[b * 4 + i] = o[i].v;
  if (gout && gdelta)
    for (int j = 0; j < 4; ++j) {
      float g = 0.f;
      for (int i = 0; i < 4; ++i) g += gout[b * 4 + i] * o[i].d[j];
      gdelta[b * 4 + j] = g;
    }
}

__device__ __forceinline__ D4 eiou_dual(const D4 (&p)[4], const float* t, float eps, D4* iou_out) {
  D4 px1 = p[0] - p[2] * 0.5f, py1 = p[1] - p[3] * 0.5f, px2 = p[0] + p[2] * 0.5f, py2 = p[1] + p[3] * 0.5f;
  D4 tx1 = dconst(t[0] - t[2] / 2), ty1 = dconst(t[1] - t[3] / 2), tx2 = dconst(t[0] + t[2] / 2), ty2 = dconst(t[1] + t[3] / 2);
  D4 inter = drelu(dmin(px2, tx2) - dmax(px1, tx1)) * drelu(dmin(py2, ty2) - dmax(py1, ty1));
  D4 uni = (px2 - px1) * (py2 - py1) + (tx2 - tx1) * (ty2 - ty1) - inter;
  D4 iou = inter / (uni + eps);
  if (iou_out) *iou_out = iou;
  D4 ew = dmax(px2, tx2) - dmin(px1, tx1), eh = dmax(py2, ty2) - dmin(py1, ty1);
  D4 dx = p[0] + (-t[0]), dy = p[1] + (-t[1]), dw = p[2] + (-t[2]), dh = p[3] + (-t[3]);
  return iou - (dx * dx + dy * dy) / (ew * ew + eh * eh + eps) - (dw * dw) / (ew * ew + eps) - (dh * dh) / (eh * eh + eps);
}

// loss[0] = 1 - mean(eiou).  If gpred: gpred[b][j] = gscale * d loss / d pred[b][j].  Single block.
__global__ void eiou_kernel(const float* __restrict__ pred, const float* __restrict__ tgt, float eps, float* __restrict__ loss,
                            const float* __restrict__ gscale, float* __restrict__ gpred, float* __restrict__ iou_out, int B) {
  __shared__ float sh[32];
  float acc = 0.f;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    D4 p[4];
    float t[4];
    for (int i = 0; i < 4; ++i) { p[i] = dvar(pred[b * 4 + i], i); t[i] = tgt[b * 4 + i]; }
    D4 iou;
    D4 e = eiou_dual(p, t, eps, &iou);
    acc += e.v;
    if (iou_out) iou_out[b] = iou.v;
    if (gpred) {
      const float gs = (gscale ? *gscale : 1.f) * (-1.f / (float)B);
      for (int j = 0; j < 4; ++j) gpred[b * 4 + j] = gs * e.d[j];
    }
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0 && loss) {
    float s = 0.f;
    for (int w = 0; w < (blockDim.x + 31) / 32; ++w) s += sh[w];
    loss[0] = 1.f - s / (float)B;
  }
}

extern "C" int b2c_apply_delta(const float* bbox, const float* delta, int training, float* out, const float* gout,
                               float* gdelta, int B, void* stream) {
  apply_delta_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(bbox, delta, training, out, gout, gdelta, B);
  return b2c_check_launch("apply_delta");
}
extern "C" int b2c_eiou(const float* pred, const float* tgt, float eps, float* loss, const float* gscale, float* gpred,
                        float* iou_out, int B, void* stream) {
  eiou_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(pred, tgt, eps, loss, gscale, gpred, iou_out, B);
  return b2c_check_launch("eiou");
}

// ---------------------------------------------------------------- reductions
// out[b] += scale * sum_i op(x[b*n + i]), op = square or identity.  grid = (splits, B).
template <int SQUARE>
__global__ void __launch_bounds__(256) rowsum_kernel(const float* __restrict__ x, long long n, float* __restrict__ out, float scale) {
  __shared__ float sh[8];
  const float* row = x + (long long)blockIdx.y * n;
  float acc = 0.f;
  const long long n4 = n / 4;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(row) + i);
    acc += SQUARE ? (v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w) : (v.x + v.y + v.z + v.w);
  }
  if (blockIdx.x == 0)
    for (long long i = n4 * 4 + threadIdx.x; i < n; i += 256) acc += SQUARE ? row[i] * row[i] : row[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += sh[w];
    atomicAdd(out + blockIdx.y, s * scale);
  }
}
// rows of length n (16-byte aligned rows: n % 4 == 0 or B == 1).  out[B] is ACCUMULATED into.
extern "C" int b2c_rowsum(const float* x, long long n, int B, int square, float scale, float* out, void* stream) {
  B2C_REQUIRE(B == 1 || n % 4 == 0, "rowsum: n=%lld must be a multiple of 4 when B>1", n);
  long long splits = (n / 4 + 2047) / 2048;
  const long long cap = 1184 / (B > 0 ? B : 1) + 1;
  if (splits > cap) splits = cap;
  if (splits < 1) splits = 1;
  dim3 grid((unsigned)splits, B);
  if (square) rowsum_kernel<1><<<grid, 256, 0, (cudaStream_t)stream>>>(x, n, out, scale);
  else rowsum_kernel<0><<<grid, 256, 0, (cudaStream_t)stream>>>(x, n, out, scale);
  return b2c_check_launch("rowsum");
}

// Per-sample sum of squares of the two input-gradient tensors (cgan/losses.py:223-226), stage 1: part[b][split], no atomics.
__global__ void __launch_bounds__(256) gp_sumsq_kernel(const float* __restrict__ g0, const float* __restrict__ g1, long long n4,
                                                        float* __restrict__ part) {
  __shared__ float sh[8];
  const float4* r0 = reinterpret_cast<const float4*>(g0) + (long long)blockIdx.y * n4;
  const float4* r1 = g1 ? reinterpret_cast<const float4*>(g1) + (long long)blockIdx.y * n4 : nullptr;
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 v = __ldg(r0 + i);
    acc += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
    if (r1) {
      const float4 w = __ldg(r1 + i);
      acc += (w.x * w.x + w.y * w.y) + (w.z * w.z + w.w * w.w);
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
// g0, g1 (nullable): [B][n] fp32, n % 4 == 0; part: [B][splits] with splits = *splits_host (<= 32; room for 32 * B floats).
extern "C" int b2c_gp_sumsq(const float* g0, const float* g1, long long n, int B, float* part, int* splits_host, void* stream) {
  B2C_REQUIRE(n % 4 == 0, "gp_sumsq: n=%lld must be a multiple of 4", n);
  long long splits = (n / 4 + 4095) / 4096;
  const long long cap = 1184 / (B > 0 ? B : 1) + 1;
  if (splits > cap) splits = cap;
  if (splits > 32) splits = 32;
  if (splits < 1) splits = 1;
  if (splits_host) *splits_host = (int)splits;
  gp_sumsq_kernel<<<dim3((unsigned)splits, B), 256, 0, (cudaStream_t)stream>>>(g0, g1, n / 4, part);
  return b2c_check_launch("gp_sumsq");
}

// Gradient penalty tail (losses.py:226-231): r_b = sqrt(ss_b + 1e-12); gp = mean((r_b-1)^2); coef_b = (2/B)(1-1/r_b);
// ss_b = fixed-order sum of ss[b][0..splits).
__global__ void gp_tail_kernel(const float* __restrict__ ss, int splits, float* __restrict__ gp, float* __restrict__ coef, int B) {
  __shared__ float sh[32];
  float acc = 0.f;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    float t = 0.f;
    for (int s = 0; s < splits; ++s) t += ss[b * splits + s];
    const float r = sqrtf(t + 1e-12f);
    acc += (r - 1.f) * (r - 1.f);
    coef[b] = (2.f / (float)B) * (1.f - 1.f / r);
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < (blockDim.x + 31) / 32; ++w) s += sh[w];
    gp[0] = s / (float)B;
  }
}
extern "C" int b2c_gp_tail(const float* ss, int splits, float* gp, float* coef, int B, void* stream) {
  gp_tail_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(ss, splits, gp, coef, B);
  return b2c_check_launch("gp_tail");
}

// out[b*n + i] = x[b*n + i] * coef[b] * (*gscale)      (seed of the GP double backward)
__global__ void scale_rows_kernel(const float* __restrict__ x, const float* __restrict__ coef, const float* __restrict__ gscale,
                                  float* __restrict__ out, long long n4, int B) {
  const float gs = gscale ? *gscale : 1.f;
  const long long total = n4 * B;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const float c = __ldg(coef + (int)(i / n4)) * gs;
    float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
    v.x *= c; v.y *= c; v.z *= c; v.w *= c;
    reinterpret_cast<float4*>(out)[i] = v;
  }
}
extern "C" int b2c_scale_rows(const float* x, const float* coef, const float* gscale, float* out, long long n, int B, void* stream) {
  B2C_REQUIRE(n % 4 == 0, "scale_rows: n=%lld must be a multiple of 4", n);
  long long blocks = (n / 4 * B + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  scale_rows_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, coef, gscale, out, n / 4, B);
  return b2c_check_launch("scale_rows");
}

// out[i] = (*gscale) * value : gradient of a mean over a score map.
__global__ void fill_scaled_kernel(float* __restrict__ out, long long n, const float* __restrict__ gscale, float value) {
  const float v = (gscale ? *gscale : 1.f) * value;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) out[i] = v;
}
extern "C" int b2c_fill_scaled(float* out, long long n, const float* gscale, float value, void* stream) {
  long long blocks = (n + 255) / 256;
  if (blocks > 592) blocks = 592;
  fill_scaled_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(out, n, gscale, value);
  return b2c_check_launch("fill_scaled");
}

// ---------------------------------------------------------------- Wasserstein terms (cgan/cgan_train_enhanced.py:327-328,362)
__device__ __forceinline__ float block_sum_1024(float v, float* sh) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = 0.f;
  if (threadIdx.x < 32) {
    r = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
    r = warp_sum(r);
    if (threadIdx.x == 0) sh[0] = r;
  }
  __syncthreads();
  r = sh[0];
  __syncthreads();
  return r;
}
// out[0] = scale * sum(x) (single block, fixed order: deterministic): torch.mean over a PatchGAN score map.
__global__ void __launch_bounds__(1024) map_mean_kernel(const float* __restrict__ x, long long n, float scale, float* __restrict__ out) {
  __shared__ float sh[32];
  float acc = 0.f;
  for (long long i = threadIdx.x; i < n; i += 1024) acc += __ldg(x + i);
  const float t = block_sum_1024(acc, sh);
  if (threadIdx.x == 0) out[0] = t * scale;
}
extern "C" int b2c_map_mean(const float* x, long long n, float scale, float* out, void* stream) {
  map_mean_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(x, n, scale, out);
  return b2c_check_launch("map_mean");
}
// The critic objective in one launch: wd = mean(real) - mean(fake); d_loss = -wd + lambda_gp * gp.  out[0..2] = (d_loss, wd, gp).
__global__ void __launch_bounds__(1024) critic_loss_kernel(const float* __restrict__ real, long long n_real,
                                                           const float* __restrict__ fake, long long n_fake,
                                                           const float* __restrict__ gp, float lambda_gp, float* __restrict__ out) {
  __shared__ float sh[32];
  float a = 0.f, b = 0.f;
  for (long long i = threadIdx.x; i < n_real; i += 1024) a += __ldg(real + i);
  for (long long i = threadIdx.x; i < n_fake; i += 1024) b += __ldg(fake + i);
  const float sr = block_sum_1024(a, sh), sf = block_sum_1024(b, sh);
  if (threadIdx.x == 0) {
    const float wd = sr / (float)n_real - sf / (float)n_fake, g = gp ? gp[0] : 0.f;
    out[0] = -wd + lambda_gp * g; out[1] = wd; out[2] = g;
  }
}
extern "C" int b2c_critic_loss(const float* real, long long n_real, const float* fake, long long n_fake, const float* gp,
                               float lambda_gp, float* out3, void* stream) {
  critic_loss_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(real, n_real, fake, n_fake, gp, lambda_gp, out3);
  return b2c_check_launch("critic_loss");
}
// Its backward for the two score maps: g_real[i] = -(*gscale) / n_real, g_fake[i] = +(*gscale) / n_fake.
__global__ void critic_loss_bwd_kernel(const float* __restrict__ gscale, float* __restrict__ g_real, long long n_real,
                                       float* __restrict__ g_fake, long long n_fake) {
  const float g = gscale ? *gscale : 1.f;
  const float vr = -g / (float)n_real, vf = g / (float)n_fake;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_real + n_fake; i += (long long)gridDim.x * blockDim.x) {
    if (i < n_real) g_real[i] = vr; else g_fake[i - n_real] = vf;
  }
}
extern "C" int b2c_critic_loss_bwd(const float* gscale, float* g_real, long long n_real, float* g_fake, long long n_fake,
                                   void* stream) {
  long long blocks = (n_real + n_fake + 255) / 256;
  if (blocks > 592) blocks = 592;
  critic_loss_bwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(gscale, g_real, n_real, g_fake, n_fake);
  return b2c_check_launch("critic_loss_bwd");
}
// cudaMemsetAsync through the C-ABI (gradient buckets are zeroed with one memset node instead of a fill kernel).
extern "C" int b2c_zero(void* p, long long bytes, void* stream) {
  cudaError_t e = cudaMemsetAsync(p, 0, (size_t)bytes, (cudaStream_t)stream);
  if (e != cudaSuccess) { b2c_set_error("b2c_zero: %s", cudaGetErrorString(e)); return B2C_ERR_CUDA; }
  return B2C_OK;
}

// ---------------------------------------------------------------- G head: pooled -> Linear(C,4) -> tanh -> * scale
// pooled_sums[N][C] are plane SUMS of the activated up4 output (norm.cu M_POOL) in 64-bit fixed point (stat_get); HW divides them.
__global__ void pool_fc_fwd_kernel(const long long* __restrict__ ps, const float* __restrict__ W, const float* __restrict__ bias,
                                   float scale, float inv_hw, float* __restrict__ delta, float* __restrict__ tanh_out, int N, int C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * 4) return;
  const int n = i >> 2, j = i & 3;
  float acc = bias[j];
  for (int c = 0; c < C; ++c) acc += W[j * C + c] * stat_get(ps[n * C + c]) * inv_hw;
  const float t = tanhf(acc);
  tanh_out[i] = t;
  delta[i] = t * scale;
}
// gdelta[N][4] -> dW[4][C], db[4] (overwritten), gvec[N][C] = d loss / d (each element of the activated plane)
__global__ void pool_fc_bwd_kernel(const float* __restrict__ gdelta, const float* __restrict__ tanh_v, const long long* __restrict__ ps,
                                   const float* __restrict__ W, float scale, float inv_hw, float* __restrict__ dW,
                                   float* __restrict__ db, float* __restrict__ gvec, int N, int C) {
  extern __shared__ float dpre[];  // [N][4]
  for (int i = threadIdx.x; i < N * 4; i += blockDim.x) dpre[i] = gdelta[i] * scale * (1.f - tanh_v[i] * tanh_v[i]);
  __syncthreads();
  for (int i = threadIdx.x; i < 4 * C; i += blockDim.x) {
    const int j = i / C, c = i % C;
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc += dpre[n * 4 + j] * stat_get(ps[n * C + c]) * inv_hw;
    dW[i] = acc;
  }
  if (threadIdx.x < 4) {
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc += dpre[n * 4 + threadIdx.x];
    db[threadIdx.x] = acc;
  }
  for (int i = threadIdx.x; i < N * C; i += blockDim.x) {
    const int n = i / C, c = i % C;
    float acc = 0.f;
    for (int j = 0; j < 4; ++j) acc += dpre[n * 4 + j] * W[j * C + c];
    gvec[i] = acc * inv_hw;
  }
}
extern "C" int b2c_pool_fc_fwd(const long long* pooled_sums, const float* W, const float* bias, float scale, int HW,
                               float* delta, float* tanh_out, int N, int C, void* stream) {
  pool_fc_fwd_kernel<<<(N * 4 + 127) / 128, 128, 0, (cudaStream_t)stream>>>(pooled_sums, W, bias, scale, 1.f / (float)HW, delta, tanh_out, N, C);
  return b2c_check_launch("pool_fc_fwd");
}
extern "C" int b2c_pool_fc_bwd(const float* gdelta, const float* tanh_v, const long long* pooled_sums, const float* W,
                               float scale, int HW, float* dW, float* db, float* gvec, int N, int C, void* stream) {
  B2C_REQUIRE(N * 16 <= 48 * 1024, "pool_fc_bwd: N=%d too large", N);
  pool_fc_bwd_kernel<<<1, 512, N * 4 * sizeof(float), (cudaStream_t)stream>>>(gdelta, tanh_v, pooled_sums, W, scale, 1.f / (float)HW, dW, db, gvec, N, C);
  return b2c_check_launch("pool_fc_bwd");
}
