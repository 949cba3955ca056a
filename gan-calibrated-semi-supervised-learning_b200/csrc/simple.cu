// Kernels that only the alternative generator GeneratorSimpleRegressor needs (cgan/models.py:147-216, selected by
// generator_type: "simple", cgan/config.yaml:21): weight packing for the 3x3 tensor-core conv (tapgemm.cu mode C3S1), the
// first 3x3 conv on the NCHW fp32 patch (Cin = 3, K = 27: bandwidth-bound, CUDA cores), MaxPool2d(2,2) forward / backward, and
// the regression head AdaptiveAvgPool2d(1) -> Linear(512,256) -> ReLU -> Linear(256,64) -> ReLU -> Linear(64,4) -> Tanh.
#include "common.cuh"

using namespace b2c;

// out[co][(ky*3+kx)*Cin + ci] = w[co][ci][ky][kx]  (K-major B operand of mode C3S1); flip != 0 packs the data-gradient form
// out[ci][(ky*3+kx)*Cout + co] = w[co][ci][2-ky][2-kx] (dx = conv3s1(dy, flipped / transposed weights)).
__global__ void pack_w_k3_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int Cout, int Cin, int flip) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)Cout * Cin) return;
  const int ci = (int)(i % Cin), co = (int)(i / Cin);
  const float* src = w + i * 9;
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    if (!flip) out[((long long)co * 9 + t) * Cin + ci] = __float2bfloat16(src[t]);
    else out[((long long)ci * 9 + (8 - t)) * Cout + co] = __float2bfloat16(src[t]);
  }
}
extern "C" int b2c_pack_w_k3(const float* w, void* out_bf16, int Cout, int Cin, int flip, void* stream) {
  const long long n = (long long)Cout * Cin;
  pack_w_k3_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(w, (__nv_bfloat16*)out_bf16, Cout, Cin, flip);
  return b2c_check_launch("pack_w_k3");
}

// y[n,oy,ox,co] = bias[co] + sum_{c,ky,kx} x[n,c,oy+ky-1,ox+kx-1] w[co][c][ky][kx]: one thread per output pixel, 64 channels
// in registers (2 passes of 32), weights (bf16-rounded, like every tensor-core operand of this library) in shared memory.
__global__ void __launch_bounds__(128) conv3x3_first_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                            const float* __restrict__ bias, __nv_bfloat16* __restrict__ y,
                                                            int N, int H, int W) {
  __shared__ float ws[27][64];
  __shared__ float bs[64];
  for (int i = threadIdx.x; i < 27 * 64; i += blockDim.x) {
    const int k = i / 64, co = i % 64;
    ws[k][co] = __bfloat162float(__float2bfloat16(w[co * 27 + k]));
  }
  if (threadIdx.x < 64) bs[threadIdx.x] = bias ? bias[threadIdx.x] : 0.f;
  __syncthreads();
  const long long npix = (long long)N * H * W;
  for (long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += (long long)gridDim.x * blockDim.x) {
    const int ox = (int)(pix % W), oy = (int)((pix / W) % H), n = (int)(pix / ((long long)W * H));
    float f[27];
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const int iy = oy + ky - 1, ix = ox + kx - 1;
          const bool ok = iy >= 0 && iy < H && ix >= 0 && ix < W;
          const float v = ok ? __ldg(x + (((long long)n * 3 + c) * H + iy) * W + ix) : 0.f;
          f[(c * 3 + ky) * 3 + kx] = __bfloat162float(__float2bfloat16(v));  // bf16 operand storage
        }
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float acc[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) acc[j] = bs[half * 32 + j];
#pragma unroll
      for (int k = 0; k < 27; ++k)
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = fmaf(f[k], ws[k][half * 32 + j], acc[j]);
      uint32_t o[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) o[j] = pack_bf16(acc[2 * j], acc[2 * j + 1]);
      __nv_bfloat16* dst = y + pix * 64 + half * 32;
      st_global_v8(dst, o);
      st_global_v8(dst + 16, o + 8);
    }
  }
}
// x: (N,3,H,W) fp32 NCHW; w: [64][3][3][3] fp32; y: dense (N,H,W,64) bf16 (32-byte aligned).
extern "C" int b2c_conv3x3_first(const float* x, const float* w, const float* bias, void* y, int N, int H, int W, void* stream) {
  B2C_REQUIRE(((uintptr_t)y & 31) == 0, "conv3x3_first: output must be 32-byte aligned");
  const long long npix = (long long)N * H * W;
  long long blocks = (npix + 127) / 128;
  if (blocks > 148 * 32) blocks = 148 * 32;
  conv3x3_first_kernel<<<(unsigned)blocks, 128, 0, (cudaStream_t)stream>>>(x, w, bias, (__nv_bfloat16*)y, N, H, W);
  return b2c_check_launch("conv3x3_first");
}

// MaxPool2d(2,2) on NHWC bf16 (cgan/models.py:169,178,187,196): out (N,H/2,W/2,C).  16-byte vectors of 8 channels.
__device__ __forceinline__ uint32_t max_bf16x2(uint32_t a, uint32_t b) {
  const float2 x = unpack_bf16(a), y = unpack_bf16(b);
  return pack_bf16(fmaxf(x.x, y.x), fmaxf(x.y, y.y));
}
__global__ void maxpool2_fwd_kernel(const __nv_bfloat16* __restrict__ a, __nv_bfloat16* __restrict__ out, long long total,
                                    int Ho, int Wo, int cvecs) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(i % cvecs);
    const long long op = i / cvecs;
    const int ox = (int)(op % Wo), oy = (int)((op / Wo) % Ho);
    const long long n = op / ((long long)Wo * Ho);
    const long long C = (long long)cvecs * 8, W = 2ll * Wo;
    const __nv_bfloat16* p = a + ((n * 2 * Ho + 2 * oy) * W + 2 * ox) * C + cv * 8;
    const uint4 v00 = __ldg(reinterpret_cast<const uint4*>(p)), v01 = __ldg(reinterpret_cast<const uint4*>(p + C));
    const uint4 v10 = __ldg(reinterpret_cast<const uint4*>(p + W * C)), v11 = __ldg(reinterpret_cast<const uint4*>(p + W * C + C));
    uint4 r;
    r.x = max_bf16x2(max_bf16x2(v00.x, v01.x), max_bf16x2(v10.x, v11.x));
    r.y = max_bf16x2(max_bf16x2(v00.y, v01.y), max_bf16x2(v10.y, v11.y));
    r.z = max_bf16x2(max_bf16x2(v00.z, v01.z), max_bf16x2(v10.z, v11.z));
    r.w = max_bf16x2(max_bf16x2(v00.w, v01.w), max_bf16x2(v10.w, v11.w));
    *reinterpret_cast<uint4*>(out + op * C + cv * 8) = r;
  }
}
extern "C" int b2c_maxpool2_fwd(const void* a, void* out, int N, int H, int W, int C, void* stream) {
  B2C_REQUIRE(H % 2 == 0 && W % 2 == 0 && C % 8 == 0, "maxpool2: H=%d W=%d must be even, C=%d a multiple of 8", H, W, C);
  const long long total = (long long)N * (H / 2) * (W / 2) * (C / 8);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  maxpool2_fwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)a, (__nv_bfloat16*)out, total, H / 2, W / 2, C / 8);
  return b2c_check_launch("maxpool2_fwd");
}
// Backward: the gradient goes to the FIRST maximum of each 2x2 window in row-major order (torch's max_pool2d tie rule).
__global__ void maxpool2_bwd_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ g,
                                    __nv_bfloat16* __restrict__ ga, long long total, int Ho, int Wo, int C) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long op = i / C;
    const int ox = (int)(op % Wo), oy = (int)((op / Wo) % Ho);
    const long long n = op / ((long long)Wo * Ho);
    const long long W = 2ll * Wo;
    const long long base = ((n * 2 * Ho + 2 * oy) * W + 2 * ox) * C + c;
    const float v[4] = {__bfloat162float(a[base]), __bfloat162float(a[base + C]), __bfloat162float(a[base + W * C]),
                        __bfloat162float(a[base + W * C + C])};
    int best = 0;
#pragma unroll
    for (int k = 1; k < 4; ++k)
      if (v[k] > v[best]) best = k;
    const __nv_bfloat16 gv = g[i], z = __float2bfloat16(0.f);
    ga[base] = best == 0 ? gv : z; ga[base + C] = best == 1 ? gv : z;
    ga[base + W * C] = best == 2 ? gv : z; ga[base + W * C + C] = best == 3 ? gv : z;
  }
}
extern "C" int b2c_maxpool2_bwd(const void* a, const void* g, void* ga, int N, int H, int W, int C, void* stream) {
  B2C_REQUIRE(H % 2 == 0 && W % 2 == 0, "maxpool2_bwd: H=%d W=%d must be even", H, W);
  const long long total = (long long)N * (H / 2) * (W / 2) * C;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  maxpool2_bwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)g, (__nv_bfloat16*)ga, total, H / 2, W / 2, C);
  return b2c_check_launch("maxpool2_bwd");
}

// Regression head, eval mode (dropout is the identity): one block per sample.
//   f = mean_hw a[n,:,:,c] (C = 512); h1 = relu(W1 f + b1) (256); h2 = relu(W2 h1 + b2) (64); delta = tanh(W3 h2 + b3) * scale.
// m1 / m2 (nullable): Dropout(0.5) keep-masks {0,1} of h1 / h2 for train mode (kept values x2).  acts (nullable): [N][512+256+64+4]
// = (f, h1, h2, tanh) saved for a backward pass.
__global__ void __launch_bounds__(256) simple_head_fwd_kernel(const __nv_bfloat16* __restrict__ a, int HW, const float* __restrict__ W1,
                                                              const float* __restrict__ b1, const float* __restrict__ W2,
                                                              const float* __restrict__ b2, const float* __restrict__ W3,
                                                              const float* __restrict__ b3, const uint8_t* __restrict__ m1,
                                                              const uint8_t* __restrict__ m2, float scale,
                                                              float* __restrict__ delta, float* __restrict__ acts) {
  __shared__ float f[512], h1[256], h2[64];
  const int n = blockIdx.x, t = threadIdx.x;
  for (int c = t; c < 512; c += 256) {
    float s = 0.f;
    for (int p = 0; p < HW; ++p) s += __bfloat162float(a[((long long)n * HW + p) * 512 + c]);
    f[c] = s / (float)HW;
  }
  __syncthreads();
  {
    float s = b1[t];
    for (int c = 0; c < 512; ++c) s = fmaf(W1[t * 512 + c], f[c], s);
    s = fmaxf(s, 0.f);
    if (m1) s *= m1[n * 256 + t] ? 2.f : 0.f;
    h1[t] = s;
  }
  __syncthreads();
  if (t < 64) {
    float s = b2[t];
    for (int c = 0; c < 256; ++c) s = fmaf(W2[t * 256 + c], h1[c], s);
    s = fmaxf(s, 0.f);
    if (m2) s *= m2[n * 64 + t] ? 2.f : 0.f;
    h2[t] = s;
  }
  __syncthreads();
  if (t < 4) {
    float s = b3[t];
    for (int c = 0; c < 64; ++c) s = fmaf(W3[t * 64 + c], h2[c], s);
    const float th = tanhf(s);
    delta[n * 4 + t] = th * scale;
    if (acts) acts[(long long)n * 836 + 832 + t] = th;
  }
  if (acts) {
    for (int c = t; c < 512; c += 256) acts[(long long)n * 836 + c] = f[c];
    acts[(long long)n * 836 + 512 + t] = h1[t];
    if (t < 64) acts[(long long)n * 836 + 768 + t] = h2[t];
  }
}
// a: dense (N,h,h,512) bf16; W1 [256][512], W2 [64][256], W3 [4][64] fp32 (+ biases); delta: [N][4].
extern "C" int b2c_simple_head_fwd(const void* a, int HW, const float* W1, const float* b1, const float* W2, const float* b2,
                                   const float* W3, const float* b3, const uint8_t* m1, const uint8_t* m2, float scale,
                                   float* delta, float* acts, int N, void* stream) {
  simple_head_fwd_kernel<<<N, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)a, HW, W1, b1, W2, b2, W3, b3, m1, m2, scale, delta, acts);
  return b2c_check_launch("simple_head_fwd");
}

// ================================================================================================ backward (training)
// First conv (nn.Conv2d(3, 64, 3, padding=1), cgan/models.py:163): dW[co][ci][ky][kx] = sum gz[n,y,x,co] x[n,ci,y+ky-1,x+kx-1].
// K = 27 cannot feed a tensor-core tile; a thread owns 4 output channels x all 27 taps (108 fp32 accumulators kept over all the
// image rows its persistent block walks), so every shared-memory x value it reads feeds 4 FMAs.  256 threads = 16 channel
// quads x 16 column strips; one atomic flush per block.  gz: dense (N,H,W,64) bf16, x: (N,3,H,W) fp32.
__global__ void __launch_bounds__(256) conv3x3_first_wgrad_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ gz,
                                                                  float* __restrict__ dw, int N, int H, int W) {
  extern __shared__ float xs[];  // [3 ci][3 rows][W + 2]
  const int t = threadIdx.x, cq = t & 15, strip = t >> 4;
  const int P = W + 2;
  float acc[4][27];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int k = 0; k < 27; ++k) acc[j][k] = 0.f;
  for (int row = blockIdx.x; row < N * H; row += gridDim.x) {
    const int n = row / H, y = row % H;
    __syncthreads();
    for (int i = t; i < 9 * P; i += 256) {
      const int ci = i / (3 * P), r = (i / P) % 3, xx = i % P - 1;
      const int yy = y + r - 1;
      xs[i] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(x + (((long long)n * 3 + ci) * H + yy) * W + xx) : 0.f;
    }
    __syncthreads();
    const __nv_bfloat16* grow = gz + ((long long)row * W) * 64 + cq * 4;
    for (int px = strip; px < W; px += 16) {
      const uint2 u = __ldg(reinterpret_cast<const uint2*>(grow + (long long)px * 64));
      const float2 g01 = unpack_bf16(u.x), g23 = unpack_bf16(u.y);
      const float g[4] = {g01.x, g01.y, g23.x, g23.y};
#pragma unroll
      for (int k = 0; k < 27; ++k) {
        const float xv = xs[(k / 9) * 3 * P + ((k / 3) % 3) * P + px + (k % 3)];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j][k] = fmaf(g[j], xv, acc[j][k]);
      }
    }
  }
  // strips 2s / 2s+1 share a warp (lanes 0-15 / 16-31): pair them, then one atomic per (co, tap) and warp
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int k = 0; k < 27; ++k) {
      const float v = acc[j][k] + __shfl_xor_sync(0xffffffffu, acc[j][k], 16);
      if ((t & 16) == 0) atomicAdd(dw + (cq * 4 + j) * 27 + k, v);
    }
}
// dw fp32 [64][3][3][3] is ACCUMULATED into (caller zeroes).
extern "C" int b2c_conv3x3_first_wgrad(const float* x, const void* gz, float* dw, int N, int H, int W, void* stream) {
  B2C_REQUIRE(W >= 16 && W <= 2048, "conv3x3_first_wgrad: W=%d out of range", W);
  const size_t smem = sizeof(float) * 9 * (size_t)(W + 2);
  static bool attr_done = false;
  if (!attr_done) {
    cudaFuncSetAttribute(conv3x3_first_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    attr_done = true;
  }
  long long blocks = (long long)N * H;
  if (blocks > 148 * 2) blocks = 148 * 2;
  conv3x3_first_wgrad_kernel<<<(unsigned)blocks, 256, smem, (cudaStream_t)stream>>>(x, (const __nv_bfloat16*)gz, dw, N, H, W);
  return b2c_check_launch("conv3x3_first_wgrad");
}

// Regression head backward (cgan/models.py:199-211), one block per sample: from gdelta (N,4) and the activations the forward
// saved (acts [N][836] = f, h1, h2, tanh) to the per-sample pre-activation gradients gv [N][836] = (gf 512, g1 256, g2 64, g3 4)
// and the gradient of the pooled map, ga[n,p,c] = gf[c] / HW (bf16, dense (N,HW,512)).
__global__ void __launch_bounds__(256) simple_head_bwd_kernel(const float* __restrict__ gdelta, const float* __restrict__ acts,
                                                              const float* __restrict__ W1, const float* __restrict__ W2,
                                                              const float* __restrict__ W3, const uint8_t* __restrict__ m1,
                                                              const uint8_t* __restrict__ m2, float scale, int HW,
                                                              float* __restrict__ gv, __nv_bfloat16* __restrict__ ga) {
  __shared__ float g3[4], g2[64], g1[256], gf[512];
  const int n = blockIdx.x, t = threadIdx.x;
  const float* A = acts + (long long)n * 836;
  if (t < 4) {
    const float th = A[832 + t];
    g3[t] = gdelta[n * 4 + t] * scale * (1.f - th * th);
  }
  __syncthreads();
  if (t < 64) {
    float s = 0.f;
#pragma unroll
    for (int o = 0; o < 4; ++o) s = fmaf(W3[o * 64 + t], g3[o], s);
    const float keep = m2 ? (m2[n * 64 + t] ? 2.f : 0.f) : 1.f;
    g2[t] = A[768 + t] > 0.f ? s * keep : 0.f;
  }
  __syncthreads();
  {
    float s = 0.f;
    for (int o = 0; o < 64; ++o) s = fmaf(W2[o * 256 + t], g2[o], s);
    const float keep = m1 ? (m1[n * 256 + t] ? 2.f : 0.f) : 1.f;
    g1[t] = A[512 + t] > 0.f ? s * keep : 0.f;
  }
  __syncthreads();
  for (int c = t; c < 512; c += 256) {
    float s = 0.f;
    for (int o = 0; o < 256; ++o) s = fmaf(W1[o * 512 + c], g1[o], s);
    gf[c] = s;
  }
  __syncthreads();
  float* G = gv + (long long)n * 836;
  for (int c = t; c < 512; c += 256) G[c] = gf[c];
  G[512 + t] = g1[t];
  if (t < 64) G[768 + t] = g2[t];
  if (t < 4) G[832 + t] = g3[t];
  const float inv = 1.f / (float)HW;
  for (int i = t; i < HW * 256; i += 256) {  // two channels per thread and store
    const int p = i >> 8, c = (i & 255) * 2;
    *reinterpret_cast<uint32_t*>(ga + ((long long)n * HW + p) * 512 + c) = pack_bf16(gf[c] * inv, gf[c + 1] * inv);
  }
}
// Weight / bias gradients of the three Linear layers as fixed-order sums over the batch: dW[o][c] (+)= sum_n g[n][o] in[n][c].
__global__ void __launch_bounds__(256) simple_head_wgrad_kernel(const float* __restrict__ acts, const float* __restrict__ gv,
                                                                float* dW1, float* db1, float* dW2, float* db2, float* dW3,
                                                                float* db3, int N, int accumulate) {
  const int total = 256 * 512 + 256 + 64 * 256 + 64 + 4 * 64 + 4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    int j = i, go, io, stride_o;
    float* dst;
    // (offset of the layer's g inside gv, offset of its input inside acts, -1: bias)
    if (j < 256 * 512) { go = 512 + j / 512; io = j % 512; dst = dW1 + j; }
    else if ((j -= 256 * 512) < 256) { go = 512 + j; io = -1; dst = db1 + j; }
    else if ((j -= 256) < 64 * 256) { go = 768 + j / 256; io = 512 + j % 256; dst = dW2 + j; }
    else if ((j -= 64 * 256) < 64) { go = 768 + j; io = -1; dst = db2 + j; }
    else if ((j -= 64) < 4 * 64) { go = 832 + j / 64; io = 768 + j % 64; dst = dW3 + j; }
    else { j -= 4 * 64; go = 832 + j; io = -1; dst = db3 + j; }
    (void)stride_o;
    float s = 0.f;
    for (int n = 0; n < N; ++n) s = fmaf(gv[(long long)n * 836 + go], io >= 0 ? acts[(long long)n * 836 + io] : 1.f, s);
    *dst = accumulate ? *dst + s : s;
  }
}
// gdelta (N,4) fp32; acts / gv: [N][836] fp32 (gv is scratch + output of the per-sample gradients); ga: dense (N,HW,512) bf16;
// dW1 [256][512], dW2 [64][256], dW3 [4][64] and biases: overwritten or (accumulate) added to.
extern "C" int b2c_simple_head_bwd(const float* gdelta, const float* acts, const float* W1, const float* W2, const float* W3,
                                   const uint8_t* m1, const uint8_t* m2, float scale, int HW, float* gv, void* ga, float* dW1,
                                   float* db1, float* dW2, float* db2, float* dW3, float* db3, int accumulate, int N,
                                   void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  simple_head_bwd_kernel<<<N, 256, 0, st>>>(gdelta, acts, W1, W2, W3, m1, m2, scale, HW, gv, (__nv_bfloat16*)ga);
  int rc = b2c_check_launch("simple_head_bwd");
  if (rc) return rc;
  simple_head_wgrad_kernel<<<148 * 2, 256, 0, st>>>(acts, gv, dW1, db1, dW2, db2, dW3, db3, N, accumulate);
  return b2c_check_launch("simple_head_wgrad");
}
