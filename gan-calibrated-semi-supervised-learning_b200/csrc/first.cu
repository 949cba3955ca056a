// First-layer convolutions (Cin = 3 for G.down1, 6 for D.model.0; cgan/models.py:103,245,257) on tcgen05.
//
// K = 16*Cin is only 48/96, so these layers are HBM-bound: the job of the kernels is to stream the
// caller's NCHW fp32 patches exactly once and keep the arithmetic off the CUDA cores.  Each thread owns one output
// pixel of a 128-pixel tile: it gathers its 4x4xCin receptive field with coalesced 8-byte loads (+ warp shuffles for
// the two halo columns), applies the optional GP interpolation x = alpha*a + (1-alpha)*b (cgan/losses.py:203-204) and
// the channel concat (cgan/models.py:257), and writes one im2col row k = (c*4 + ky)*4 + kx (bf16, the weight tensor's
// own order, K = 16*Cin, no padding) straight into the 128B-swizzled UMMA layout in shared memory.  The loads of tile
// i+1 are issued before the epilogue of tile i and consumed after it, so their latency hides behind it.  That tile is
//   fprop : the K-major  A operand of  y[128 pix x 64 co]  = A . W^T      (+bias, LeakyReLU, bf16 NHWC store)
//   wgrad : the MN-major A operand of  dW[128 k  x 64 co] += A^T . gz     (gz tile by TMA; fp32 TMEM accumulation over
//           all tiles of the CTA; one atomic flush per CTA; bias gradient = the all-ones im2col column k = 127)
// dgrad (needed only under the gradient penalty) lives in tapgemm.cu (mode D1).
#include "common.cuh"
#include "tmap.h"

using namespace b2c;

struct FirstIn {
  const float* a[2];   // up to two (N,3,H,W) fp32 tensors = channels 0-2 and 3-5
  const float* b[2];   // optional interpolation partners
  const float* alpha;  // optional (N) : x = alpha*a + (1-alpha)*b
};

static constexpr int FT_THREADS = 128;
static constexpr int FT_A_BYTES = 2 * 128 * 128;   // [2 k-blocks][128 rows][64 bf16]
static constexpr int FT_B_BYTES = 128 * 128;       // fprop: [2 k-blocks][64 co][64 k] ; wgrad: [128 pix][64 co]
static constexpr int FT_SMEM = FT_A_BYTES + 2 * FT_B_BYTES + 1024 + 64;   // wgrad (double-buffered gz tile)
static constexpr int FT_SMEM_FPROP = FT_A_BYTES + FT_B_BYTES + 1024 + 64;

// Raw receptive field of one output pixel as it comes back from global memory: v[c][ky] = (x[2ox], x[2ox+1]);
// h[c][ky] = the halo element only the warp-edge lanes fetch (lane 0: x[2ox-1], lane 31: x[2ox+2]; 0 elsewhere) -
// every other lane takes its two halo columns from its neighbours by shuffle.
template <int CIN, bool INTERP>
struct Raw {
  float2 v[CIN][4];
  float h[CIN][4];
  float2 u[INTERP ? CIN : 1][4];
  float hb[INTERP ? CIN : 1][4];
  float al;
  int ox;
};

// Phase 1: issue every global load of the field (4*CIN 8-byte row loads + one predicated halo load each); nothing here
// depends on a loaded value, so the whole field is in flight while the caller runs the previous tile's epilogue.
template <int CIN, bool INTERP>
__device__ __forceinline__ void issue_field(const FirstIn& in, int pix, int npix, int Ho, int Wo, int H, int W,
                                            Raw<CIN, INTERP>& f) {
  const int lane = threadIdx.x & 31;
  const bool live = pix < npix;
  int ox = 0, oy = 0, n = 0;
  if (live) {
    ox = pix % Wo;
    const int q = pix / Wo;
    oy = q % Ho;
    n = q / Ho;
  }
  f.ox = ox;
  f.al = (INTERP && live) ? __ldg(in.alpha + n) : 1.f;
  const bool want = live && ((lane == 0 && ox > 0) || (lane == 31 && ox < Wo - 1));
  const int hoff = lane == 0 ? -1 : 2;
  const long long img = (long long)n * 3 * H * W + 2 * ox;
#pragma unroll
  for (int c = 0; c < CIN; ++c) {
    const int part = c / 3, cc = c % 3;
    const float* pa0 = in.a[part] + img;
    const float* pb0 = INTERP ? in.b[part] + img : nullptr;
#pragma unroll
    for (int ky = 0; ky < 4; ++ky) {
      const int iy = 2 * oy + ky - 1;
      const bool row_ok = live && iy >= 0 && iy < H;
      const int off = row_ok ? (cc * H + iy) * W : 0;
      f.v[c][ky] = row_ok ? __ldg(reinterpret_cast<const float2*>(pa0 + off)) : make_float2(0.f, 0.f);
      f.h[c][ky] = (row_ok && want) ? __ldg(pa0 + off + hoff) : 0.f;
      if (INTERP) {
        f.u[c][ky] = row_ok ? __ldg(reinterpret_cast<const float2*>(pb0 + off)) : make_float2(0.f, 0.f);
        f.hb[c][ky] = (row_ok && want) ? __ldg(pb0 + off + hoff) : 0.f;
      }
    }
  }
}

// Phase 2: optional GP interpolation, halo exchange, bf16 packing and the swizzled im2col row of pixel `row`:
// k = (c*4 + ky)*4 + kx, i.e. the weight tensor's own [c][ky][kx] order, so a 16-byte chunk is two (c,ky) groups of
// (left, x0, x1, right).  K = 16*CIN: 2*CIN chunks, no channel padding.
template <int CIN, bool INTERP>
__device__ __forceinline__ void pack_store_field(uint8_t* sA, int row, int Wo, const Raw<CIN, INTERP>& f) {
  const int lane = threadIdx.x & 31;
  const bool edge_l = lane == 0 || f.ox == 0, edge_r = lane == 31 || f.ox == Wo - 1;
  const float al = f.al, be = 1.f - f.al;
  uint8_t* base = sA + (row >> 3) * 1024 + (row & 7) * 128;
  uint32_t p[4];
#pragma unroll
  for (int c = 0; c < CIN; ++c) {
#pragma unroll
    for (int ky = 0; ky < 4; ++ky) {
      float2 t = f.v[c][ky];
      float hh = f.h[c][ky];
      if (INTERP) {
        t.x = al * t.x + be * f.u[c][ky].x;
        t.y = al * t.y + be * f.u[c][ky].y;
        hh = al * hh + be * f.hb[c][ky];
      }
      const float lft = __shfl_up_sync(0xffffffffu, t.y, 1);
      const float rgt = __shfl_down_sync(0xffffffffu, t.x, 1);
      const float l = edge_l ? (lane == 0 ? hh : 0.f) : lft;
      const float r = edge_r ? (lane == 31 ? hh : 0.f) : rgt;
      const int g = c * 4 + ky;
      p[(g & 1) * 2] = pack_bf16(l, t.x);
      p[(g & 1) * 2 + 1] = pack_bf16(t.y, r);
      if (g & 1) {
        const int j = g >> 1;
        *reinterpret_cast<uint4*>(base + (j >> 3) * 16384 + (((j & 7) ^ (row & 7)) << 4)) = make_uint4(p[0], p[1], p[2], p[3]);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ fprop
template <int CIN, bool INTERP>
__global__ void __launch_bounds__(FT_THREADS, INTERP ? 2 : 4) first_fprop_tc_kernel(FirstIn in, const float* __restrict__ w,
                                                                    const float* __restrict__ inv_sigma,
                                                                    const float* __restrict__ bias,
                                                                    __nv_bfloat16* __restrict__ y, int N, int H, int W,
                                                                    float slope) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + FT_A_BYTES;
  uint64_t* mma_bar = reinterpret_cast<uint64_t*>(sB + FT_B_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mma_bar + 1);
  __shared__ float bias_s[64];

  const int t = threadIdx.x, warp = t >> 5;
  const int Ho = H / 2, Wo = W / 2;
  const int npix = N * Ho * Wo;   // host checks < 2^31 - 128
  const int ntiles = (npix + 127) / 128;

  if (t == 0) {
    mbar_init(mma_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<64>(tmem_slot);
  // weights: B[co][k] = bf16(w[co][c][ky][kx]) in the tensor's own order, K-major, k-blocks of [64 rows x 128 B]; the
  // spectral-norm 1/sigma multiplies the fp32 accumulators in the epilogue (same rounding point as the tapgemm kernels)
  const float osc = inv_sigma ? __ldg(inv_sigma) : 1.f;
  {
    const float s = 1.f;
    constexpr int NCH = 2 * CIN;
    for (int i = t; i < 64 * NCH; i += FT_THREADS) {
      const int co = i / NCH, j = i % NCH;
      const float4 lo = __ldg(reinterpret_cast<const float4*>(w + (long long)co * (16 * CIN) + j * 8));
      const float4 hi = __ldg(reinterpret_cast<const float4*>(w + (long long)co * (16 * CIN) + j * 8 + 4));
      const uint4 u = make_uint4(pack_bf16(lo.x * s, lo.y * s), pack_bf16(lo.z * s, lo.w * s), pack_bf16(hi.x * s, hi.y * s),
                                 pack_bf16(hi.z * s, hi.w * s));
      *reinterpret_cast<uint4*>(sB + (j >> 3) * 8192 + (co >> 3) * 1024 + (co & 7) * 128 + (((j & 7) ^ (co & 7)) << 4)) = u;
    }
    if (t < 64) bias_s[t] = bias ? __ldg(bias + t) : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);

  Raw<CIN, INTERP> f;
  int tile = blockIdx.x;
  if (tile < ntiles) issue_field<CIN, INTERP>(in, tile * 128 + t, npix, Ho, Wo, H, W, f);
  uint32_t phase = 0;
  for (; tile < ntiles; tile += gridDim.x) {
    pack_store_field<CIN, INTERP>(sA, t, Wo, f);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (t < 32 && elect_one()) {
      tc_fence_after();
      const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
#pragma unroll
      for (int k = 0; k < CIN; ++k)  // K = 16*CIN
        umma_bf16(tmem_base, umma_desc_sw128(a0 + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                  umma_desc_sw128(b0 + (k >> 2) * 8192 + (k & 3) * 32, 16, 1024), idesc, (uint32_t)(k != 0));
      umma_commit(mma_bar);
    }
    const int pix = tile * 128 + t;
    const int next = tile + gridDim.x;
    // the next tile's loads fly during this tile's MMA + epilogue; they are consumed at the top of the loop
    if (next < ntiles) issue_field<CIN, INTERP>(in, next * 128 + t, npix, Ho, Wo, H, W, f);
    mbar_wait(mma_bar, phase);
    phase ^= 1;
    tc_fence_after();
#pragma unroll
    for (int c0 = 0; c0 < 64; c0 += 32) {
      uint32_t v0[16], v1[16];
      const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + c0;
      tmem_ld16(taddr, v0);
      tmem_ld16(taddr + 16, v1);
      tmem_ld_wait();
      if (pix < npix) {
        uint32_t o[16];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float a = fmaf(__uint_as_float(v0[2 * j]), osc, bias_s[c0 + 2 * j]), b = fmaf(__uint_as_float(v0[2 * j + 1]), osc, bias_s[c0 + 2 * j + 1]);
          float c = fmaf(__uint_as_float(v1[2 * j]), osc, bias_s[c0 + 16 + 2 * j]), d = fmaf(__uint_as_float(v1[2 * j + 1]), osc, bias_s[c0 + 17 + 2 * j]);
          a = a > 0.f ? a : a * slope; b = b > 0.f ? b : b * slope;
          c = c > 0.f ? c : c * slope; d = d > 0.f ? d : d * slope;
          o[j] = pack_bf16(a, b);
          o[8 + j] = pack_bf16(c, d);
        }
        // whole 32-byte sectors per store (STG.256)
        __nv_bfloat16* dst = y + (long long)pix * 64 + c0;
        st_global_v8(dst, o);
        st_global_v8(dst + 16, o + 8);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<64>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ wgrad
// dw[co][c][ky][kx] += sum_pix gz[pix,co] * x[n,c,2oy+ky-1,2ox+kx-1];  db[co] += sum_pix gz[pix,co].
template <int CIN, bool INTERP>
__global__ void __launch_bounds__(FT_THREADS) first_wgrad_tc_kernel(FirstIn in, const __grid_constant__ CUtensorMap tmG,
                                                                    float* __restrict__ dw, float* __restrict__ db,
                                                                    const float* __restrict__ scale,
                                                                    const float* __restrict__ wdot, float* __restrict__ part,
                                                                    int N, int H, int W) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + FT_A_BYTES;                      // two gz tiles (double-buffered TMA destination)
  uint64_t* mma_bar = reinterpret_cast<uint64_t*>(sB + 2 * FT_B_BYTES);
  uint64_t* tma_bar = mma_bar + 1;                      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tma_bar + 2);

  const int t = threadIdx.x, warp = t >> 5;
  const int Ho = H / 2, Wo = W / 2;
  const int npix = N * Ho * Wo;   // host checks < 2^31 - 128
  const int ntiles = (npix + 127) / 128;

  int tile = blockIdx.x;
  if (t == 0) {
    tma_prefetch_desc(&tmG);
    mbar_init(mma_bar, 1);
    mbar_init(&tma_bar[0], 1);
    mbar_init(&tma_bar[1], 1);
    fence_barrier_init();
    if (tile < ntiles) {
      mbar_expect_tx(&tma_bar[0], FT_B_BYTES);
      tma_load_2d(sB, &tmG, &tma_bar[0], 0, tile * 128);  // rows past npix are zero-filled
    }
  }
  if (warp == 1) tmem_alloc<64>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 1, 1);

  // k = 127 (last element of chunk 15) is the all-ones im2col column whose dW row is the bias gradient; rows past npix
  // meet zero-filled gz rows, so it never changes.  Chunks 2*CIN..14 are never written: their dW rows are not flushed.
  *reinterpret_cast<uint4*>(sA + 16384 + (t >> 3) * 1024 + (t & 7) * 128 + ((7 ^ (t & 7)) << 4)) =
      make_uint4(0u, 0u, 0u, pack_bf16(0.f, 1.f));
  Raw<CIN, INTERP> f;
  if (tile < ntiles) issue_field<CIN, INTERP>(in, tile * 128 + t, npix, Ho, Wo, H, W, f);
  uint32_t it = 0;
  for (; tile < ntiles; tile += gridDim.x, ++it) {
    if (it > 0) mbar_wait(mma_bar, (it - 1) & 1);  // the previous tile's MMAs are done reading sA and sB[(it-1)&1]
    const int next = tile + gridDim.x;
    if (t == 0 && next < ntiles) {
      mbar_expect_tx(&tma_bar[(it + 1) & 1], FT_B_BYTES);
      tma_load_2d(sB + ((it + 1) & 1) * FT_B_BYTES, &tmG, &tma_bar[(it + 1) & 1], 0, next * 128);
    }
    pack_store_field<CIN, INTERP>(sA, t, Wo, f);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (t < 32 && elect_one()) {
      mbar_wait(&tma_bar[it & 1], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB + (it & 1) * FT_B_BYTES);
#pragma unroll
      for (int k = 0; k < 8; ++k)  // 16 pixels of the reduction per MMA
        umma_bf16(tmem_base, umma_desc_sw128(a0 + k * 2048, 16384, 1024), umma_desc_sw128(b0 + k * 2048, 16384, 1024),
                  idesc, (uint32_t)((it | (uint32_t)k) != 0));
      umma_commit(mma_bar);
    }
    if (next < ntiles) issue_field<CIN, INTERP>(in, next * 128 + t, npix, Ho, Wo, H, W, f);
  }
  float dot = 0.f;  // <this CTA's partial dW (unscaled), wdot>: deferred rank-1 term of the spectral-norm backward
  if (it > 0) {
    const float sc = scale ? __ldg(scale) : 1.f;
    mbar_wait(mma_bar, (it - 1) & 1);
    tc_fence_after();
    // TMEM lane t = im2col column k = t = (c*4 + ky)*4 + kx: dw[co][k] in the weight tensor's own layout.
    // CTAs finish together: rotate the column order per CTA so concurrent flushes hit different addresses
    const int rot = (blockIdx.x * 16) & 63;
#pragma unroll 1
    for (int cc = 0; cc < 64; cc += 16) {
      const int c0 = (cc + rot) & 63;
      uint32_t v[16];
      tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + c0, v);
      tmem_ld_wait();
      if (t < 16 * CIN) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const long long idx = (long long)(c0 + j) * (16 * CIN) + t;
          if (wdot) dot += __uint_as_float(v[j]) * __ldg(wdot + idx);
          atomicAdd(dw + idx, __uint_as_float(v[j]) * sc);
        }
      } else if (t == 127 && db) {
#pragma unroll
        for (int j = 0; j < 16; ++j) atomicAdd(db + c0 + j, __uint_as_float(v[j]));
      }
    }
  }
  if (part) {
    __shared__ float dsh[4];
    dot = warp_sum(dot);
    if ((t & 31) == 0) dsh[warp] = dot;
    __syncthreads();
    if (t == 0) part[blockIdx.x] = (dsh[0] + dsh[1]) + (dsh[2] + dsh[3]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<64>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ host side
static FirstIn make_in(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha) {
  FirstIn in;
  in.a[0] = a0; in.a[1] = a1 ? a1 : a0; in.b[0] = b0; in.b[1] = b1 ? b1 : b0; in.alpha = alpha;
  return in;
}

template <typename K>
static int first_attr(K kernel) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FT_SMEM);
  if (e != cudaSuccess) {
    b2c_set_error("cudaFuncSetAttribute(first layer): %s", cudaGetErrorString(e));
    return B2C_ERR_CUDA;
  }
  return B2C_OK;
}

static int first_check(const float* a0, const float* b0, const float* alpha, int N, int H, int W) {
  B2C_REQUIRE(H % 2 == 0 && W % 2 == 0, "first layer: odd size %dx%d", H, W);
  B2C_REQUIRE((long long)N * (H / 2) * (W / 2) < (1ll << 31) - 128 * 1024, "first layer: %d x %d x %d patches exceed 2^31 output pixels", N, H, W);
  B2C_REQUIRE(((uintptr_t)a0 & 7) == 0, "first layer: input must be 8-byte aligned");
  B2C_REQUIRE(!alpha || b0, "first layer: alpha given without interpolation partners");
  return B2C_OK;
}

template <int CIN, bool INTERP>
static int launch_first_fprop(const FirstIn& in, const float* w, const float* inv_sigma, const float* bias, void* y,
                              int N, int H, int W, float slope, unsigned blocks, cudaStream_t st) {
  static bool attr_done = false;
  if (!attr_done) {
    int rc = first_attr(first_fprop_tc_kernel<CIN, INTERP>);
    if (rc) return rc;
    attr_done = true;
  }
  first_fprop_tc_kernel<CIN, INTERP><<<blocks, FT_THREADS, FT_SMEM_FPROP, st>>>(in, w, inv_sigma, bias, (__nv_bfloat16*)y, N,
                                                                          H, W, slope);
  return b2c_check_launch("first_fprop");
}

template <int CIN, bool INTERP>
static int launch_first_wgrad(const FirstIn& in, const CUtensorMap& tmG, float* dw, float* db, const float* scale,
                              const float* wdot, float* part, int N, int H, int W, unsigned blocks, cudaStream_t st) {
  static bool attr_done = false;
  if (!attr_done) {
    int rc = first_attr(first_wgrad_tc_kernel<CIN, INTERP>);
    if (rc) return rc;
    attr_done = true;
  }
  first_wgrad_tc_kernel<CIN, INTERP><<<blocks, FT_THREADS, FT_SMEM, st>>>(in, tmG, dw, db, scale, wdot, part, N, H, W);
  return b2c_check_launch("first_wgrad");
}

// a0/a1: (N,3,H,W) fp32 NCHW halves (a1 null => Cin=3).  b0/b1/alpha: optional interpolation (all or none).
extern "C" int b2c_first_fprop(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                               const float* w, const float* inv_sigma, const float* bias, void* y, int N, int H, int W,
                               float slope, void* stream) {
  int rc = first_check(a0, b0, alpha, N, H, W);
  if (rc) return rc;
  FirstIn in = make_in(a0, a1, b0, b1, alpha);
  const long long npix = (long long)N * (H / 2) * (W / 2);
  long long blocks = (npix + 127) / 128;
  if (blocks > 148 * 4) blocks = 148 * 4;
  cudaStream_t st = (cudaStream_t)stream;
  if (a1) {
    if (alpha) return launch_first_fprop<6, true>(in, w, inv_sigma, bias, y, N, H, W, slope, (unsigned)blocks, st);
    return launch_first_fprop<6, false>(in, w, inv_sigma, bias, y, N, H, W, slope, (unsigned)blocks, st);
  }
  if (alpha) return launch_first_fprop<3, true>(in, w, inv_sigma, bias, y, N, H, W, slope, (unsigned)blocks, st);
  return launch_first_fprop<3, false>(in, w, inv_sigma, bias, y, N, H, W, slope, (unsigned)blocks, st);
}

// dw: fp32 [64][Cin][4][4], db: fp32 [64] or null; both are ACCUMULATED into (caller zeroes).  gz: dense NHWC bf16.
// scale (device scalar, nullable) multiplies what is added to dw (spectral norm: 1/sigma); wdot/part/nparts_host (nullable,
// together): part[0..*nparts_host) <- per-CTA partials of <dW_unscaled, wdot> for b2c_sn_rank1 (room for 1024 floats).
extern "C" int b2c_first_wgrad(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                               const void* gz, float* dw, float* db, const float* scale, const float* wdot, float* part,
                               int* nparts_host, int N, int H, int W, void* stream) {
  int rc = first_check(a0, b0, alpha, N, H, W);
  if (rc) return rc;
  B2C_REQUIRE((wdot == nullptr) == (part == nullptr), "first_wgrad: wdot and part go together");
  FirstIn in = make_in(a0, a1, b0, b1, alpha);
  const long long npix = (long long)N * (H / 2) * (W / 2);
  CUtensorMap tmG;
  {
    uint64_t dims[2] = {64, (uint64_t)npix};
    uint64_t str[1] = {128};
    uint32_t box[2] = {64, 128};
    rc = b2c_make_tmap_bf16(&tmG, gz, 2, dims, str, box);
    if (rc) return rc;
  }
  long long blocks = (npix + 127) / 128;
  if (blocks > 148 * 4) blocks = 148 * 4;
  if (nparts_host) *nparts_host = (int)blocks;
  cudaStream_t st = (cudaStream_t)stream;
  if (a1) {
    if (alpha) return launch_first_wgrad<6, true>(in, tmG, dw, db, scale, wdot, part, N, H, W, (unsigned)blocks, st);
    return launch_first_wgrad<6, false>(in, tmG, dw, db, scale, wdot, part, N, H, W, (unsigned)blocks, st);
  }
  if (alpha) return launch_first_wgrad<3, true>(in, tmG, dw, db, scale, wdot, part, N, H, W, (unsigned)blocks, st);
  return launch_first_wgrad<3, false>(in, tmG, dw, db, scale, wdot, part, N, H, W, (unsigned)blocks, st);
}
