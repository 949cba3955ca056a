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

// Sign bits of 32 consecutive channels from their 16 packed bf16 pairs (word j = channels 2j, 2j+1): bit c = (value > 0).
__device__ __forceinline__ uint32_t sign_bits32(const uint32_t (&o)[16]) {
  uint32_t bits = 0;
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    bits |= (uint32_t)((int16_t)(o[j] & 0xffffu) > 0) << (2 * j);
    bits |= (uint32_t)((int32_t)o[j] >= 0x10000) << (2 * j + 1);
  }
  return bits;
}

// ------------------------------------------------------------------------------------------------ fprop
template <int CIN, bool INTERP>
__global__ void __launch_bounds__(FT_THREADS, INTERP ? 2 : 4) first_fprop_tc_kernel(FirstIn in, const float* __restrict__ w,
                                                                    const float* __restrict__ inv_sigma,
                                                                    const float* __restrict__ bias,
                                                                    __nv_bfloat16* __restrict__ y, int N, int H, int W,
                                                                    float slope, unsigned long long* __restrict__ sign) {
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
        if (sign) reinterpret_cast<uint32_t*>(sign + pix)[c0 >> 5] = sign_bits32(o);
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
                              int N, int H, int W, float slope, unsigned long long* sign, unsigned blocks, cudaStream_t st) {
  static bool attr_done = false;
  if (!attr_done) {
    int rc = first_attr(first_fprop_tc_kernel<CIN, INTERP>);
    if (rc) return rc;
    attr_done = true;
  }
  first_fprop_tc_kernel<CIN, INTERP><<<blocks, FT_THREADS, FT_SMEM_FPROP, st>>>(in, w, inv_sigma, bias, (__nv_bfloat16*)y, N,
                                                                          H, W, slope, sign);
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

// ================================================================================================ TMA-staged variants
// Round-2 profile of the register-gather kernels above (profiles/r02_bw_step_ncu.txt, ncu_first): 1230 instructions per
// output pixel, L1TEX the busiest unit (every input row is fetched by two output rows, plus predicated halo loads and
// shuffles), 15 resident warps waiting on their own global loads -> 2.1 TB/s.  Here the caller's NCHW fp32 rows are
// staged by TMA instead: a box of [3 channels][4 input rows][136 columns] (x = 2*ox0-4 .. 2*ox0+131: the innermost TMA
// coordinate must be 16-byte aligned - x = 2*ox0-1 traps as an illegal instruction; rows / columns outside the image
// are zero-filled by the TMA unit = the conv's padding) per 64 output pixels of one output row and source tensor.
//
// ONE CTA per SM: a producer warp fills a ring of NS half-tile stages that takes all the shared memory the fixed
// buffers leave (>= 100 KB in flight per SM: the first version, two CTAs with a private 2-3 stage ring each, spent a
// third of its time waiting for data), and NG = 2 or 3 groups of 128 packer threads take tiles from it in turn, each with its
// own im2col tile and TMEM accumulator, so one group's MMA / epilogue latency is covered by the other groups' packing
// (with two groups the 8 warps' dependent instruction chains were the limit: IPC 1.2).  The fprop output staging tile
// aliases the group's im2col tile, which is dead once the MMAs have completed.  Shared-memory wavefronts are the next limit (the first version ran L1TEX at 80 %), so
// a lane pair (2m, 2m+1) shares one aligned LDS.128 - staged columns 4m+4..4m+7 hold three of the four taps of both
// pixels - plus one LDS.32 for the fourth (4 wavefronts per (channel, ky) row instead of 6), the bias is read with
// LDS.128, and the fprop output leaves through a swizzled staging tile and ONE TMA store per tile instead of 16 strided
// STG.256 per warp.  Needs Wo % 64 == 0 (patches >= 128) and 16-byte aligned inputs; smaller patches keep the kernels above.
static constexpr int FS_BOXW = 136;
static constexpr int FS_BOX_BYTES = 3 * 4 * FS_BOXW * 4;  // 6528 = 51 * 128
static constexpr int FS_SLOT = FS_BOX_BYTES;
// threads = NG * 128 + 32: warps 4g..4g+3 = packer group g (NG = 2 or 3), last warp = TMA producer
static constexpr int FS_GZ_BYTES = 64 * 128;              // wgrad: gz rows of one half tile

struct FirstMaps {
  CUtensorMap m[4];  // a0, a1, b0, b1 as (W, H, 3N) fp32
};

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(src)), "r"(c0), "r"(c1)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// One half tile (64 output pixels) -> 64 im2col rows of sA.  `st` = this half's stage; slot order: a-parts, then the
// b-parts that are really interpolated (lerp_mask bit p).  Staged column j <-> x = 2*ox0 - 4 + j, so pixel oxl needs
// j = 2*oxl+3 .. 2*oxl+6: for the lane pair (2m, 2m+1) the aligned chunk B = j 4m+4..4m+7 holds (x0, x1, r) of the even
// pixel and (l, x0, x1) of the odd one; the even lane's l (j = 4m+3) / the odd lane's r (j = 4m+8) come from one LDS.32.
template <int CIN, bool INTERP>
__device__ __forceinline__ void pack_from_stage(uint8_t* sA, int row, const uint8_t* st, int oxl, float al, int lerp_mask) {
  uint8_t* base = sA + (row >> 3) * 1024 + (row & 7) * 128;
  const float be = 1.f - al;
  const bool odd = oxl & 1;
  const int cb = (oxl >> 1) * 4 + 4, ce = odd ? cb + 4 : cb - 1;
  uint32_t p[4];
  int bslot = CIN / 3;
#pragma unroll
  for (int part = 0; part < CIN / 3; ++part) {
    const float* ra = reinterpret_cast<const float*>(st + part * FS_SLOT);
    const bool lerp = INTERP && ((lerp_mask >> part) & 1);
    const float* rb = reinterpret_cast<const float*>(st + bslot * FS_SLOT);
    if (lerp) ++bslot;
#pragma unroll
    for (int r = 0; r < 12; ++r) {  // r = cc*4 + ky: the box's own row order
      float4 B = *reinterpret_cast<const float4*>(ra + r * FS_BOXW + cb);
      float e = ra[r * FS_BOXW + ce];
      if (lerp) {
        const float4 Bb = *reinterpret_cast<const float4*>(rb + r * FS_BOXW + cb);
        const float eb = rb[r * FS_BOXW + ce];
        B.x = al * B.x + be * Bb.x; B.y = al * B.y + be * Bb.y;
        B.z = al * B.z + be * Bb.z; B.w = al * B.w + be * Bb.w;
        e = al * e + be * eb;
      }
      const float l = odd ? B.y : e, x0 = odd ? B.z : B.x, x1 = odd ? B.w : B.y, rr = odd ? e : B.z;
      const int g = part * 12 + r;
      p[(g & 1) * 2] = pack_bf16(l, x0);
      p[(g & 1) * 2 + 1] = pack_bf16(x1, rr);
      if (g & 1) {
        const int j = g >> 1;
        *reinterpret_cast<uint4*>(base + (j >> 3) * 16384 + (((j & 7) ^ (row & 7)) << 4)) = make_uint4(p[0], p[1], p[2], p[3]);
      }
    }
  }
}

// Producer: lane 0 of warp 8 walks this CTA's tiles, two half-tile stages per tile.
template <int CIN, bool INTERP, bool WGRAD>
__device__ __forceinline__ void first_producer(const FirstMaps& maps, const CUtensorMap* tmG, uint8_t* raw, uint8_t* gzs,
                                               uint64_t* full, uint64_t* empty, int lerp_mask, int NS, int stage_bytes,
                                               int ntiles, int Ho, int Wo) {
  int nslots = CIN / 3;
  if (INTERP) nslots += __popc(lerp_mask & ((1 << (CIN / 3)) - 1));
  const uint32_t tx = (uint32_t)nslots * FS_BOX_BYTES + (WGRAD ? FS_GZ_BYTES : 0);
  int s = 0, use = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    for (int h = 0; h < 2; ++h) {
      if (use > 0) mbar_wait(&empty[s], (use - 1) & 1);
      const int p0 = tile * 128 + h * 64;
      const int ox0 = p0 % Wo, q = p0 / Wo, oy = q % Ho, n = q / Ho;
      mbar_expect_tx(&full[s], tx);
      uint8_t* dst = raw + (size_t)s * stage_bytes;
      int slot = 0;
#pragma unroll
      for (int part = 0; part < CIN / 3; ++part)
        tma_load_3d(dst + (slot++) * FS_SLOT, &maps.m[part], &full[s], 2 * ox0 - 4, 2 * oy - 1, 3 * n);
      if (INTERP) {
#pragma unroll
        for (int part = 0; part < CIN / 3; ++part)
          if ((lerp_mask >> part) & 1) tma_load_3d(dst + (slot++) * FS_SLOT, &maps.m[2 + part], &full[s], 2 * ox0 - 4, 2 * oy - 1, 3 * n);
      }
      if (WGRAD) tma_load_2d(gzs + (size_t)s * FS_GZ_BYTES, tmG, &full[s], 0, p0);
      if (++s == NS) { s = 0; ++use; }
    }
  }
}

// Ring position of packer group `grp`, half `half`, at the group's j-th tile: the CTA's i-th tile (i = NG*j + grp) uses
// half-stage counts 2i and 2i+1.  The groups share one ring and skip each other's stages, so a group can reach a stage
// whose PREVIOUS fill has not even landed yet, and a bare parity wait for "its" fill would then pass on the wrong phase.
// wait_full() therefore first waits until the previous fill was consumed (=> it landed), which is well defined because
// the host guarantees NS >= 2*NG: the previous occupant of the stage is then a half tile the in-order producer issued
// before the tile this group has just finished.
struct RingPos {
  int s, use;
  __device__ __forceinline__ RingPos(int grp, int half, int NS) {
    const int cnt = 2 * grp + half;
    s = cnt % NS;
    use = cnt / NS;
  }
  __device__ __forceinline__ void next_tile(int NS, int NG) {
    s += 2 * NG;
    while (s >= NS) { s -= NS; ++use; }
  }
};
__device__ __forceinline__ void wait_full(uint64_t* full, uint64_t* empty, int s, int use) {
  if (use > 0) mbar_wait(&empty[s], (use - 1) & 1);
  mbar_wait(&full[s], use & 1);
}

// ------------------------------------------------------------------------------------------------ fprop (TMA-staged)
template <int CIN, bool INTERP, int NG>
__global__ void __launch_bounds__(NG * 128 + 32, 1) first_fprop_tma_kernel(const __grid_constant__ FirstMaps maps,
                                                                      const __grid_constant__ CUtensorMap tmY,
                                                                      const float* __restrict__ alpha, int lerp_mask,
                                                                      const float* __restrict__ w,
                                                                      const float* __restrict__ inv_sigma,
                                                                      const float* __restrict__ bias, int N, int H, int W,
                                                                      float slope, int NS, int stage_bytes,
                                                                      unsigned long long* __restrict__ sign) {
  constexpr int KB = (16 * CIN + 63) / 64;  // k-blocks of 64
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                            // [NG groups][KB][128 rows][128 B]; the first 16 KB double as output staging
  uint8_t* sB = sA + NG * KB * 16384;         // [KB][64 co][128 B]
  // output staging: with three groups it aliases the group's im2col tile (dead once the MMAs have completed; costs one more
  // barrier per tile), with two there is room for a tile of its own
  constexpr bool ALIAS_O = NG == 3;
  uint8_t* sO = sB + KB * 8192;                  // [NG groups][128 rows][128 B] unless aliased
  uint8_t* raw = sO + (ALIAS_O ? 0 : NG * 16384);  // [NS][stage_bytes]
  uint64_t* full = reinterpret_cast<uint64_t*>(raw + (size_t)NS * stage_bytes);
  uint64_t* empty = full + 32;
  uint64_t* mma_bar = empty + 32;                // [NG groups]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mma_bar + 4);
  __shared__ __align__(16) float bias_s[64];

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int Ho = H / 2, Wo = W / 2;
  const int ntiles = N * Ho * Wo / 128;

  if (t == 0) {
    for (int i = 0; i < CIN / 3; ++i) tma_prefetch_desc(&maps.m[i]);
    tma_prefetch_desc(&tmY);
    for (int s = 0; s < NS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 2);  // the two packer warps of a half
    }
    for (int g = 0; g < NG; ++g) mbar_init(&mma_bar[g], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<256>(tmem_slot);
  const float osc = inv_sigma ? __ldg(inv_sigma) : 1.f;
  {
    constexpr int NCH = 2 * CIN;
    for (int i = t; i < 64 * NCH; i += NG * 128 + 32) {
      const int co = i / NCH, j = i % NCH;
      const float4 lo = __ldg(reinterpret_cast<const float4*>(w + (long long)co * (16 * CIN) + j * 8));
      const float4 hi = __ldg(reinterpret_cast<const float4*>(w + (long long)co * (16 * CIN) + j * 8 + 4));
      const uint4 u = make_uint4(pack_bf16(lo.x, lo.y), pack_bf16(lo.z, lo.w), pack_bf16(hi.x, hi.y), pack_bf16(hi.z, hi.w));
      *reinterpret_cast<uint4*>(sB + (j >> 3) * 8192 + (co >> 3) * 1024 + (co & 7) * 128 + (((j & 7) ^ (co & 7)) << 4)) = u;
    }
    if (t < 64) bias_s[t] = bias ? __ldg(bias + t) : 0.f;
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);

  if (warp == 4 * NG) {
    if (lane == 0)
      first_producer<CIN, INTERP, false>(maps, nullptr, raw, nullptr, full, empty, lerp_mask, NS, stage_bytes, ntiles, Ho, Wo);
  } else {
    const int grp = warp >> 2, tg = t & 127, half = tg >> 6, oxl = tg & 63;
    uint8_t* sAg = sA + grp * (KB * 16384);
    uint8_t* sOg = ALIAS_O ? sAg : sO + grp * 16384;
    const uint32_t tmem_g = tmem_base + (uint32_t)grp * 64;
    RingPos rp(grp, half, NS);
    int j = 0;
    for (int tile = blockIdx.x + grp * gridDim.x; tile < ntiles; tile += NG * gridDim.x, ++j) {
      if (ALIAS_O) {
        if (tg == 0) tma_store_wait_read();  // the previous tile's store has read the staging tile = this im2col tile
        named_bar_sync(1 + grp, 128);
      }
      float al = 1.f;
      if (INTERP) al = __ldg(alpha + (tile * 128 + half * 64) / (Ho * Wo));
      wait_full(full, empty, rp.s, rp.use);
      pack_from_stage<CIN, INTERP>(sAg, tg, raw + (size_t)rp.s * stage_bytes, oxl, al, lerp_mask);
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[rp.s]);
      rp.next_tile(NS, NG);
      if (!ALIAS_O && tg == 0) tma_store_wait_read();  // the previous tile's store has read the staging tile
      fence_proxy_async();
      tc_fence_before();
      named_bar_sync(1 + grp, 128);
      if (tg == 0) {
        tc_fence_after();
        const uint32_t a0 = smem_u32(sAg), b0 = smem_u32(sB);
#pragma unroll
        for (int k = 0; k < CIN; ++k)  // K = 16*CIN
          umma_bf16(tmem_g, umma_desc_sw128(a0 + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                    umma_desc_sw128(b0 + (k >> 2) * 8192 + (k & 3) * 32, 16, 1024), idesc, (uint32_t)(k != 0));
        umma_commit(&mma_bar[grp]);
      }
      __syncwarp();
      mbar_wait(&mma_bar[grp], j & 1);
      tc_fence_after();
      uint8_t* orow = sOg + (tg >> 3) * 1024 + (tg & 7) * 128;
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t v0[16], v1[16];
        const uint32_t taddr = tmem_g + ((uint32_t)((warp & 3) * 32) << 16) + c0;
        tmem_ld16(taddr, v0);
        tmem_ld16(taddr + 16, v1);
        tmem_ld_wait();
        uint32_t o[16];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 b0v = *reinterpret_cast<const float4*>(bias_s + c0 + 4 * q);
          const float4 b1v = *reinterpret_cast<const float4*>(bias_s + c0 + 16 + 4 * q);
          float a = fmaf(__uint_as_float(v0[4 * q]), osc, b0v.x), b = fmaf(__uint_as_float(v0[4 * q + 1]), osc, b0v.y);
          float c = fmaf(__uint_as_float(v0[4 * q + 2]), osc, b0v.z), d = fmaf(__uint_as_float(v0[4 * q + 3]), osc, b0v.w);
          a = fmaxf(a, a * slope); b = fmaxf(b, b * slope);  // LeakyReLU for 0 <= slope <= 1 (host-checked)
          c = fmaxf(c, c * slope); d = fmaxf(d, d * slope);
          o[2 * q] = pack_bf16(a, b);
          o[2 * q + 1] = pack_bf16(c, d);
          a = fmaf(__uint_as_float(v1[4 * q]), osc, b1v.x); b = fmaf(__uint_as_float(v1[4 * q + 1]), osc, b1v.y);
          c = fmaf(__uint_as_float(v1[4 * q + 2]), osc, b1v.z); d = fmaf(__uint_as_float(v1[4 * q + 3]), osc, b1v.w);
          a = fmaxf(a, a * slope); b = fmaxf(b, b * slope);  // LeakyReLU for 0 <= slope <= 1 (host-checked)
          c = fmaxf(c, c * slope); d = fmaxf(d, d * slope);
          o[8 + 2 * q] = pack_bf16(a, b);
          o[8 + 2 * q + 1] = pack_bf16(c, d);
        }
        // channels c0 .. c0+31 = 16-byte chunks c0/8 .. c0/8+3 of this pixel's 128-byte row (128B-swizzled like a TMA tile)
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<uint4*>(orow + ((((c0 >> 3) + q) ^ (tg & 7)) << 4)) = make_uint4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
        if (sign) reinterpret_cast<uint32_t*>(sign + ((long long)tile * 128 + tg))[c0 >> 5] = sign_bits32(o);
      }
      fence_proxy_async();
      tc_fence_before();
      named_bar_sync(1 + grp, 128);
      if (tg == 0) tma_store_2d(&tmY, sOg, 0, tile * 128);
      __syncwarp();
    }
    if (tg == 0) tma_store_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<256>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ wgrad (TMA-staged)
// The gz rows of a half tile travel in the same stage as its input boxes; a stage is released by the commit of the MMAs
// that read its gz rows.  Each group accumulates its tiles in its own TMEM columns; one atomic flush per group.
template <int CIN, bool INTERP, int NG>
__global__ void __launch_bounds__(NG * 128 + 32, 1) first_wgrad_tma_kernel(const __grid_constant__ FirstMaps maps,
                                                                      const __grid_constant__ CUtensorMap tmG,
                                                                      const float* __restrict__ alpha, int lerp_mask,
                                                                      float* __restrict__ dw, float* __restrict__ db,
                                                                      const float* __restrict__ scale,
                                                                      const float* __restrict__ wdot, float* __restrict__ part,
                                                                      int N, int H, int W, int NS, int stage_bytes) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                              // [NG groups][2][128 rows][128 B]
  uint8_t* gzs = sA + NG * FT_A_BYTES;          // [NS][64 pix][128 B]
  uint8_t* raw = gzs + (size_t)NS * FS_GZ_BYTES;   // [NS][stage_bytes]
  uint64_t* full = reinterpret_cast<uint64_t*>(raw + (size_t)NS * stage_bytes);
  uint64_t* empty = full + 32;
  uint64_t* mma_bar = empty + 32;                  // [NG groups]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mma_bar + 4);
  __shared__ float dsh[4 * NG];

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int Ho = H / 2, Wo = W / 2;
  const int ntiles = N * Ho * Wo / 128;

  if (t == 0) {
    for (int i = 0; i < CIN / 3; ++i) tma_prefetch_desc(&maps.m[i]);
    tma_prefetch_desc(&tmG);
    for (int s = 0; s < NS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);  // the commit of the MMAs that read the stage's gz rows
    }
    for (int g = 0; g < NG; ++g) mbar_init(&mma_bar[g], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<256>(tmem_slot);
  // k = 127 (last element of chunk 15) is the all-ones im2col column whose dW row is the bias gradient
  if (t < NG * 128) {
    const int g = t >> 7, r = t & 127;
    *reinterpret_cast<uint4*>(sA + g * FT_A_BYTES + 16384 + (r >> 3) * 1024 + (r & 7) * 128 + ((7 ^ (r & 7)) << 4)) =
        make_uint4(0u, 0u, 0u, pack_bf16(0.f, 1.f));
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 1, 1);

  float dot = 0.f;  // <this group's partial dW (unscaled), wdot>: deferred rank-1 term of the spectral-norm backward
  if (warp == 4 * NG) {
    if (lane == 0)
      first_producer<CIN, INTERP, true>(maps, &tmG, raw, gzs, full, empty, lerp_mask, NS, stage_bytes, ntiles, Ho, Wo);
  } else {
    const int grp = warp >> 2, tg = t & 127, half = tg >> 6, oxl = tg & 63;
    uint8_t* sAg = sA + grp * FT_A_BYTES;
    const uint32_t tmem_g = tmem_base + (uint32_t)grp * 64;
    RingPos rp(grp, half, NS);
    int j = 0;
    for (int tile = blockIdx.x + grp * gridDim.x; tile < ntiles; tile += NG * gridDim.x, ++j) {
      if (j > 0) mbar_wait(&mma_bar[grp], (j - 1) & 1);  // the group's previous MMAs are done reading sAg
      float al = 1.f;
      if (INTERP) al = __ldg(alpha + (tile * 128 + half * 64) / (Ho * Wo));
      wait_full(full, empty, rp.s, rp.use);
      pack_from_stage<CIN, INTERP>(sAg, tg, raw + (size_t)rp.s * stage_bytes, oxl, al, lerp_mask);
      fence_proxy_async();
      tc_fence_before();
      named_bar_sync(1 + grp, 128);
      if (tg == 0) {
        // this thread packed half 0; half 1's stage is the next one in the ring
        int s1 = rp.s + 1, use1 = rp.use;
        if (s1 == NS) { s1 = 0; ++use1; }
        wait_full(full, empty, s1, use1);
        tc_fence_after();
        const uint32_t a0 = smem_u32(sAg);
        const uint32_t g0 = smem_u32(gzs + (size_t)rp.s * FS_GZ_BYTES), g1 = smem_u32(gzs + (size_t)s1 * FS_GZ_BYTES);
#pragma unroll
        for (int k = 0; k < 8; ++k)  // 16 pixels of the reduction per MMA
          umma_bf16(tmem_g, umma_desc_sw128(a0 + k * 2048, 16384, 1024),
                    umma_desc_sw128((k < 4 ? g0 : g1) + (k & 3) * 2048, 16384, 1024), idesc, (uint32_t)((j | k) != 0));
        umma_commit(&empty[rp.s]);
        umma_commit(&empty[s1]);
        umma_commit(&mma_bar[grp]);
      }
      __syncwarp();
      rp.next_tile(NS, NG);
    }
    if (j > 0) {
      const float sc = scale ? __ldg(scale) : 1.f;
      mbar_wait(&mma_bar[grp], (j - 1) & 1);
      tc_fence_after();
      // TMEM lane tg = im2col column k = (c*4 + ky)*4 + kx: dw[co][k] in the weight tensor's own layout.
      // CTAs finish together: rotate the column order per CTA so concurrent flushes hit different addresses
      const int rot = ((blockIdx.x * NG + grp) * 16) & 63;
#pragma unroll 1
      for (int cc = 0; cc < 64; cc += 16) {
        const int c0 = (cc + rot) & 63;
        uint32_t v[16];
        tmem_ld16(tmem_g + ((uint32_t)((warp & 3) * 32) << 16) + c0, v);
        tmem_ld_wait();
        if (tg < 16 * CIN) {
#pragma unroll
          for (int q = 0; q < 16; ++q) {
            const long long idx = (long long)(c0 + q) * (16 * CIN) + tg;
            if (wdot) dot += __uint_as_float(v[q]) * __ldg(wdot + idx);
            atomicAdd(dw + idx, __uint_as_float(v[q]) * sc);
          }
        } else if (tg == 127 && db) {
#pragma unroll
          for (int q = 0; q < 16; ++q) atomicAdd(db + c0 + q, __uint_as_float(v[q]));
        }
      }
    }
  }
  if (part) {
    dot = warp_sum(dot);
    if (lane == 0 && warp < 4 * NG) dsh[warp] = dot;
    __syncthreads();
    if (t == 0) {
      float d = 0.f;
      for (int i = 0; i < 4 * NG; ++i) d += dsh[i];
      part[blockIdx.x] = d;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<256>(tmem_base);
}

// ---- host side of the TMA-staged variants
static bool first_tma_ok(const float* a0, const float* a1, const float* b0, const float* b1, int N, int H, int W) {
  static int enabled = -1;
  if (enabled < 0) {
    const char* e = getenv("B2C_FIRST_TMA");
    enabled = (e && e[0] == '0') ? 0 : 1;
  }
  if (!enabled) return false;
  const int Ho = H / 2, Wo = W / 2;
  if (Wo % 64 != 0 || ((long long)Ho * Wo) % 128 != 0) return false;
  if ((long long)N * 3 >= (1ll << 31)) return false;
  const uintptr_t al = (uintptr_t)a0 | (uintptr_t)a1 | (uintptr_t)b0 | (uintptr_t)b1;
  return (al & 15) == 0;
}

static int first_maps(FirstMaps* maps, const FirstIn& in, bool two, bool interp, int N, int H, int W, int* lerp_mask) {
  uint64_t dims[3] = {(uint64_t)W, (uint64_t)H, (uint64_t)N * 3};
  uint64_t str[2] = {(uint64_t)W * 4, (uint64_t)W * H * 4};
  uint32_t box[3] = {(uint32_t)FS_BOXW, 4, 3};
  *lerp_mask = 0;
  const int nparts = two ? 2 : 1;
  for (int p = 0; p < nparts; ++p) {
    int rc = b2c_make_tmap_f32(&maps->m[p], in.a[p], 3, dims, str, box);
    if (rc) return rc;
    // x = alpha*a + (1-alpha)*a = a: a part whose two sources are the same tensor (the critic's `pred` channels under the
    // gradient penalty, cgan/losses.py:203-204 with real = (pred, gt), fake = (pred, refined)) is staged once
    if (interp && in.b[p] != in.a[p]) {
      rc = b2c_make_tmap_f32(&maps->m[2 + p], in.b[p], 3, dims, str, box);
      if (rc) return rc;
      *lerp_mask |= 1 << p;
    } else {
      maps->m[2 + p] = maps->m[p];
    }
  }
  for (int p = nparts; p < 2; ++p) maps->m[p] = maps->m[0], maps->m[2 + p] = maps->m[0];
  return B2C_OK;
}

// Packer groups and ring depth: three groups when the shared memory one CTA per SM has left (227 KB) still holds a ring of
// NS >= 2*NG half-tile stages (the condition RingPos needs), else two; NG = 0: use the register-gather kernels.
static void first_plan(int per_group_bytes, int shared_bytes, int stage_bytes_total, int max_ng, int* NG, int* NS) {
  static int forced_ng = -1, forced_ns = -1;
  if (forced_ng < 0) {
    const char* e = getenv("B2C_FIRST_NG");
    forced_ng = e ? atoi(e) : 0;
    e = getenv("B2C_FIRST_NS");
    forced_ns = e ? atoi(e) : 0;
  }
  *NG = 0;
  for (int ng = max_ng; ng >= 2; --ng) {
    if (forced_ng && ng != forced_ng) continue;
    // two fprop groups carry an output staging tile of their own (per_group_bytes is the im2col tile: 16 KB per k-block)
    const int extra = (max_ng == 3 && ng == 2) ? 2 * 16384 : 0;
    int ns = (226 * 1024 - ng * per_group_bytes - extra - shared_bytes - 2048) / stage_bytes_total;
    if (forced_ns && forced_ns < ns) ns = forced_ns;
    if (ns > 16) ns = 16;
    if (ns >= 2 * ng) {
      *NG = ng;
      *NS = ns;
      return;
    }
  }
}

template <typename K>
static int first_smem_attr(K kernel, int smem, int* have) {
  if (smem > *have) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) {
      b2c_set_error("cudaFuncSetAttribute(first layer, %d): %s", smem, cudaGetErrorString(e));
      return B2C_ERR_CUDA;
    }
    *have = smem;
  }
  return B2C_OK;
}

// returns B2C_OK + *done = false when the shapes do not fit the TMA-staged kernels
template <int CIN, bool INTERP>
static int launch_first_fprop_tma(const FirstIn& in, const float* w, const float* inv_sigma, const float* bias, void* y, int N,
                                  int H, int W, float slope, unsigned long long* sign, cudaStream_t st, bool* done) {
  *done = false;
  FirstMaps maps;
  int lerp_mask = 0;
  int rc = first_maps(&maps, in, CIN == 6, INTERP, N, H, W, &lerp_mask);
  if (rc) return rc;
  const long long npix = (long long)N * (H / 2) * (W / 2);
  CUtensorMap tmY;
  {
    uint64_t dims[2] = {64, (uint64_t)npix};
    uint64_t str[1] = {128};
    uint32_t box[2] = {64, 128};
    rc = b2c_make_tmap_bf16(&tmY, y, 2, dims, str, box);
    if (rc) return rc;
  }
  constexpr int KB = (16 * CIN + 63) / 64;
  const int nslots = CIN / 3 + __builtin_popcount(lerp_mask);
  const int stage_bytes = nslots * FS_SLOT;
  int NG, NS;
  first_plan(KB * 16384, KB * 8192, stage_bytes, 3, &NG, &NS);
  if (!NG) return B2C_OK;
  const int smem = NG * KB * 16384 + (NG == 2 ? 2 * 16384 : 0) + KB * 8192 + 2048 + NS * stage_bytes;
  const long long ntiles = npix / 128;
  const unsigned blocks = (unsigned)(ntiles < 148 ? ntiles : 148);
  static int have2 = 0, have3 = 0;
  if (NG == 3) {
    rc = first_smem_attr(first_fprop_tma_kernel<CIN, INTERP, 3>, smem, &have3);
    if (rc) return rc;
    first_fprop_tma_kernel<CIN, INTERP, 3><<<blocks, 3 * 128 + 32, smem, st>>>(maps, tmY, in.alpha, lerp_mask, w, inv_sigma, bias,
                                                                           N, H, W, slope, NS, stage_bytes, sign);
  } else {
    rc = first_smem_attr(first_fprop_tma_kernel<CIN, INTERP, 2>, smem, &have2);
    if (rc) return rc;
    first_fprop_tma_kernel<CIN, INTERP, 2><<<blocks, 2 * 128 + 32, smem, st>>>(maps, tmY, in.alpha, lerp_mask, w, inv_sigma, bias,
                                                                           N, H, W, slope, NS, stage_bytes, sign);
  }
  *done = true;
  return b2c_check_launch("first_fprop_tma");
}

template <int CIN, bool INTERP>
static int launch_first_wgrad_tma(const FirstIn& in, const void* gz, float* dw, float* db, const float* scale,
                                  const float* wdot, float* part, int* nparts_host, int N, int H, int W, cudaStream_t st,
                                  bool* done) {
  *done = false;
  FirstMaps maps;
  int lerp_mask = 0;
  int rc = first_maps(&maps, in, CIN == 6, INTERP, N, H, W, &lerp_mask);
  if (rc) return rc;
  const long long npix = (long long)N * (H / 2) * (W / 2);
  CUtensorMap tmG;
  {
    uint64_t dims[2] = {64, (uint64_t)npix};
    uint64_t str[1] = {128};
    uint32_t box[2] = {64, 64};
    rc = b2c_make_tmap_bf16(&tmG, gz, 2, dims, str, box);
    if (rc) return rc;
  }
  const int nslots = CIN / 3 + __builtin_popcount(lerp_mask);
  const int stage_bytes = nslots * FS_SLOT;
  int NG, NS;
  first_plan(FT_A_BYTES, 0, stage_bytes + FS_GZ_BYTES, 2, &NG, &NS);  // wgrad has no epilogue to hide: two groups, deeper ring
  if (!NG) return B2C_OK;
  const int smem = NG * FT_A_BYTES + 2048 + NS * (stage_bytes + FS_GZ_BYTES);
  const long long ntiles = npix / 128;
  const unsigned blocks = (unsigned)(ntiles < 148 ? ntiles : 148);
  if (nparts_host) *nparts_host = (int)blocks;
  static int have2 = 0, have3 = 0;
  if (NG == 3) {
    rc = first_smem_attr(first_wgrad_tma_kernel<CIN, INTERP, 3>, smem, &have3);
    if (rc) return rc;
    first_wgrad_tma_kernel<CIN, INTERP, 3><<<blocks, 3 * 128 + 32, smem, st>>>(maps, tmG, in.alpha, lerp_mask, dw, db, scale, wdot,
                                                                           part, N, H, W, NS, stage_bytes);
  } else {
    rc = first_smem_attr(first_wgrad_tma_kernel<CIN, INTERP, 2>, smem, &have2);
    if (rc) return rc;
    first_wgrad_tma_kernel<CIN, INTERP, 2><<<blocks, 2 * 128 + 32, smem, st>>>(maps, tmG, in.alpha, lerp_mask, dw, db, scale, wdot,
                                                                           part, N, H, W, NS, stage_bytes);
  }
  *done = true;
  return b2c_check_launch("first_wgrad_tma");
}

// a0/a1: (N,3,H,W) fp32 NCHW halves (a1 null => Cin=3).  b0/b1/alpha: optional interpolation (all or none).
static int first_fprop_impl(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                            const float* w, const float* inv_sigma, const float* bias, void* y, int N, int H, int W, float slope,
                            unsigned long long* sign, void* stream) {
  int rc = first_check(a0, b0, alpha, N, H, W);
  if (rc) return rc;
  FirstIn in = make_in(a0, a1, b0, b1, alpha);
  const long long npix = (long long)N * (H / 2) * (W / 2);
  long long blocks = (npix + 127) / 128;
  if (blocks > 148 * 4) blocks = 148 * 4;
  cudaStream_t st = (cudaStream_t)stream;
  if (slope >= 0.f && slope <= 1.f && first_tma_ok(a0, a1, b0, b1, N, H, W)) {
    bool done = false;
    if (a1) {
      rc = alpha ? launch_first_fprop_tma<6, true>(in, w, inv_sigma, bias, y, N, H, W, slope, sign, st, &done)
                 : launch_first_fprop_tma<6, false>(in, w, inv_sigma, bias, y, N, H, W, slope, sign, st, &done);
    } else {
      rc = alpha ? launch_first_fprop_tma<3, true>(in, w, inv_sigma, bias, y, N, H, W, slope, sign, st, &done)
                 : launch_first_fprop_tma<3, false>(in, w, inv_sigma, bias, y, N, H, W, slope, sign, st, &done);
    }
    if (rc || done) return rc;
  }
  if (a1) {
    if (alpha) return launch_first_fprop<6, true>(in, w, inv_sigma, bias, y, N, H, W, slope, sign, (unsigned)blocks, st);
    return launch_first_fprop<6, false>(in, w, inv_sigma, bias, y, N, H, W, slope, sign, (unsigned)blocks, st);
  }
  if (alpha) return launch_first_fprop<3, true>(in, w, inv_sigma, bias, y, N, H, W, slope, sign, (unsigned)blocks, st);
  return launch_first_fprop<3, false>(in, w, inv_sigma, bias, y, N, H, W, slope, sign, (unsigned)blocks, st);
}

extern "C" int b2c_first_fprop(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                               const float* w, const float* inv_sigma, const float* bias, void* y, int N, int H, int W,
                               float slope, void* stream) {
  return first_fprop_impl(a0, a1, b0, b1, alpha, w, inv_sigma, bias, y, N, H, W, slope, nullptr, stream);
}
// The same, and sign_bits[pixel] (one 64-bit word per output pixel, N*(H/2)*(W/2) of them) <- bit c = (y[pixel][c] > 0): what
// the LeakyReLU backward of this layer needs, 1/16 of the activation's bytes (b2c_convt4s2_signbits_bf16 consumes it).
extern "C" int b2c_first_fprop_bits(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                                    const float* w, const float* inv_sigma, const float* bias, void* y, int N, int H, int W,
                                    float slope, unsigned long long* sign_bits, void* stream) {
  B2C_REQUIRE(sign_bits != nullptr && ((uintptr_t)sign_bits & 7) == 0, "first_fprop_bits: sign_bits must be given, 8-byte aligned");
  return first_fprop_impl(a0, a1, b0, b1, alpha, w, inv_sigma, bias, y, N, H, W, slope, sign_bits, stream);
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
  if (first_tma_ok(a0, a1, b0, b1, N, H, W)) {
    cudaStream_t st = (cudaStream_t)stream;
    bool done = false;
    if (a1) {
      rc = alpha ? launch_first_wgrad_tma<6, true>(in, gz, dw, db, scale, wdot, part, nparts_host, N, H, W, st, &done)
                 : launch_first_wgrad_tma<6, false>(in, gz, dw, db, scale, wdot, part, nparts_host, N, H, W, st, &done);
    } else {
      rc = alpha ? launch_first_wgrad_tma<3, true>(in, gz, dw, db, scale, wdot, part, nparts_host, N, H, W, st, &done)
                 : launch_first_wgrad_tma<3, false>(in, gz, dw, db, scale, wdot, part, nparts_host, N, H, W, st, &done);
    }
    if (rc || done) return rc;
  }
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

// ================================================================================================ first-layer DATA gradient
// gx[n,c,Y,X] = sum_{co,ky,kx : 2oy+ky-1 = Y, 2ox+kx-1 = X} gz[n,oy,ox,co] * w[co][c][ky][kx]   (needed only under the gradient
// penalty: d D(x_hat) / d x_hat, cgan/losses.py:213-220).  tapgemm.cu mode D1 computes it as four parity GEMMs over NINE
// tap-shifted boxes of gz: 1.2 GB through the L2 for 134 MB of input (142 us, 25 % of HBM).  Here the shifts move to the
// OUTPUT side: one GEMM  P[pixel][(c,ky,kx)] = gz[pixel][co] . Wd[co][(c,ky,kx)]  (M = pixels, N = 16*Cin, K = 64) over a
// tile of 8 x 16 gz pixels plus a one-pixel halo (ONE TMA box of 10 x 18 pixels, zero-filled outside the image), P parked in
// shared memory as fp32, and every output pixel of the tile's 16 x 32 block is the sum of its four (pixel, tap) entries,
// written as coalesced fp32 rows of the caller's NCHW tensors.  gz crosses the L2 1.4 times instead of 9.
namespace fd2 {
constexpr int TH = 8, TW = 16, HR = TH + 2, HC = TW + 2, NPIX = HR * HC;  // 180 halo pixels, rows 180..255 of A are never read
constexpr int A_BYTES = 256 * 128;       // two M = 128 tiles of [pixel][64 co] bf16
constexpr int PSTR = 100;                // floats per pixel row of P (96 used; 100 keeps float4 stores conflict-free)
constexpr int P_FLOATS = NPIX * PSTR;    // P is double-buffered: the output pass of tile i overlaps the TMEM pass of tile i + 1
constexpr int THREADS = 448;             // warp 0: TMA producer, 1: MMA, 2-5: TMEM -> P (one per lane quarter), 6-13: P -> output rows
constexpr int SMEM = 2 * A_BYTES + 96 * 128 + 2 * P_FLOATS * 4 + 256 + 1024;
// Measured at (64, 256, 256): 86 us (tapgemm mode D1: 142 us).  What did not help: 8 instead of 4 TMEM-pass warps, two TMEM
// accumulator sets, a 6-stage ring of 4 x 16 tiles (100 us: the per-tile hand-offs TMA -> MMA -> P -> rows cost more than the
// larger halo saves).
}  // namespace fd2

template <int CIN>
__global__ void __launch_bounds__(fd2::THREADS, 1) first_dgrad2_kernel(const __grid_constant__ CUtensorMap tmG,
                                                                     const float* __restrict__ w, const float* __restrict__ inv_sigma,
                                                                     float* __restrict__ gx0, float* __restrict__ gx1, int N, int Ho,
                                                                     int Wo, int ntiles) {
  using namespace fd2;
  constexpr int NCOL = 16 * CIN;  // 48 / 96
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                                        // [2 stages][256 rows][128 B], 128B-swizzled by TMA
  uint8_t* sB = sA + 2 * A_BYTES;                            // [NCOL rows (c,ky,kx)][64 co] bf16, K-major, 128B-swizzled
  float* Pbuf = reinterpret_cast<float*>(sB + 96 * 128);     // [2][NPIX][PSTR]
  uint64_t* full = reinterpret_cast<uint64_t*>(Pbuf + 2 * P_FLOATS);
  uint64_t* empty = full + 2;
  uint64_t* tmem_full = empty + 2;
  uint64_t* tmem_empty = tmem_full + 1;
  uint64_t* p_full = tmem_empty + 1;   // [2]
  uint64_t* p_empty = p_full + 2;      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(p_empty + 2);

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int tiles_x = Wo / TW, tiles_y = Ho / TH;

  if (t == 0) {
    tma_prefetch_desc(&tmG);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, 4);  // one arrival per TMEM-pass warp
    for (int b = 0; b < 2; ++b) {
      mbar_init(&p_full[b], 4);
      mbar_init(&p_empty[b], 8);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<256>(tmem_slot);
  // Wd[j = (c*4 + ky)*4 + kx][co] = bf16(w[co][c][ky][kx]): one 16-byte chunk (8 co) per thread and step
  for (int i = t; i < NCOL * 8; i += THREADS) {
    const int j = i >> 3, kc = i & 7;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = __ldg(w + (long long)(kc * 8 + e) * NCOL + j);
    *reinterpret_cast<uint4*>(sB + (j >> 3) * 1024 + (j & 7) * 128 + ((kc ^ (j & 7)) << 4)) =
        make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int s = 0, use = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int tx = tile % tiles_x, ty = (tile / tiles_x) % tiles_y, n = tile / (tiles_x * tiles_y);
        if (use > 0) mbar_wait(&empty[s], (use - 1) & 1);
        mbar_expect_tx(&full[s], NPIX * 128);
        tma_load_4d(sA + s * A_BYTES, &tmG, &full[s], 0, tx * TW - 1, ty * TH - 1, n);
        if (++s == 2) { s = 0; ++use; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, NCOL, 0, 0);
      constexpr uint32_t DHI = umma_desc_hi_sw128(1024);
      const uint32_t a_lo0 = umma_desc_lo(smem_u32(sA), 16), b_lo = umma_desc_lo(smem_u32(sB), 16);
      int s = 0, ph = 0, it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        if (it > 0) mbar_wait(tmem_empty, (it - 1) & 1);  // the previous P has been moved out of TMEM
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t alo = a_lo0 + s * (A_BYTES >> 4);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem_base + mt * 128, umma_desc_join(DHI, alo + mt * (16384 >> 4) + 2 * k), umma_desc_join(DHI, b_lo + 2 * k),
                      idesc, (uint32_t)(k != 0));
        umma_commit(&empty[s]);
        umma_commit(tmem_full);
        if (++s == 2) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp < 6) {
    // ---- pass 1 warps: P (both M tiles of this warp's TMEM lane quarter) from TMEM to shared memory
    const int q = warp & 3;
    const float sc = inv_sigma ? __ldg(inv_sigma) : 1.f;
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      float* P = Pbuf + (it & 1) * P_FLOATS;
      if (it >= 2) mbar_wait(&p_empty[it & 1], ((it >> 1) - 1) & 1);  // the output pass of tile it - 2 has read this buffer
      mbar_wait(tmem_full, it & 1);
      tc_fence_after();
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const int r = mt * 128 + q * 32 + lane;  // halo pixel
#pragma unroll
        for (int c0 = 0; c0 < NCOL; c0 += 48) {
          uint32_t v[3][16];
#pragma unroll
          for (int u = 0; u < 3; ++u) tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + mt * 128 + c0 + 16 * u, v[u]);
          tmem_ld_wait();
          if (r < NPIX) {
#pragma unroll
            for (int u = 0; u < 3; ++u) {
              float4* d = reinterpret_cast<float4*>(P + r * PSTR + c0 + 16 * u);
#pragma unroll
              for (int jj = 0; jj < 4; ++jj)
                d[jj] = make_float4(__uint_as_float(v[u][4 * jj]) * sc, __uint_as_float(v[u][4 * jj + 1]) * sc,
                                    __uint_as_float(v[u][4 * jj + 2]) * sc, __uint_as_float(v[u][4 * jj + 3]) * sc);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(tmem_empty);
        mbar_arrive(&p_full[it & 1]);
      }
    }
  } else {
    // ---- pass 2 warps: output rows.  Lane = output column X of the tile's 32; warp ew owns rows yl = ew and ew + 8 of every
    // channel.  Y + 1 = 2 oy + ky has the two solutions (oy1 = (Y+1) >> 1, ky1 = (Y+1) & 1) and (oy1 - 1, ky1 + 2); same along X.
    const int ew = warp - 6;
    const int H = 2 * Ho, W = 2 * Wo;
    const int hc1 = ((lane + 1) >> 1) + 1, kx1 = (lane + 1) & 1;  // halo column / tap of the first solution
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int tx = tile % tiles_x, ty = (tile / tiles_x) % tiles_y, n = tile / (tiles_x * tiles_y);
      const float* pl = Pbuf + (it & 1) * P_FLOATS + hc1 * PSTR + kx1;
      mbar_wait(&p_full[it & 1], (it >> 1) & 1);
#pragma unroll
      for (int cg = 0; cg < CIN; cg += 3) {  // all loads of a group of 3 channels are issued before the sums
        float a[6][4];
#pragma unroll
        for (int i = 0; i < 6; ++i) {
          const int c = cg + (i >> 1), yl = ew + 8 * (i & 1);
          const int hr1 = ((yl + 1) >> 1) + 1, ky1 = (yl + 1) & 1;
          const float* p11 = pl + hr1 * HC * PSTR + c * 16 + ky1 * 4;
          a[i][0] = p11[0]; a[i][1] = p11[-PSTR + 2]; a[i][2] = p11[-HC * PSTR + 8]; a[i][3] = p11[-HC * PSTR - PSTR + 10];
        }
        if (cg + 3 >= CIN) {  // last reads of this buffer
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_empty[it & 1]);
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) {
          const int c = cg + (i >> 1), yl = ew + 8 * (i & 1);
          float* plane = (c < 3 ? gx0 : gx1) + ((long long)n * 3 + (c % 3)) * H * W;
          plane[(long long)(2 * ty * TH + yl) * W + 2 * tx * TW + lane] = (a[i][0] + a[i][1]) + (a[i][2] + a[i][3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<256>(tmem_base);
}

static int launch_first_dgrad2(const void* gz, const float* w, const float* inv_sigma, float* gx0, float* gx1, int N, int Ho, int Wo,
                               cudaStream_t st) {
  CUtensorMap tmG;
  {
    uint64_t dims[4] = {64, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)N};
    uint64_t str[3] = {128, (uint64_t)128 * Wo, (uint64_t)128 * Wo * Ho};
    uint32_t box[4] = {64, (uint32_t)fd2::HC, (uint32_t)fd2::HR, 1};
    if (int rc = b2c_make_tmap_bf16(&tmG, gz, 4, dims, str, box)) return rc;
  }
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(first_dgrad2_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, fd2::SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(first_dgrad2_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, fd2::SMEM);
    if (e != cudaSuccess) { b2c_set_error("cudaFuncSetAttribute(first_dgrad2): %s", cudaGetErrorString(e)); return B2C_ERR_CUDA; }
    attr_done = true;
  }
  const long long ntiles = (long long)N * (Ho / fd2::TH) * (Wo / fd2::TW);
  const unsigned blocks = (unsigned)(ntiles < 148 ? ntiles : 148);
  if (gx1) first_dgrad2_kernel<6><<<blocks, fd2::THREADS, fd2::SMEM, st>>>(tmG, w, inv_sigma, gx0, gx1, N, Ho, Wo, (int)ntiles);
  else first_dgrad2_kernel<3><<<blocks, fd2::THREADS, fd2::SMEM, st>>>(tmG, w, inv_sigma, gx0, gx1, N, Ho, Wo, (int)ntiles);
  return b2c_check_launch("first_dgrad2");
}

// Called by b2c_first_dgrad (tapgemm.cu).  Returns B2C_OK with *done = false when the shape does not fit this kernel.
int b2c_first_dgrad2(const void* gz, const float* w, const float* inv_sigma, float* gx0, float* gx1, int N, int H, int W,
                     void* stream, bool* done) {
  static int enabled = -1;
  if (enabled < 0) {
    const char* e = getenv("B2C_FIRST_DGRAD2");
    enabled = (e && e[0] == '0') ? 0 : 1;
  }
  *done = false;
  const int Ho = H / 2, Wo = W / 2;
  if (!enabled || Ho % fd2::TH || Wo % fd2::TW || ((uintptr_t)gz & 15) || (long long)N * (Ho / fd2::TH) * (Wo / fd2::TW) >= (1ll << 31))
    return B2C_OK;
  *done = true;
  return launch_first_dgrad2(gz, w, inv_sigma, gx0, gx1, N, Ho, Wo, (cudaStream_t)stream);
}
