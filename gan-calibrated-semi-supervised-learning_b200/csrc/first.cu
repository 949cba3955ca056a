// First-layer convolutions (Cin = 3 for G.down1, 6 for D.model.0; cgan/models.py:103,245,257) on tcgen05.
//
// K = 16*Cin is only 48/96, so these layers are HBM-bound: the job of the kernels is to stream the
// caller's NCHW fp32 patches exactly once and keep the arithmetic off the CUDA cores.  Each thread owns one output
// pixel of a 128-pixel tile: it gathers its 4x4xCin receptive field with coalesced 8-byte loads (+ warp shuffles for
// the two halo columns), applies the optional GP interpolation x = alpha*a + (1-alpha)*b (cgan/losses.py:203-204) and
// the channel concat (cgan/models.py:257), and writes one 256-byte im2col row k = tap*8 + c (bf16, channels padded
// to 8) straight into the 128B-swizzled UMMA layout in shared memory.  That tile is
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

// Receptive field of one output pixel: v[c][ky] = (x[2ox], x[2ox+1]), l[c][ky] = x[2ox-1], r[c][ky] = x[2ox+2].
template <int CIN>
struct Field {
  float2 v[CIN][4];
  float l[CIN][4], r[CIN][4];
};

// Two phases so that every global load of the field is in flight before the first dependent instruction: phase 1
// issues the 4*CIN row loads (+ the predicated halo loads of the warp-edge lanes), phase 2 exchanges halo columns.
template <int CIN, bool INTERP>
__device__ __forceinline__ void gather_field(const FirstIn& in, long long pix, long long npix, int Ho, int Wo, int H,
                                             int W, Field<CIN>& f) {
  const int lane = threadIdx.x & 31;
  const bool live = pix < npix;
  int ox = 0, oy = 0, n = 0;
  if (live) {
    ox = (int)(pix % Wo);
    oy = (int)((pix / Wo) % Ho);
    n = (int)(pix / ((long long)Wo * Ho));
  }
  constexpr bool interp = INTERP;
  const float al = (interp && live) ? __ldg(in.alpha + n) : 1.f;
  const float be = 1.f - al;
  const bool want_l = lane == 0 && ox > 0, want_r = lane == 31 && ox < Wo - 1;
  float2 u[CIN][4];
  float lb[CIN][4], rb[CIN][4];
#pragma unroll
  for (int c = 0; c < CIN; ++c) {
    const int part = c / 3, cc = c % 3;
    const long long plane = ((long long)n * 3 + cc) * H;
#pragma unroll
    for (int ky = 0; ky < 4; ++ky) {
      const int iy = 2 * oy + ky - 1;
      const bool row_ok = live && iy >= 0 && iy < H;
      const long long off = row_ok ? (plane + iy) * W + 2 * ox : 0;
      const float* pa = in.a[part] + off;
      f.v[c][ky] = row_ok ? __ldg(reinterpret_cast<const float2*>(pa)) : make_float2(0.f, 0.f);
      f.l[c][ky] = (row_ok && want_l) ? __ldg(pa - 1) : 0.f;
      f.r[c][ky] = (row_ok && want_r) ? __ldg(pa + 2) : 0.f;
      if (interp) {
        const float* pb = in.b[part] + off;
        u[c][ky] = row_ok ? __ldg(reinterpret_cast<const float2*>(pb)) : make_float2(0.f, 0.f);
        lb[c][ky] = (row_ok && want_l) ? __ldg(pb - 1) : 0.f;
        rb[c][ky] = (row_ok && want_r) ? __ldg(pb + 2) : 0.f;
      }
    }
  }
#pragma unroll
  for (int c = 0; c < CIN; ++c) {
#pragma unroll
    for (int ky = 0; ky < 4; ++ky) {
      float2 t = f.v[c][ky];
      float le = f.l[c][ky], re = f.r[c][ky];
      if (interp) {
        t.x = al * t.x + be * u[c][ky].x;
        t.y = al * t.y + be * u[c][ky].y;
        le = al * le + be * lb[c][ky];
        re = al * re + be * rb[c][ky];
        f.v[c][ky] = t;
      }
      // halo columns come from the neighbouring lanes (same image row) except at warp / image-row edges
      const float lft = __shfl_up_sync(0xffffffffu, t.y, 1);
      const float rgt = __shfl_down_sync(0xffffffffu, t.x, 1);
      f.l[c][ky] = (lane == 0 || ox == 0) ? le : lft;
      f.r[c][ky] = (lane == 31 || ox == Wo - 1) ? re : rgt;
    }
  }
}

// Write im2col row `row` (k = tap*8 + c) into the swizzled tile; `ones` puts 1.0 into the padding slot k = 127.
template <int CIN>
__device__ __forceinline__ void store_field(uint8_t* sA, int row, const Field<CIN>& f, bool ones) {
  uint8_t* base = sA + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
  for (int ky = 0; ky < 4; ++ky) {
#pragma unroll
    for (int kx = 0; kx < 4; ++kx) {
      float e[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        if (c < CIN) e[c] = kx == 0 ? f.l[c][ky] : kx == 1 ? f.v[c][ky].x : kx == 2 ? f.v[c][ky].y : f.r[c][ky];
        else e[c] = 0.f;
      }
      const int tap = ky * 4 + kx;
      if (tap == 15 && ones) e[7] = 1.f;
      const uint4 u = make_uint4(pack_bf16(e[0], e[1]), pack_bf16(e[2], e[3]), pack_bf16(e[4], e[5]), pack_bf16(e[6], e[7]));
      *reinterpret_cast<uint4*>(base + (tap >> 3) * 16384 + (((tap & 7) ^ (row & 7)) << 4)) = u;
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
  const long long npix = (long long)N * Ho * Wo;
  const int ntiles = (int)((npix + 127) / 128);

  if (t == 0) {
    mbar_init(mma_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<64>(tmem_slot);
  // weights: B[co][k = tap*8 + c] = w[co][c][tap] / sigma, K-major, two 64-wide k-blocks of [64 rows x 128 B]
  {
    const float s = inv_sigma ? __ldg(inv_sigma) : 1.f;
    for (int i = t; i < 64 * 16; i += FT_THREADS) {
      const int co = i >> 4, tap = i & 15;
      float e[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) e[c] = c < CIN ? __ldg(w + ((long long)co * CIN + c) * 16 + tap) * s : 0.f;
      const uint4 u = make_uint4(pack_bf16(e[0], e[1]), pack_bf16(e[2], e[3]), pack_bf16(e[4], e[5]), pack_bf16(e[6], e[7]));
      *reinterpret_cast<uint4*>(sB + (tap >> 3) * 8192 + (co >> 3) * 1024 + (co & 7) * 128 + (((tap & 7) ^ (co & 7)) << 4)) = u;
    }
    if (t < 64) bias_s[t] = bias ? __ldg(bias + t) : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);

  Field<CIN> f;
  int tile = blockIdx.x;
  if (tile < ntiles) gather_field<CIN, INTERP>(in, (long long)tile * 128 + t, npix, Ho, Wo, H, W, f);
  uint32_t phase = 0;
  for (; tile < ntiles; tile += gridDim.x) {
    store_field<CIN>(sA, t, f, false);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (t < 32 && elect_one()) {
      tc_fence_after();
      const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        umma_bf16(tmem_base, umma_desc_sw128(a0 + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                  umma_desc_sw128(b0 + (k >> 2) * 8192 + (k & 3) * 32, 16, 1024), idesc, (uint32_t)(k != 0));
      umma_commit(mma_bar);
    }
    const long long pix = (long long)tile * 128 + t;
    const int next = tile + gridDim.x;
    if (next < ntiles) gather_field<CIN, INTERP>(in, (long long)next * 128 + t, npix, Ho, Wo, H, W, f);  // overlaps the MMA
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
          float a = __uint_as_float(v0[2 * j]) + bias_s[c0 + 2 * j], b = __uint_as_float(v0[2 * j + 1]) + bias_s[c0 + 2 * j + 1];
          float c = __uint_as_float(v1[2 * j]) + bias_s[c0 + 16 + 2 * j], d = __uint_as_float(v1[2 * j + 1]) + bias_s[c0 + 17 + 2 * j];
          a = a > 0.f ? a : a * slope; b = b > 0.f ? b : b * slope;
          c = c > 0.f ? c : c * slope; d = d > 0.f ? d : d * slope;
          o[j] = pack_bf16(a, b);
          o[8 + j] = pack_bf16(c, d);
        }
        uint4* d4 = reinterpret_cast<uint4*>(y + pix * 64 + c0);
        d4[0] = make_uint4(o[0], o[1], o[2], o[3]);
        d4[1] = make_uint4(o[4], o[5], o[6], o[7]);
        d4[2] = make_uint4(o[8], o[9], o[10], o[11]);
        d4[3] = make_uint4(o[12], o[13], o[14], o[15]);
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
  const long long npix = (long long)N * Ho * Wo;
  const int ntiles = (int)((npix + 127) / 128);

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

  Field<CIN> f;
  if (tile < ntiles) gather_field<CIN, INTERP>(in, (long long)tile * 128 + t, npix, Ho, Wo, H, W, f);
  uint32_t it = 0;
  for (; tile < ntiles; tile += gridDim.x, ++it) {
    if (it > 0) mbar_wait(mma_bar, (it - 1) & 1);  // the previous tile's MMAs are done reading sA and sB[(it-1)&1]
    const int next = tile + gridDim.x;
    if (t == 0 && next < ntiles) {
      mbar_expect_tx(&tma_bar[(it + 1) & 1], FT_B_BYTES);
      tma_load_2d(sB + ((it + 1) & 1) * FT_B_BYTES, &tmG, &tma_bar[(it + 1) & 1], 0, next * 128);
    }
    store_field<CIN>(sA, t, f, (long long)tile * 128 + t < npix);
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
    if (next < ntiles) gather_field<CIN, INTERP>(in, (long long)next * 128 + t, npix, Ho, Wo, H, W, f);
  }
  if (it > 0) {
    mbar_wait(mma_bar, (it - 1) & 1);
    tc_fence_after();
    const int tap = t >> 3, c = t & 7;
    // CTAs finish together: rotate the column order per CTA so concurrent flushes hit different addresses
    const int rot = (blockIdx.x * 16) & 63;
#pragma unroll 1
    for (int cc = 0; cc < 64; cc += 16) {
      const int c0 = (cc + rot) & 63;
      uint32_t v[16];
      tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + c0, v);
      tmem_ld_wait();
      if (c < CIN) {
#pragma unroll
        for (int j = 0; j < 16; ++j) atomicAdd(dw + ((long long)(c0 + j) * CIN + c) * 16 + tap, __uint_as_float(v[j]));
      } else if (t == 127 && db) {
#pragma unroll
        for (int j = 0; j < 16; ++j) atomicAdd(db + c0 + j, __uint_as_float(v[j]));
      }
    }
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

static int first_check(const float* a0, const float* b0, const float* alpha, int H, int W) {
  B2C_REQUIRE(H % 2 == 0 && W % 2 == 0, "first layer: odd size %dx%d", H, W);
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
static int launch_first_wgrad(const FirstIn& in, const CUtensorMap& tmG, float* dw, float* db, int N, int H, int W,
                              unsigned blocks, cudaStream_t st) {
  static bool attr_done = false;
  if (!attr_done) {
    int rc = first_attr(first_wgrad_tc_kernel<CIN, INTERP>);
    if (rc) return rc;
    attr_done = true;
  }
  first_wgrad_tc_kernel<CIN, INTERP><<<blocks, FT_THREADS, FT_SMEM, st>>>(in, tmG, dw, db, N, H, W);
  return b2c_check_launch("first_wgrad");
}

// a0/a1: (N,3,H,W) fp32 NCHW halves (a1 null => Cin=3).  b0/b1/alpha: optional interpolation (all or none).
extern "C" int b2c_first_fprop(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                               const float* w, const float* inv_sigma, const float* bias, void* y, int N, int H, int W,
                               float slope, void* stream) {
  int rc = first_check(a0, b0, alpha, H, W);
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
extern "C" int b2c_first_wgrad(const float* a0, const float* a1, const float* b0, const float* b1, const float* alpha,
                               const void* gz, float* dw, float* db, int N, int H, int W, void* stream) {
  int rc = first_check(a0, b0, alpha, H, W);
  if (rc) return rc;
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
  cudaStream_t st = (cudaStream_t)stream;
  if (a1) {
    if (alpha) return launch_first_wgrad<6, true>(in, tmG, dw, db, N, H, W, (unsigned)blocks, st);
    return launch_first_wgrad<6, false>(in, tmG, dw, db, N, H, W, (unsigned)blocks, st);
  }
  if (alpha) return launch_first_wgrad<3, true>(in, tmG, dw, db, N, H, W, (unsigned)blocks, st);
  return launch_first_wgrad<3, false>(in, tmG, dw, db, N, H, W, (unsigned)blocks, st);
}
