// tcgen05 weight-gradient kernel for every k4/s2/p1 conv / convT of the cGAN step (sm_100a only).
//
//   dW[a][b][ky][kx] = sum_{n,sy,sx} small[n,sy,sx,a] * large[n,2sy+ky-1,2sx+kx-1,b]
// conv   (cgan/models.py:57,236): small = grad of the output (Cout), large = input (Cin)  -> dW[Cout][Cin][4][4]
// convT  (cgan/models.py:72,113): small = input (Cin), large = grad of the output (Cout)  -> dW[Cin][Cout][4][4]
// and the double-backward term of the gradient penalty (cgan/losses.py:213-220): small = g_z, large = cotangent h.
//
// GEMM per tap: D[128 a x NB b] += A^T B with the reduction over pixels.  Both operands sit in NHWC, i.e. the
// reduction index (pixel) is the slow one: they are MN-major UMMA operands, loaded as [32 pixels x 64 channels]
// 128B-swizzled TMA boxes (the "large" one through the same strided 5-D view the fprop kernel uses).
// The kernel is L2->SM bandwidth bound, so one CTA owns TPC = 512/NB taps of its (128 x NB) weight tile: the `small`
// box of a pixel chunk is loaded ONCE and multiplied against TPC tap-shifted `large` boxes, with TPC accumulators side
// by side in the 512 TMEM columns (105 FLOP per staged byte instead of 64 / 44 with one tap per CTA).  Split-K over
// pixel chunks; partial tiles are reduced with vector fp32 atomics into a tap-major workspace ws[16][Cs][Cl].
#include "common.cuh"
#include "tmap.h"

using namespace b2c;

struct WgradParams {
  int N, Cs, Cl;
  int BW, BH, BI;          // 32-pixel box over the small tensor
  int tiles_x, tiles_y;    // box grid over (Hs, Ws)
  int nchunks;             // total pixel chunks
  int mtiles, ntiles;      // Cs/128, Cl/NB
  float* ws;               // [16][Cs][Cl]
};

template <int NB>
struct WgradCfg {
  static constexpr int TPC = 512 / NB;                       // taps per CTA
  static constexpr int STAGES = 5;
  static constexpr int PIX = 32;                             // pixels (K) per stage = 2 MMAs of K=16
  static constexpr int BOX_BYTES = PIX * 128;                // one [32 px x 64 ch] box
  static constexpr int A_BYTES = 2 * BOX_BYTES;              // 128 channels of `small`
  static constexpr int B_TAP_BYTES = (NB / 64) * BOX_BYTES;  // NB channels of `large` for one tap
  static constexpr int B_BYTES = TPC * B_TAP_BYTES;
  static constexpr int SMEM = STAGES * (A_BYTES + B_BYTES) + 256 + 1024;
};

template <int NB>
__global__ void __launch_bounds__(320) wgrad_kernel(const __grid_constant__ CUtensorMap tmS,
                                                    const __grid_constant__ CUtensorMap tmL, WgradParams p) {
  using Cfg = WgradCfg<NB>;
  constexpr int STAGES = Cfg::STAGES, TPC = Cfg::TPC;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * Cfg::A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + STAGES * Cfg::B_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int t = blockIdx.x;
  const int nt = t % p.ntiles; t /= p.ntiles;
  const int mt = t % p.mtiles; t /= p.mtiles;
  const int tap0 = t * TPC;
  const int a0 = mt * 128, b0 = nt * NB;
  const int splits = gridDim.y;
  const int c_begin = (int)((long long)p.nchunks * blockIdx.y / splits);
  const int c_end = (int)((long long)p.nchunks * (blockIdx.y + 1) / splits);
  const int niters = c_end - c_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmS);
    tma_prefetch_desc(&tmL);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (niters > 0) {
    if (warp == 0) {
      if (elect_one()) {
        int stage = 0, phase = 0;
        int tx = c_begin % p.tiles_x, ty = (c_begin / p.tiles_x) % p.tiles_y, tn = c_begin / (p.tiles_x * p.tiles_y);
        for (int c = c_begin; c < c_end; ++c) {
          const int x0 = tx * p.BW, y0 = ty * p.BH, n0 = tn * p.BI;
          if (++tx == p.tiles_x) { tx = 0; if (++ty == p.tiles_y) { ty = 0; ++tn; } }
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_expect_tx(&full[stage], Cfg::A_BYTES + Cfg::B_BYTES);
          uint8_t* dA = sA + stage * Cfg::A_BYTES;
          uint8_t* dB = sB + stage * Cfg::B_BYTES;
          tma_load_4d(dA, &tmS, &full[stage], a0, x0, y0, n0);
          tma_load_4d(dA + Cfg::BOX_BYTES, &tmS, &full[stage], a0 + 64, x0, y0, n0);
#pragma unroll
          for (int tt = 0; tt < TPC; ++tt) {
            const int ky = (tap0 + tt) >> 2, kx = (tap0 + tt) & 3;
            const int fy = (ky - 1) >> 1, di = (ky - 1) & 1, fx = (kx - 1) >> 1, dj = (kx - 1) & 1;
#pragma unroll
            for (int j = 0; j < NB / 64; ++j)
              tma_load_5d(dB + tt * Cfg::B_TAP_BYTES + j * Cfg::BOX_BYTES, &tmL, &full[stage],
                          dj * p.Cl + b0 + 64 * j, x0 + fx, di, y0 + fy, n0);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    } else if (warp == 1) {
      if (elect_one()) {
        // taps sit side by side in shared memory (64-channel blocks BOX_BYTES apart) and in TMEM (NB columns each), so
        // 256 / NB taps go into ONE N = 256 instruction: 4 MMAs per stage instead of 8 or 16
        constexpr uint32_t idesc = umma_idesc_bf16(128, 256, 1, 1);
        constexpr int TPM = 256 / NB;  // taps per MMA
        int stage = 0, phase = 0;
        for (int it = 0; it < niters; ++it) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(sA + stage * Cfg::A_BYTES), sb = smem_u32(sB + stage * Cfg::B_BYTES);
#pragma unroll
          for (int g = 0; g < TPC / TPM; ++g)
#pragma unroll
            for (int k = 0; k < 2; ++k)  // 16 pixels (= 2 swizzle atoms of 8 k-rows) per MMA
              umma_bf16(tmem_base + g * 256, umma_desc_sw128(sa + k * 2048, Cfg::BOX_BYTES, 1024),
                        umma_desc_sw128(sb + g * TPM * Cfg::B_TAP_BYTES + k * 2048, Cfg::BOX_BYTES, 1024), idesc,
                        (uint32_t)((it | k) != 0));
          umma_commit(&empty[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(tmem_full);
      }
    } else {
      const int q = warp & 3;
      const int a = a0 + q * 32 + lane;
      mbar_wait(tmem_full, 0);
      tc_fence_after();
#pragma unroll 1
      for (int tt = ((warp - 2) >> 2) * (TPC / 2); tt < (((warp - 2) >> 2) + 1) * (TPC / 2); ++tt) {  // 2 warps per lane quarter
        float* dst = p.ws + ((long long)(tap0 + tt) * p.Cs + a) * p.Cl + b0;
#pragma unroll 1
        for (int c0 = 0; c0 < NB; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + tt * NB + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float4 f = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                   __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
            atomicAdd(reinterpret_cast<float4*>(dst + c0 + 4 * j), f);
          }
        }
      }
      tc_fence_before();
    }
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// ws[tap][a][b] -> dw[a][b][ky][kx] (torch weight layout), optionally accumulating.
__global__ void wgrad_unpack_kernel(const float* __restrict__ ws, float* __restrict__ dw, int Cs, int Cl, int accumulate) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // over Cs*Cl
  if (i >= (long long)Cs * Cl) return;
  float v[16];
#pragma unroll
  for (int t = 0; t < 16; ++t) v[t] = ws[(long long)t * Cs * Cl + i];
  float4* o = reinterpret_cast<float4*>(dw + i * 16);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float4 f = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    if (accumulate) { float4 g = o[j]; f.x += g.x; f.y += g.y; f.z += g.z; f.w += g.w; }
    o[j] = f;
  }
}

// ------------------------------------------------------------------------------------------------ host side
static bool w_is_pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }

template <int NB>
static int launch_wgrad(const CUtensorMap& tmS, const CUtensorMap& tmL, const WgradParams& p, dim3 grid, cudaStream_t st) {
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, WgradCfg<NB>::SMEM);
    if (e != cudaSuccess) {
      b2c_set_error("cudaFuncSetAttribute(wgrad): %s", cudaGetErrorString(e));
      return B2C_ERR_CUDA;
    }
    attr_done = true;
  }
  wgrad_kernel<NB><<<grid, 320, WgradCfg<NB>::SMEM, st>>>(tmS, tmL, p);
  return b2c_check_launch("wgrad_kernel");
}

// small: (N,Hs,Ws,Cs) bf16 with pixel stride lds; large: dense (N,2Hs,2Ws,Cl) bf16; ws: fp32 [16][Cs][Cl] workspace
// (zeroed here); dw: fp32 [Cs][Cl][4][4].  sm_count sizes the split-K factor.
extern "C" int b2c_wgrad4s2_bf16(const void* small, long long lds, const void* large, long long ldl, float* ws,
                                 float* dw, int accumulate, int N, int Hs, int Ws, int Cs, int Cl, int sm_count,
                                 void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B2C_REQUIRE(Cs % 128 == 0 && Cl % 64 == 0, "wgrad4s2: Cs=%d must be a multiple of 128, Cl=%d of 64", Cs, Cl);
  B2C_REQUIRE(ldl == Cl, "wgrad4s2: large operand must be dense NHWC (ldl=%lld, Cl=%d)", ldl, Cl);
  B2C_REQUIRE(lds % 8 == 0, "wgrad4s2: lds=%lld must be a multiple of 8", lds);
  B2C_REQUIRE(w_is_pow2(Hs) && w_is_pow2(Ws), "wgrad4s2: unsupported spatial size %dx%d", Hs, Ws);
  WgradParams p;
  p.N = N; p.Cs = Cs; p.Cl = Cl; p.ws = ws;
  p.BW = Ws < 32 ? Ws : 32;
  int rem = 32 / p.BW;
  p.BH = Hs < rem ? Hs : rem;
  p.BI = rem / p.BH;
  p.tiles_x = Ws / p.BW; p.tiles_y = Hs / p.BH;
  const int tiles_n = (N + p.BI - 1) / p.BI;
  p.nchunks = p.tiles_x * p.tiles_y * tiles_n;
  const int NB = (Cl % 128 == 0) ? 128 : 64;
  p.mtiles = Cs / 128; p.ntiles = Cl / NB;
  CUtensorMap tmS, tmL;
  {
    uint64_t dims[4] = {(uint64_t)Cs, (uint64_t)Ws, (uint64_t)Hs, (uint64_t)N};
    uint64_t str[3] = {(uint64_t)2 * lds, (uint64_t)2 * Ws * lds, (uint64_t)2 * Hs * Ws * lds};
    uint32_t box[4] = {64, (uint32_t)p.BW, (uint32_t)p.BH, (uint32_t)p.BI};
    int rc = b2c_make_tmap_bf16(&tmS, small, 4, dims, str, box);
    if (rc) return rc;
  }
  {
    const int H = 2 * Hs, W = 2 * Ws;
    uint64_t dims[5] = {(uint64_t)2 * Cl, (uint64_t)Ws, 2, (uint64_t)Hs, (uint64_t)N};
    uint64_t str[4] = {(uint64_t)4 * Cl, (uint64_t)2 * W * Cl, (uint64_t)4 * W * Cl, (uint64_t)2 * H * W * Cl};
    uint32_t box[5] = {64, (uint32_t)p.BW, 1, (uint32_t)p.BH, (uint32_t)p.BI};
    int rc = b2c_make_tmap_bf16(&tmL, large, 5, dims, str, box);
    if (rc) return rc;
  }
  // One CTA per SM (it owns all 512 TMEM columns).  Pick the split-K factor by a small cost model: waves of CTAs times
  // (pixel chunks per CTA + a fixed prologue/epilogue cost of about 12 chunk-times).
  const int tiles = (16 / (512 / NB)) * p.mtiles * p.ntiles;
  const int sms = sm_count > 0 ? sm_count : 148;
  int splits = 1;
  {
    double best = 1e30;
    for (int s = 1; s <= 2 * sms && s <= p.nchunks; ++s) {
      const int waves = (tiles * s + sms - 1) / sms;
      const double cost = waves * ((double)((p.nchunks + s - 1) / s) + 12.0);
      if (cost < best - 1e-9) { best = cost; splits = s; }
    }
  }
  if (splits > p.nchunks) splits = p.nchunks;
  if (splits < 1) splits = 1;
  cudaError_t e = cudaMemsetAsync(ws, 0, sizeof(float) * 16 * (size_t)Cs * Cl, st);
  if (e != cudaSuccess) {
    b2c_set_error("wgrad4s2 memset: %s", cudaGetErrorString(e));
    return B2C_ERR_CUDA;
  }
  dim3 grid(tiles, splits);
  int rc = NB == 128 ? launch_wgrad<128>(tmS, tmL, p, grid, st) : launch_wgrad<64>(tmS, tmL, p, grid, st);
  if (rc) return rc;
  const long long n = (long long)Cs * Cl;
  wgrad_unpack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(ws, dw, Cs, Cl, accumulate);
  return b2c_check_launch("wgrad_unpack_kernel");
}
