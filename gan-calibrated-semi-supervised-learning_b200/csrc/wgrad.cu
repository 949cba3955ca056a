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
  float* ws;               // [16][Cs][Cl] (atomic mode) or [splits][16][Cs][Cl] (deterministic mode)
  long long slab;          // 0: split-K partials are reduced with fp32 vector atomics into one slab (ws zeroed by the host);
                           // else: floats per slab, split s stores its partial tile into slab s (no atomics, no memset)
  int k3;                  // 1: the k3/s1/p1 convs of GeneratorSimpleRegressor (cgan/models.py:163-196): 9 taps, `large` at the
                           // resolution of `small` through a plain 4-D map, dW[a][b][ky][kx] = sum small[n,y,x,a] large[n,y+ky-1,x+kx-1,b]
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
__global__ void __launch_bounds__(352) wgrad_kernel(const __grid_constant__ CUtensorMap tmS,
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
      mbar_init(&full[s], 2);  // two producer threads (warps 0 and 10) split the stage's TMA loads
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
    if (warp == 0 || warp == 10) {
      // A stage is 2 + TPC small boxes (4-8 KB each): one issuing thread needs about as long per stage as the MMAs take, so
      // two threads share it - warp 0: the `small` boxes and the first half of the taps, warp 10: the second half.
      const int half = warp == 10;
      if (elect_one()) {
        int stage = 0, phase = 0;
        int tx = c_begin % p.tiles_x, ty = (c_begin / p.tiles_x) % p.tiles_y, tn = c_begin / (p.tiles_x * p.tiles_y);
        for (int c = c_begin; c < c_end; ++c) {
          const int x0 = tx * p.BW, y0 = ty * p.BH, n0 = tn * p.BI;
          if (++tx == p.tiles_x) { tx = 0; if (++ty == p.tiles_y) { ty = 0; ++tn; } }
          mbar_wait(&empty[stage], phase ^ 1);
          const int live = p.k3 ? (9 - tap0 < TPC ? 9 - tap0 : TPC) : TPC;  // taps of this CTA that exist
          const int t_lo = half ? TPC / 2 : 0, t_hi = half ? TPC : TPC / 2;
          const int mine = (live < t_hi ? live : t_hi) - t_lo;  // live taps in this thread's half (may be <= 0)
          mbar_expect_tx(&full[stage], (half ? 0 : Cfg::A_BYTES) + (mine > 0 ? mine : 0) * Cfg::B_TAP_BYTES);
          uint8_t* dA = sA + stage * Cfg::A_BYTES;
          uint8_t* dB = sB + stage * Cfg::B_BYTES;
          if (!half) {
            tma_load_4d(dA, &tmS, &full[stage], a0, x0, y0, n0);
            tma_load_4d(dA + Cfg::BOX_BYTES, &tmS, &full[stage], a0 + 64, x0, y0, n0);
          }
#pragma unroll
          for (int tt = 0; tt < TPC; ++tt) {
            if (tt < t_lo || tt >= t_hi) continue;
            if (p.k3) {
              if (tt < live) {
                const int ky = (tap0 + tt) / 3, kx = (tap0 + tt) % 3;
#pragma unroll
                for (int j = 0; j < NB / 64; ++j)
                  tma_load_4d(dB + tt * Cfg::B_TAP_BYTES + j * Cfg::BOX_BYTES, &tmL, &full[stage], b0 + 64 * j, x0 + kx - 1,
                              y0 + ky - 1, n0);
              }
              continue;
            }
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
      // The split-K CTAs of a tile finish together and flush to the SAME addresses: each starts its sweep over (tap, 16-column
      // chunk) at a different position, so concurrent atomics spread over the L2 instead of queueing on one line after another
      constexpr int CHUNKS = (TPC / 2) * (NB / 16);
      const int rot = p.slab ? 0 : (int)((blockIdx.y * 5u) % CHUNKS);
#pragma unroll 1
      for (int ci = 0; ci < CHUNKS; ++ci) {
        const int cr = (ci + rot) % CHUNKS;
        const int tt = ((warp - 2) >> 2) * (TPC / 2) + cr / (NB / 16), c0 = (cr % (NB / 16)) * 16;  // 2 warps per lane quarter
        if (p.k3 && tap0 + tt >= 9) continue;  // dead taps of the last group (their columns hold products of stale shared memory)
        const bool row_ok = a < p.Cs;          // Cs = 64: rows 64..127 are the zero-filled half of the `small` box
        float* dst = p.ws + (long long)blockIdx.y * p.slab + ((long long)(tap0 + tt) * p.Cs + a) * p.Cl + b0;
        {
          uint32_t v[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + tt * NB + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float4 f = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                   __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
            if (!row_ok) continue;
            if (p.slab) *reinterpret_cast<float4*>(dst + c0 + 4 * j) = f;
            else atomicAdd(reinterpret_cast<float4*>(dst + c0 + 4 * j), f);
          }
        }
      }
      tc_fence_before();
    }
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ CTA pairs
// The kernel above is L2 -> SM fill bound: per 32-pixel stage a CTA stages 8 KB of `small` and 32 KB of tap-shifted `large`
// boxes.  For Cs >= 256 two CTAs of a cluster take adjacent 128-channel tiles of `small` and issue ONE M = 256
// tcgen05.mma.cta_group::2 per tap pair: the N = 256 operand (two taps x 128 channels of `large`) is split between the two
// CTAs' shared memories, so each CTA stages only HALF of the tap boxes (16 KB + its own 8 KB: 24 instead of 40 KB per stage,
// and the ring deepens from 5 to 8 stages).  Protocol as in tapgemm_pair_kernel: both CTAs' TMA loads complete on the
// LEADER's mbarrier (cta_group::2 TMA), the leader's elected thread issues the MMAs, its commits are multicast to both CTAs.
struct WgradP2Cfg {
  static constexpr int NB = 128, TPC = 4, STAGES = 9, PIX = 32;
  static constexpr int BOX_BYTES = PIX * 128;
  static constexpr int A_BYTES = 2 * BOX_BYTES;              // this CTA's 128 channels of `small`
  static constexpr int B_TAP_BYTES = 2 * BOX_BYTES;          // 128 channels of `large` for one tap
  static constexpr int B_BYTES = (TPC / 2) * B_TAP_BYTES;    // this CTA's half of the taps
  static constexpr int SMEM = STAGES * (A_BYTES + B_BYTES) + 256 + 1024;
};

__global__ void __launch_bounds__(320, 1) wgrad_pair_kernel(const __grid_constant__ CUtensorMap tmS,
                                                            const __grid_constant__ CUtensorMap tmL, WgradParams p) {
  using Cfg = WgradP2Cfg;
  constexpr int STAGES = Cfg::STAGES, TPC = Cfg::TPC, NB = Cfg::NB;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * Cfg::A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + STAGES * Cfg::B_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank_u32();
  int t = blockIdx.x >> 1;                  // tile of the pair
  const int nt = t % p.ntiles; t /= p.ntiles;
  const int mt2 = t % (p.mtiles / 2); t /= (p.mtiles / 2);
  const int tap0 = t * TPC;
  const int a0 = (2 * mt2 + (int)rank) * 128, b0 = nt * NB;
  const int splits = gridDim.y;
  const int c_begin = (int)((long long)p.nchunks * blockIdx.y / splits);
  const int c_end = (int)((long long)p.nchunks * (blockIdx.y + 1) / splits);
  const int niters = c_end - c_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmS);
    tma_prefetch_desc(&tmL);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 2);   // (leader's) one arrive per producer of the pair + the transaction bytes of both
      mbar_init(&empty[s], 1);  // multicast commit from the leader's MMA thread
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc2<512>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0, phase = 0;
      int tx = c_begin % p.tiles_x, ty = (c_begin / p.tiles_x) % p.tiles_y, tn = c_begin / (p.tiles_x * p.tiles_y);
      for (int c = c_begin; c < c_end; ++c) {
        const int x0 = tx * p.BW, y0 = ty * p.BH, n0 = tn * p.BI;
        if (++tx == p.tiles_x) { tx = 0; if (++ty == p.tiles_y) { ty = 0; ++tn; } }
        mbar_wait(&empty[stage], phase ^ 1);
        const uint32_t lead_full = mapa_u32(smem_u32(&full[stage]), 0);
        if (rank == 0) mbar_expect_tx(&full[stage], 2 * (Cfg::A_BYTES + Cfg::B_BYTES));
        else mbar_arrive_cluster(lead_full);
        uint8_t* dA = sA + stage * Cfg::A_BYTES;
        uint8_t* dB = sB + stage * Cfg::B_BYTES;
        tma2_load_4d(dA, &tmS, lead_full, a0, x0, y0, n0);
        tma2_load_4d(dA + Cfg::BOX_BYTES, &tmS, lead_full, a0 + 64, x0, y0, n0);
#pragma unroll
        for (int g = 0; g < TPC / 2; ++g) {  // columns [256 g + 128 rank, +128) of the accumulators = tap tap0 + 2 g + rank
          const int tap = tap0 + 2 * g + (int)rank;
          const int ky = tap >> 2, kx = tap & 3;
          const int fy = (ky - 1) >> 1, di = (ky - 1) & 1, fx = (kx - 1) >> 1, dj = (kx - 1) & 1;
#pragma unroll
          for (int j = 0; j < 2; ++j)
            tma2_load_5d(dB + g * Cfg::B_TAP_BYTES + j * Cfg::BOX_BYTES, &tmL, lead_full, dj * p.Cl + b0 + 64 * j, x0 + fx, di,
                         y0 + fy, n0);
        }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(256, 256, 1, 1);
      int stage = 0, phase = 0;
      for (int it = 0; it < niters; ++it) {
        mbar_wait_cluster(&full[stage], phase);  // both CTAs' boxes have landed
        tc_fence_after();
        const uint32_t sa = smem_u32(sA + stage * Cfg::A_BYTES), sb = smem_u32(sB + stage * Cfg::B_BYTES);
#pragma unroll
        for (int g = 0; g < TPC / 2; ++g)
#pragma unroll
          for (int k = 0; k < 2; ++k)  // 16 pixels (= 2 swizzle atoms of 8 k-rows) per MMA
            umma2_bf16(tmem_base + g * 256, umma_desc_sw128(sa + k * 2048, Cfg::BOX_BYTES, 1024),
                       umma_desc_sw128(sb + g * Cfg::B_TAP_BYTES + k * 2048, Cfg::BOX_BYTES, 1024), idesc,
                       (uint32_t)((it | k) != 0));
        umma2_commit(&empty[stage]);  // frees the ring slot in BOTH CTAs
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      umma2_commit(tmem_full);
    }
  } else {
    const int q = warp & 3;
    const int a = a0 + q * 32 + lane;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    constexpr int CHUNKS = (TPC / 2) * (NB / 16);
    const int rot = p.slab ? 0 : (int)((blockIdx.y * 5u) % CHUNKS);  // see wgrad_kernel
#pragma unroll 1
    for (int ci = 0; ci < CHUNKS; ++ci) {
      const int cr = (ci + rot) % CHUNKS;
      const int tt = ((warp - 2) >> 2) * (TPC / 2) + cr / (NB / 16), c0 = (cr % (NB / 16)) * 16;  // 2 warps per lane quarter
      float* dst = p.ws + (long long)blockIdx.y * p.slab + ((long long)(tap0 + tt) * p.Cs + a) * p.Cl + b0;
      {
        uint32_t v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + tt * NB + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float4 f = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                 __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
          if (p.slab) *reinterpret_cast<float4*>(dst + c0 + 4 * j) = f;
          else atomicAdd(reinterpret_cast<float4*>(dst + c0 + 4 * j), f);
        }
      }
    }
    tc_fence_before();
  }
  __syncwarp();
  cluster_sync_all();  // the peer's shared memory and barriers stay alive until the leader's last MMA / commit has landed
  if (warp == 1) tmem_dealloc2<512>(tmem_base);
}

// ws[split][tap][a][b] -> dw[a][b][ky][kx] (torch weight layout): fixed-order sum over the split-K slabs, optional
// device-scalar scale (spectral norm: dW_orig gets dWe / sigma), optional accumulation into dw, and optionally the block
// partials of <dWe, wdot> (unscaled; wdot in dw's layout) for the deferred rank-1 term of the spectral-norm backward
// (b2c_sn_rank1): part[blockIdx.x], summed later in a fixed order - no atomics anywhere.
__global__ void __launch_bounds__(256) wgrad_unpack_kernel(const float* __restrict__ ws, long long slab, int nslabs,
                                                           float* __restrict__ dw, int Cs, int Cl, int accumulate,
                                                           const float* __restrict__ scale, const float* __restrict__ wdot,
                                                           float* __restrict__ part) {
  __shared__ float sh[8];
  const long long n = (long long)Cs * Cl;
  const float sc = scale ? __ldg(scale) : 1.f;
  float dot = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float v[16];
#pragma unroll
    for (int t = 0; t < 16; ++t) v[t] = ws[(long long)t * n + i];
    for (int sl = 1; sl < nslabs; ++sl) {
      const float* w2 = ws + (long long)sl * slab;
#pragma unroll
      for (int t = 0; t < 16; ++t) v[t] += w2[(long long)t * n + i];
    }
    float4* o = reinterpret_cast<float4*>(dw + i * 16);
    const float4* wd = reinterpret_cast<const float4*>(wdot + i * 16);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (wdot) {
        const float4 w4 = __ldg(wd + j);
        dot += v[4 * j] * w4.x + v[4 * j + 1] * w4.y + v[4 * j + 2] * w4.z + v[4 * j + 3] * w4.w;
      }
      float4 f = make_float4(v[4 * j] * sc, v[4 * j + 1] * sc, v[4 * j + 2] * sc, v[4 * j + 3] * sc);
      if (accumulate) { float4 g = o[j]; f.x += g.x; f.y += g.y; f.z += g.z; f.w += g.w; }
      o[j] = f;
    }
  }
  if (part) {
    dot = warp_sum(dot);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = dot;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) t += sh[w];
      part[blockIdx.x] = t;
    }
  }
}

// ------------------------------------------------------------------------------------------------ host side

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
  wgrad_kernel<NB><<<grid, 352, WgradCfg<NB>::SMEM, st>>>(tmS, tmL, p);
  return b2c_check_launch("wgrad_kernel");
}

struct WgradPlan { int BW, BH, BI, tiles_x, tiles_y, nchunks, NB, tiles, splits; };
static WgradPlan wgrad_plan(int N, int Hs, int Ws, int Cs, int Cl, int sm_count) {
  WgradPlan q;
  const int lw = Ws & -Ws, lh = Hs & -Hs;  // largest powers of two dividing the extents: any spatial size works
  q.BW = lw < 32 ? lw : 32;
  int rem = 32 / q.BW;
  q.BH = lh < rem ? lh : rem;
  q.BI = rem / q.BH;
  q.tiles_x = Ws / q.BW; q.tiles_y = Hs / q.BH;
  const int tiles_n = (N + q.BI - 1) / q.BI;
  q.nchunks = q.tiles_x * q.tiles_y * tiles_n;
  q.NB = (Cl % 128 == 0) ? 128 : 64;
  // One CTA per SM (it owns all 512 TMEM columns).  Pick the split-K factor by a small cost model: waves of CTAs times
  // (pixel chunks per CTA + a fixed prologue/epilogue cost of about 12 chunk-times).
  q.tiles = (16 / (512 / q.NB)) * (Cs / 128) * (Cl / q.NB);
  const int sms = sm_count > 0 ? sm_count : 148;
  int splits = 1;
  double best = 1e30;
  for (int s = 1; s <= 2 * sms && s <= q.nchunks; ++s) {
    const int waves = (q.tiles * s + sms - 1) / sms;
    const double cost = waves * ((double)((q.nchunks + s - 1) / s) + 12.0);
    if (cost < best - 1e-9) { best = cost; splits = s; }
  }
  if (splits > q.nchunks) splits = q.nchunks;
  if (splits < 1) splits = 1;
  q.splits = splits;
  return q;
}

// Workspace (in floats) that makes b2c_wgrad4s2_bf16 deterministic for this shape: one [16][Cs][Cl] slab per split-K
// partition.  With less (but >= 16*Cs*Cl) the partitions are reduced with fp32 atomics instead (run-to-run rounding noise).
extern "C" long long b2c_wgrad4s2_ws_floats(int N, int Hs, int Ws, int Cs, int Cl, int sm_count) {
  if (Hs <= 0 || Ws <= 0 || Cs <= 0 || Cl <= 0 || N <= 0) return 0;
  const WgradPlan q = wgrad_plan(N, Hs, Ws, Cs, Cl, sm_count);
  return (long long)q.splits * 16 * Cs * Cl;
}

// small: (N,Hs,Ws,Cs) bf16 with pixel stride lds; large: dense (N,2Hs,2Ws,Cl) bf16; ws: fp32 workspace of ws_floats
// floats (see b2c_wgrad4s2_ws_floats); dw: fp32 [Cs][Cl][4][4], overwritten or (accumulate) added to, times *scale
// (device scalar, nullable).  wdot/part/nparts_host (all nullable): part[0..*nparts_host) receive block partials of
// <dW_unscaled, wdot> (spectral-norm backward, see b2c_sn_rank1); part needs room for 1024 floats.
// sm_count sizes the split-K factor.
extern "C" int b2c_wgrad4s2_bf16(const void* small, long long lds, const void* large, long long ldl, float* ws,
                                 long long ws_floats, float* dw, int accumulate, const float* scale, const float* wdot,
                                 float* part, int* nparts_host, int N, int Hs, int Ws, int Cs, int Cl, int sm_count,
                                 void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B2C_REQUIRE(Cs % 128 == 0 && Cl % 64 == 0, "wgrad4s2: Cs=%d must be a multiple of 128, Cl=%d of 64", Cs, Cl);
  B2C_REQUIRE(ldl == Cl, "wgrad4s2: large operand must be dense NHWC (ldl=%lld, Cl=%d)", ldl, Cl);
  B2C_REQUIRE(lds % 8 == 0, "wgrad4s2: lds=%lld must be a multiple of 8", lds);
  B2C_REQUIRE(Hs > 0 && Ws > 0, "wgrad4s2: unsupported spatial size %dx%d", Hs, Ws);
  B2C_REQUIRE((wdot == nullptr) == (part == nullptr), "wgrad4s2: wdot and part go together");
  const WgradPlan q = wgrad_plan(N, Hs, Ws, Cs, Cl, sm_count);
  const long long slab = (long long)16 * Cs * Cl;
  B2C_REQUIRE(ws_floats >= slab, "wgrad4s2: workspace of %lld floats < %lld", ws_floats, slab);
  const bool det = ws_floats >= slab * q.splits;
  WgradParams p;
  p.N = N; p.Cs = Cs; p.Cl = Cl; p.ws = ws; p.slab = det ? slab : 0; p.k3 = 0;
  p.BW = q.BW; p.BH = q.BH; p.BI = q.BI; p.tiles_x = q.tiles_x; p.tiles_y = q.tiles_y; p.nchunks = q.nchunks;
  const int NB = q.NB;
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
  if (!det) {
    cudaError_t e = cudaMemsetAsync(ws, 0, sizeof(float) * (size_t)slab, st);
    if (e != cudaSuccess) {
      b2c_set_error("wgrad4s2 memset: %s", cudaGetErrorString(e));
      return B2C_ERR_CUDA;
    }
  }
  dim3 grid(q.tiles, q.splits);
  static const bool pair_ok = !(getenv("B2C_WGRAD_2CTA") && getenv("B2C_WGRAD_2CTA")[0] == '0');
  int rc;
  if (pair_ok && NB == 128 && Cs % 256 == 0) {
    static bool attr_done = false;
    if (!attr_done) {
      cudaError_t e = cudaFuncSetAttribute(wgrad_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WgradP2Cfg::SMEM);
      if (e != cudaSuccess) { b2c_set_error("cudaFuncSetAttribute(wgrad_pair): %s", cudaGetErrorString(e)); return B2C_ERR_CUDA; }
      attr_done = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;  // x = 2 * (tap groups x channel-tile pairs x Cl tiles): CTAs 2i, 2i+1 form a cluster
    cfg.blockDim = dim3(320);
    cfg.dynamicSmemBytes = WgradP2Cfg::SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, wgrad_pair_kernel, tmS, tmL, p);
    if (e != cudaSuccess) { b2c_set_error("wgrad_pair_kernel launch: %s", cudaGetErrorString(e)); return B2C_ERR_CUDA; }
    rc = b2c_check_launch("wgrad_pair_kernel");
  } else {
    rc = NB == 128 ? launch_wgrad<128>(tmS, tmL, p, grid, st) : launch_wgrad<64>(tmS, tmL, p, grid, st);
  }
  if (rc) return rc;
  const long long n = (long long)Cs * Cl;
  long long blocks = (n + 255) / 256;
  if (blocks > 1024) blocks = 1024;
  if (nparts_host) *nparts_host = (int)blocks;
  wgrad_unpack_kernel<<<(unsigned)blocks, 256, 0, st>>>(ws, slab, det ? q.splits : 1, dw, Cs, Cl, accumulate, scale, wdot, part);
  return b2c_check_launch("wgrad_unpack_kernel");
}

// ------------------------------------------------------------------------------------------------ k3 / s1 / p1
// ws[split][9][a][b] -> dw[a][b][3][3], fixed-order sum over the slabs, optional accumulation.
__global__ void __launch_bounds__(256) wgrad_unpack9_kernel(const float* __restrict__ ws, long long slab, int nslabs,
                                                            float* __restrict__ dw, long long n, int accumulate) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float v[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) v[t] = ws[(long long)t * n + i];
    for (int sl = 1; sl < nslabs; ++sl) {
      const float* w2 = ws + (long long)sl * slab;
#pragma unroll
      for (int t = 0; t < 9; ++t) v[t] += w2[(long long)t * n + i];
    }
#pragma unroll
    for (int t = 0; t < 9; ++t) dw[i * 9 + t] = accumulate ? dw[i * 9 + t] + v[t] : v[t];
  }
}

static WgradPlan wgrad3_plan(int N, int H, int W, int Cs, int Cl, int sm_count) {
  WgradPlan q = wgrad_plan(N, H, W, Cs >= 128 ? Cs : 128, Cl, sm_count);  // box geometry and NB
  const int tpc = 512 / q.NB;
  q.tiles = ((9 + tpc - 1) / tpc) * ((Cs + 127) / 128) * (Cl / q.NB);
  const int sms = sm_count > 0 ? sm_count : 148;
  int splits = 1;
  double best = 1e30;
  for (int s = 1; s <= 2 * sms && s <= q.nchunks; ++s) {
    const int waves = (q.tiles * s + sms - 1) / sms;
    const double cost = waves * ((double)((q.nchunks + s - 1) / s) + 12.0);
    if (cost < best - 1e-9) { best = cost; splits = s; }
  }
  q.splits = splits < 1 ? 1 : splits;
  return q;
}

extern "C" long long b2c_wgrad3s1_ws_floats(int N, int H, int W, int Cs, int Cl, int sm_count) {
  if (H <= 0 || W <= 0 || Cs <= 0 || Cl <= 0 || N <= 0) return 0;
  return (long long)wgrad3_plan(N, H, W, Cs, Cl, sm_count).splits * 9 * Cs * Cl;
}

// Weight gradient of nn.Conv2d(Cl, Cs, 3, padding=1) (cgan/models.py:163-196): small = grad of the conv output (N,H,W,Cs),
// large = the conv input (N,H,W,Cl), both dense NHWC bf16; dw fp32 [Cs][Cl][3][3] overwritten or (accumulate) added to.
// ws: >= 9*Cs*Cl floats (atomic split-K) or b2c_wgrad3s1_ws_floats (deterministic).  Cs: 64 or a multiple of 128.
extern "C" int b2c_wgrad3s1_bf16(const void* small, const void* large, float* ws, long long ws_floats, float* dw,
                                 int accumulate, int N, int H, int W, int Cs, int Cl, int sm_count, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B2C_REQUIRE((Cs == 64 || Cs % 128 == 0) && Cl % 64 == 0, "wgrad3s1: Cs=%d must be 64 or a multiple of 128, Cl=%d of 64", Cs, Cl);
  B2C_REQUIRE(H > 0 && W > 0, "wgrad3s1: unsupported spatial size %dx%d", H, W);
  const WgradPlan q = wgrad3_plan(N, H, W, Cs, Cl, sm_count);
  const long long slab = (long long)9 * Cs * Cl;
  B2C_REQUIRE(ws_floats >= slab, "wgrad3s1: workspace of %lld floats < %lld", ws_floats, slab);
  const bool det = ws_floats >= slab * q.splits;
  WgradParams p;
  p.N = N; p.Cs = Cs; p.Cl = Cl; p.ws = ws; p.slab = det ? slab : 0; p.k3 = 1;
  p.BW = q.BW; p.BH = q.BH; p.BI = q.BI; p.tiles_x = q.tiles_x; p.tiles_y = q.tiles_y; p.nchunks = q.nchunks;
  const int NB = q.NB;
  p.mtiles = (Cs + 127) / 128; p.ntiles = Cl / NB;
  CUtensorMap tmS, tmL;
  for (int which = 0; which < 2; ++which) {
    const int C = which ? Cl : Cs;
    uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t str[3] = {(uint64_t)2 * C, (uint64_t)2 * W * C, (uint64_t)2 * H * W * C};
    uint32_t box[4] = {64, (uint32_t)p.BW, (uint32_t)p.BH, (uint32_t)p.BI};
    int rc = b2c_make_tmap_bf16(which ? &tmL : &tmS, which ? large : small, 4, dims, str, box);
    if (rc) return rc;
  }
  if (!det) {
    cudaError_t e = cudaMemsetAsync(ws, 0, sizeof(float) * (size_t)slab, st);
    if (e != cudaSuccess) {
      b2c_set_error("wgrad3s1 memset: %s", cudaGetErrorString(e));
      return B2C_ERR_CUDA;
    }
  }
  dim3 grid(q.tiles, q.splits);
  int rc = NB == 128 ? launch_wgrad<128>(tmS, tmL, p, grid, st) : launch_wgrad<64>(tmS, tmL, p, grid, st);
  if (rc) return rc;
  const long long n = (long long)Cs * Cl;
  long long blocks = (n + 255) / 256;
  if (blocks > 1024) blocks = 1024;
  wgrad_unpack9_kernel<<<(unsigned)blocks, 256, 0, st>>>(ws, slab, det ? q.splits : 1, dw, n, accumulate);
  return b2c_check_launch("wgrad_unpack9_kernel");
}
