// Parity-merged transposed convolution for Cout = 64 (sm_100a only): ConvTranspose2d(k4,s2,p1) fprop of the
// generator's up3 / up4 (cgan/models.py:111-113) and the data gradient of every Conv2d(64 -> 128) (models.py:104,239).
//
// tapgemm.cu runs the four output parities (py,px) as four independent GEMMs with N = 64.  Those are bound by the
// L2 -> shared-memory stream of the ACTIVATION operand: 16 boxes of [128 pixels x 64 channels] per 64-channel slice of
// Cin, although only 9 distinct shifted boxes exist (parity (py,px) with tap (a,b) reads x[i+py+a-1, j+px+b-1], and
// py+a-1, px+b-1 take the values -1, 0, +1).  Here one CTA owns all four parities of its 128 input-resolution pixels:
// it walks the 9 shifts, loads each activation box ONCE and multiplies it against the 1, 2 or 4 weight slices
// (parity, tap) that use that shift, into four 64-column accumulators side by side in TMEM (256 columns).
// Staged bytes per pixel tile and 64-channel slice drop from 16*(16+8) = 384 KB to 9*16 + 16*8 = 272 KB.
// Warp roles: warp 0 = TMA producer of the activation ring, warp 6 = TMA producer of the weight ring, warp 1 = tcgen05.mma
// issuer, warps 2-5 = epilogue.
// Activation boxes and weight slices travel through two separate mbarrier rings (a box is used by a variable number of
// slices), produced and consumed in the same order, so both rings stay full: 208 KB in flight per SM.
#include <cstdlib>
#include "common.cuh"
#include "tmap.h"

using namespace b2c;

struct ConvT64Params {
  int N, H, W, Cin;      // input (N,H,W,Cin); output (N,2H,2W,64)
  int BW, BH, BI;        // pixel box over the input grid, BW*BH*BI == 128
  int tiles_x, tiles_y;
  int kchunks;           // Cin / 64
  __nv_bfloat16* y;
  long long ldy;
  long long* stats;      // nullable: [N][64][2] += fixed-point (stat_add) per-(image, channel) (sum, sum of squares) of y; needs BI == 1
  int f32out;            // fp32 check mode (b2c_set_f32_output): y is float*, accumulators stored unrounded
  const __nv_bfloat16* act;  // nullable: dense (N,2H,2W,64) activation a = LeakyReLU(z); y *= (a > 0 ? 1 : slope), i.e. the
  float slope;               // backward of the LeakyReLU that produced this conv's input, fused into its data gradient
  int st256;                 // y and ldy allow 32-byte stores
  const float* scale;        // nullable device scalar multiplying the accumulators (spectral norm 1/sigma)
  const unsigned long long* act_bits;  // alternative to act: one 64-bit word per output pixel, bit c = (a[c] > 0) (first_fprop)
};

namespace {
constexpr int CT_NA = 7;                   // activation ring: [128 pixels][64 channels] bf16 boxes
constexpr int CT_NB = 12;                  // weight ring: [64 cout][64 cin] bf16 slices (1, 2 or 4 consumed per box)
constexpr int CT_A_BYTES = 128 * 128;
constexpr int CT_SLICE_BYTES = 64 * 128;
constexpr int CT_RED_BYTES = 4 * 64 * 8;   // statistics scratch of the 4 epilogue warps
constexpr int CT_SMEM = CT_NA * CT_A_BYTES + CT_NB * CT_SLICE_BYTES + CT_RED_BYTES + 512 + 1024;
}  // namespace

// Shift visiting order.  Raster: s = 0..8.  MERGE: centre, the two vertical neighbours, then the rest - the centre box
// feeds all four parities and a vertical neighbour feeds parities (0,1) or (2,3), whose accumulators are adjacent TMEM
// columns, so their 4 / 2 weight slices (consecutive ring slots: a tile consumes 16*kchunks slots and these groups start at
// multiples of 4 / 2, so with CT_NB = 12 they never wrap) are multiplied by ONE N = 256 / N = 128 MMA per k-step.  Centre
// first makes the accumulate flag uniform over a merged group.
template <bool MERGE>
__device__ __forceinline__ constexpr int ct_shift(int i) {
  return MERGE ? (i == 0 ? 4 : i == 1 ? 1 : i == 2 ? 7 : i == 3 ? 3 : i == 4 ? 5 : i == 5 ? 0 : i == 6 ? 2 : i == 7 ? 6 : 8) : i;
}

// Weight slices that read shift d in {-1,0,+1} along one axis: (parity bit, tap bit) pairs with parity + tap - 1 == d.
__device__ __forceinline__ int axis_pairs(int d, int (&par)[2], int (&tap)[2]) {
  if (d < 0) { par[0] = 0; tap[0] = 0; return 1; }
  if (d > 0) { par[0] = 1; tap[0] = 1; return 1; }
  par[0] = 0; tap[0] = 1; par[1] = 1; tap[1] = 0;
  return 2;
}

// PERSISTENT: one CTA per SM walks pixel tiles blockIdx.x, blockIdx.x + gridDim.x, ...  The TMA producer runs ahead
// across tile boundaries (the rings never drain), and the 512 TMEM columns hold two 256-column accumulator sets, so the
// epilogue of tile i (TMEM -> bf16 NHWC stores + InstanceNorm statistics) overlaps the MMAs of tile i + 1.
// ACT = the fused LeakyReLU-backward variant (never combined with statistics): the 4 x 128-byte activation rows a thread
// needs for its tile are fetched with 256-bit loads BEFORE it waits for the accumulator, so their latency hides behind the
// MMAs instead of being paid once per 32-channel chunk.
// ACT: 0 = none, 1 = activation tensor `act`, 2 = its sign bits `act_bits` (8 bytes instead of 128 per output pixel).
template <int ACT, bool MERGE>
__global__ void __launch_bounds__(352, 1) convt64_kernel(const __grid_constant__ CUtensorMap tmA,
                                                         const __grid_constant__ CUtensorMap tmB, ConvT64Params p,
                                                         int ntiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + CT_NA * CT_A_BYTES;
  float2* red = reinterpret_cast<float2*>(sB + CT_NB * CT_SLICE_BYTES);
  uint64_t* fullA = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(red) + CT_RED_BYTES);
  uint64_t* emptyA = fullA + CT_NA;
  uint64_t* fullB = emptyA + CT_NA;
  uint64_t* emptyB = fullB + CT_NB;
  uint64_t* tmem_full = emptyB + CT_NB;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < CT_NA; ++s) {
      mbar_init(&fullA[s], 1);
      mbar_init(&emptyA[s], 1);
    }
    for (int s = 0; s < CT_NB; ++s) {
      mbar_init(&fullB[s], 1);
      mbar_init(&emptyB[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 8);  // one arrival per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      // Activation-box producer.  25 TMA loads per 64-channel slice of Cin and pixel tile would leave ONE thread ~80
      // tensor-pipe cycles per load; the two rings are independent, so this thread feeds the activation ring (9 boxes)
      // and warp 6 the weight ring (16 slices).  Shifts are unrolled with compile-time offsets, like the MMA loop.
      int sa = 0, pa = 0;
      const int kchunks = p.kchunks;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int tx = tile % p.tiles_x, ty = (tile / p.tiles_x) % p.tiles_y, tn = tile / (p.tiles_x * p.tiles_y);
        const int x0 = tx * p.BW, y0 = ty * p.BH, n0 = tn * p.BI;
#pragma unroll
        for (int si = 0; si < 9; ++si) {
          const int s = ct_shift<MERGE>(si);
          const int dy = s / 3 - 1, dx = s % 3 - 1;
          for (int cc = 0; cc < kchunks; ++cc) {
            mbar_wait(&emptyA[sa], pa ^ 1);
            mbar_expect_tx(&fullA[sa], CT_A_BYTES);
            tma_load_4d(sA + sa * CT_A_BYTES, &tmA, &fullA[sa], cc * 64, x0 + dx, y0 + dy, n0);
            if (++sa == CT_NA) { sa = 0; pa ^= 1; }
          }
        }
      }
    }
  } else if (warp == 6) {
    if (elect_one()) {
      // Weight-slice producer: same walk, the 1 / 2 / 4 (parity, tap) slices of every (shift, channel slice).
      int sb = 0, pb = 0;
      const int kchunks = p.kchunks, Cin = p.Cin;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
#pragma unroll
        for (int si = 0; si < 9; ++si) {
          const int s = ct_shift<MERGE>(si);
          const int dy = s / 3 - 1, dx = s % 3 - 1;
          for (int cc = 0; cc < kchunks; ++cc) {
#pragma unroll
            for (int py = (dy > 0 ? 1 : 0); py <= (dy < 0 ? 0 : 1); ++py)
#pragma unroll
              for (int px = (dx > 0 ? 1 : 0); px <= (dx < 0 ? 0 : 1); ++px) {
                // parity (py,px) reads shift (dy,dx) through tap (a,b) with py + a - 1 == dy, px + b - 1 == dx
                const int parity = py * 2 + px, tap = (dy - py + 1) * 2 + (dx - px + 1);
                mbar_wait(&emptyB[sb], pb ^ 1);
                mbar_expect_tx(&fullB[sb], CT_SLICE_BYTES);
                tma_load_2d(sB + sb * CT_SLICE_BYTES, &tmB, &fullB[sb], tap * Cin + cc * 64, parity * 64);
                if (++sb == CT_NB) { sb = 0; pb ^= 1; }
              }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      // This ONE thread paces the tensor pipe: an N = 64 MMA is 32 cycles of tensor work, so a slice (4 MMAs) must be
      // issued in < 128 cycles.  The 9 shifts are unrolled with compile-time parities / accumulate flags (no local
      // arrays, no divisions) and the descriptors are a constant high word + (base >> 4) + slot * size + 2k.
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);
      constexpr uint32_t DHI = (1024u >> 4) | (1u << 14) | (2u << 29);  // SBO = 1024 B, descriptor version, SWIZZLE_128B
      const uint32_t a_lo0 = (smem_u32(sA) >> 4) | (1u << 16), b_lo0 = (smem_u32(sB) >> 4) | (1u << 16);  // LBO = 16 B
      auto desc = [](uint32_t lo) { return ((uint64_t)DHI << 32) | lo; };
      int sa = 0, pa = 0, sb = 0, pb = 0, lt = 0;
      const int kchunks = p.kchunks;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
        const int buf = lt & 1;
        const uint32_t acc = tmem_base + buf * 256;
        mbar_wait(&tmem_empty[buf], ((lt >> 1) & 1) ^ 1);  // the epilogue has drained this accumulator set
        tc_fence_after();
#pragma unroll
        for (int si = 0; si < 9; ++si) {
          const int s = ct_shift<MERGE>(si);
          const int dy = s / 3 - 1, dx = s % 3 - 1;
          for (int cc = 0; cc < kchunks; ++cc) {
            mbar_wait(&fullA[sa], pa);
            const uint32_t alo = a_lo0 + sa * (CT_A_BYTES >> 4);
            if (MERGE && dx == 0) {
              // one N = 256 (centre) / N = 128 (vertical neighbour) MMA per k-step over g adjacent slices and parities
              constexpr uint32_t idesc128 = umma_idesc_bf16(128, 128, 0, 0), idesc256 = umma_idesc_bf16(128, 256, 0, 0);
              const int g = dy == 0 ? 4 : 2, parity0 = dy > 0 ? 2 : 0;
#pragma unroll
              for (int i = 0; i < g; ++i) mbar_wait(&fullB[sb + i], pb);
              tc_fence_after();
              const uint32_t blo = b_lo0 + sb * (CT_SLICE_BYTES >> 4);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma_bf16(acc + parity0 * 64, desc(alo + 2 * k), desc(blo + 2 * k), g == 4 ? idesc256 : idesc128,
                          (si != 0 || k != 0) ? 1u : (uint32_t)(cc != 0));
#pragma unroll
              for (int i = 0; i < g; ++i) umma_commit(&emptyB[sb + i]);
              sb += g;
              if (sb == CT_NB) { sb = 0; pb ^= 1; }
            } else {
#pragma unroll
              for (int py = (dy > 0 ? 1 : 0); py <= (dy < 0 ? 0 : 1); ++py)
#pragma unroll
                for (int px = (dx > 0 ? 1 : 0); px <= (dx < 0 ? 0 : 1); ++px) {
                  const int parity = py * 2 + px;
                  // raster order: parity 0 is first written by shift 0, parity 1 by shift 1, parity 2 by 3, parity 3 by 4;
                  // MERGE: the centre (first) writes all four
                  const bool later = MERGE ? true : s != (parity == 0 ? 0 : parity == 1 ? 1 : parity == 2 ? 3 : 4);
                  mbar_wait(&fullB[sb], pb);
                  tc_fence_after();
                  const uint32_t blo = b_lo0 + sb * (CT_SLICE_BYTES >> 4);
#pragma unroll
                  for (int k = 0; k < 4; ++k)
                    umma_bf16(acc + parity * 64, desc(alo + 2 * k), desc(blo + 2 * k), idesc,
                              (later || k != 0) ? 1u : (uint32_t)(cc != 0));
                  umma_commit(&emptyB[sb]);
                  if (++sb == CT_NB) { sb = 0; pb ^= 1; }
                }
            }
            umma_commit(&emptyA[sa]);
            if (++sa == CT_NA) { sa = 0; pa ^= 1; }
          }
        }
        umma_commit(&tmem_full[buf]);
      }
    }
  } else {
    // EIGHT epilogue warps (2-5 and 7-10), two per TMEM lane quarter: with four, one warp per scheduler had ~1600 (2100 with the
    // fused LeakyReLU backward) dependent instructions per tile against ~4100 cycles of MMAs - the epilogue, not the tensor pipe,
    // set the pace (38 % tensor-pipe active with the sign-bit variant).  The two warps of a quarter take 32 channels each.
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    const int c0 = warp >= 7 ? 32 : 0;
    const int r = q * 32 + lane;
    const int bw = r % p.BW, bh = (r / p.BW) % p.BH, bi = r / (p.BW * p.BH);
    const int Ho = 2 * p.H, Wo = 2 * p.W;
    const float osc = p.scale ? __ldg(p.scale) : 1.f;
    int lt = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++lt) {
      const int tx = tile % p.tiles_x, ty = (tile / p.tiles_x) % p.tiles_y, tn = tile / (p.tiles_x * p.tiles_y);
      const int x0 = tx * p.BW, y0 = ty * p.BH, n0 = tn * p.BI;
      const int n = n0 + bi;
      const bool valid = n < p.N;
      const int buf = lt & 1;
      const uint32_t acc = tmem_base + buf * 256 + ((uint32_t)(q * 32) << 16);
      uint32_t am[ACT == 1 ? 4 : 1][ACT == 1 ? 16 : 1];  // this warp's 32 channels of the 4 parity pixels
      uint2 ab[ACT == 2 ? 4 : 1];
      if (ACT == 2 && valid) {
#pragma unroll
        for (int parity = 0; parity < 4; ++parity) {
          const int oy = 2 * (y0 + bh) + (parity >> 1), ox = 2 * (x0 + bw) + (parity & 1);
          ab[ACT == 2 ? parity : 0] = __ldg(reinterpret_cast<const uint2*>(p.act_bits + (((long long)n * Ho + oy) * Wo + ox)));
        }
      }
      if (ACT == 1 && valid) {
#pragma unroll
        for (int parity = 0; parity < 4; ++parity) {
          const int oy = 2 * (y0 + bh) + (parity >> 1), ox = 2 * (x0 + bw) + (parity & 1);
          const __nv_bfloat16* a = p.act + (((long long)n * Ho + oy) * Wo + ox) * 64;
#pragma unroll
          for (int j = 0; j < 2; ++j) ld_global_nc_v8(a + c0 + 16 * j, &am[ACT == 1 ? parity : 0][ACT == 1 ? 8 * j : 0]);
        }
      }
      mbar_wait(&tmem_full[buf], (lt >> 1) & 1);
      tc_fence_after();
      {
        float f[32], sq[32];  // per-channel partial statistics of this pixel over the 4 parities
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = sq[j] = 0.f;
#pragma unroll(ACT ? 4 : 1)
        for (int parity = 0; parity < 4; ++parity) {
          const int oy = 2 * (y0 + bh) + (parity >> 1), ox = 2 * (x0 + bw) + (parity & 1);
          __nv_bfloat16* dst = p.y + (((long long)n * Ho + oy) * Wo + ox) * p.ldy;
          uint32_t v0[16], v1[16];
          tmem_ld16(acc + parity * 64 + c0, v0);
          tmem_ld16(acc + parity * 64 + c0 + 16, v1);
          tmem_ld_wait();
          if (parity == 3) {  // this warp's last read of the accumulator set: hand it back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[buf]);
          }
          if (p.scale) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              v0[i] = __float_as_uint(__uint_as_float(v0[i]) * osc);
              v1[i] = __float_as_uint(__uint_as_float(v1[i]) * osc);
            }
          }
          if (ACT == 2 && valid) {  // the same from the sign bits: v0 = channels c0 .. c0+15, v1 = c0+16 .. c0+31
            const uint32_t bits = c0 == 0 ? ab[ACT == 2 ? parity : 0].x : ab[ACT == 2 ? parity : 0].y;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              if (!((bits >> i) & 1u)) v0[i] = __float_as_uint(__uint_as_float(v0[i]) * p.slope);
              if (!((bits >> (16 + i)) & 1u)) v1[i] = __float_as_uint(__uint_as_float(v1[i]) * p.slope);
            }
          }
          if (ACT == 1 && valid) {  // fused LeakyReLU backward: scale by the slope where the forward activation was <= 0
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float2 t = unpack_bf16(am[ACT == 1 ? parity : 0][ACT == 1 ? i : 0]);
              uint32_t* v = i < 8 ? v0 : v1;
              const int i0 = (2 * i) & 15;
              if (!(t.x > 0.f)) v[i0] = __float_as_uint(__uint_as_float(v[i0]) * p.slope);
              if (!(t.y > 0.f)) v[i0 + 1] = __float_as_uint(__uint_as_float(v[i0 + 1]) * p.slope);
            }
          }
          uint32_t o[16];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            o[j] = pack_bf16(__uint_as_float(v0[2 * j]), __uint_as_float(v0[2 * j + 1]));
            o[8 + j] = pack_bf16(__uint_as_float(v1[2 * j]), __uint_as_float(v1[2 * j + 1]));
          }
          if (valid && p.f32out) {
            float4* d4 = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.y) + (dst - p.y) + c0);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              d4[j] = make_float4(__uint_as_float(v0[4 * j]), __uint_as_float(v0[4 * j + 1]), __uint_as_float(v0[4 * j + 2]),
                                  __uint_as_float(v0[4 * j + 3]));
              d4[4 + j] = make_float4(__uint_as_float(v1[4 * j]), __uint_as_float(v1[4 * j + 1]),
                                      __uint_as_float(v1[4 * j + 2]), __uint_as_float(v1[4 * j + 3]));
            }
          } else if (valid) {
            if (p.st256) {  // whole 32-byte sectors per store
              st_global_v8(dst + c0, o);
              st_global_v8(dst + c0 + 16, o + 8);
            } else {
              uint4* d4 = reinterpret_cast<uint4*>(dst + c0);
              d4[0] = make_uint4(o[0], o[1], o[2], o[3]);
              d4[1] = make_uint4(o[4], o[5], o[6], o[7]);
              d4[2] = make_uint4(o[8], o[9], o[10], o[11]);
              d4[3] = make_uint4(o[12], o[13], o[14], o[15]);
            }
          }
          if (!ACT && p.stats) {  // statistics of the bf16-ROUNDED outputs (what the normalisation kernels will read)
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float2 t = unpack_bf16(o[j]);
              f[2 * j] += t.x; f[2 * j + 1] += t.y;
              sq[2 * j] += t.x * t.x; sq[2 * j + 1] += t.y * t.y;
            }
          }
        }
        if (!ACT && p.stats) {  // lane L ends up with channel c0 + L of this warp's 32 pixels x 4 parities
          const float s1 = warp_transpose_sum32(f), s2 = warp_transpose_sum32(sq);
          red[q * 64 + c0 + lane] = make_float2(s1, s2);
        }
      }
      if (!ACT && p.stats) {  // combine the 4 epilogue warps, then one atomic per (channel, statistic) and tile
        named_bar_sync(1, 256);
        if (warp < 7 && r < 64) {
          const float2 a = red[r], b = red[64 + r], c = red[128 + r], d = red[192 + r];
          long long* st = p.stats + ((long long)n0 * 64 + r) * 2;
          stat_add(st, (a.x + b.x) + (c.x + d.x));
          stat_add(st + 1, (a.y + b.y) + (c.y + d.y));
        }
        named_bar_sync(1, 256);  // scratch may be overwritten by the next tile
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// x: (N,H,W,Cin) bf16 with pixel stride ldx; wpack_t: [4 parities][64][4*Cin] bf16 (pack.cu); y: (N,2H,2W,64) bf16
// with pixel stride ldy; stats as in b2c_convt4s2_bf16.  Called by b2c_convt4s2_bf16 (tapgemm.cu) when Cout == 64.
extern "C" int b2c_in_stats(const void* z, long long ldz, long long* zs, int N, int HW, int C, void* stream);  // norm.cu

int b2c_run_convt64(const void* x, long long ldx, const void* wpack_t, const float* scale, void* y, long long ldy,
                    long long* stats, const void* act, const unsigned long long* act_bits, float slope, int N, int H, int W,
                    int Cin, void* stream) {
  extern int g_f32_output;  // tapgemm.cu
  ConvT64Params p;
  p.stats = nullptr;
  p.f32out = g_f32_output;
  p.act = (const __nv_bfloat16*)act; p.act_bits = act_bits; p.slope = slope; p.scale = scale;
  p.st256 = (((uintptr_t)y & 31) == 0 && ldy % 16 == 0) ? 1 : 0;
  p.N = N; p.H = H; p.W = W; p.Cin = Cin; p.kchunks = Cin / 64; p.y = (__nv_bfloat16*)y; p.ldy = ldy;
  const int lw = W & -W, lh = H & -H;  // largest powers of two dividing the extents (tapgemm.cu pick_box)
  p.BW = lw < 128 ? lw : 128;
  int rem = 128 / p.BW;
  p.BH = lh < rem ? lh : rem;
  p.BI = rem / p.BH;
  B2C_REQUIRE(p.BW * p.BH * p.BI == 128 && p.BI <= 256, "convt64: unsupported spatial size %dx%d", H, W);
  p.tiles_x = W / p.BW; p.tiles_y = H / p.BH;
  const int tiles_n = (N + p.BI - 1) / p.BI;
  CUtensorMap tmA, tmB;
  {
    uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t str[3] = {(uint64_t)2 * ldx, (uint64_t)2 * W * ldx, (uint64_t)2 * H * W * ldx};
    uint32_t box[4] = {64, (uint32_t)p.BW, (uint32_t)p.BH, (uint32_t)p.BI};
    if (int rc = b2c_make_tmap_bf16(&tmA, x, 4, dims, str, box)) return rc;
  }
  {
    const uint64_t K = (uint64_t)4 * Cin;
    uint64_t dims[2] = {K, 256};
    uint64_t str[1] = {2 * K};
    uint32_t box[2] = {64, 64};
    if (int rc = b2c_make_tmap_bf16(&tmB, wpack_t, 2, dims, str, box)) return rc;
  }
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(convt64_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CT_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(convt64_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CT_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(convt64_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CT_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(convt64_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, CT_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(convt64_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, CT_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(convt64_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, CT_SMEM);
    if (e != cudaSuccess) {
      b2c_set_error("cudaFuncSetAttribute(convt64): %s", cudaGetErrorString(e));
      return B2C_ERR_CUDA;
    }
    attr_done = true;
  }
  B2C_REQUIRE(!(act && act_bits), "convt64: act and act_bits are alternatives");
  B2C_REQUIRE(!((act || act_bits) && stats), "convt64: the fused LeakyReLU backward and the statistics epilogue are exclusive");
  B2C_REQUIRE(((uintptr_t)act_bits & 7) == 0, "convt64: act_bits must be 8-byte aligned");
  B2C_REQUIRE(!act || ((uintptr_t)act & 31) == 0, "convt64: act must be 32-byte aligned");
  const bool fuse_stats = stats && p.BI == 1;
  if (fuse_stats) {
    cudaError_t e = cudaMemsetAsync(stats, 0, sizeof(long long) * 2 * (size_t)N * 64, (cudaStream_t)stream);
    if (e != cudaSuccess) { b2c_set_error("convt64 stats memset: %s", cudaGetErrorString(e)); return B2C_ERR_CUDA; }
    p.stats = stats;
  }
  const int ntiles = p.tiles_x * p.tiles_y * tiles_n;
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  }
  static const bool merge = getenv("B2C_CONVT_NMERGE") && getenv("B2C_CONVT_NMERGE")[0] == '1';
  const int ctas = ntiles < sms ? ntiles : sms;
  const cudaStream_t st = (cudaStream_t)stream;
  if (merge) {
    if (act_bits) convt64_kernel<2, true><<<ctas, 352, CT_SMEM, st>>>(tmA, tmB, p, ntiles);
    else if (act) convt64_kernel<1, true><<<ctas, 352, CT_SMEM, st>>>(tmA, tmB, p, ntiles);
    else convt64_kernel<0, true><<<ctas, 352, CT_SMEM, st>>>(tmA, tmB, p, ntiles);
  } else {
    if (act_bits) convt64_kernel<2, false><<<ctas, 352, CT_SMEM, st>>>(tmA, tmB, p, ntiles);
    else if (act) convt64_kernel<1, false><<<ctas, 352, CT_SMEM, st>>>(tmA, tmB, p, ntiles);
    else convt64_kernel<0, false><<<ctas, 352, CT_SMEM, st>>>(tmA, tmB, p, ntiles);
  }
  int rc = b2c_check_launch("convt64_kernel");
  if (rc == 0 && stats && !fuse_stats) rc = b2c_in_stats(y, ldy, stats, N, 4 * H * W, 64, stream);
  return rc;
}
