// Dataset indexing: the many-to-one greedy matching of pseudo-label boxes to ground-truth boxes that the reference's
// CalibratorDataset._prepare_index runs as a Python double loop per image (cgan/dataset.py:128-205, `_bbox_iou` :57-70,
// `_hungarian_matching` :176-205; O(Np * Ng) tensor-scalar operations, about a minute of CPU for a real dataset).  One
// thread per predicted box walks the ground-truth boxes of its image.  The arithmetic is the reference's fp32 sequence with
// every operation rounded on its own (no FMA contraction), so the IoU values - and with them argmax and threshold decisions -
// are bit-identical to the Python loop's.
#include "common.cuh"

__global__ void match_boxes_kernel(const float* __restrict__ pred, const float* __restrict__ gt,
                                   const int* __restrict__ gt_begin, const int* __restrict__ gt_end,
                                   float* __restrict__ best_iou, int* __restrict__ best_idx, int P) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const float4 p = __ldg(reinterpret_cast<const float4*>(pred) + i);
  const float hw = __fdiv_rn(p.z, 2.f), hh = __fdiv_rn(p.w, 2.f);
  const float ax1 = __fsub_rn(p.x, hw), ay1 = __fsub_rn(p.y, hh), ax2 = __fadd_rn(p.x, hw), ay2 = __fadd_rn(p.y, hh);
  const float area_a = __fmul_rn(__fsub_rn(ax2, ax1), __fsub_rn(ay2, ay1));
  float best = 0.f;
  int arg = 0;
  const int j0 = gt_begin[i], j1 = gt_end[i];
  for (int j = j0; j < j1; ++j) {
    const float4 g = __ldg(reinterpret_cast<const float4*>(gt) + j);
    const float gw = __fdiv_rn(g.z, 2.f), gh = __fdiv_rn(g.w, 2.f);
    const float bx1 = __fsub_rn(g.x, gw), by1 = __fsub_rn(g.y, gh), bx2 = __fadd_rn(g.x, gw), by2 = __fadd_rn(g.y, gh);
    // max(0, t) of the Python builtin: 0 unless t > 0
    float dx = __fsub_rn(fminf(ax2, bx2), fmaxf(ax1, bx1)), dy = __fsub_rn(fminf(ay2, by2), fmaxf(ay1, by1));
    dx = dx > 0.f ? dx : 0.f;
    dy = dy > 0.f ? dy : 0.f;
    const float inter = __fmul_rn(dx, dy);
    const float uni = __fsub_rn(__fadd_rn(area_a, __fmul_rn(__fsub_rn(bx2, bx1), __fsub_rn(by2, by1))), inter);
    const float iou = uni > 0.f ? __fdiv_rn(inter, uni) : 0.f;
    if (j == j0 || iou > best) {  // torch.max(dim): the first maximal value
      best = iou;
      arg = j - j0;
    }
  }
  best_iou[i] = best;
  best_idx[i] = j1 > j0 ? arg : -1;
}

// pred (P,4), gt (G,4): normalised (cx, cy, w, h) fp32, 16-byte aligned; gt_begin/gt_end (P): the row range in `gt` of the
// image pred box i belongs to.  best_iou (P) <- max IoU, best_idx (P) <- its index within that range (-1: empty range).
extern "C" int b2c_match_boxes(const float* pred, const float* gt, const int* gt_begin, const int* gt_end, float* best_iou,
                               int* best_idx, int P, void* stream) {
  if (P <= 0) return B2C_OK;
  B2C_REQUIRE((((uintptr_t)pred | (uintptr_t)gt) & 15) == 0, "match_boxes: boxes must be 16-byte aligned");
  match_boxes_kernel<<<(P + 127) / 128, 128, 0, (cudaStream_t)stream>>>(pred, gt, gt_begin, gt_end, best_iou, best_idx, P);
  return b2c_check_launch("match_boxes");
}
