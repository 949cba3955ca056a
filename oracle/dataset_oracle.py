"""TEST INFRASTRUCTURE (never imported by the product): CPU restatement of the reference's dataset indexing
(cgan/dataset.py:57-101,128-205: `_bbox_iou`, `_bbox2delta`, `_prepare_index`, `_hungarian_matching`) as vectorised fp32
torch, plus the synthetic YOLO-style tree the dataset tests are run on.  Pinned to the UNMODIFIED reference by
oracle/make_golden_dataset.py -> tests/golden/dataset.json."""
import math
from pathlib import Path

import numpy as np
import torch

import refine_oracle as RO

SIZES = [(97, 150), (240, 320), (333, 211), (64, 64), (480, 640), (200, 200), (120, 360)]


def make_tree(root, seed=7):
    """images/imgK.jpg, labels_gt/imgK.txt (cls cx cy w h), labels_pred/imgK.txt (cls cx cy w h conf).  Covers: several preds
    on one gt, preds below the threshold, exact IoU ties between two identical gt boxes, an empty pred file, a pred file
    without a gt file, a 5-field line in a pred file (skipped by the reference), an image without any match."""
    root = Path(root)
    for d in ("images", "labels_gt", "labels_pred"):
        (root / d).mkdir(parents=True, exist_ok=True)
    images = RO.synth_images(seed, SIZES)
    g = torch.Generator().manual_seed(seed)
    for k, im in enumerate(images):
        im.save(root / "images" / f"img{k}.jpg", quality=92)
        ng = 1 + k % 4
        gt = torch.cat([torch.rand(ng, 2, generator=g) * 0.6 + 0.2, torch.rand(ng, 2, generator=g) * 0.3 + 0.08], 1)
        if k == 2:
            gt = torch.cat([gt, gt[:1]])  # duplicate gt: IoU tie, the first index must win
        preds = []
        for j in range(len(gt)):
            for _ in range(1 + (j + k) % 3):
                jitter = torch.cat([torch.randn(2, generator=g) * 0.02, torch.randn(2, generator=g) * 0.03])
                preds.append(gt[j] + jitter * (1.0 if (len(preds) % 4) else 6.0))  # every fourth one is far off
        preds.append(torch.tensor([0.05, 0.05, 0.04, 0.04]))  # matches nothing
        pred = torch.stack(preds).clamp(0.01, 0.99)
        if k == 5:
            pred = pred[-1:]  # an image whose only pred matches nothing
        (root / "labels_gt" / f"img{k}.txt").write_text("".join("0 %.6f %.6f %.6f %.6f\n" % tuple(b.tolist()) for b in gt))
        lines = ["0 %.6f %.6f %.6f %.6f %.3f\n" % (*b.tolist(), 0.5 + 0.01 * i) for i, b in enumerate(pred)]
        if k == 1:
            lines.insert(1, "0 0.5 0.5 0.2 0.2\n")  # 5 fields: not a pred line
        (root / "labels_pred" / f"img{k}.txt").write_text("".join(lines) if k != 3 else "")
    (root / "labels_pred" / "orphan.txt").write_text("0 0.5 0.5 0.2 0.2 0.9\n")
    return root


def iou_matrix(pred: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """cgan/dataset.py:57-70 for all pairs, the same fp32 operations in the same order."""
    p, g = pred.float()[:, None, :], gt.float()[None, :, :]
    ax1, ay1, ax2, ay2 = p[..., 0] - p[..., 2] / 2, p[..., 1] - p[..., 3] / 2, p[..., 0] + p[..., 2] / 2, p[..., 1] + p[..., 3] / 2
    bx1, by1, bx2, by2 = g[..., 0] - g[..., 2] / 2, g[..., 1] - g[..., 3] / 2, g[..., 0] + g[..., 2] / 2, g[..., 1] + g[..., 3] / 2
    dx = torch.minimum(ax2, bx2) - torch.maximum(ax1, bx1)
    dy = torch.minimum(ay2, by2) - torch.maximum(ay1, by1)
    dx, dy = torch.where(dx > 0, dx, torch.zeros_like(dx)), torch.where(dy > 0, dy, torch.zeros_like(dy))
    inter = dx * dy
    union = ((ax2 - ax1) * (ay2 - ay1) + (bx2 - bx1) * (by2 - by1)) - inter
    return torch.where(union > 0, inter / union, torch.zeros_like(inter))


def bbox2delta(gt, pred):
    nf = max(math.sqrt(float(pred[2]) * float(pred[3])), 0.05)
    eps = 1e-6
    wr = max(0.1, min(10.0, max(float(gt[2]), eps) / max(float(pred[2]), eps)))
    hr = max(0.1, min(10.0, max(float(gt[3]), eps) / max(float(pred[3]), eps)))
    return torch.tensor([(float(gt[0]) - float(pred[0])) / nf, (float(gt[1]) - float(pred[1])) / nf, math.log(wr), math.log(hr)],
                        dtype=torch.float32)


def read_boxes(path: Path, min_fields: int) -> torch.Tensor:
    if path.stat().st_size == 0:
        return torch.empty((0, 4))
    rows = [[float(x) for x in ln.split()[1:5]] for ln in path.read_text().strip().splitlines() if len(ln.split()) >= min_fields]
    return torch.tensor(rows) if rows else torch.empty((0, 4))


def prepare_index(root, iou_thr):
    """-> list of (image name, pred_box, delta, gt_box) in the reference's order."""
    root = Path(root)
    out = []
    for txt_pred in sorted((root / "labels_pred").glob("*.txt")):
        name = txt_pred.stem
        txt_gt, img = root / "labels_gt" / f"{name}.txt", root / "images" / f"{name}.jpg"
        if not txt_gt.exists() or not img.exists():
            continue
        gt, pred = read_boxes(txt_gt, 5), read_boxes(txt_pred, 6)
        if len(gt) == 0 or len(pred) == 0:
            continue
        best_iou, best_idx = iou_matrix(pred, gt).max(dim=1)
        for i in range(len(pred)):
            if best_iou[i] >= iou_thr:
                out.append((name, pred[i], bbox2delta(gt[int(best_idx[i])], pred[i]), gt[int(best_idx[i])]))
    return out


def patch_u8(t: torch.Tensor) -> np.ndarray:
    return ((t.detach().float().cpu() * 0.5 + 0.5) * 255).round().to(torch.uint8).numpy()
