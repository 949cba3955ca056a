"""Drop-in for the reference's cgan/dataset.py `CalibratorDataset`: same constructor, directory layout, `samples` tuples and
`__getitem__` 5-tuple, with the two expensive parts on the GPU.

* `_prepare_index` (cgan/dataset.py:128-205): the reference matches every pseudo-label box to its best ground-truth box with a
  Python double loop of tensor-scalar IoUs per image.  Here the label files of the whole dataset are parsed once and ONE
  kernel launch (`b2c_match_boxes`, the reference's fp32 arithmetic operation for operation) returns best IoU / best index
  for every predicted box; threshold, many-to-one semantics, sample order and `_bbox2delta` (python-float maths, kept as is)
  are the reference's.
* `__getitem__` / `batch` (cgan/dataset.py:207-262): images are decoded once (PIL, as in the reference), kept in HBM as uint8
  (`refine.DeviceImageBank`) and the crop -> grey pad -> BICUBIC resize -> ToTensor -> Normalize chain runs as
  `b2c_refine_patches` (bit-identical to Pillow, tests/test_refine_gpu.py).  `__getitem__` returns the reference's tuple with
  the two patches already on the device (the trainer's `.to(device)` is then a no-op); `batch(indices)` does a whole batch
  in one launch, and `bank` / `sample_image_ids` are what `refine.make_refine_fn` needs for the in-loop refined patches.
"""
from __future__ import annotations

import math
import os
from pathlib import Path
from typing import List, Sequence, Tuple

import torch
import yaml
from torch.utils.data import Dataset

from . import _lib
from .refine import DeviceImageBank, letterbox_patches


def match_boxes(pred: torch.Tensor, gt: torch.Tensor, gt_begin: torch.Tensor, gt_end: torch.Tensor):
    """(best_iou (P,) fp32, best_idx (P,) int32) of every pred box over gt[gt_begin[i]:gt_end[i]] - on the device."""
    dev = "cuda"
    P = pred.shape[0]
    best_iou = torch.zeros(P, dtype=torch.float32, device=dev)
    best_idx = torch.full((P,), -1, dtype=torch.int32, device=dev)
    if P:
        pred_d = pred.to(dev, torch.float32).contiguous()
        gt_d = gt.to(dev, torch.float32).contiguous() if gt.numel() else torch.zeros(1, 4, device=dev)
        _lib.call("b2c_match_boxes", pred_d, gt_d, gt_begin.to(dev, torch.int32).contiguous(),
                  gt_end.to(dev, torch.int32).contiguous(), best_iou, best_idx, P)
    return best_iou, best_idx


class CalibratorDataset(Dataset):
    """(pred_patch, gt_patch, delta_true, pred_bbox, img_path) tuples from a YOLO-style tree under *root*:
    images/*.jpg, labels_gt/*.txt (class cx cy w h), labels_pred/*.txt (class cx cy w h conf); all normalised."""

    def __init__(self, root: str | os.PathLike, img_size: int = None, iou_thr: float = None, transform=None, device="cuda"):
        super().__init__()
        if transform is not None:
            raise ValueError("b200cgan.dataset.CalibratorDataset applies the reference's default transform "
                             "(ToTensor + Normalize(0.5, 0.5)) inside the device kernel; custom transforms are not supported")
        self.root = Path(root)
        self.img_dir = self.root / "images"
        self.gt_dir = self.root / "labels_gt"
        self.pred_dir = self.root / "labels_pred"
        with open(Path(__file__).parent / "config.yaml", "r", encoding="utf-8") as f:
            config = yaml.safe_load(f)
        self.img_size = int(config["img_size"] if img_size is None else img_size)
        self.iou_thr = float(config["iou_threshold"] if iou_thr is None else iou_thr)
        self.device = device
        self.samples: List[Tuple[Path, int, torch.Tensor, torch.Tensor, torch.Tensor]] = []
        self._image_paths: List[Path] = []      # distinct images, in first-use order
        self._sample_image: List[int] = []      # per sample: index into _image_paths
        self._bank = None
        self._prepare_index()

    # --------------------- the reference's static helpers (same names and results)
    @staticmethod
    def _bbox_iou(box1: torch.Tensor, box2: torch.Tensor) -> float:
        """IoU of two normalised (cx, cy, w, h) boxes (cgan/dataset.py:57-70), through the matching kernel."""
        iou, _ = match_boxes(box1.reshape(1, 4), box2.reshape(1, 4), torch.zeros(1, dtype=torch.int32),
                             torch.ones(1, dtype=torch.int32))
        return float(iou[0])

    @staticmethod
    def _bbox2delta(gt: torch.Tensor, pred: torch.Tensor) -> torch.Tensor:
        """(dx, dy, log dw, log dh), centre offsets normalised by max(sqrt(pred area), 0.05), size ratios clamped to
        [0.1, 10] (cgan/dataset.py:72-101); python-float maths as in the reference."""
        norm_factor = max(math.sqrt(float(pred[2]) * float(pred[3])), 0.05)
        dx = (float(gt[0]) - float(pred[0])) / norm_factor
        dy = (float(gt[1]) - float(pred[1])) / norm_factor
        eps = 1e-6
        w_ratio = max(0.1, min(10.0, max(float(gt[2]), eps) / max(float(pred[2]), eps)))
        h_ratio = max(0.1, min(10.0, max(float(gt[3]), eps) / max(float(pred[3]), eps)))
        return torch.tensor([dx, dy, math.log(w_ratio), math.log(h_ratio)], dtype=torch.float32)

    @staticmethod
    def _apply_delta_to_bbox(bbox: torch.Tensor, delta: torch.Tensor) -> torch.Tensor:
        """cgan/dataset.py:208-215."""
        return torch.tensor([bbox[0] + delta[0] * bbox[2], bbox[1] + delta[1] * bbox[3], bbox[2] * torch.exp(delta[2]),
                             bbox[3] * torch.exp(delta[3])], dtype=torch.float32)

    @staticmethod
    def _read_boxes(txt_path: Path, min_fields: int) -> torch.Tensor:
        """cgan/dataset.py:152-174: fields 1..4 of every line with at least `min_fields` fields."""
        if txt_path.stat().st_size == 0:
            return torch.empty((0, 4))
        boxes = []
        for line in txt_path.read_text().strip().splitlines():
            parts = line.strip().split()
            if len(parts) >= min_fields:
                boxes.append([float(x) for x in parts[1:5]])
        return torch.tensor(boxes) if boxes else torch.empty((0, 4))

    def _load_boxes(self, txt_path: Path) -> torch.Tensor:
        return self._read_boxes(txt_path, 5)

    def _load_pred_boxes(self, txt_path: Path) -> torch.Tensor:
        return self._read_boxes(txt_path, 6)

    def _hungarian_matching(self, pred_boxes: torch.Tensor, gt_boxes: torch.Tensor) -> List[Tuple[int, int]]:
        """Many-to-one matching of one image (cgan/dataset.py:176-205): every pred takes its max-IoU gt if IoU >= iou_thr."""
        if len(pred_boxes) == 0 or len(gt_boxes) == 0:
            return []
        Np, Ng = len(pred_boxes), len(gt_boxes)
        best_iou, best_idx = match_boxes(pred_boxes, gt_boxes, torch.zeros(Np, dtype=torch.int32),
                                         torch.full((Np,), Ng, dtype=torch.int32))
        keep = (best_iou.cpu() >= self.iou_thr).tolist()
        idx = best_idx.cpu().tolist()
        return [(i, idx[i]) for i in range(Np) if keep[i]]

    # --------------------- index: all images in one launch
    def _prepare_index(self) -> None:
        entries = []  # (img_path, pred_boxes, gt_boxes)
        for txt_pred in sorted(self.pred_dir.glob("*.txt")):
            name = txt_pred.stem
            txt_gt = self.gt_dir / f"{name}.txt"
            img_path = self.img_dir / f"{name}.jpg"
            if not txt_gt.exists() or not img_path.exists():
                continue
            gt_boxes = self._load_boxes(txt_gt)
            pred_boxes = self._load_pred_boxes(txt_pred)
            if len(gt_boxes) == 0 or len(pred_boxes) == 0:
                continue
            entries.append((img_path, pred_boxes, gt_boxes))
        if not entries:
            return
        pred_all = torch.cat([e[1] for e in entries]).float()
        gt_all = torch.cat([e[2] for e in entries]).float()
        begins, ends, g0 = [], [], 0
        for _, pb, gb in entries:
            begins += [g0] * len(pb)
            ends += [g0 + len(gb)] * len(pb)
            g0 += len(gb)
        best_iou, best_idx = match_boxes(pred_all, gt_all, torch.tensor(begins, dtype=torch.int32),
                                         torch.tensor(ends, dtype=torch.int32))
        keep = (best_iou.cpu() >= self.iou_thr).tolist()
        best_idx = best_idx.cpu().tolist()
        p0 = 0
        for img_path, pred_boxes, gt_boxes in entries:
            img_id = None
            for i in range(len(pred_boxes)):
                if not keep[p0 + i]:
                    continue
                if img_id is None:
                    img_id = len(self._image_paths)
                    self._image_paths.append(img_path)
                pred_box, gt_box = pred_boxes[i], gt_boxes[best_idx[p0 + i]]
                self.samples.append((img_path, 0, pred_box, self._bbox2delta(gt_box, pred_box), gt_box))
                self._sample_image.append(img_id)
            p0 += len(pred_boxes)

    # --------------------- device-side patches
    @property
    def bank(self) -> DeviceImageBank:
        """All images that own at least one sample, decoded once and resident in HBM as uint8."""
        if self._bank is None:
            import numpy as np
            from PIL import Image
            imgs = [torch.from_numpy(np.asarray(Image.open(p).convert("RGB")).copy()) for p in self._image_paths]
            self._bank = DeviceImageBank(imgs, device=self.device)
        return self._bank

    @property
    def sample_image_ids(self) -> torch.Tensor:
        """(len(self),) int32: the bank index of every sample's image (for refine.make_refine_fn)."""
        return torch.tensor(self._sample_image, dtype=torch.int32)

    def __len__(self) -> int:
        return len(self.samples)

    def batch(self, indices: Sequence[int]):
        """(pred_patch (B,3,S,S), gt_patch, delta_true (B,4), pred_box (B,4), img_paths, sample_img) - device tensors, two
        launches for the whole batch."""
        idx = [int(i) for i in indices]
        sample_img = torch.tensor([self._sample_image[i] for i in idx], dtype=torch.int32)
        pred_box = torch.stack([self.samples[i][2] for i in idx]).float()
        gt_box = torch.stack([self.samples[i][4] for i in idx]).float()
        delta = torch.stack([self.samples[i][3] for i in idx]).to(self.device)
        pred_patch = letterbox_patches(self.bank, sample_img, pred_box, self.img_size)
        gt_patch = letterbox_patches(self.bank, sample_img, gt_box, self.img_size)
        return pred_patch, gt_patch, delta, pred_box.to(self.device), [str(self.samples[i][0]) for i in idx], sample_img

    def __getitem__(self, idx: int):
        img_path, _, pred_box, delta_true, gt_box = self.samples[idx]
        sample_img = torch.tensor([self._sample_image[idx]] * 2, dtype=torch.int32)
        patches = letterbox_patches(self.bank, sample_img, torch.stack([pred_box, gt_box]).float(), self.img_size)
        return patches[0], patches[1], delta_true, pred_box, str(img_path)
