"""Device-side patch generation: the "refined" (fake) patch of the cGAN loop and the dataset / inference letterbox.

The reference re-crops the fake patch on the CPU with PIL from the generator's current delta in every critic step and in the
generator step (cgan/cgan_train_enhanced.py:37-137,311-315,358-360: device -> host -> PIL crop / pad / BICUBIC resize ->
device, 1.7 ms per patch), and builds pred / gt patches the same way in its DataLoader (cgan/dataset.py:103-124,222-236).
Here the images live in HBM as uint8 and one C-ABI call (b2c_refine_patches: a restatement of Pillow's BICUBIC resample that
is bit-identical to it) produces the (B,3,S,S) fp32 patches.

    bank = DeviceImageBank([uint8 HWC tensors ...])
    pred = letterbox_patches(bank, sample_img, pred_boxes, S)                       # dataset.py _letterbox + transform
    trainer = CGANTrainer(G, D, refine_fn=make_refine_fn(bank, sample_img, pred_boxes, S, fallback=pred))
    trainer.step(pred, gt, None, pred_boxes, delta_true)                            # refined recomputed from G's delta
"""
from __future__ import annotations

from typing import Sequence

import torch

from . import _lib

MODE_DATASET, MODE_REFINED, MODE_INFERENCE = 0, 1, 2


class DeviceImageBank:
    """uint8 HWC RGB images resident in HBM, back to back, with their (offset, H, W) tables."""

    def __init__(self, images: Sequence[torch.Tensor], device="cuda"):
        offs, hs, ws, total = [], [], [], 0
        for im in images:
            assert im.dtype == torch.uint8 and im.dim() == 3 and im.shape[2] == 3, "images must be uint8 (H, W, 3)"
            offs.append(total); hs.append(im.shape[0]); ws.append(im.shape[1])
            total += (im.numel() + 15) // 16 * 16
        self.data = torch.empty(max(total, 16), dtype=torch.uint8, device=device)
        for im, o in zip(images, offs):
            self.data[o:o + im.numel()].copy_(im.reshape(-1), non_blocking=True)
        self.offsets = torch.tensor(offs, dtype=torch.int64, device=device)
        self.heights = torch.tensor(hs, dtype=torch.int32, device=device)
        self.widths = torch.tensor(ws, dtype=torch.int32, device=device)
        self.qmax = max(max(hs), max(ws))
        self.count = len(offs)


def crop_patches(bank: DeviceImageBank, sample_img: torch.Tensor, boxes: torch.Tensor, img_size: int, *, delta=None,
                 mode=MODE_DATASET, fallback=None, want_u8=False, want_boxes=False):
    """(B,3,S,S) fp32 patches in [-1,1] for normalised (cx,cy,w,h) boxes; see include/b2c.h b2c_refine_patches."""
    dev = bank.data.device
    B, S = boxes.shape[0], int(img_size)
    sample_img = sample_img.to(dev, torch.int32).contiguous()
    boxes = boxes.to(dev, torch.float32).contiguous()
    delta = None if delta is None else delta.detach().to(dev, torch.float32).contiguous()
    fallback = None if fallback is None else fallback.detach().to(dev, torch.float32).contiguous()
    nws = _lib.query("b2c_refine_ws_bytes", B, S, bank.qmax)
    ws = torch.empty(nws, dtype=torch.uint8, device=dev)
    out = torch.empty(B, 3, S, S, dtype=torch.float32, device=dev)
    u8 = torch.empty(B, S, S, 3, dtype=torch.uint8, device=dev) if want_u8 else None
    bx = torch.empty(B, 4, dtype=torch.float32, device=dev) if want_boxes else None
    _lib.call("b2c_refine_patches", bank.data, bank.offsets, bank.heights, bank.widths, sample_img, boxes, delta, int(mode),
              fallback, ws, nws, out, u8, bx, B, S, bank.qmax)
    extra = tuple(t for t in (u8, bx) if t is not None)
    return (out, *extra) if extra else out


def letterbox_patches(bank, sample_img, boxes, img_size):
    """CalibratorDataset._letterbox + ToTensor/Normalize for a batch (cgan/dataset.py:103-124,50-53)."""
    return crop_patches(bank, sample_img, boxes, img_size, mode=MODE_DATASET)


def refined_patches(bank, sample_img, pred_boxes, deltas, img_size, fallback=None):
    """get_refined_patch_batch (cgan/cgan_train_enhanced.py:37-137) on the device."""
    return crop_patches(bank, sample_img, pred_boxes, img_size, delta=deltas, mode=MODE_REFINED, fallback=fallback)


def make_refine_fn(bank, sample_img, pred_boxes, img_size, fallback=None):
    """-> refine_fn(delta) for CGANTrainer: the fake patch from the generator's current delta, as the reference computes it
    at :311-315 (critic steps) and :358-360 (generator step).  All arguments are fixed per batch; under CUDA-graph capture the
    tensors must be the captured static buffers (pass views of them)."""
    def refine_fn(delta):
        return refined_patches(bank, sample_img, pred_boxes, delta, img_size, fallback)
    return refine_fn
