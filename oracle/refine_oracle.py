"""CPU oracle for the patch-generation path (crop -> grey-pad -> PIL BICUBIC resize -> normalise).  TEST INFRASTRUCTURE ONLY
(imported by tests/ and bench.py's CPU legs; the product path b200cgan.refine never imports it).

Restates, over in-memory PIL images instead of file paths:
  refined_patches   cgan/cgan_train_enhanced.py:37-137  get_refined_patch_batch (image cache and file I/O left out)
  letterbox         cgan/dataset.py:103-124             CalibratorDataset._letterbox + transform (:50-53)
The arithmetic lives in Pillow (Image.crop / ImageOps.expand / Image.resize(BICUBIC)) and torchvision-equivalent tensor
ops, exactly as in the reference.  Pinned by tests/golden/refine.json, which oracle/make_golden_refine.py produced by calling
the UNMODIFIED reference function on image files."""
from __future__ import annotations

import numpy as np
import torch

import cgan_oracle as O


def to_tensor_normalized(pil_img) -> torch.Tensor:
    """transforms.ToTensor() + Normalize([.5]*3, [.5]*3)."""
    arr = torch.from_numpy(np.asarray(pil_img, dtype=np.uint8).copy())
    return (arr.permute(2, 0, 1).to(torch.float32).div(255) - 0.5) / 0.5


def _pad_square(crop):
    from PIL import ImageOps
    pad_w = max(crop.height - crop.width, 0)
    pad_h = max(crop.width - crop.height, 0)
    padding = (pad_w // 2, pad_h // 2, pad_w - pad_w // 2, pad_h - pad_h // 2)
    return ImageOps.expand(crop, padding, fill=(128, 128, 128))


def letterbox(img, bbox, out_size: int):
    """cgan/dataset.py:103-124."""
    from PIL import Image
    W, H = img.size
    cx, cy, w, h = (float(v) for v in bbox)
    px, py, pw, ph = cx * W, cy * H, w * W, h * H
    x1, y1 = max(0, px - pw / 2), max(0, py - ph / 2)
    x2, y2 = min(W, px + pw / 2), min(H, py + ph / 2)
    crop = img.crop((int(x1), int(y1), int(x2), int(y2)))
    return _pad_square(crop).resize((out_size, out_size), Image.BICUBIC)


def refined_patches(images, sample_img, pred_bboxes, deltas_pred, img_size, fallback_patches=None):
    """cgan/cgan_train_enhanced.py:37-137 with `images[sample_img[i]]` in place of Image.open(path)."""
    from PIL import Image
    out = []
    refined_boxes = O.apply_delta_to_bbox(pred_bboxes.cpu().float(), deltas_pred.cpu().float(), training=False)
    for i in range(len(sample_img)):
        img = images[int(sample_img[i])]
        pred_box, refined_box = pred_bboxes[i], refined_boxes[i]
        try:
            W, H = img.size
            cx, cy, w, h = refined_box
            cx = float(torch.clamp(cx, 0.1, 0.9)); cy = float(torch.clamp(cy, 0.1, 0.9))
            w = float(torch.clamp(w, 0.05, 0.8)); h = float(torch.clamp(h, 0.05, 0.8))
            px, py, pw, ph = cx * W, cy * H, w * W, h * H
            x1, y1 = max(0, px - pw / 2), max(0, py - ph / 2)
            x2, y2 = min(W, px + pw / 2), min(H, py + ph / 2)
            if x2 <= x1 or y2 <= y1 or (x2 - x1) < 10 or (y2 - y1) < 10:
                ocx, ocy, ow, oh = (float(v) for v in pred_box)
                opx, opy, opw, oph = ocx * W, ocy * H, ow * W, oh * H
                ox1, oy1 = max(0, opx - opw / 2), max(0, opy - oph / 2)
                ox2, oy2 = min(W, opx + opw / 2), min(H, opy + oph / 2)
                crop = img.crop((int(ox1), int(oy1), int(ox2), int(oy2)))
            else:
                crop = img.crop((int(x1), int(y1), int(x2), int(y2)))
            if crop.width != crop.height:
                crop = _pad_square(crop)
            if crop.size != (img_size, img_size):
                crop = crop.resize((img_size, img_size), Image.BICUBIC)
            out.append(to_tensor_normalized(crop))
        except Exception:
            out.append(fallback_patches[i].cpu() if fallback_patches is not None else torch.zeros(3, img_size, img_size))
    return torch.stack(out)


def synth_images(seed: int, sizes):
    """Smooth random RGB test images (low-frequency noise + gradients + edges: interpolation-sensitive content)."""
    from PIL import Image
    rng = np.random.default_rng(seed)
    ims = []
    for (h, w) in sizes:
        base = rng.integers(0, 256, size=(max(h // 8, 2), max(w // 8, 2), 3), dtype=np.uint8)
        im = Image.fromarray(base, "RGB").resize((w, h), Image.BILINEAR)
        arr = np.asarray(im, dtype=np.int32).copy()
        arr[:: max(h // 7, 1), :, :] = 255 - arr[:: max(h // 7, 1), :, :]          # sharp horizontal lines
        arr[:, :: max(w // 9, 1), 1] = rng.integers(0, 256)                          # vertical lines in one channel
        arr += rng.integers(-6, 7, size=arr.shape)
        ims.append(Image.fromarray(np.clip(arr, 0, 255).astype(np.uint8), "RGB"))
    return ims
