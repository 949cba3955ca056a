"""Drop-in for the reference's cgan/inference.py: calibrate the YOLO pseudo-labels of one image with a trained generator.

    python -m b200cgan.inference --weights runs/G_best.pth --image demo.jpg --pred_txt demo_pred.txt \
                                 --out_txt demo_calib.txt [--img_size 128]

Same CLI flags, file formats, checkpoint handling and box arithmetic as cgan/inference.py:95-178.  What changes is the
device work: every box of the image is cropped/letterboxed on the host exactly as the reference does
(cgan/inference.py:51-66) and the whole stack of patches goes through ONE generator forward on the B200 kernels
(the reference loops box by box with batch 1, cgan/inference.py:161-171).  InstanceNorm statistics are per sample, so
batching does not change any output.  Both generator types of the reference (`unet`, `simple`; cgan/inference.py:131-135)
are supported; there is no CPU path.
"""
from __future__ import annotations

import argparse
import math
from pathlib import Path
from typing import List, Sequence

import torch

from .models import GeneratorSimpleRegressor, GeneratorUNet


# ------------------------------------------------------------------------------------------------ YOLO txt I/O
def load_yolo_txt(txt_path: Path) -> List[List[float]]:
    """Rows of `cls cx cy w h [conf ...]`, normalised to [0, 1]; missing file -> [] (cgan/inference.py:29-41)."""
    txt_path = Path(txt_path)
    if not txt_path.exists():
        return []
    rows = []
    for line in txt_path.read_text().splitlines():
        if line.strip():
            rows.append([float(tok) for tok in line.split()])
    return rows


def save_yolo_txt(txt_path: Path, rows: Sequence[Sequence[float]]) -> None:
    """One `cls cx cy w h [extras]` line per row, class written as an int (cgan/inference.py:43-49)."""
    with open(txt_path, "w") as f:
        for r in rows:
            f.write(" ".join(str(v) for v in [int(r[0]), *r[1:]]) + "\n")


# ------------------------------------------------------------------------------------------------ host-side patch prep
def crop_patch(img, bbox):
    """Crop the (cx, cy, w, h)-normalised box out of a PIL image, clipped to the image (cgan/inference.py:59-66)."""
    W, H = img.size
    cx, cy, w, h = bbox[:4]
    half_w, half_h = w * W / 2, h * H / 2
    return img.crop((max(0, cx * W - half_w), max(0, cy * H - half_h), min(W, cx * W + half_w), min(H, cy * H + half_h)))


def letterbox(img, out_size: int):
    """Grey-pad (128) to a square, then bicubic resize to out_size (cgan/inference.py:51-57)."""
    from PIL import Image, ImageOps
    extra_w, extra_h = max(img.height - img.width, 0), max(img.width - img.height, 0)
    border = (extra_w // 2, extra_h // 2, extra_w - extra_w // 2, extra_h - extra_h // 2)
    return ImageOps.expand(img, border, fill=(128, 128, 128)).resize((out_size, out_size), Image.BICUBIC)


def patch_to_tensor(pil_img) -> torch.Tensor:
    """ToTensor + Normalize(0.5, 0.5): uint8 HWC -> fp32 CHW in [-1, 1] (cgan/inference.py:154-157)."""
    import numpy as np
    arr = torch.from_numpy(np.asarray(pil_img, dtype=np.uint8).copy())
    return (arr.permute(2, 0, 1).to(torch.float32) / 255.0 - 0.5) / 0.5


def apply_delta_to_bbox_inference(bbox: Sequence[float], delta: torch.Tensor) -> List[float]:
    """Inference-time box update: deltas hard-clamped to +-2, raw exp on w/h, centre in [.05,.95], size in [.01,.9]
    (cgan/inference.py:69-89; NOT the training-time smooth clamp of cgan/losses.py:108-150)."""
    cx, cy, w, h = bbox[:4]
    d = [min(2.0, max(-2.0, float(v))) for v in delta.tolist()]
    clip = lambda v, lo, hi: max(lo, min(hi, v))
    return [clip(cx + d[0] * w, 0.05, 0.95), clip(cy + d[1] * h, 0.05, 0.95),
            clip(w * math.exp(d[2]), 0.01, 0.9), clip(h * math.exp(d[3]), 0.01, 0.9)]


# ------------------------------------------------------------------------------------------------ checkpoint handling
def resolve_checkpoint(weights: str, device):
    """-> (state_dict, delta_scale, generator_type) with the reference's fallbacks (cgan/inference.py:104-128,139-147):
    a dict with 'config'/'generator' keys, else a bare state dict with delta_scale parsed from `name=0.3.pth`, else 0.25."""
    ckpt = torch.load(weights, map_location=device)
    delta_scale, gen_type = 0.25, "unet"
    if isinstance(ckpt, dict) and "config" in ckpt:
        delta_scale = ckpt["config"].get("delta_scale", 0.25)
        gen_type = ckpt["config"].get("generator_type", "unet")
    else:
        try:
            delta_scale = float(Path(weights).stem.split("=")[-1])
        except ValueError:
            pass
    state = ckpt["generator"] if isinstance(ckpt, dict) and "generator" in ckpt else ckpt
    return state, delta_scale, gen_type


@torch.no_grad()
def calibrate_patches(netG, patches: torch.Tensor, max_batch: int = 256) -> torch.Tensor:
    """Generator-only eval forward over a stack of (B,3,S,S) patches -> (B,4) deltas on the host
    (BASELINE configs[4]: batch 256 per launch)."""
    netG.eval()
    out = []
    for i in range(0, patches.shape[0], max_batch):
        out.append(netG(patches[i:i + max_batch].cuda(non_blocking=True)).float().cpu())
    return torch.cat(out) if out else torch.zeros(0, 4)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--weights", type=str, required=True)
    ap.add_argument("--image", type=str, required=True)
    ap.add_argument("--pred_txt", type=str, required=True)
    ap.add_argument("--out_txt", type=str, required=True)
    ap.add_argument("--img_size", type=int, default=128)
    args = ap.parse_args(argv)
    if not torch.cuda.is_available():
        raise SystemExit("b200cgan.inference: no CUDA device (this package has no CPU path)")
    from PIL import Image
    state, delta_scale, gen_type = resolve_checkpoint(args.weights, "cpu")
    if gen_type == "simple":  # cgan/inference.py:131-135
        netG = GeneratorSimpleRegressor(delta_scale=delta_scale)
    else:
        netG = GeneratorUNet(delta_scale=delta_scale)
    netG.load_state_dict(state)
    netG = netG.cuda()
    img = Image.open(args.image).convert("RGB")
    preds = load_yolo_txt(Path(args.pred_txt))
    rows = []
    if preds:
        patches = torch.stack([patch_to_tensor(letterbox(crop_patch(img, r[1:5]), args.img_size)) for r in preds])
        deltas = calibrate_patches(netG, patches.pin_memory())
        for r, d in zip(preds, deltas):
            rows.append([int(r[0])] + apply_delta_to_bbox_inference(r[1:5], d) + list(r[5:]))
    save_yolo_txt(Path(args.out_txt), rows)
    print(f"saved {len(rows)} calibrated boxes -> {args.out_txt}")


if __name__ == "__main__":
    main()
