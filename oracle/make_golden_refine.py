"""Golden vectors for the patch-generation path, produced by the UNMODIFIED reference: get_refined_patch_batch
(cgan/cgan_train_enhanced.py:37-137) and CalibratorDataset._letterbox (cgan/dataset.py:103-124) run on image FILES.
Run in the build container only (needs /root/reference):  WANDB_MODE=disabled python oracle/make_golden_refine.py
TEST INFRASTRUCTURE: writes tests/golden/refine.json; nothing in the product imports this."""
import hashlib
import json
import os
import sys
import tempfile
from pathlib import Path

import numpy as np
import torch

os.environ.setdefault("WANDB_MODE", "disabled")
sys.dont_write_bytecode = True
HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, "/root/reference/cgan")
import refine_oracle as RO  # noqa: E402
import cgan_train_enhanced as T  # noqa: E402  (the reference trainer module)
from dataset import CalibratorDataset  # noqa: E402

SIZES = [(97, 150), (240, 320), (333, 211), (64, 64), (480, 640)]
SEED, B = 11, 24
images = RO.synth_images(SEED, SIZES)
g = torch.Generator().manual_seed(SEED)
sample_img = torch.randint(0, len(SIZES), (B,), generator=g)
pred_box = torch.cat([torch.rand(B, 2, generator=g) * 0.7 + 0.15, torch.rand(B, 2, generator=g) * 0.5 + 0.04], 1)
pred_box[3, 2:] = torch.tensor([0.01, 0.01])     # degenerate: "< 10 px" fallback to pred_box, which is itself tiny
pred_box[5] = torch.tensor([0.5, 0.5, 0.9, 0.9])  # clamped to 0.8
delta = torch.randn(B, 4, generator=g) * torch.tensor([0.3, 0.3, 0.6, 0.6])
delta[7] = torch.tensor([2.5, -2.5, 3.0, -3.0])   # saturates the hard clamps
cases = {}
with tempfile.TemporaryDirectory() as td:
    paths = []
    for i, im in enumerate(images):
        p = Path(td) / f"img{i}.png"
        im.save(p)
        paths.append(str(p))
    for S in (64, 128):
        ref = T.get_refined_patch_batch([paths[int(i)] for i in sample_img], pred_box, delta, S, "cpu")
        lb = torch.stack([RO.to_tensor_normalized(CalibratorDataset._letterbox(images[int(i)], pred_box[b], S))
                          for b, i in enumerate(sample_img) if pred_box[b, 2] > 0.02])
        mine = RO.refined_patches(images, sample_img, pred_box, delta, S)
        assert torch.equal(ref, mine), "oracle restatement != reference"
        u8 = lambda t: ((t * 0.5 + 0.5) * 255).round().to(torch.uint8).numpy()
        cases[str(S)] = {"refined_sha256": hashlib.sha256(u8(ref).tobytes()).hexdigest(),
                         "letterbox_sha256": hashlib.sha256(u8(lb).tobytes()).hexdigest(),
                         "refined_sum": [int(v) for v in u8(ref).reshape(B, -1).astype(np.int64).sum(1)],
                         "refined_patch0": u8(ref)[0].reshape(-1).tolist() if S == 64 else None}
dst = HERE.parent / "tests" / "golden" / "refine.json"
dst.write_text(json.dumps({"seed": SEED, "sizes": SIZES, "B": B, "sample_img": sample_img.tolist(),
                           "pred_box": pred_box.tolist(), "delta": delta.tolist(), "cases": cases}))
print("wrote", dst, {k: v["refined_sha256"][:12] for k, v in cases.items()})
