"""The patch-generation oracle (oracle/refine_oracle.py) against golden vectors produced by the UNMODIFIED reference
functions on image files (oracle/make_golden_refine.py: get_refined_patch_batch, cgan/cgan_train_enhanced.py:37-137, and
CalibratorDataset._letterbox, cgan/dataset.py:103-124).  CPU only."""
import hashlib
import json
from pathlib import Path

import numpy as np
import torch

import refine_oracle as RO

GOLD = json.loads((Path(__file__).parent / "golden" / "refine.json").read_text())


def u8(t):
    return ((t * 0.5 + 0.5) * 255).round().to(torch.uint8).numpy()


def test_refine_oracle_matches_reference_golden():
    images = RO.synth_images(GOLD["seed"], [tuple(s) for s in GOLD["sizes"]])
    si, pb, dl = torch.tensor(GOLD["sample_img"]), torch.tensor(GOLD["pred_box"]), torch.tensor(GOLD["delta"])
    for S, c in GOLD["cases"].items():
        S = int(S)
        ref = RO.refined_patches(images, si, pb, dl, S)
        assert hashlib.sha256(u8(ref).tobytes()).hexdigest() == c["refined_sha256"]
        assert [int(v) for v in u8(ref).reshape(len(si), -1).astype(np.int64).sum(1)] == c["refined_sum"]
        lb = torch.stack([RO.to_tensor_normalized(RO.letterbox(images[int(i)], pb[b], S)) for b, i in enumerate(si) if pb[b, 2] > 0.02])
        assert hashlib.sha256(u8(lb).tobytes()).hexdigest() == c["letterbox_sha256"]
        if c["refined_patch0"] is not None:
            assert u8(ref)[0].reshape(-1).tolist() == c["refined_patch0"]
