"""Golden vectors for the host-side helpers of cgan/inference.py, produced by the UNMODIFIED reference functions.
Run in the build container only (needs /root/reference):  python oracle/make_golden_inference.py
TEST INFRASTRUCTURE: writes tests/golden/inference.json; nothing in the product imports this."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference/cgan")
import inference as R  # noqa: E402  (the reference module)
from PIL import Image  # noqa: E402

rng = np.random.default_rng(7)
img = Image.fromarray(rng.integers(0, 256, size=(97, 150, 3), dtype=np.uint8), "RGB")
cases = []
for _ in range(12):
    bbox = [float(rng.uniform(.1, .9)), float(rng.uniform(.1, .9)), float(rng.uniform(.05, .6)), float(rng.uniform(.05, .6))]
    delta = (rng.normal(size=4) * 1.5).astype(np.float32)
    out = R.apply_delta_to_bbox_inference(bbox, torch.from_numpy(delta))
    crop = R.crop_patch(img, bbox)
    lb = R.letterbox(crop, 64)
    arr = np.asarray(lb, dtype=np.int64)
    cases.append({"bbox": bbox, "delta": [float(v) for v in delta], "out": [float(v) for v in out],
                  "crop_size": list(crop.size), "lb_sum": int(arr.sum()), "lb_weighted": int((arr * np.arange(arr.size).reshape(arr.shape) % 251).sum())})
dst = Path(__file__).resolve().parent.parent / "tests" / "golden" / "inference.json"
dst.write_text(json.dumps({"img_seed": 7, "img_hw": [97, 150], "cases": cases}, indent=1))
print("wrote", dst)
