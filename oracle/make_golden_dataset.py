"""Golden vectors for the dataset path, produced by the UNMODIFIED reference CalibratorDataset (cgan/dataset.py) on the
synthetic tree of oracle/dataset_oracle.make_tree.  Run in the build container only (needs /root/reference):
    python oracle/make_golden_dataset.py
TEST INFRASTRUCTURE: writes tests/golden/dataset.json; nothing in the product imports this."""
import hashlib
import json
import sys
import tempfile
from pathlib import Path

import torch

sys.dont_write_bytecode = True
HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, "/root/reference/cgan")
import dataset_oracle as DO  # noqa: E402
from dataset import CalibratorDataset  # noqa: E402  (the reference)

out = {"seed": 7, "cases": {}}
with tempfile.TemporaryDirectory() as td:
    DO.make_tree(td, 7)
    for thr, S in ((0.5, 64), (0.3, 128)):
        ds = CalibratorDataset(td, img_size=S, iou_thr=thr)
        mine = DO.prepare_index(td, thr)
        assert len(mine) == len(ds.samples)
        for (name, pb, dl, gb), (p, _, rpb, rdl, rgb) in zip(mine, ds.samples):
            assert name == p.stem and torch.equal(pb, rpb) and torch.equal(dl, rdl) and torch.equal(gb, rgb), "oracle != reference"
        items = [ds[i] for i in range(len(ds))]
        out["cases"][f"{thr}_{S}"] = {
            "iou_thr": thr, "img_size": S,
            "samples": [{"img": p.stem, "pred_box": pb.tolist(), "delta": dl.tolist(), "gt_box": gb.tolist()}
                        for p, _, pb, dl, gb in ds.samples],
            "pred_sha256": [hashlib.sha256(DO.patch_u8(it[0]).tobytes()).hexdigest() for it in items],
            "gt_sha256": [hashlib.sha256(DO.patch_u8(it[1]).tobytes()).hexdigest() for it in items],
        }
dst = HERE.parent / "tests" / "golden" / "dataset.json"
dst.write_text(json.dumps(out))
print("wrote", dst, {k: len(v["samples"]) for k, v in out["cases"].items()})
