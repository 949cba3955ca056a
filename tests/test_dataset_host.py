"""CPU: the dataset-indexing oracle (oracle/dataset_oracle.py) against golden vectors produced by the UNMODIFIED reference
CalibratorDataset (tests/golden/dataset.json, oracle/make_golden_dataset.py): sample order, boxes and deltas bit for bit."""
import json
from pathlib import Path

import torch

import dataset_oracle as DO

GOLD = json.loads((Path(__file__).parent / "golden" / "dataset.json").read_text())


def test_index_oracle_matches_reference_golden(tmp_path):
    DO.make_tree(tmp_path, GOLD["seed"])
    for case in GOLD["cases"].values():
        mine = DO.prepare_index(tmp_path, case["iou_thr"])
        assert len(mine) == len(case["samples"])
        for (name, pb, dl, gb), ref in zip(mine, case["samples"]):
            assert name == ref["img"]
            assert torch.equal(pb, torch.tensor(ref["pred_box"])) and torch.equal(gb, torch.tensor(ref["gt_box"]))
            assert torch.equal(dl, torch.tensor(ref["delta"]))


def test_tree_has_the_edge_cases(tmp_path):
    DO.make_tree(tmp_path, GOLD["seed"])
    case = GOLD["cases"]["0.3_128"]
    imgs = [s["img"] for s in case["samples"]]
    assert "img3" not in imgs and "img5" not in imgs and "orphan" not in imgs  # empty pred file / no match / no gt file
    assert len(set(imgs)) < len(imgs)                                            # several samples per image
    gts = [tuple(s["gt_box"]) for s in case["samples"]]
    assert len(set(gts)) < len(gts)                                              # many-to-one: a gt chosen more than once
