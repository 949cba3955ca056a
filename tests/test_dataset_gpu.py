"""GPU: b200cgan.dataset.CalibratorDataset (index built by b2c_match_boxes, patches by b2c_refine_patches) against golden
vectors of the UNMODIFIED reference dataset: same samples in the same order, boxes and deltas bit for bit, pred / gt patches
bit-identical to PIL's; the matching kernel against the oracle on random boxes with exact ties."""
import hashlib
import json
from pathlib import Path

import pytest
import torch

import dataset_oracle as DO

pytestmark = pytest.mark.gpu
GOLD = json.loads((Path(__file__).parent / "golden" / "dataset.json").read_text())


@pytest.mark.parametrize("key", list(GOLD["cases"]))
def test_dataset_matches_reference(tmp_path, key):
    from b200cgan.dataset import CalibratorDataset
    case = GOLD["cases"][key]
    DO.make_tree(tmp_path, GOLD["seed"])
    ds = CalibratorDataset(tmp_path, img_size=case["img_size"], iou_thr=case["iou_thr"])
    assert len(ds) == len(case["samples"])
    for (p, zero, pb, dl, gb), ref in zip(ds.samples, case["samples"]):
        assert p.stem == ref["img"] and zero == 0
        assert torch.equal(pb, torch.tensor(ref["pred_box"])) and torch.equal(gb, torch.tensor(ref["gt_box"]))
        assert torch.equal(dl, torch.tensor(ref["delta"]))
    sha = lambda t: hashlib.sha256(DO.patch_u8(t).tobytes()).hexdigest()
    for i in range(len(ds)):
        pred_patch, gt_patch, delta, pred_box, path = ds[i]
        assert pred_patch.is_cuda and pred_patch.shape == (3, case["img_size"], case["img_size"])
        assert sha(pred_patch) == case["pred_sha256"][i] and sha(gt_patch) == case["gt_sha256"][i], i
        assert torch.equal(delta, torch.tensor(case["samples"][i]["delta"])) and Path(path).stem == case["samples"][i]["img"]
    # one batch = the same patches
    idx = list(range(len(ds)))[::-1]
    pp, gp, dl, pb, paths, sample_img = ds.batch(idx)
    for k, i in enumerate(idx):
        assert sha(pp[k]) == case["pred_sha256"][i] and sha(gp[k]) == case["gt_sha256"][i]
    assert sample_img.tolist() == [int(ds.sample_image_ids[i]) for i in idx] and len(paths) == len(idx)


def test_match_kernel_vs_oracle_with_ties():
    from b200cgan.dataset import CalibratorDataset, match_boxes
    g = torch.Generator().manual_seed(3)
    n_img, begins, ends, preds, gts = 40, [], [], [], []
    g0 = 0
    for k in range(n_img):
        ng, npred = int(torch.randint(0, 9, (1,), generator=g)), int(torch.randint(1, 30, (1,), generator=g))
        gt = torch.cat([torch.rand(ng, 2, generator=g) * 0.8 + 0.1, torch.rand(ng, 2, generator=g) * 0.4 + 0.02], 1)
        if ng >= 2:
            gt[-1] = gt[0]  # exact tie between the first and the last gt
        pr = torch.cat([torch.rand(npred, 2, generator=g) * 0.8 + 0.1, torch.rand(npred, 2, generator=g) * 0.4 + 0.02], 1)
        if ng:
            pr[0] = gt[0]
        preds.append(pr); gts.append(gt)
        begins += [g0] * npred; ends += [g0 + ng] * npred
        g0 += ng
    iou, idx = match_boxes(torch.cat(preds), torch.cat(gts), torch.tensor(begins, dtype=torch.int32), torch.tensor(ends, dtype=torch.int32))
    iou, idx = iou.cpu(), idx.cpu()
    p0 = 0
    for pr, gt in zip(preds, gts):
        if len(gt):
            ref_iou, ref_idx = DO.iou_matrix(pr, gt).max(dim=1)
            assert torch.equal(iou[p0:p0 + len(pr)], ref_iou) and torch.equal(idx[p0:p0 + len(pr)].long(), ref_idx)
        else:
            assert (idx[p0:p0 + len(pr)] == -1).all() and (iou[p0:p0 + len(pr)] == 0).all()
        p0 += len(pr)
    assert CalibratorDataset._bbox_iou(gts[1][0], gts[1][0]) == 1.0 if len(gts[1]) else True
