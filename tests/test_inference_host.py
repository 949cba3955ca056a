"""Host-side helpers of the inference entry point vs golden vectors produced by the reference's own functions
(oracle/make_golden_inference.py ran cgan/inference.py:51-89 unmodified).  CPU only; the generator forward itself is
covered by tests/test_models_gpu.py and tests/test_inference_gpu.py."""
import json
from pathlib import Path

import numpy as np
import torch

GOLD = json.loads((Path(__file__).parent / "golden" / "inference.json").read_text())


def _image():
    from PIL import Image
    rng = np.random.default_rng(GOLD["img_seed"])
    h, w = GOLD["img_hw"]
    return Image.fromarray(rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8), "RGB")


def test_box_update_and_patch_prep_match_reference():
    from b200cgan import inference as I
    img = _image()
    rng = np.random.default_rng(GOLD["img_seed"])
    for c in GOLD["cases"]:
        out = I.apply_delta_to_bbox_inference(c["bbox"], torch.tensor(c["delta"]))
        # the reference multiplies np.float32 scalars by Python floats, so its cx/cy carry float32 rounding under NumPy 2
        # promotion rules (float64 under NumPy 1): agree to float32 precision
        assert np.allclose(out, c["out"], rtol=0, atol=2e-7), (out, c["out"])
        crop = I.crop_patch(img, c["bbox"])
        assert list(crop.size) == c["crop_size"]
        arr = np.asarray(I.letterbox(crop, 64), dtype=np.int64)
        assert int(arr.sum()) == c["lb_sum"]
        assert int((arr * np.arange(arr.size).reshape(arr.shape) % 251).sum()) == c["lb_weighted"]
        t = I.patch_to_tensor(I.letterbox(crop, 64))
        assert t.shape == (3, 64, 64) and float(t.min()) >= -1 and float(t.max()) <= 1
        assert torch.equal(t, (torch.from_numpy(arr.astype(np.float32)).permute(2, 0, 1) / 255.0 - 0.5) / 0.5)


def test_yolo_txt_roundtrip(tmp_path):
    from b200cgan import inference as I
    p = tmp_path / "a.txt"
    rows = [[0, 0.5, 0.25, 0.125, 0.3], [3, 0.1, 0.2, 0.3, 0.4, 0.9]]
    I.save_yolo_txt(p, rows)
    back = I.load_yolo_txt(p)
    assert back == [[float(v) for v in r] for r in rows]
    assert p.read_text().splitlines()[0].split()[0] == "0"
    assert I.load_yolo_txt(tmp_path / "missing.txt") == []


def test_checkpoint_resolution(tmp_path):
    from b200cgan import inference as I
    sd = {"w": torch.ones(2)}
    torch.save({"generator": sd, "config": {"delta_scale": 0.3, "generator_type": "unet"}}, tmp_path / "G_best.pth")
    st, ds, gt = I.resolve_checkpoint(str(tmp_path / "G_best.pth"), "cpu")
    assert ds == 0.3 and gt == "unet" and torch.equal(st["w"], sd["w"])
    torch.save(sd, tmp_path / "G_scale=0.4.pth")
    st, ds, gt = I.resolve_checkpoint(str(tmp_path / "G_scale=0.4.pth"), "cpu")
    assert ds == 0.4 and gt == "unet" and "w" in st
    torch.save(sd, tmp_path / "plain.pth")
    assert I.resolve_checkpoint(str(tmp_path / "plain.pth"), "cpu")[1] == 0.25
