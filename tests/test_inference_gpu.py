"""The generator-only paths around the hot loop on the GPU:
  * b200cgan.fit.validate (cgan/cgan_train_enhanced.py:395-420) against the oracle's eval forward + hard-clamp box update + IoU;
  * the G_best.pth dict written by fit.save_best_checkpoint (:481-489) read back through b200cgan.inference.resolve_checkpoint
    (cgan/inference.py:104-147) into a fresh generator: bit-identical deltas, and the same weights drive the CPU oracle;
  * the b200cgan.inference CLI end to end on a synthetic image: every box batched through one generator call, against the
    reference's per-box arithmetic (host helpers pinned to the reference in tests/test_inference_host.py + the oracle forward);
  * fit(): two epochs in CUDA-graph mode, the device-side learning rate (ReduceLROnPlateau under a captured graph)."""
import json

import numpy as np
import pytest
import torch

import cgan_oracle as O
from test_models_gpu import build, cuda_batch, rel

pytestmark = pytest.mark.gpu


def test_validation_loop_matches_oracle():
    from b200cgan import fit
    st = O.init_state(42)
    G, _ = build(st)
    batches, ref_b, ref_a, n = [], 0.0, 0.0, 0
    for i, B in enumerate((5, 3)):
        b, _, _ = O.make_batch(B, 64, 10 + i)
        batches.append((b["pred"].cuda(), b["delta_true"].cuda(), b["pred_box"].cuda()))
        with torch.no_grad():
            d = O.g_forward(st["G"], b["pred"], 0.3, train=False)
            cal = O.apply_delta_to_bbox(b["pred_box"], d, training=False)
            gtb = O.apply_delta_to_bbox(b["pred_box"], b["delta_true"], training=False)
            ref_b += float(O.iou_metric(b["pred_box"], gtb).sum()); ref_a += float(O.iou_metric(cal, gtb).sum()); n += B
    G.train()
    val = fit.validate(G, batches)
    assert G.training  # mode restored
    assert abs(val["mean_iou_before"] - ref_b / n) < 1e-5
    assert abs(val["mean_iou_after"] - ref_a / n) < 2e-3
    assert abs(val["delta_iou"] - (ref_a - ref_b) / n) < 2e-3


def test_best_checkpoint_round_trip(tmp_path):
    from b200cgan import fit, inference
    from b200cgan.models import GeneratorUNet
    from b200cgan.train_step import CGANTrainer
    st = O.init_state(42)
    G, D = build(st)
    tr = CGANTrainer(G, D)
    b, _, _ = O.make_batch(2, 64, 0)
    cb = cuda_batch(b)
    tr.step(cb["pred"], cb["gt"], cb["refined"], cb["pred_box"], cb["delta_true"])  # weights now live in the flat buckets
    path = tmp_path / "G_best.pth"
    fit.save_best_checkpoint(path, G, D, 3, 0.0123, {"delta_scale": 0.3, "generator_type": "unet", "img_size": 64})
    ck = torch.load(str(path), map_location="cpu")
    assert set(ck) == {"generator", "discriminator", "epoch", "delta_iou", "config"} and ck["epoch"] == 3
    assert set(ck["generator"]) == set(st["G"]) and set(ck["discriminator"]) == set(st["D"])
    state, delta_scale, gen_type = inference.resolve_checkpoint(str(path), "cpu")
    assert delta_scale == 0.3 and gen_type == "unet"
    G2 = GeneratorUNet(delta_scale=delta_scale)
    G2.load_state_dict(state)
    G2 = G2.cuda().eval()
    G.eval()
    with torch.no_grad():
        d1, d2 = G(cb["pred"]), G2(cb["pred"])
        ref = O.g_forward({k: v.float() for k, v in state.items()}, b["pred"], 0.3, train=False)
    assert torch.equal(d1, d2)  # a forward pass is bit-reproducible: fixed-point statistics, no float atomics
    assert rel(d2, ref) < 2e-2


def test_inference_cli_end_to_end(tmp_path):
    from PIL import Image
    from b200cgan import inference as I
    st = O.init_state(42)
    rng = np.random.default_rng(5)
    img = Image.fromarray(rng.integers(0, 256, size=(120, 176, 3), dtype=np.uint8), "RGB")
    img.save(tmp_path / "demo.png")
    rows = [[0, float(rng.uniform(.2, .8)), float(rng.uniform(.2, .8)), float(rng.uniform(.1, .5)), float(rng.uniform(.1, .5)),
             float(rng.uniform(.3, .9))] for _ in range(7)]
    I.save_yolo_txt(tmp_path / "pred.txt", rows)
    torch.save({"generator": st["G"], "config": {"delta_scale": 0.3, "generator_type": "unet"}}, tmp_path / "G_best.pth")
    I.main(["--weights", str(tmp_path / "G_best.pth"), "--image", str(tmp_path / "demo.png"), "--pred_txt",
            str(tmp_path / "pred.txt"), "--out_txt", str(tmp_path / "out.txt"), "--img_size", "64"])
    got = I.load_yolo_txt(tmp_path / "out.txt")
    assert len(got) == len(rows)
    # the reference's per-box path (cgan/inference.py:161-178) with the oracle as the generator
    img2 = Image.open(tmp_path / "demo.png").convert("RGB")
    for g, r in zip(got, I.load_yolo_txt(tmp_path / "pred.txt")):
        patch = I.patch_to_tensor(I.letterbox(I.crop_patch(img2, r[1:5]), 64)).unsqueeze(0)
        with torch.no_grad():
            d = O.g_forward(st["G"], patch, 0.3, train=False)[0]
        want = [int(r[0])] + I.apply_delta_to_bbox_inference(r[1:5], d) + list(r[5:])
        assert g[0] == want[0] and g[5:] == [float(v) for v in want[5:]]
        assert np.allclose(g[1:5], want[1:5], atol=3e-3), (g, want)


def test_fit_two_epochs_graph_mode_and_device_lr(tmp_path):
    from b200cgan import fit
    from b200cgan.train_step import CGANTrainer
    st = O.init_state(42)
    G, D = build(st)
    tr = CGANTrainer(G, D, use_cuda_graph=True)
    data = [cuda_batch(O.make_batch(2, 64, i)[0]) for i in range(4)]
    keys = ("pred", "gt", "refined", "pred_box", "delta_true")
    train = lambda: [tuple(b[k] for k in keys) for b in data]
    val = lambda: [(b["pred"], b["delta_true"], b["pred_box"]) for b in data[:2]]
    hist = fit.fit(tr, train, val, n_epochs=2, save_dir=tmp_path, config={"delta_scale": 0.3, "generator_type": "unet"},
                   log=lambda *_: None)
    assert len(hist) == 2 and tr._graph is not None
    assert all(np.isfinite(h["loss_G"]) and np.isfinite(h["loss_D"]) and 0.0 <= h["mean_iou_after"] <= 1.0 for h in hist)
    assert (tmp_path / "G_best.pth").exists() and json.loads((tmp_path / "training_history.json").read_text())[1]["epoch"] == 2
    # a scheduler lowering the rate must reach the CAPTURED graph: with lr = 0 a replayed step leaves the weights unchanged
    tr.optG.lr = 0.0
    tr.optD.lr = 0.0
    before = [p.detach().clone() for p in list(G.parameters()) + list(D.parameters())]
    tr.step(*train()[0])
    for p, q in zip(list(G.parameters()) + list(D.parameters()), before):
        assert torch.equal(p.detach(), q)
    tr.optG.lr = 2e-4
    tr.step(*train()[0])
    assert not torch.equal(G.fc_delta[1].weight.detach(), before[8]) or not torch.equal(G.down2.model[0].weight.detach(), before[1])
