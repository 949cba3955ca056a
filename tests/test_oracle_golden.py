"""The CPU oracle (oracle/cgan_oracle.py) against known answers produced by the unmodified
reference modules (tests/golden/*.json, written by oracle/make_golden.py in the build container)."""
import json
from pathlib import Path

import pytest
import torch

import cgan_oracle as O

GOLD = Path(__file__).parent / "golden"


def stats(t):
    t = t.detach().double()
    return [float(t.sum()), float(t.abs().sum()), float((t * t).sum())]


def close(a, b, rtol=2e-4, atol=1e-6):
    a, b = torch.as_tensor(a, dtype=torch.float64), torch.as_tensor(b, dtype=torch.float64)
    assert torch.allclose(a, b, rtol=rtol, atol=atol), (a, b)


def test_param_counts_and_init_match_reference():
    g = json.loads((GOLD / "step_B2_S64.json").read_text())
    st = O.init_state(42)
    assert sum(v.numel() for v in st["G"].values()) == g["n_params"]["G"] == 6294788
    assert sum(st["D"][k].numel() for k in O.d_param_names(st["D"])) == g["n_params"]["D"] == 2767808
    for net in ("G", "D"):
        assert set(st[net]) == set(g["init"][net])
        for k, v in st[net].items():
            close(stats(v), g["init"][net][k], rtol=1e-6)


def test_loss_helpers_match_reference():
    g = json.loads((GOLD / "losses.json").read_text())
    box, delta, dt = (torch.tensor(g[k]) for k in ("box", "delta", "delta_true"))
    for tr in (True, False):
        a = O.apply_delta_to_bbox(box, delta, tr)
        b = O.apply_delta_to_bbox(box, dt, tr)
        close(a, g["apply_%s" % tr], rtol=1e-6)
        close(O.eiou_loss(a, b), g["eiou_%s" % tr], rtol=1e-6)
        close(O.iou_metric(a, b), g["iou_%s" % tr], rtol=1e-5)
    d = delta.clone().requires_grad_(True)
    O.eiou_loss(O.apply_delta_to_bbox(box, d, True), O.apply_delta_to_bbox(box, dt, True)).backward()
    close(d.grad, g["eiou_grad_delta"], rtol=1e-5, atol=1e-8)
    close(O.smooth_clamp(torch.linspace(-3, 3, 13), -1.5, 1.5), g["smooth_clamp"], rtol=1e-6)


@pytest.mark.parametrize("name", ["step_B2_S64", "step_B2_S128"])
def test_train_steps_match_reference(name):
    g = json.loads((GOLD / f"{name}.json").read_text())
    B, S = g["B"], g["S"]
    torch.set_num_threads(8)
    st = O.init_state(42)
    for step, ref in enumerate(g["steps"]):
        batch, masks, alphas = O.make_batch(B, S, step)
        if step == 0:
            with torch.no_grad():
                close(O.g_forward(st["G"], batch["pred"], 0.3, train=False), ref["g_eval_delta"], rtol=1e-4, atol=1e-7)
                d2 = {k: v.clone() for k, v in st["D"].items()}
                o = O.d_forward(d2, batch["pred"], batch["gt"], True)
            assert list(o.shape) == ref["d_out_shape"] == [B, 1, S // 16 - 1, S // 16 - 1]
            close(o.flatten()[:8], ref["d_out_first"], rtol=1e-4, atol=1e-6)
            close(stats(o), ref["d_out_stats"], rtol=1e-4, atol=1e-5)
        out = O.train_step(st, batch, masks, alphas, n_critic=g["n_critic"], lambda_gp=g["lambda_gp"])
        for k in ("d_loss", "gp", "wd", "loss_G", "loss_iou", "loss_wgan", "g_gradnorm", "d_gradnorm_0", "d_gradnorm_1"):
            close(out[k], ref[k], rtol=2e-3, atol=2e-5)
        close(out["delta"], ref["delta"], rtol=1e-3, atol=1e-6)
    for net in ("G", "D"):
        for k, v in st[net].items():
            if k in O.D_BIAS_UNOBSERVABLE:
                continue  # true gradient is 0; Adam amplifies fp noise
            close(stats(v)[1:], g["final"][net][k][1:], rtol=1e-4, atol=1e-5)
