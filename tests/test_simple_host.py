"""GeneratorSimpleRegressor (cgan/models.py:147-216): the drop-in module draws the reference's initial weights under the same
seed and the oracle restatement reproduces the reference's eval forward (tests/golden/simple.json was written by the
UNMODIFIED reference module, oracle/make_golden_simple.py).  CPU only."""
import json
from pathlib import Path

import torch

import cgan_oracle as O

GOLD = json.loads((Path(__file__).parent / "golden" / "simple.json").read_text())


def build_cpu():
    from b200cgan.models import GeneratorSimpleRegressor, weights_init_normal
    torch.manual_seed(GOLD["seed"])
    net = GeneratorSimpleRegressor(delta_scale=0.3)
    net.apply(weights_init_normal)
    return net


def test_simple_regressor_init_and_oracle_match_reference():
    net = build_cpu()
    assert sum(p.numel() for p in net.parameters()) == GOLD["params"] == 4833412
    sd = net.state_dict()
    assert set(sd) == set(GOLD["weight_stats"])
    for k, (s, a) in GOLD["weight_stats"].items():  # same draws in the same order as the reference constructor + init
        assert abs(float(sd[k].double().sum()) - s) <= 1e-9 * max(1.0, abs(s)) and abs(float(sd[k].double().abs().sum()) - a) <= 1e-9 * a, k
    x = torch.rand(*GOLD["shape"], generator=torch.Generator().manual_seed(GOLD["x_seed"])) * 2 - 1
    with torch.no_grad():
        out = O.simple_forward({k: v.float() for k, v in sd.items()}, x, 0.3)
    assert torch.allclose(out, torch.tensor(GOLD["delta"]), rtol=1e-4, atol=1e-6)


def test_simple_regressor_oracle_gradients_match_reference():
    """loss = sum(delta * r) in eval mode: autograd through the oracle restatement reproduces the reference module's parameter
    gradients (norm and a strided partial sum per tensor; conv biases feed an InstanceNorm: rounding noise, compared absolutely)."""
    net = build_cpu()
    sd = {k: v.float().clone().requires_grad_(True) for k, v in net.state_dict().items()}
    x = torch.rand(*GOLD["shape"], generator=torch.Generator().manual_seed(GOLD["x_seed"])) * 2 - 1
    (O.simple_forward(sd, x, 0.3) * torch.tensor(GOLD["r"])).sum().backward()
    for k, (nrm, part) in GOLD["grad_stats"].items():
        g = sd[k].grad.double()
        if k.startswith("features") and k.endswith("bias"):
            assert float(g.norm()) < 1e-6 and nrm < 1e-6, k
            continue
        assert abs(float(g.norm()) - nrm) <= 2e-3 * nrm + 1e-9, (k, float(g.norm()), nrm)
        ps = float(g.flatten()[:: max(1, g.numel() // 7)][:7].sum())
        assert abs(ps - part) <= 2e-3 * nrm + 1e-9, (k, ps, part)
