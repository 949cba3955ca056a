"""Golden vectors for GeneratorSimpleRegressor from the UNMODIFIED reference module (cgan/models.py:147-216).
Run in the build container only (needs /root/reference):  python oracle/make_golden_simple.py
TEST INFRASTRUCTURE: writes tests/golden/simple.json; nothing in the product imports this."""
import json
import sys
from pathlib import Path

import torch

sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference/cgan")
import models as R  # noqa: E402

torch.manual_seed(42)
net = R.GeneratorSimpleRegressor(delta_scale=0.3)
net.apply(R.weights_init_normal)
net.eval()
g = torch.Generator().manual_seed(9)
x = torch.rand(3, 3, 64, 64, generator=g) * 2 - 1
with torch.no_grad():
    out = net(x)
sd = net.state_dict()
# first-order training gradients (eval mode: no dropout draws) of loss = sum(delta * r)
r = torch.randn(3, 4, generator=g)
net.zero_grad()
(net(x) * r).sum().backward()
grads = {k: p.grad for k, p in net.named_parameters()}
grad_stats = {k: [float(v.double().norm()), float(v.double().flatten()[:: max(1, v.numel() // 7)][:7].sum())] for k, v in grads.items()}
dst = Path(__file__).resolve().parent.parent / "tests" / "golden" / "simple.json"
dst.write_text(json.dumps({"seed": 42, "x_seed": 9, "shape": [3, 3, 64, 64], "delta": out.tolist(),
                           "params": sum(p.numel() for p in net.parameters()), "r": r.tolist(), "grad_stats": grad_stats,
                           "weight_stats": {k: [float(v.double().sum()), float(v.double().abs().sum())] for k, v in sd.items()}}))
print("wrote", dst, out[0].tolist())
