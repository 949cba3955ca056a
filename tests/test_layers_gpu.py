"""Teacher-forced per-layer parity of the generator on the GPU.

End-to-end weight gradients of G cannot agree to 2e-2 for ANY bf16-storage implementation at random init: the loss
sees the trunk only through InstanceNorm -> ReLU -> mean, whose gradient is an indicator function, so the 0.3 % of
pre-activations that bf16 rounding moves across the kink change the gradient by sqrt(2*0.003) ~ 8 % in relative L2
(oracle fp32 vs oracle bf16-storage, both on the CPU, differ by 13-20 %: DESIGN.md "Parity").  What CAN be pinned at
the north_star tolerance is every layer on its own: each block of the B200 model is re-run on the inputs and incoming
gradients it saw inside a full forward/backward, and compared with the oracle block on those same tensors."""
import pytest
import torch

import cgan_oracle as O
from test_models_gpu import build, cuda_batch, inject_masks, rel

pytestmark = pytest.mark.gpu
TOL = 2e-2


def _capture(G, x):
    """Full kernel forward/backward; returns per-block (inputs, output, grad_output)."""
    rec = {}
    hooks = []
    for name in ("down1", "down2", "down3", "down4", "up1", "up2", "up3"):
        def fh(mod, inp, out, name=name):
            rec[name] = {"in": [t.detach().clone() for t in inp], "out": out.detach().clone()}
            out.register_hook(lambda g, name=name: rec[name].__setitem__("gout", g.detach().clone()))
        hooks.append(getattr(G, name).register_forward_hook(fh))
    delta = G(x)
    gd = torch.randn(delta.shape, generator=torch.Generator().manual_seed(3)).cuda()
    delta.backward(gd)
    for h in hooks:
        h.remove()
    rec["tail"] = {"in": [rec["up3"]["out"]], "out": delta.detach().clone(), "gout": gd}
    return rec


@pytest.mark.parametrize("B,S", [(2, 64), (2, 128)])
def test_generator_blocks_teacher_forced(B, S):
    st = O.init_state(42)
    G, _ = build(st)
    batch, masks, _ = O.make_batch(B, S, 0)
    G.train()
    inject_masks(G, masks[0])
    rec = _capture(G, cuda_batch(batch)["pred"])
    mask_of = {"down4": masks[0][0], "up1": masks[0][1], "up2": masks[0][2]}
    worst = {}
    for name in ("down1", "down2", "down3", "down4", "up1", "up2", "up3", "tail"):
        r = rec[name]
        G.zero_grad()
        xs = [t.clone().requires_grad_(True) for t in r["in"]]
        cpu = [t.detach().float().cpu().requires_grad_(True) for t in r["in"]]
        if name == "tail":
            conv, fc = G.up4[0], G.fc_delta[1]
            from b200cgan import ops
            z = ops.ConvT4s2.apply(xs[0], conv.weight, ops.WeightCtx(conv.weight))
            y = ops.INActPoolFC.apply(z, fc.weight, fc.bias, 0.0, G.delta_scale)
            params = {"up4.0.weight": conv.weight, "fc_delta.1.weight": fc.weight, "fc_delta.1.bias": fc.bias}
            leaves = {k: st["G"][k].clone().requires_grad_(True) for k in params}
            with O.emulate_bf16():
                ref = O.g_tail(leaves["up4.0.weight"], leaves["fc_delta.1.weight"], leaves["fc_delta.1.bias"], cpu[0], 0.3)
        else:
            y = getattr(G, name)(*xs)
            key = name + ".model.0.weight"
            params = {key: getattr(G, name).model[0].weight}
            leaves = {key: st["G"][key].clone().requires_grad_(True)}
            with O.emulate_bf16():
                if name.startswith("down"):
                    ref = O.g_down_block(leaves[key], cpu[0], name != "down1", mask_of.get(name))
                else:
                    ref = O.g_up_block(leaves[key], cpu[0], cpu[1], mask_of.get(name))
        y.backward(r["gout"])
        ref.backward(r["gout"].float().cpu())
        errs = {"out": rel(y, ref)}
        for i, (a, b) in enumerate(zip(xs, cpu)):
            if a.grad is not None and name != "down1":
                errs[f"gin{i}"] = rel(a.grad, b.grad)
        for k, p in params.items():
            errs[k] = rel(p.grad, leaves[k].grad)
        worst[name] = errs
        for k, e in errs.items():
            assert e < TOL, (name, k, e, errs)
    print("teacher-forced G block errors:", {n: {k: round(v, 5) for k, v in e.items()} for n, e in worst.items()})
