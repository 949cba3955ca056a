import sys; sys.path[:0] = ["oracle", "gan-calibrated-semi-supervised-learning_b200", "tests"]
import torch, cgan_oracle as O
from test_models_gpu import build, inject_masks, cuda_batch, rel
B, S = int(sys.argv[1]), int(sys.argv[2])
st = O.init_state(42); G, D = build(st)
batch, masks, alphas = O.make_batch(B, S, 0); cb = cuda_batch(batch)
G.train(); inject_masks(G, masks[0])
outs = {}
for n in ("down1", "down2", "down3", "down4", "up1", "up2", "up3"):
    getattr(G, n).register_forward_hook(lambda m, i, o, n=n: outs.__setitem__(n, o.detach().float().cpu()))
with torch.no_grad():
    G(cb["pred"])
t32, temu = {}, {}
with torch.no_grad():
    O.g_forward(st["G"], batch["pred"], 0.3, True, masks[0], taps=t32)
    with O.emulate_bf16():
        O.g_forward(st["G"], batch["pred"], 0.3, True, masks[0], taps=temu)
for n in outs:
    d = (outs[n] - temu[n]).abs()
    print("%-6s vs fp32 %.2e vs emu %.2e emu-vs-fp32 %.2e  frac(|diff|>0 vs emu) %.4f maxdiff %.3e" % (
        n, rel(outs[n], t32[n]), rel(outs[n], temu[n]), rel(temu[n], t32[n]), float((d > 0).float().mean()), float(d.max())))
