import sys; sys.path[:0] = ["oracle", "gan-calibrated-semi-supervised-learning_b200", "tests"]
import torch, cgan_oracle as O
from test_models_gpu import build, inject_masks, cuda_batch, rel
B, S = int(sys.argv[1]), int(sys.argv[2])
st = O.init_state(42); G, D = build(st)
batch, masks, alphas = O.make_batch(B, S, 0); cb = cuda_batch(batch)
G.train(); inject_masks(G, masks[0])
delta = G(cb["pred"])
gd = torch.randn(B, 4, generator=torch.Generator().manual_seed(3))
delta.backward(gd.cuda())
res = {}
for mode in ("fp32", "emu"):
    leaves = {k: v.clone().requires_grad_(True) for k, v in st["G"].items()}
    if mode == "emu":
        with O.emulate_bf16():
            ref = O.g_forward(leaves, batch["pred"], 0.3, True, masks[0])
    else:
        ref = O.g_forward(leaves, batch["pred"], 0.3, True, masks[0])
    ref.backward(gd)
    res[mode] = (rel(delta, ref), {k: leaves[k].grad for k in leaves})
print("delta", res["fp32"][0], res["emu"][0])
for k, p in G.named_parameters():
    print("%-24s vs fp32 %.4f  vs emu %.4f  emu-vs-fp32 %.4f  |g| %.3e" % (k, rel(p.grad, res["fp32"][1][k]), rel(p.grad, res["emu"][1][k]), rel(res["emu"][1][k], res["fp32"][1][k]), float(res["fp32"][1][k].norm())))
