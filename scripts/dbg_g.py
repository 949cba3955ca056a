import sys; sys.path[:0] = ["oracle", "gan-calibrated-semi-supervised-learning_b200", "tests"]
import torch, cgan_oracle as O
from test_models_gpu import build, inject_masks, cuda_batch, rel
B, S = int(sys.argv[1]), int(sys.argv[2])
st = O.init_state(42); G, D = build(st)
batch, masks, alphas = O.make_batch(B, S, 0); cb = cuda_batch(batch)
G.train(); inject_masks(G, masks[0])
delta = G(cb["pred"])
leaves = {k: v.clone().requires_grad_(True) for k, v in st["G"].items()}
ref = O.g_forward(leaves, batch["pred"], 0.3, True, masks[0])
gd = torch.randn(B, 4, generator=torch.Generator().manual_seed(3))
delta.backward(gd.cuda()); ref.backward(gd)
print("delta", rel(delta, ref))
for k, p in G.named_parameters():
    print(k, "%.4f" % rel(p.grad, leaves[k].grad), float(leaves[k].grad.norm()))
