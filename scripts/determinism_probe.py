"""Which tensors of one critic loss (real + fake + penalty) differ between two identical runs?  python scripts/determinism_probe.py [B] [S]"""
import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200", "oracle", "."]
import torch
import cgan_oracle as O
from b200cgan import _lib, losses, ops
from b200cgan.models import Discriminator
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
S = int(sys.argv[2]) if len(sys.argv) > 2 else 64


def run():
    st = O.init_state(42)
    D = Discriminator(True); D.load_state_dict(st["D"]); D = D.cuda().train()
    batch, _, alphas = O.make_batch(B, S, 0)
    cb = {k: v.cuda() for k, v in batch.items()}
    rec = {}
    names = iter(range(10000))
    orig = _lib.backend().call
    log = []

    def spy(name, *args):
        orig(name, *args)
        outs = [a for a in args if isinstance(a, torch.Tensor)]
        log.append((name, [a.detach().clone() for a in outs]))
    _lib.backend().call = spy
    try:
        real, fake = D(cb["pred"], cb["gt"]), D(cb["pred"], cb["refined"])
        gp = losses.compute_gradient_penalty(D, (cb["pred"], cb["gt"]), (cb["pred"], cb["refined"]), "cuda", alpha=alphas[0].cuda())
        cl = ops.CriticLoss.apply(real, fake, gp, 1.0)
        torch.autograd.backward(cl, torch.tensor([1.0, 0, 0], device="cuda"))
    finally:
        _lib.backend().call = orig
    grads = {k: p.grad.detach().clone() for k, p in D.named_parameters()}
    return log, grads


a, ga = run()
b, gb = run()
assert len(a) == len(b)
first = None
for i, ((na, ta), (nb, tb)) in enumerate(zip(a, b)):
    same = all(x.shape == y.shape and torch.equal(x, y) for x, y in zip(ta, tb))
    if not same:
        bad = [j for j, (x, y) in enumerate(zip(ta, tb)) if not torch.equal(x, y)]
        d = max(float((ta[j].double() - tb[j].double()).abs().max()) for j in bad)
        print(f"call {i:4d} {na:28s} differs in tensor args {bad}  max|diff| {d:.3e}  shapes {[tuple(ta[j].shape) for j in bad]}")
        if first is None:
            first = i
        if i - first > 12:
            break
print("first differing call:", first, "of", len(a))
for k in ga:
    r = float((ga[k].double() - gb[k].double()).norm() / (ga[k].double().norm() + 1e-30))
    print(f"grad {k:24s} run-to-run rel-L2 {r:.2e}")
