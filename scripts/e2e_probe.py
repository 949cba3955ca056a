"""Where does the end-to-end leg of bench.py lose time against the device-resident leg?  Variants of the same 10 graph-replayed
steps: resident inputs / pipelined scalar read-back / H2D prefetch / both / blocking read-back; plus the bare H2D time."""
import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200", "."]
import torch
import bench
from b200cgan import rng
from b200cgan.models import Discriminator, GeneratorUNet, weights_init_normal
from b200cgan.train_step import CGANTrainer, HostPrefetcher
B, S, K = 64, 256, 10
dev = torch.device("cuda", 0)
torch.manual_seed(42)
G, D = GeneratorUNet(0.3), Discriminator(True)
G.apply(weights_init_normal); D.apply(weights_init_normal)
G, D = G.to(dev), D.to(dev)
rng.manual_seed(42)
host = bench.synth_host_batches(B, S, 0)
resident = [{k: v.to(dev) for k, v in b.items()} for b in host]
keys = ("pred", "gt", "refined", "pred_box", "delta_true")
tr = CGANTrainer(G, D, n_critic=2, lambda_gp=1.0, use_cuda_graph=True)
for i in range(4):
    tr.step(*[resident[i % 2][k] for k in keys])
pf = HostPrefetcher([host[0][k] for k in keys], dev)


def run(h2d, read):
    pending = None
    if h2d:
        pf.submit([host[0][k] for k in keys])
    for i in range(K):
        if h2d:
            if i + 1 < K:
                pf.submit([host[(i + 1) % 2][k] for k in keys])
            args = pf.take()
        else:
            args = [resident[i % 2][k] for k in keys]
        o = tr.step(*args)
        if h2d:
            pf.release()
        if read == "block":
            o.to_host()
        elif read == "pipe":
            h = o.to_host_async()
            if pending is not None:
                pending.get()
            pending = h
    if pending is not None:
        pending.get()


for name, h2d, read in (("resident", 0, None), ("resident+pipe", 0, "pipe"), ("resident+block", 0, "block"), ("h2d", 1, None),
                        ("h2d+pipe", 1, "pipe"), ("h2d+block", 1, "block"), ("resident", 0, None)):
    run(h2d, read); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(h2d, read); e1.record(); torch.cuda.synchronize()
    print(f"{name:16s} {e0.elapsed_time(e1) / K:8.3f} ms/step")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
dst = [torch.empty_like(host[0][k], device=dev) for k in keys]
torch.cuda.synchronize(); e0.record()
for d, k in zip(dst, keys):
    d.copy_(host[0][k], non_blocking=True)
e1.record(); torch.cuda.synchronize()
nb = sum(host[0][k].numel() * 4 for k in keys)
print(f"bare H2D of one batch: {e0.elapsed_time(e1):.3f} ms for {nb / 1e6:.0f} MB = {nb / e0.elapsed_time(e1) / 1e6:.1f} GB/s")
