"""Stability probe at the benchmarked configuration: python scripts/long_run.py [steps] [B] [S]
Steps a CGANTrainer (CUDA-graph mode) over 8 rotating synthetic batches and prints the logged scalars every 50 steps;
exits 1 on the first non-finite scalar or weight."""
import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200", "."]
import math
import torch
from b200cgan.models import Discriminator, GeneratorUNet, weights_init_normal
from b200cgan.train_step import CGANTrainer
import bench
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
S = int(sys.argv[3]) if len(sys.argv) > 3 else 256
torch.manual_seed(42)
G, D = GeneratorUNet(0.3), Discriminator(True)
G.apply(weights_init_normal); D.apply(weights_init_normal)
G, D = G.cuda(), D.cuda()
tr = CGANTrainer(G, D, use_cuda_graph=True)
batches = [{k: v.cuda() for k, v in b.items()} for b in bench.synth_host_batches(B, S, 0, 8)]
keys = ("pred", "gt", "refined", "pred_box", "delta_true")
for i in range(steps):
    out = tr.step(*[batches[i % 8][k] for k in keys])
    if i % 50 == 0 or i == steps - 1:
        h = out.to_host()
        print(i, {k: round(v, 4) for k, v in h.items()}, flush=True)
        if not all(math.isfinite(v) for v in h.values()):
            sys.exit(1)
for p in list(G.parameters()) + list(D.parameters()):
    if not bool(torch.isfinite(p).all()):
        print("non-finite weight"); sys.exit(1)
print("ok")
