"""One eager train step between cudaProfilerStart/Stop (for `ncu --profile-from-start off`): python scripts/one_step.py [B] [S]"""
import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200", "."]
import torch
from b200cgan.models import Discriminator, GeneratorUNet, weights_init_normal
from b200cgan.train_step import CGANTrainer
import bench
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
S = int(sys.argv[2]) if len(sys.argv) > 2 else 256
torch.manual_seed(42)
G, D = GeneratorUNet(0.3), Discriminator(True)
G.apply(weights_init_normal); D.apply(weights_init_normal)
G, D = G.cuda(), D.cuda()
tr = CGANTrainer(G, D)
b = {k: v.cuda() for k, v in bench.synth_host_batches(B, S, 0, 1)[0].items()}
keys = ("pred", "gt", "refined", "pred_box", "delta_true")
for _ in range(2):
    tr.step(*[b[k] for k in keys])
torch.cuda.synchronize()
torch.cuda.profiler.start()
tr.step(*[b[k] for k in keys])
torch.cuda.synchronize()
torch.cuda.profiler.stop()
