"""Per-C-ABI-call device time of one train step (CUDA events around every libb2c call), aggregated by entry point.
Usage: python scripts/kprof.py [B] [S] [steps]"""
import sys, collections
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200", "."]
import torch
from b200cgan import _lib
from b200cgan.models import Discriminator, GeneratorUNet, weights_init_normal
from b200cgan.train_step import CGANTrainer
import bench
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
S = int(sys.argv[2]) if len(sys.argv) > 2 else 256
K = int(sys.argv[3]) if len(sys.argv) > 3 else 3
torch.manual_seed(42)
G, D = GeneratorUNet(0.3), Discriminator(True)
G.apply(weights_init_normal); D.apply(weights_init_normal)
G, D = G.cuda(), D.cuda()
tr = CGANTrainer(G, D)
import os
if os.environ.get('KPROF_GRAPH_WARM'): pass
b = {k: v.cuda() for k, v in bench.synth_host_batches(B, S, 0, 1)[0].items()}
keys = ("pred", "gt", "refined", "pred_box", "delta_true")
for _ in range(3):
    tr.step(*[b[k] for k in keys])
torch.cuda.synchronize()
names = list(_lib.backend().protos)
timed = _lib.time_kernels(names)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(K):
    tr.step(*[b[k] for k in keys])
e1.record(); torch.cuda.synchronize()
tot = e0.elapsed_time(e1) / K
rows = []
for n, recs in timed.items():
    if recs:
        ms = sum(a.elapsed_time(c) for a, c, _ in recs) / K
        rows.append((ms, n, len(recs) / K))
rows.sort(reverse=True)
acc = sum(r[0] for r in rows)
print(f"step {tot:.2f} ms; inside b2c calls {acc:.2f} ms; outside {tot-acc:.2f} ms")
for ms, n, cnt in rows:
    print(f"{n:28s} {cnt:6.1f}/step {ms:8.3f} ms {100*ms/tot:5.1f}%  {1e3*ms/cnt:8.1f} us/call")

print("\n-- bandwidth kernels by shape (algorithmic bytes: each distinct tensor once) --")
agg = collections.defaultdict(lambda: [0, 0.0, 0])
for n, recs in timed.items():
    for a, c, sc in recs:
        nb = bench.bw_bytes(n, sc)
        if nb is None: continue
        shp = tuple(v for v in sc[-4:] if isinstance(v, int) and not isinstance(v, bool))[-3:]
        key = (n, shp, tuple(bool(v) for v in sc[:5] if isinstance(v, bool)))
        agg[key][0] += 1; agg[key][1] += a.elapsed_time(c); agg[key][2] = nb
for (n, shp, pat), (cnt, ms, nb) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n:22s} {str(shp):22s} {''.join('x' if p else '-' for p in pat):6s} x{cnt/K:5.1f}/step {ms/K:7.3f} ms/step  {1e3*ms/cnt:7.1f} us/call  {nb/(ms/cnt)/1e6:7.0f} GB/s")

print("\n-- tensor-core GEMM calls by shape --")
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for n in bench.GEMM_KERNELS:
    for a, c, sc in timed.get(n, []):
        key = (n, sc[-6:-1] if n == "b2c_wgrad4s2_bf16" else sc[-5:])
        agg[key][0] += 1; agg[key][1] += a.elapsed_time(c); agg[key][2] = bench.gemm_flops(n, sc)
for (n, shp), (cnt, ms, fl) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n:20s} {str(shp):32s} x{cnt/K:5.1f}/step {ms/K:7.3f} ms/step {1e3*ms/cnt:7.1f} us/call {fl/(ms/cnt)/1e9:7.0f} TFLOP/s")
