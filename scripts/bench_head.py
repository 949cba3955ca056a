"""Micro-benchmark of the PatchGAN head kernels at (N, h, C): python scripts/bench_head.py [N] [h] [iters]"""
import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200"]
import torch
from b200cgan import _lib as L
N = int(sys.argv[1]) if len(sys.argv) > 1 else 64
h = int(sys.argv[2]) if len(sys.argv) > 2 else 16
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 10
C, dev = 512, "cuda"
a = torch.randn(N, h, h, C, device=dev).to(torch.bfloat16)
w = torch.randn(1, C, 4, 4, device=dev) * 0.05
out = torch.empty(N, 1, h - 1, h - 1, device=dev)
gy = torch.randn(N, 1, h - 1, h - 1, device=dev)
ga = torch.empty(N, h, h, C, device=dev, dtype=torch.bfloat16)
dw = torch.zeros(1, C, 4, 4, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
nb = a.numel() * 2
cases = {
    "head_fprop": lambda: L.call("b2c_head_fprop", a, w, out, N, h, C),
    "head_dgrad": lambda: L.call("b2c_head_dgrad", gy, w, ga, N, h, C),
    "head_wgrad": lambda: L.call("b2c_head_wgrad", a, gy, dw, N, h, C),
}
for name, fn in cases.items():
    fn(); ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[len(ts) // 2]
    print(f"{name:12s} N={N} h={h}: {t*1e3:8.1f} us  {nb/t/1e6:8.0f} GB/s of {nb/1e6:.1f} MB")
