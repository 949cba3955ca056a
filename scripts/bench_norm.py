"""Micro-benchmark of the InstanceNorm kernels at one shape: python scripts/bench_norm.py N HW C [iters]"""
import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200"]
import torch
from b200cgan import _lib as L
N, HW, C = (int(v) for v in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 10
bf = lambda: torch.randn(N, HW, C, device="cuda").to(torch.bfloat16)
z, g, h, o0, o1 = bf(), bf(), bf(), bf(), bf()
zs = torch.empty(N, C, 2, dtype=torch.int64, device="cuda"); ws = torch.empty(N, C, 5, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
T = N * HW * C * 2
cases = {
    "in_stats": (lambda: L.call("b2c_in_stats", z, C, zs, N, HW, C), T),
    "in_act_fwd": (lambda: L.call("b2c_in_act_fwd", z, C, zs, None, 0.2, o0, C, None, 0, N, HW, C), 2 * T),
    "in_act_bwd": (lambda: L.call("b2c_in_act_bwd", g, C, None, None, z, C, zs, None, 0.2, ws, o0, C, N, HW, C), 3 * T),
    "in_act_bwd2": (lambda: L.call("b2c_in_act_bwd2", h, C, g, C, z, C, zs, None, 0.2, ws, o0, C, o1, C, N, HW, C), 5 * T),
}
for name, (fn, nb) in cases.items():
    fn(); ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[len(ts) // 2]
    print(f"{name:12s} ({N},{HW},{C}): {t*1e3:8.1f} us  {nb/t/1e6:8.0f} GB/s")
