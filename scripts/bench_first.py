"""Micro-benchmark of the first-layer kernels (Cin=6, D.model.0) at B x S x S: python scripts/bench_first.py [B] [S] [iters]"""
import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200"]
import torch
from b200cgan import _lib as L
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
S = int(sys.argv[2]) if len(sys.argv) > 2 else 256
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 10
dev = "cuda"
a0 = torch.randn(B, 3, S, S, device=dev); a1 = torch.randn(B, 3, S, S, device=dev)
w = torch.randn(64, 6, 4, 4, device=dev) * 0.1
isg = torch.ones(1, device=dev); bias = torch.randn(64, device=dev)
y = torch.empty(B, S // 2, S // 2, 64, device=dev, dtype=torch.bfloat16)
gz = torch.randn(B, S // 2, S // 2, 64, device=dev).to(torch.bfloat16)
dw = torch.zeros(64, 6, 4, 4, device=dev); db = torch.zeros(64, device=dev)
g0 = torch.empty_like(a0); g1 = torch.empty_like(a1)
ws = torch.empty(32 * 576, dtype=torch.bfloat16, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
nb_in, nb_map = 2 * a0.numel() * 4, y.numel() * 2
b1 = torch.randn(B, 3, S, S, device=dev); alpha = torch.rand(B, device=dev)
cases = {
    "fprop_interp": (lambda: L.call("b2c_first_fprop", a0, a1, a0, b1, alpha, w, isg, bias, y, B, S, S, 0.2), nb_in + nb_map + nb_in // 2),
    "wgrad_interp": (lambda: L.call("b2c_first_wgrad", a0, a1, a0, b1, alpha, gz, dw, db, None, None, None, None, B, S, S), nb_in + nb_map + nb_in // 2),
    "first_fprop": (lambda: L.call("b2c_first_fprop", a0, a1, None, None, None, w, isg, bias, y, B, S, S, 0.2), nb_in + nb_map),
    "first_wgrad": (lambda: L.call("b2c_first_wgrad", a0, a1, None, None, None, gz, dw, db, None, None, None, None, B, S, S), nb_in + nb_map),
    "first_dgrad": (lambda: L.call("b2c_first_dgrad", gz, w, isg, g0, g1, ws, B, S, S), nb_in + nb_map),
}
for name, (fn, nb) in cases.items():
    fn(); ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[len(ts) // 2]
    print(f"{name:12s} B={B} S={S}: {t*1e3:8.1f} us  {nb/t/1e6:8.0f} GB/s of {nb/1e6:.0f} MB")
