"""Micro-benchmark of the first-layer kernels at the bench shape (B=64, S=256): us/call and achieved GB/s."""
import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200"]
import torch
from b200cgan import _lib as L
N, S = 64, 256
a0, a1 = torch.randn(N, 3, S, S, device="cuda"), torch.randn(N, 3, S, S, device="cuda")
w = torch.randn(64, 6, 4, 4, device="cuda") * 0.1
w3 = torch.randn(64, 3, 4, 4, device="cuda") * 0.1
b = torch.randn(64, device="cuda")
y = torch.empty(N, S // 2, S // 2, 64, dtype=torch.bfloat16, device="cuda")
gz = torch.randn(N, S // 2, S // 2, 64, device="cuda").to(torch.bfloat16)
gx0, gx1 = torch.empty_like(a0), torch.empty_like(a1)
ws = torch.empty(32 * 576, dtype=torch.bfloat16, device="cuda")
dw, db = torch.zeros(64, 6, 4, 4, device="cuda"), torch.zeros(64, device="cuda")
dw3 = torch.zeros(64, 3, 4, 4, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
inb, outb = 2 * a0.numel() * 4, y.numel() * 2
cases = {
    "fprop6": (lambda: L.call("b2c_first_fprop", a0, a1, None, None, None, w, None, b, y, N, S, S, 0.2), inb + outb),
    "fprop3": (lambda: L.call("b2c_first_fprop", a0, None, None, None, None, w3, None, None, y, N, S, S, 0.2), inb // 2 + outb),
    "wgrad6": (lambda: L.call("b2c_first_wgrad", a0, a1, None, None, None, gz, dw, db, N, S, S), inb + outb),
    "wgrad3": (lambda: L.call("b2c_first_wgrad", a0, None, None, None, None, gz, dw3, None, N, S, S), inb // 2 + outb),
    "dgrad6": (lambda: L.call("b2c_first_dgrad", gz, w, None, gx0, gx1, ws, N, S, S), inb + outb),
}
for name, (fn, nbytes) in cases.items():
    for _ in range(3):
        fn()
    ts = []
    for _ in range(10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[len(ts) // 2]
    print(f"{name}: {t*1e3:8.1f} us  {nbytes/t/1e6:8.1f} GB/s (algorithmic {nbytes/1e6:.0f} MB)")
