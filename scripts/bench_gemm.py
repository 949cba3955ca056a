"""Micro-benchmark of the tensor-core entry points at the cGAN step's shapes (B=64, S=256), L2 flushed between runs.
Usage: python scripts/bench_gemm.py [conv|convt|wgrad|all] [iters]"""
import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200"]
import torch
from b200cgan import _lib as L
which = sys.argv[1] if len(sys.argv) > 1 else "all"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
B = 64
bf = lambda *s: (torch.randn(*s, device="cuda") * 0.5).to(torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn):
    fn(); ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


CONV = [(128, 128, 64, 128), (64, 64, 128, 256), (32, 32, 256, 512), (256, 256, 64, 128), (128, 128, 64, 256), (64, 64, 128, 512)]
CONVT = [(128, 128, 128, 64), (64, 64, 128, 64), (64, 64, 256, 64), (32, 32, 256, 128), (32, 32, 512, 128), (16, 16, 512, 256)]
WGRAD = [(64, 64, 128, 64), (32, 32, 256, 128), (16, 16, 512, 256), (128, 128, 128, 64), (64, 64, 256, 64), (32, 32, 512, 128),
         (16, 16, 512, 512), (32, 32, 1024, 256), (64, 64, 512, 128), (64, 64, 256, 128)]
if which in ("conv", "all"):
    for H, W, Cin, Cout in CONV:
        x, w, y = bf(B, H, W, Cin), bf(Cout, 16 * Cin), bf(B, H // 2, W // 2, Cout)
        t = timeit(lambda: L.call("b2c_conv4s2_bf16", x, Cin, w, None, None, y, Cout, None, B, H, W, Cin, Cout))
        fl = 2.0 * B * (H // 2) * (W // 2) * Cout * 16 * Cin
        print(f"conv4s2  ({B},{H},{W},{Cin},{Cout}): {t*1e3:8.1f} us {fl/t/1e9:7.0f} TFLOP/s")
if which in ("convt", "all"):
    for H, W, Cin, Cout in CONVT:
        x, w, y = bf(B, H, W, Cin), bf(4, Cout, 4 * Cin), bf(B, 2 * H, 2 * W, Cout)
        t = timeit(lambda: L.call("b2c_convt4s2_bf16", x, Cin, w, None, None, y, Cout, None, None, 1.0, B, H, W, Cin, Cout))
        fl = 2.0 * B * H * W * Cout * 16 * Cin
        print(f"convt4s2 ({B},{H},{W},{Cin},{Cout}): {t*1e3:8.1f} us {fl/t/1e9:7.0f} TFLOP/s")
        if Cout == 64:  # the fused LeakyReLU backward: from the activation tensor / from its sign bits
            act = bf(B, 2 * H, 2 * W, Cout)
            bits = torch.randint(-2**62, 2**62, (B * 4 * H * W,), dtype=torch.int64, device="cuda")
            t = timeit(lambda: L.call("b2c_convt4s2_bf16", x, Cin, w, None, None, y, Cout, None, act, 0.2, B, H, W, Cin, Cout))
            print(f"  + act tensor                   : {t*1e3:8.1f} us {fl/t/1e9:7.0f} TFLOP/s")
            t = timeit(lambda: L.call("b2c_convt4s2_signbits_bf16", x, Cin, w, None, y, Cout, bits, 0.2, B, H, W, Cin, Cout))
            print(f"  + sign bits                    : {t*1e3:8.1f} us {fl/t/1e9:7.0f} TFLOP/s")
if which in ("wgrad", "all"):
    for Hs, Ws, Cs, Cl in WGRAD:
        sm, lg = bf(B, Hs, Ws, Cs), bf(B, 2 * Hs, 2 * Ws, Cl)
        dw = torch.empty(Cs, Cl, 4, 4, device="cuda")
        fl = 2.0 * B * Hs * Ws * Cs * 16 * Cl
        ndet = L.query("b2c_wgrad4s2_ws_floats", B, Hs, Ws, Cs, Cl, 148)
        for tag, nf in (("atomic", 16 * Cs * Cl), ("det", ndet)):
            ws = torch.empty(nf, device="cuda")
            t = timeit(lambda: L.call("b2c_wgrad4s2_bf16", sm, Cs, lg, Cl, ws, nf, dw, 0, None, None, None, None, B, Hs, Ws, Cs, Cl, 148))
            print(f"wgrad4s2 ({B},{Hs},{Ws},{Cs},{Cl}) {tag:6s} splits={nf // (16 * Cs * Cl):3d}: {t*1e3:8.1f} us {fl/t/1e9:7.0f} TFLOP/s")
