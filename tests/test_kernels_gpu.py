"""Per-kernel parity of libb2c.so (through the C-ABI) against torch CPU fp32 on the same bf16-rounded inputs.
Tolerance: relative L2 <= 2e-2 for bf16 outputs (BASELINE.json north_star), <= 1e-4 for fp32-only kernels."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def nhwc(x):  # NCHW fp32 cpu -> NHWC bf16 cuda
    return x.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()


def nchw(y):  # NHWC cuda -> NCHW fp32 cpu
    return y.float().cpu().permute(0, 3, 1, 2).contiguous()


def q(x):  # bf16 rounding on cpu
    return x.to(torch.bfloat16).float()


@pytest.fixture(scope="module")
def L():
    from b200cgan import _lib
    return _lib


CONV_CASES = [(2, 16, 16, 64, 128), (3, 32, 32, 64, 64), (2, 8, 8, 256, 512), (1, 128, 128, 64, 128), (5, 4, 4, 128, 256),
              (2, 256, 256, 64, 64),
              (4, 128, 128, 64, 128), (17, 64, 64, 64, 256), (65, 32, 32, 128, 128),  # two pixel tiles per CTA (MT = 2)
              (2, 48, 48, 64, 128), (3, 24, 24, 128, 256), (2, 12, 12, 256, 512), (5, 6, 6, 512, 512), (2, 80, 40, 64, 64)]  # not powers of two


@pytest.mark.parametrize("N,H,W,Cin,Cout", CONV_CASES)
def test_conv4s2(L, N, H, W, Cin, Cout):
    g = torch.Generator().manual_seed(1)
    x = q(torch.randn(N, Cin, H, W, generator=g))
    w = torch.randn(Cout, Cin, 4, 4, generator=g) * 0.05
    b = torch.randn(Cout, generator=g)
    wp = torch.empty(Cout, 16 * Cin, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_pack_w_c", w.cuda(), None, wp, Cout, Cin)
    y = torch.empty(N, H // 2, W // 2, Cout, dtype=torch.bfloat16, device="cuda")
    zsi = torch.full((N, Cout, 2), 7, dtype=torch.int64, device="cuda")  # fused InstanceNorm statistics of the bf16 output
    sc = torch.tensor([0.75], device="cuda")  # epilogue scale (spectral-norm 1/sigma): y = acc * scale + bias
    L.call("b2c_conv4s2_bf16", nhwc(x), Cin, wp, sc, b.cuda(), y, Cout, zsi, N, H, W, Cin, Cout)
    zs = zsi.double() * 2.0 ** -30  # 64-bit fixed-point sums
    yf = y.float()
    assert float((zs[:, :, 0] - yf.sum((1, 2))).abs().max()) < 1e-3 * float(yf.abs().sum((1, 2)).max())
    assert rel(zs[:, :, 1], (yf * yf).sum((1, 2))) < 1e-4
    torch.cuda.synchronize()
    ref = F.conv2d(x, q(w) * 0.75, b, 2, 1)
    assert rel(nchw(y), ref) < 1e-2


@pytest.mark.parametrize("N,H,W,Cin,Cout", [(2, 8, 8, 128, 64), (3, 16, 16, 64, 128), (2, 4, 4, 512, 256),
                                            (1, 64, 64, 128, 64), (2, 128, 128, 128, 64), (5, 2, 2, 64, 64),
                                            (3, 32, 32, 256, 64), (130, 1, 1, 64, 64),
                                            (16, 32, 32, 128, 128), (65, 16, 16, 64, 256),  # MT = 2
                                            # sizes that are not powers of two (img_size 96 / 160 / 192: any multiple of 16)
                                            (2, 24, 24, 128, 64), (3, 12, 12, 256, 128), (2, 6, 6, 512, 256), (5, 3, 3, 512, 512),
                                            (2, 40, 40, 128, 64), (1, 48, 24, 64, 64)])
def test_convt4s2(L, N, H, W, Cin, Cout):
    g = torch.Generator().manual_seed(2)
    x = q(torch.randn(N, Cin, H, W, generator=g))
    w = torch.randn(Cin, Cout, 4, 4, generator=g) * 0.05
    wp = torch.empty(4, Cout, 4 * Cin, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_pack_w_t", w.cuda(), None, wp, Cin, Cout)
    y = torch.empty(N, 2 * H, 2 * W, Cout, dtype=torch.bfloat16, device="cuda")
    zsi = torch.full((N, Cout, 2), 7, dtype=torch.int64, device="cuda")
    sc = torch.tensor([1.5], device="cuda")
    L.call("b2c_convt4s2_bf16", nhwc(x), Cin, wp, sc, None, y, Cout, zsi, None, 1.0, N, H, W, Cin, Cout)
    zs = zsi.double() * 2.0 ** -30
    yf = y.float()
    assert rel(zs[:, :, 1], (yf * yf).sum((1, 2))) < 1e-4
    assert float((zs[:, :, 0] - yf.sum((1, 2))).abs().max()) < 1e-3 * float(yf.abs().sum((1, 2)).max())
    torch.cuda.synchronize()
    ref = F.conv_transpose2d(x, q(w) * 1.5, None, 2, 1)
    assert rel(nchw(y), ref) < 1e-2


@pytest.mark.parametrize("mode,N,H,W,Cin,Cout", [("c", 2, 32, 32, 64, 128), ("c", 2, 16, 16, 256, 512), ("c", 17, 64, 64, 64, 256),
                                                 ("t", 2, 16, 16, 256, 128), ("t", 2, 32, 32, 128, 64), ("t", 3, 8, 8, 512, 256)])
def test_fp32_check_mode(L, mode, N, H, W, Cin, Cout):
    """BASELINE north_star fp32 check (<= 1e-4): the same tensor-core kernels with fp32 stores instead of bf16, against an
    fp64 reference on the same bf16 operands - isolates the arithmetic (tcgen05, fp32 TMEM accumulation over K up to
    4096) from the bf16 storage rounding that the 2e-2 criterion covers."""
    g = torch.Generator().manual_seed(9)
    x = q(torch.randn(N, Cin, H, W, generator=g))
    L.call("b2c_set_f32_output", 1)
    try:
        if mode == "c":
            w = torch.randn(Cout, Cin, 4, 4, generator=g) * 0.05
            wp = torch.empty(Cout, 16 * Cin, dtype=torch.bfloat16, device="cuda")
            L.call("b2c_pack_w_c", w.cuda(), None, wp, Cout, Cin)
            y = torch.empty(N, H // 2, W // 2, Cout, dtype=torch.float32, device="cuda")
            L.call("b2c_conv4s2_bf16", nhwc(x), Cin, wp, None, None, y, Cout, None, N, H, W, Cin, Cout)
            ref = F.conv2d(x.double(), q(w).double(), None, 2, 1)
        else:
            w = torch.randn(Cin, Cout, 4, 4, generator=g) * 0.05
            wp = torch.empty(4, Cout, 4 * Cin, dtype=torch.bfloat16, device="cuda")
            L.call("b2c_pack_w_t", w.cuda(), None, wp, Cin, Cout)
            y = torch.empty(N, 2 * H, 2 * W, Cout, dtype=torch.float32, device="cuda")
            L.call("b2c_convt4s2_bf16", nhwc(x), Cin, wp, None, None, y, Cout, None, None, 1.0, N, H, W, Cin, Cout)
            ref = F.conv_transpose2d(x.double(), q(w).double(), None, 2, 1)
        torch.cuda.synchronize()
    finally:
        L.call("b2c_set_f32_output", 0)
    assert rel(y.permute(0, 3, 1, 2), ref) < 1e-4


@pytest.mark.parametrize("kind,N,H,W,Cin,Cout,slope", [("D", 3, 32, 32, 64, 128, 0.2), ("D", 2, 16, 16, 256, 512, 0.2),
                                                       ("Gdown", 2, 32, 32, 128, 256, 0.2), ("Gup", 2, 8, 8, 512, 256, 0.0)])
def test_fp32_check_mode_whole_block(L, kind, N, H, W, Cin, Cout, slope):
    """North-star fp32 check on a whole block - D's SN-conv -> InstanceNorm -> LeakyReLU (cgan/models.py:236-242), UNetDown
    (:57-62), UNetUp's ConvTranspose -> InstanceNorm -> ReLU (:72-77) - on fp32 storage: tensor-core conv with unrounded
    fp32 stores, fp32 InstanceNorm + activation, against fp64 torch on the same bf16-exact input / weights: output <= 1e-4;
    and the block's backward from that fp32 pre-activation: d/dz <= 1e-4, then data / weight gradients of the conv on the
    bf16-rounded d/dz (the tensor cores' operand format) <= 1e-4 against fp64 on the same rounded tensor."""
    g = torch.Generator().manual_seed(12)
    x = q(torch.randn(N, Cin, H, W, generator=g))
    up = kind == "Gup"
    w = q(torch.randn(*((Cin, Cout) if up else (Cout, Cin)), 4, 4, generator=g) * 0.05)
    Ho, Wo = (2 * H, 2 * W) if up else (H // 2, W // 2)
    xd, wd = x.double().requires_grad_(True), w.double().requires_grad_(True)
    zr = F.conv_transpose2d(xd, wd, None, 2, 1) if up else F.conv2d(xd, wd, None, 2, 1)
    ar = F.leaky_relu(F.instance_norm(zr, eps=1e-5), slope)
    L.call("b2c_set_f32_output", 1)
    try:
        z = torch.empty(N, Ho, Wo, Cout, dtype=torch.float32, device="cuda")
        if up:
            wp = torch.empty(4, Cout, 4 * Cin, dtype=torch.bfloat16, device="cuda")
            L.call("b2c_pack_w_t", w.cuda(), None, wp, Cin, Cout)
            L.call("b2c_convt4s2_bf16", nhwc(x), Cin, wp, None, None, z, Cout, None, None, 1.0, N, H, W, Cin, Cout)
        else:
            wp = torch.empty(Cout, 16 * Cin, dtype=torch.bfloat16, device="cuda")
            L.call("b2c_pack_w_c", w.cuda(), None, wp, Cout, Cin)
            L.call("b2c_conv4s2_bf16", nhwc(x), Cin, wp, None, None, z, Cout, None, N, H, W, Cin, Cout)
        a = torch.empty_like(z)
        L.call("b2c_in_act_fwd_f32", z, slope, a, N, Ho * Wo, Cout)
        assert rel(a.permute(0, 3, 1, 2), ar) < 1e-4
        # backward
        gy = q(torch.randn(N, Cout, Ho, Wo, generator=g))
        gzr, = torch.autograd.grad(ar, zr, gy.double(), retain_graph=True)
        gz = torch.empty_like(z)
        L.call("b2c_in_act_bwd_f32", gy.permute(0, 2, 3, 1).contiguous().cuda(), z, slope, gz, N, Ho * Wo, Cout)
        assert rel(gz.permute(0, 3, 1, 2), gzr) < 1e-4
        gzq = q(gz.permute(0, 3, 1, 2).cpu())                       # tensor-core operand format
        dxr, dwr = torch.autograd.grad(zr, (xd, wd), gzq.double())
        dx = torch.empty(N, H, W, Cin, dtype=torch.float32, device="cuda")
        if up:   # dgrad of a transposed conv = conv with the same weights, repacked
            wc = torch.empty(Cin, 16 * Cout, dtype=torch.bfloat16, device="cuda")
            L.call("b2c_pack_w_c", w.cuda(), None, wc, Cin, Cout)
            L.call("b2c_conv4s2_bf16", nhwc(gzq), Cout, wc, None, None, dx, Cin, None, N, Ho, Wo, Cout, Cin)
        else:    # dgrad of a conv = transposed conv with the same weights
            wt = torch.empty(4, Cin, 4 * Cout, dtype=torch.bfloat16, device="cuda")
            L.call("b2c_pack_w_t", w.cuda(), None, wt, Cout, Cin)
            L.call("b2c_convt4s2_bf16", nhwc(gzq), Cout, wt, None, None, dx, Cin, None, None, 1.0, N, Ho, Wo, Cout, Cin)
        assert rel(dx.permute(0, 3, 1, 2), dxr) < 1e-4
    finally:
        L.call("b2c_set_f32_output", 0)
    # weight gradient: fp32 accumulate and fp32 output in every mode
    small, large, Cs, Cl, Hs, Ws = (x, gzq, Cin, Cout, H, W) if up else (gzq, x, Cout, Cin, Ho, Wo)
    nws = 16 * Cs * Cl
    ws = torch.empty(nws, dtype=torch.float32, device="cuda")
    dw = torch.empty(Cs, Cl, 4, 4, device="cuda")
    L.call("b2c_wgrad4s2_bf16", nhwc(small), Cs, nhwc(large), Cl, ws, nws, dw, 0, None, None, None, None, N, Hs, Ws, Cs, Cl, 148)
    assert rel(dw, dwr) < 1e-4


def test_convt64_fused_lrelu_backward(L):
    """convt4s2 with `act`: y = convT(x) * (act > 0 ? 1 : slope), the LeakyReLU backward fused into the dgrad epilogue."""
    g = torch.Generator().manual_seed(5)
    N, H, W, Cin, Cout = 3, 16, 16, 128, 64
    x = q(torch.randn(N, Cin, H, W, generator=g))
    w = torch.randn(Cin, Cout, 4, 4, generator=g) * 0.05
    act = q(torch.randn(N, Cout, 2 * H, 2 * W, generator=g))
    wp = torch.empty(4, Cout, 4 * Cin, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_pack_w_t", w.cuda(), None, wp, Cin, Cout)
    y = torch.empty(N, 2 * H, 2 * W, Cout, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_convt4s2_bf16", nhwc(x), Cin, wp, None, None, y, Cout, None, nhwc(act), 0.2, N, H, W, Cin, Cout)
    torch.cuda.synchronize()
    ref = F.conv_transpose2d(x, q(w), None, 2, 1) * torch.where(act > 0, 1.0, 0.2)
    assert rel(nchw(y), ref) < 1e-2
    # the same from the activation's SIGN BITS (one 64-bit word per output pixel): bit-identical output
    pos = (act.permute(0, 2, 3, 1) > 0).to(torch.int64)  # (N, 2H, 2W, 64)
    sh = torch.arange(64, dtype=torch.int64)
    bits = (pos << sh).sum(-1)  # bit 63 wraps into the sign bit of int64: the same 64-bit pattern
    y2 = torch.empty_like(y)
    L.call("b2c_convt4s2_signbits_bf16", nhwc(x), Cin, wp, None, y2, Cout, bits.contiguous().cuda(), 0.2, N, H, W, Cin, Cout)
    torch.cuda.synchronize()
    assert torch.equal(y2, y)


def test_conv_dgrad_is_convt(L):
    """dgrad of Conv2d(k4,s2,p1) == convt4s2 with the conv weight packed in T-form."""
    g = torch.Generator().manual_seed(3)
    N, H, W, Cin, Cout = 2, 16, 16, 64, 128
    gy = q(torch.randn(N, Cout, H // 2, W // 2, generator=g))
    w = torch.randn(Cout, Cin, 4, 4, generator=g) * 0.05
    wp = torch.empty(4, Cin, 4 * Cout, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_pack_w_t", w.cuda(), None, wp, Cout, Cin)
    dx = torch.empty(N, H, W, Cin, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_convt4s2_bf16", nhwc(gy), Cout, wp, None, None, dx, Cin, None, None, 1.0, N, H // 2, W // 2, Cout, Cin)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_input((N, Cin, H, W), q(w), gy, 2, 1)
    assert rel(nchw(dx), ref) < 1e-2


@pytest.mark.parametrize("N,Hs,Ws,Cs,Cl", [(2, 8, 8, 128, 64), (3, 16, 16, 128, 128), (2, 4, 4, 512, 256),
                                           (1, 64, 64, 128, 64), (5, 2, 2, 256, 128), (4, 32, 32, 256, 128),
                                           (2, 24, 24, 128, 64), (3, 12, 12, 256, 128), (2, 6, 6, 512, 256), (5, 3, 3, 512, 512),
                                           (2, 40, 20, 128, 64)])  # not powers of two
def test_wgrad4s2(L, N, Hs, Ws, Cs, Cl):
    g = torch.Generator().manual_seed(4)
    small = q(torch.randn(N, Cs, Hs, Ws, generator=g))
    large = q(torch.randn(N, Cl, 2 * Hs, 2 * Ws, generator=g))
    ref = torch.nn.grad.conv2d_weight(large.double(), (Cs, Cl, 4, 4), small.double(), 2, 1)
    nws = L.query("b2c_wgrad4s2_ws_floats", N, Hs, Ws, Cs, Cl, 148)
    assert nws >= 16 * Cs * Cl and nws % (16 * Cs * Cl) == 0
    dw = torch.empty(Cs, Cl, 4, 4, dtype=torch.float32, device="cuda")
    runs = []
    for nf in (nws, 16 * Cs * Cl):  # deterministic split-K slabs, then the atomic fallback with a one-slab workspace
        ws = torch.empty(nf, dtype=torch.float32, device="cuda")
        L.call("b2c_wgrad4s2_bf16", nhwc(small), Cs, nhwc(large), Cl, ws, nf, dw, 0, None, None, None, None, N, Hs, Ws, Cs,
               Cl, 148)
        torch.cuda.synchronize()
        assert rel(dw, ref) < 1e-4  # fp32 output: the fp32 check criterion applies directly
        runs.append(dw.clone())
    ws = torch.empty(nws, dtype=torch.float32, device="cuda")
    L.call("b2c_wgrad4s2_bf16", nhwc(small), Cs, nhwc(large), Cl, ws, nws, dw, 0, None, None, None, None, N, Hs, Ws, Cs, Cl, 148)
    assert torch.equal(dw, runs[0])  # bit-identical from run to run
    # accumulate + scale + the <dW, W> partials of the spectral-norm backward
    wdot = torch.randn(Cs, Cl, 4, 4, generator=g).cuda()
    part, nparts = torch.full((1024,), 9.0, device="cuda"), ctypes.c_int(0)
    sc = torch.tensor([0.5], device="cuda")
    L.call("b2c_wgrad4s2_bf16", nhwc(small), Cs, nhwc(large), Cl, ws, nws, dw, 1, sc, wdot, part, ctypes.byref(nparts), N, Hs,
           Ws, Cs, Cl, 148)
    torch.cuda.synchronize()
    assert rel(dw, 1.5 * ref) < 1e-4
    dot = float(part[:nparts.value].double().sum())
    assert abs(dot - float((ref * wdot.cpu().double()).sum())) <= 1e-4 * float(ref.norm() * wdot.norm())


@pytest.mark.parametrize("cin,N,S,interp", [(3, 3, 32, True), (6, 3, 32, True), (6, 2, 128, False), (3, 5, 64, False),
                                             (6, 1, 256, True), (3, 3, 128, True), (6, 5, 256, "same0"), (3, 5, 256, False),
                                             (6, 3, 512, False)])
def test_first_layer(L, cin, N, S, interp):
    """First-layer kernels (tcgen05, bf16 operands, fp32 accumulate) against torch on bf16-rounded x and w.  S >= 128 takes
    the TMA-staged kernels (5 x 256^2 = 640 tiles: several per CTA); "same0": the first three channels of both
    interpolation sources are the same tensor (the critic's `pred` under the gradient penalty) and are staged once."""
    g = torch.Generator().manual_seed(5)
    x = torch.randn(N, cin, S, S, generator=g)
    xb = torch.randn(N, cin, S, S, generator=g)
    if interp == "same0":
        xb[:, :3] = x[:, :3]
    alpha = torch.rand(N, generator=g)
    w = torch.randn(64, cin, 4, 4, generator=g) * 0.1
    b = torch.randn(64, generator=g) * 0.1
    sig = torch.tensor([2.0, 0.5])
    parts = lambda t: [t[:, :3].contiguous().cuda(), t[:, 3:].contiguous().cuda() if cin == 6 else None]
    a0, a1 = parts(x)
    b0, b1 = parts(xb) if interp else (None, None)
    if interp == "same0":
        b0 = a0
    al_d = alpha.cuda() if interp else None
    y = torch.empty(N, S // 2, S // 2, 64, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_first_fprop", a0, a1, b0, b1, al_d, w.cuda(), sig.cuda()[1:], b.cuda(), y, N, S, S, 0.2)
    al = alpha.view(N, 1, 1, 1)
    xi = q(al * x + (1 - al) * xb) if interp else q(x)
    ref = F.leaky_relu(F.conv2d(xi, q(w * 0.5), b, 2, 1), 0.2)
    assert rel(nchw(y), ref) < 5e-3
    # the variant that also leaves the sign bits of its output (input of the fused LeakyReLU backward one layer up)
    y2 = torch.empty_like(y)
    bits = torch.zeros(N * (S // 2) * (S // 2), dtype=torch.int64, device="cuda")
    L.call("b2c_first_fprop_bits", a0, a1, b0, b1, al_d, w.cuda(), sig.cuda()[1:], b.cuda(), y2, N, S, S, 0.2, bits)
    assert torch.equal(y2, y)
    want = ((y.float().reshape(-1, 64) > 0).to(torch.int64) << torch.arange(64, dtype=torch.int64, device="cuda")).sum(-1)
    assert torch.equal(bits, want)
    gz = q(torch.randn(N, 64, S // 2, S // 2, generator=g))
    gx0 = torch.empty(N, 3, S, S, device="cuda")
    gx1 = torch.empty(N, 3, S, S, device="cuda") if cin == 6 else None
    ws = torch.empty(32 * 576, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_first_dgrad", nhwc(gz), w.cuda(), sig.cuda()[1:], gx0, gx1, ws, N, S, S)
    refdx = torch.nn.grad.conv2d_input((N, cin, S, S), q(w * 0.5), gz, 2, 1)
    got = torch.cat([gx0.cpu()] + ([gx1.cpu()] if cin == 6 else []), 1)
    assert rel(got, refdx) < 1e-4
    dw = torch.zeros(64, cin, 4, 4, device="cuda")
    db = torch.zeros(64, device="cuda")
    L.call("b2c_first_wgrad", a0, a1, b0, b1, al_d, nhwc(gz), dw, db, None, None, None, None, N, S, S)
    refdw = torch.nn.grad.conv2d_weight(xi, (64, cin, 4, 4), gz, 2, 1)
    assert rel(dw, refdw) < 1e-4
    assert rel(db, gz.sum((0, 2, 3))) < 1e-4
    # accumulate with a scale (1/sigma) and leave the <dW, W> partials of the spectral-norm backward
    part, nparts = torch.full((1024,), 9.0, device="cuda"), ctypes.c_int(0)
    L.call("b2c_first_wgrad", a0, a1, b0, b1, al_d, nhwc(gz), dw, None, sig.cuda()[1:], w.cuda(), part,
           ctypes.byref(nparts), N, S, S)
    assert rel(dw, 1.5 * refdw) < 1e-4
    dot = float(part[:nparts.value].double().sum())
    assert abs(dot - float((refdw.double() * w.double()).sum())) <= 1e-4 * float(refdw.norm() * w.norm())


@pytest.mark.parametrize("N,h", [(3, 8), (5, 16), (2, 4), (3, 2), (64, 16), (3, 3), (2, 5), (2, 6)])
def test_head(L, N, h):
    g = torch.Generator().manual_seed(6)
    C = 512
    a = q(torch.randn(N, C, h, h, generator=g))
    w = torch.randn(1, C, 4, 4, generator=g) * 0.05
    out = torch.empty(N, 1, h - 1, h - 1, device="cuda")
    L.call("b2c_head_fprop", nhwc(a), w.cuda(), out, N, h, C)
    assert rel(out, F.conv2d(a, w, None, 1, 1)) < 1e-4
    gy = torch.randn(N, 1, h - 1, h - 1, generator=g)
    ga = torch.empty(N, h, h, C, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_head_dgrad", gy.cuda(), w.cuda(), ga, N, h, C)
    assert rel(nchw(ga), torch.nn.grad.conv2d_input((N, C, h, h), w, gy, 1, 1)) < 5e-3
    dw = torch.zeros(1, C, 4, 4, device="cuda")
    L.call("b2c_head_wgrad", nhwc(a), gy.cuda(), dw, N, h, C)
    assert rel(dw, torch.nn.grad.conv2d_weight(a, (1, C, 4, 4), gy, 1, 1)) < 1e-4


def test_pack_w_multi_equals_single_packs(L):
    """b2c_pack_w_multi (all stale packs of a network in one launch) writes exactly what b2c_pack_w_c / _t write."""
    g = torch.Generator().manual_seed(11)
    shapes = [(128, 64), (256, 128), (512, 512), (64, 128), (512, 1024)]
    ws = [torch.randn(d0, d1, 4, 4, generator=g).cuda() for d0, d1 in shapes]
    ref, jobs = [], []
    for w, (d0, d1) in zip(ws, shapes):
        for kind, name, shape in ((0, "b2c_pack_w_c", (d0, 16 * d1)), (1, "b2c_pack_w_t", (4, d1, 4 * d0))):
            r = torch.empty(shape, dtype=torch.bfloat16, device="cuda")
            L.call(name, w, None, r, d0, d1)
            o = torch.full(shape, float("nan"), dtype=torch.bfloat16, device="cuda")
            ref.append(r)
            jobs.append((w, o, d0, d1, kind))
    n = len(jobs)
    P, I = ctypes.c_void_p, ctypes.c_int
    L.call("b2c_pack_w_multi", (P * n)(*[j[0].data_ptr() for j in jobs]), (P * n)(*[j[1].data_ptr() for j in jobs]),
           (I * n)(*[j[2] for j in jobs]), (I * n)(*[j[3] for j in jobs]), (I * n)(*[j[4] for j in jobs]), n)
    torch.cuda.synchronize()
    for r, j in zip(ref, jobs):
        assert torch.equal(r.view(torch.int16), j[1].view(torch.int16))
