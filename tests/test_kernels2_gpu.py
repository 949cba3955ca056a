"""Bandwidth / loss / optimizer kernels of libb2c.so against torch CPU autograd (fp64 where it is cheap)."""
import ctypes
import json
from pathlib import Path

import pytest
import torch
import torch.nn.functional as F

from test_kernels_gpu import nchw, nhwc, q, rel

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    from b200cgan import _lib
    return _lib


def mask_nhwc(m):  # (N,C,H,W) {0,1} float cpu -> dense u8 NHWC cuda
    return m.permute(0, 2, 3, 1).contiguous().to(torch.uint8).cuda()


@pytest.mark.parametrize("N,C,H,slope,drop", [(2, 64, 16, 0.2, False), (3, 128, 8, 0.0, True), (2, 512, 2, 0.2, True),
                                              (1, 64, 64, 0.0, False),
                                              # planes small enough for the fused single-pass backward (HW = 1024 / 256 / 64)
                                              (2, 256, 32, 0.0, True), (2, 512, 16, 0.2, True), (3, 512, 8, 0.2, False)])
def test_instance_norm_act_fwd_bwd_bwd2(L, N, C, H, slope, drop):
    g = torch.Generator().manual_seed(11)
    HW = H * H
    z = q(torch.randn(N, C, H, H, generator=g) * 1.5 + 0.3)
    m = (torch.rand(N, C, H, H, generator=g) < 0.5).float() if drop else None
    mk = mask_nhwc(m) if drop else None
    zd = nhwc(z)
    zs = torch.empty(N, C, 2, device="cuda")
    L.call("b2c_in_stats", zd, C, zs, N, HW, C)
    out = torch.empty(N, H, H, C, dtype=torch.bfloat16, device="cuda")
    out2 = torch.empty(N, H, H, 2 * C, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_in_act_fwd", zd, C, zs, mk, slope, out, C, out2[..., C:], 2 * C, N, HW, C)

    def fwd(zz):
        a = F.leaky_relu(F.instance_norm(zz, eps=1e-5), slope)
        return a * m * 2 if drop else a

    zz = z.double().requires_grad_(True)
    ref = fwd(zz)
    assert rel(nchw(out), ref) < 6e-3
    assert torch.equal(out2[..., C:].contiguous(), out)
    # first backward
    ga = q(torch.randn(N, C, H, H, generator=g))
    gad = ga.double().requires_grad_(True)
    (gz_ref,) = torch.autograd.grad(ref, zz, gad, create_graph=True)
    ws = torch.empty(N, C, 5, device="cuda")
    gz = torch.empty(N, H, H, C, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_in_act_bwd", nhwc(ga), C, None, zd, C, zs, mk, slope, ws, gz, C, N, HW, C)
    assert rel(nchw(gz), gz_ref) < 1e-2
    # backward of backward
    h = q(torch.randn(N, C, H, H, generator=g))
    gg_ref, gx_ref = torch.autograd.grad(gz_ref, [gad, zz], h.double())
    gg = torch.empty_like(gz)
    gx = torch.empty_like(gz)
    L.call("b2c_in_act_bwd2", nhwc(h), C, nhwc(ga), C, zd, C, zs, mk, slope, ws, gg, C, gx, C, N, HW, C)
    assert rel(nchw(gg), gg_ref) < 1e-2
    assert rel(nchw(gx), gx_ref) < 1.5e-2
    # pooled variant (G head): pooled sums and broadcast-gradient backward
    if not drop:
        ps = torch.empty(N, C, device="cuda")
        L.call("b2c_in_act_pool_fwd", zd, C, zs, slope, ps, N, HW, C)
        assert rel(ps.cpu() / HW, ref.mean((2, 3))) < 1e-3
        gv = torch.randn(N, C, generator=g)
        (gzp_ref,) = torch.autograd.grad(fwd(zz), zz, gv.double()[:, :, None, None].expand(N, C, H, H))
        L.call("b2c_in_act_bwd", None, 0, gv.cuda(), zd, C, zs, None, slope, ws, gz, C, N, HW, C)
        # 8x8 planes: 64-sample statistics amplify the bf16 rounding of z (two-pass and fused kernels agree to 1e-13)
        assert rel(nchw(gz), gzp_ref) < (1e-2 if HW >= 256 else 2e-2)


def test_lrelu_bwd_copy_dropout(L):
    g = torch.Generator().manual_seed(12)
    a = torch.randn(4, 8, 8, 64, generator=g).to(torch.bfloat16).cuda()
    gr = torch.randn(4, 8, 8, 64, generator=g).to(torch.bfloat16).cuda()
    out = torch.empty_like(a)
    L.call("b2c_lrelu_bwd", gr, a, out, a.numel(), 0.2)
    ref = gr.float() * torch.where(a.float() > 0, 1.0, 0.2)
    assert rel(out, ref) < 5e-3
    dst = torch.zeros(4, 8, 8, 128, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_copy_strided", a, 64, dst[..., 64:], 128, 4 * 64, 64)
    assert torch.equal(dst[..., 64:], a) and float(dst[..., :64].abs().sum()) == 0
    # gradient sum of a skip tensor: dense tensor + channel slice of a concat gradient
    cat = torch.randn(4, 8, 8, 128, generator=g).to(torch.bfloat16).cuda()
    tot = torch.empty_like(a)
    L.call("b2c_add_strided", a, 64, cat[..., 64:], 128, tot, 64, 4 * 64, 64)
    assert torch.equal(tot, (a.float() + cat[..., 64:].float()).to(torch.bfloat16))
    m = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
    L.call("b2c_dropout_mask", m, m.numel(), 7, 0, None)
    assert set(m.unique().tolist()) == {0, 1} and abs(float(m.float().mean()) - 0.5) < 5e-3
    m2 = torch.empty_like(m)
    L.call("b2c_dropout_mask", m2, m.numel(), 7, 0, None)
    assert torch.equal(m, m2)
    L.call("b2c_dropout_mask", m2, m.numel(), 8, 0, None)
    assert not torch.equal(m, m2)
    # stream position = host offset + device base (what a replayed CUDA graph advances)
    L.call("b2c_dropout_mask", m, m.numel(), 7, 4096, None)
    L.call("b2c_dropout_mask", m2, m.numel(), 7, 1024, torch.tensor([3072], dtype=torch.int64, device="cuda"))
    assert torch.equal(m, m2)


@pytest.mark.parametrize("R,K", [(64, 96), (128, 1024), (512, 4096)])
def test_spectral_norm(L, R, K):
    g = torch.Generator().manual_seed(13)
    W = torch.randn(R, K, generator=g) * 0.02
    u = F.normalize(torch.randn(R, generator=g), dim=0)
    v = F.normalize(torch.randn(K, generator=g), dim=0)
    ud, vd, sig, ws = u.cuda(), v.cuda(), torch.empty(2, device="cuda"), torch.empty((R + 31) // 32 * K + R, device="cuda")
    L.call("b2c_sn_power_iter", W.cuda(), ud, vd, sig, ws, R, K, 1)
    v1 = F.normalize(W.t() @ u, dim=0, eps=1e-12)
    u1 = F.normalize(W @ v1, dim=0, eps=1e-12)
    s1 = torch.dot(u1, W @ v1)
    assert rel(vd, v1) < 1e-5 and rel(ud, u1) < 1e-5 and rel(sig[:1], s1.view(1)) < 1e-5
    ud2, vd2 = u.cuda(), v.cuda()  # atomic-free reductions: a second run is bit-identical (ranks stay in lock step)
    L.call("b2c_sn_power_iter", W.cuda(), ud2, vd2, sig, ws, R, K, 1)
    assert torch.equal(ud, ud2) and torch.equal(vd, vd2)
    L.call("b2c_sn_power_iter", W.cuda(), ud, vd, sig, ws, R, K, 0)  # eval: no iteration
    assert rel(ud, u1) < 1e-5 and rel(sig[:1], s1.view(1)) < 1e-5 and rel(sig[1:], 1 / s1.view(1)) < 1e-5
    dWe = torch.randn(R, K, generator=g)
    Wl = W.clone().requires_grad_(True)
    (Wl / torch.dot(u1, Wl @ v1) * dWe).sum().backward()
    dW = torch.ones(R, K, device="cuda")
    L.call("b2c_sn_bwd", dWe.cuda(), W.cuda(), ud, vd, sig, ws, dW, R, K, 1)
    assert rel(dW.cpu() - 1, Wl.grad) < 1e-4


def test_losses_match_reference_golden(L):
    g = json.loads((Path(__file__).parent / "golden" / "losses.json").read_text())
    box, delta, dt = (torch.tensor(g[k]).cuda() for k in ("box", "delta", "delta_true"))
    B = box.shape[0]
    for tr in (True, False):
        a, b = torch.empty(B, 4, device="cuda"), torch.empty(B, 4, device="cuda")
        L.call("b2c_apply_delta", box, delta, int(tr), a, None, None, B)
        L.call("b2c_apply_delta", box, dt, int(tr), b, None, None, B)
        assert rel(a, torch.tensor(g["apply_%s" % tr])) < 1e-5
        loss, iou = torch.empty(1, device="cuda"), torch.empty(B, device="cuda")
        L.call("b2c_eiou", a, b, 1e-6, loss, None, None, iou, B)
        assert abs(float(loss) - g["eiou_%s" % tr]) < 1e-5
        assert rel(iou, torch.tensor(g["iou_%s" % tr])) < 1e-4
    a, b = torch.empty(B, 4, device="cuda"), torch.empty(B, 4, device="cuda")
    L.call("b2c_apply_delta", box, delta, 1, a, None, None, B)
    L.call("b2c_apply_delta", box, dt, 1, b, None, None, B)
    gp_, gd = torch.empty(B, 4, device="cuda"), torch.empty(B, 4, device="cuda")
    L.call("b2c_eiou", a, b, 1e-6, None, None, gp_, None, B)
    L.call("b2c_apply_delta", box, delta, 1, None, gp_, gd, B)
    assert rel(gd, torch.tensor(g["eiou_grad_delta"])) < 1e-4


def test_gp_and_reductions(L):
    g = torch.Generator().manual_seed(14)
    B, n = 5, 3 * 32 * 32
    x0, x1 = torch.randn(B, n, generator=g), torch.randn(B, n, generator=g)
    ss = torch.zeros(B, device="cuda")
    L.call("b2c_rowsum", x0.cuda(), n, B, 1, 1.0, ss)
    L.call("b2c_rowsum", x1.cuda(), n, B, 1, 1.0, ss)
    assert rel(ss, (x0 ** 2).sum(1) + (x1 ** 2).sum(1)) < 1e-5
    gp, coef = torch.empty(1, device="cuda"), torch.empty(B, device="cuda")
    L.call("b2c_gp_tail", ss, gp, coef, B)
    xx = torch.cat([x0, x1], 1).double().requires_grad_(True)
    ref = ((torch.sqrt((xx ** 2).sum(1) + 1e-12) - 1) ** 2).mean()
    ref.backward()
    assert abs(float(gp) - float(ref)) / float(ref) < 1e-5
    h0 = torch.empty(B, n, device="cuda")
    gs = torch.tensor([0.7], device="cuda")
    L.call("b2c_scale_rows", x0.cuda(), coef, gs, h0, n, B)
    assert rel(h0, 0.7 * xx.grad[:, :n]) < 1e-4
    m = torch.zeros(1, device="cuda")
    L.call("b2c_rowsum", x0.cuda(), B * n, 1, 0, 1.0 / (B * n), m)
    assert abs(float(m) - float(x0.mean())) < 1e-5
    f = torch.empty(100, device="cuda")
    L.call("b2c_fill_scaled", f, 100, gs, -0.01)
    assert rel(f, torch.full((100,), -0.007)) < 1e-6


def test_pool_fc_tanh(L):
    g = torch.Generator().manual_seed(15)
    N, C, HW = 6, 64, 256
    ps = torch.randn(N, C, generator=g) * HW
    W, b = torch.randn(4, C, generator=g) * 0.1, torch.randn(4, generator=g) * 0.1
    delta, th = torch.empty(N, 4, device="cuda"), torch.empty(N, 4, device="cuda")
    L.call("b2c_pool_fc_fwd", ps.cuda(), W.cuda(), b.cuda(), 0.3, HW, delta, th, N, C)
    pl = (ps / HW).double().requires_grad_(True)
    Wd, bd = W.double().requires_grad_(True), b.double().requires_grad_(True)
    ref = torch.tanh(F.linear(pl, Wd, bd)) * 0.3
    assert rel(delta, ref) < 1e-5
    gd = torch.randn(N, 4, generator=g)
    ref.backward(gd.double())
    dW, db, gv = torch.empty(4, C, device="cuda"), torch.empty(4, device="cuda"), torch.empty(N, C, device="cuda")
    L.call("b2c_pool_fc_bwd", gd.cuda(), th, ps.cuda(), W.cuda(), 0.3, HW, dW, db, gv, N, C)
    assert rel(dW, Wd.grad) < 1e-4 and rel(db, bd.grad) < 1e-4 and rel(gv.cpu() * HW, pl.grad) < 1e-4


def test_clip_adam_matches_torch(L):
    g = torch.Generator().manual_seed(16)
    shapes = [(64, 3, 4, 4), (128, 64, 4, 4), (4, 64), (4,)]
    ps = [torch.randn(*s, generator=g) for s in shapes]
    ref_p = [torch.nn.Parameter(p.clone()) for p in ps]
    opt = torch.optim.Adam(ref_p, lr=2e-4, betas=(0.5, 0.999))
    dp = [p.clone().cuda() for p in ps]
    dm, dv = [torch.zeros_like(p) for p in dp], [torch.zeros_like(p) for p in dp]
    n = len(ps)
    arr = lambda ts: (ctypes.c_void_p * n)(*[t.data_ptr() for t in ts])
    cnt = (ctypes.c_longlong * n)(*[p.numel() for p in ps])
    nrm = torch.empty(1, device="cuda")
    for step in range(1, 4):
        gs = [torch.randn(*s, generator=g) * (10.0 if step < 3 else 1e-3) for s in shapes]
        for p, gr in zip(ref_p, gs):
            p.grad = gr.clone()
        tn = torch.nn.utils.clip_grad_norm_(ref_p, 1.0)
        opt.step()
        dg = [x.cuda() for x in gs]
        L.call("b2c_multi_sumsq", arr(dg), cnt, n, nrm)
        assert abs(float(nrm.sqrt()) - float(tn)) / float(tn) < 1e-5
        L.call("b2c_multi_adam", arr(dp), arr(dg), arr(dm), arr(dv), cnt, n, nrm, 1.0, 2e-4, 0.5, 0.999, 1e-8,
               0 if step == 2 else step, torch.tensor([2], dtype=torch.int64, device="cuda") if step == 2 else None)
        for a, b in zip(dp, ref_p):
            assert rel(a, b) < 1e-6
