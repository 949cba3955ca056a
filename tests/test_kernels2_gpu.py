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
    zs = torch.empty(N, C, 2, dtype=torch.int64, device="cuda")
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
    L.call("b2c_in_act_bwd", nhwc(ga), C, None, None, zd, C, zs, mk, slope, ws, gz, C, N, HW, C)
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
        ps = torch.empty(2, N, C, dtype=torch.int64, device="cuda")  # planar fixed-point (sum act(x_hat), sum act'(x_hat))
        L.call("b2c_in_act_pool_fwd", zd, C, zs, slope, ps, N, HW, C)
        psf = ps.double().cpu() * 2.0 ** -30
        assert rel(psf[0] / HW, ref.mean((2, 3))) < 1e-3
        xh = F.instance_norm(z.double(), eps=1e-5)
        assert rel(psf[1], torch.where(xh > 0, 1.0, slope).sum((2, 3))) < 2e-3
        gv = torch.randn(N, C, generator=g)
        (gzp_ref,) = torch.autograd.grad(fwd(zz), zz, gv.double()[:, :, None, None].expand(N, C, H, H))
        for fs in (None, ps):  # two-pass reduction, then the single apply pass fed by the forward's plane sums
            gz.zero_()
            L.call("b2c_in_act_bwd", None, 0, gv.cuda(), fs, zd, C, zs, None, slope, ws, gz, C, N, HW, C)
            # 8x8 planes: 64-sample statistics amplify the bf16 rounding of z (two-pass and fused kernels agree to 1e-13)
            assert rel(nchw(gz), gzp_ref) < (1e-2 if HW >= 256 else 2e-2)


@pytest.mark.parametrize("N,C,H,ld", [(3, 128, 64, 128), (2, 256, 32, 256), (5, 512, 16, 512), (2, 64, 8, 64),
                                      (4, 128, 32, 192), (65, 128, 16, 128)])
def test_instance_norm_cluster_kernels(L, N, C, H, ld):
    """The single-launch cluster kernels (TMA ring + DSMEM exchange; CS = 1..8, resident and streaming rings) at plane sizes
    of the step's D layers, incl. a channel-slice gradient (pixel stride ld > C), against fp64 autograd - and against the
    two-pass kernels they replace (B2C_NORM_CLUSTER=0 is read once per process, so the reference here is torch)."""
    g = torch.Generator().manual_seed(21)
    HW, slope = H * H, 0.2
    z = q(torch.randn(N, C, H, H, generator=g) * 1.3 - 0.2)
    zd = nhwc(z)
    zs = torch.empty(N, C, 2, dtype=torch.int64, device="cuda")
    L.call("b2c_in_stats", zd, C, zs, N, HW, C)
    zz = z.double().requires_grad_(True)
    ref = F.leaky_relu(F.instance_norm(zz, eps=1e-5), slope)
    ga = q(torch.randn(N, C, H, H, generator=g))
    gad = ga.double().requires_grad_(True)
    (gz_ref,) = torch.autograd.grad(ref, zz, gad, create_graph=True)
    gwide = torch.zeros(N, H, H, ld, dtype=torch.bfloat16, device="cuda")
    gwide[..., ld - C:] = nhwc(ga)
    gsl = gwide[..., ld - C:]
    ws = torch.empty(N, C, 5, device="cuda")
    gz = torch.empty(N, H, H, C, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_in_act_bwd", gsl, ld, None, None, zd, C, zs, None, slope, ws, gz, C, N, HW, C)
    assert rel(nchw(gz), gz_ref) < 1e-2
    gz2 = torch.empty_like(gz)
    L.call("b2c_in_act_bwd", gsl, ld, None, None, zd, C, zs, None, slope, ws, gz2, C, N, HW, C)
    assert torch.equal(gz, gz2)  # no atomics: bit-identical from run to run
    h = q(torch.randn(N, C, H, H, generator=g))
    gg_ref, gx_ref = torch.autograd.grad(gz_ref, [gad, zz], h.double())
    gg, gx = torch.empty_like(gz), torch.empty_like(gz)
    L.call("b2c_in_act_bwd2", nhwc(h), C, gsl, ld, zd, C, zs, None, slope, ws, gg, C, gx, C, N, HW, C)
    assert rel(nchw(gg), gg_ref) < 1e-2
    assert rel(nchw(gx), gx_ref) < 1.5e-2


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
    # uniform draws of the same stream (GP alpha): in [0,1), mean 1/2, a shifted window is the same numbers
    u1, u2 = torch.empty(1 << 16, device="cuda"), torch.empty(1 << 16, device="cuda")
    L.call("b2c_uniform", u1, u1.numel(), 7, 0, None)
    L.call("b2c_uniform", u2, u2.numel() - 100, 7, 60, torch.tensor([40], dtype=torch.int64, device="cuda"))
    assert float(u1.min()) >= 0.0 and float(u1.max()) < 1.0 and abs(float(u1.mean()) - 0.5) < 5e-3
    assert torch.equal(u1[100:], u2[:-100])


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


def test_spectral_norm_batched_and_deferred_rank1(L):
    """b2c_sn_power_iter_multi (all layers of a D forward in three launches, with the per-forward u / v snapshots) equals
    the per-layer kernel, and the deferred form of the backward - gradient sink += dWe / sigma (weight-gradient kernels),
    then ONE b2c_sn_rank1 for every layer and forward instance - equals autograd through W / sigma(u, v)."""
    g = torch.Generator().manual_seed(17)
    shapes = [(64, 96), (128, 1024), (256, 2048), (512, 4096)]
    Ws = [torch.randn(R, K, generator=g) * 0.02 for R, K in shapes]
    us = [F.normalize(torch.randn(R, generator=g), dim=0) for R, _ in shapes]
    vs = [F.normalize(torch.randn(K, generator=g), dim=0) for _, K in shapes]
    n = len(shapes)
    P, I = ctypes.c_void_p, ctypes.c_int
    arr = lambda ts: (P * n)(*[t.data_ptr() for t in ts])
    Wd = [w.cuda() for w in Ws]
    insts = []  # per forward instance: (u_snap, v_snap, sig) per layer
    ud, vd = [u.cuda() for u in us], [v.cuda() for v in vs]
    ur, vr = [u.clone() for u in us], [v.clone() for v in vs]
    for it in range(2):  # two forwards: u / v advance, each instance keeps its own snapshot
        sig = [torch.empty(2, device="cuda") for _ in shapes]
        ws = [torch.empty(K + (K + 31) // 32 + R, device="cuda") for R, K in shapes]
        usn = [torch.empty(R, device="cuda") for R, _ in shapes]
        vsn = [torch.empty(K, device="cuda") for _, K in shapes]
        L.call("b2c_sn_power_iter_multi", arr(Wd), arr(ud), arr(vd), arr(sig), arr(ws), arr(usn), arr(vsn),
               (I * n)(*[r for r, _ in shapes]), (I * n)(*[k for _, k in shapes]), n, 1)
        for l in range(n):
            vr[l] = F.normalize(Ws[l].t() @ ur[l], dim=0, eps=1e-12)
            ur[l] = F.normalize(Ws[l] @ vr[l], dim=0, eps=1e-12)
            s1 = torch.dot(ur[l], Ws[l] @ vr[l])
            assert rel(vd[l], vr[l]) < 1e-5 and rel(ud[l], ur[l]) < 1e-5 and rel(sig[l][:1], s1.view(1)) < 1e-5
            assert torch.equal(usn[l], ud[l]) and torch.equal(vsn[l], vd[l])
        insts.append((usn, vsn, sig, [u.clone() for u in ur], [v.clone() for v in vr]))
    # deferred backward over both instances
    grads = [torch.zeros(R, K, device="cuda") for R, K in shapes]
    refs = [torch.zeros(R, K, dtype=torch.float64) for R, K in shapes]
    fu, fv, fs, fp, fn = [], [], [], [], []
    for l, (R, K) in enumerate(shapes):
        for usn, vsn, sig, uref, vref in insts:
            dWe = torch.randn(R, K, generator=g)
            Wl = Ws[l].double().clone().requires_grad_(True)
            (Wl / torch.dot(uref[l].double(), Wl @ vref[l].double()) * dWe.double()).sum().backward()
            refs[l] += Wl.grad
            grads[l] += dWe.cuda() * sig[l][1]                      # what the wgrad kernels leave in the sink
            prod = (dWe.double() * Ws[l].double()).reshape(-1)
            parts = torch.stack([c.sum() for c in prod.chunk(7)]).float().cuda()  # block partials of <dWe, W>
            fu.append(usn[l]); fv.append(vsn[l]); fs.append(sig[l]); fp.append(parts); fn.append(parts.numel())
    t = len(fu)
    L.call("b2c_sn_rank1", arr(grads), (I * n)(*[r for r, _ in shapes]), (I * n)(*[k for _, k in shapes]), (I * n)(*[2] * n), n,
           (P * t)(*[x.data_ptr() for x in fu]), (P * t)(*[x.data_ptr() for x in fv]), (P * t)(*[x.data_ptr() for x in fs]),
           (P * t)(*[x.data_ptr() for x in fp]), (I * t)(*fn))
    for l in range(n):
        assert rel(grads[l], refs[l]) < 1e-4, l


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
    part, splits = torch.full((32 * B,), 3.0, device="cuda"), ctypes.c_int(0)
    L.call("b2c_gp_sumsq", x0.cuda(), x1.cuda(), n, B, part, ctypes.byref(splits))  # two-stage, no atomics
    assert rel(part[:B * splits.value].view(B, -1).sum(1), ss) < 1e-5
    gp, coef = torch.empty(1, device="cuda"), torch.empty(B, device="cuda")
    L.call("b2c_gp_tail", part, splits.value, gp, coef, B)
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
    L.call("b2c_map_mean", x0.cuda(), B * n, -1.0 / (B * n), m)
    assert abs(float(m) + float(x0.mean())) < 1e-5
    # the critic objective in one launch and its backward
    real, fake = torch.randn(B, 1, 15, 15, generator=g), torch.randn(B, 1, 15, 15, generator=g)
    out3 = torch.empty(3, device="cuda")
    L.call("b2c_critic_loss", real.cuda(), real.numel(), fake.cuda(), fake.numel(), gp, 10.0, out3)
    wd = float(real.double().mean() - fake.double().mean())
    assert abs(float(out3[1]) - wd) < 1e-5 and abs(float(out3[0]) - (-wd + 10.0 * float(gp))) < 1e-4 * (1 + 10 * float(gp))
    gr, gf = torch.empty(B * 225, device="cuda"), torch.empty(B * 225, device="cuda")
    L.call("b2c_critic_loss_bwd", gs, gr, gr.numel(), gf, gf.numel())
    assert rel(gr, torch.full((B * 225,), -0.7 / (B * 225))) < 1e-6 and rel(gf, torch.full((B * 225,), 0.7 / (B * 225))) < 1e-6
    zz = torch.ones(1000, device="cuda")
    L.call("b2c_zero", zz, 4000)
    assert float(zz.abs().sum()) == 0.0


def test_pool_fc_tanh(L):
    g = torch.Generator().manual_seed(15)
    N, C, HW = 6, 64, 256
    ps = torch.randn(N, C, generator=g) * HW
    W, b = torch.randn(4, C, generator=g) * 0.1, torch.randn(4, generator=g) * 0.1
    delta, th = torch.empty(N, 4, device="cuda"), torch.empty(N, 4, device="cuda")
    psi = (ps.double() * 2.0 ** 30).round().to(torch.int64).cuda()  # pooled sums travel as 64-bit fixed point
    ps = psi.double().cpu() * 2.0 ** -30
    L.call("b2c_pool_fc_fwd", psi, W.cuda(), b.cuda(), 0.3, HW, delta, th, N, C)
    pl = (ps / HW).double().requires_grad_(True)
    Wd, bd = W.double().requires_grad_(True), b.double().requires_grad_(True)
    ref = torch.tanh(F.linear(pl, Wd, bd)) * 0.3
    assert rel(delta, ref) < 1e-5
    gd = torch.randn(N, 4, generator=g)
    ref.backward(gd.double())
    dW, db, gv = torch.empty(4, C, device="cuda"), torch.empty(4, device="cuda"), torch.empty(N, C, device="cuda")
    L.call("b2c_pool_fc_bwd", gd.cuda(), th, psi, W.cuda(), 0.3, HW, dW, db, gv, N, C)
    assert rel(dW, Wd.grad) < 1e-4 and rel(db, bd.grad) < 1e-4 and rel(gv.cpu() * HW, pl.grad) < 1e-4


def test_clip_adam_matches_torch(L):
    g = torch.Generator().manual_seed(16)
    shapes = [(64, 3, 4, 4), (128, 64, 4, 4), (4, 64), (4,), (7,)]  # the last one forces the scalar (non-float4) path
    ps = [torch.randn(*s, generator=g) for s in shapes]
    ref_p = [torch.nn.Parameter(p.clone()) for p in ps]
    opt = torch.optim.Adam(ref_p, lr=2e-4, betas=(0.5, 0.999))
    dp = [p.clone().cuda() for p in ps]
    dm, dv = [torch.zeros_like(p) for p in dp], [torch.zeros_like(p) for p in dp]
    n = len(ps)
    arr = lambda ts: (ctypes.c_void_p * n)(*[t.data_ptr() for t in ts])
    cnt = (ctypes.c_longlong * n)(*[p.numel() for p in ps])
    part, nrm = torch.empty(1024, device="cuda"), torch.empty(1, device="cuda")
    lr_dev = torch.tensor([2e-4], device="cuda")
    for step in range(1, 5):
        gs = [torch.randn(*s, generator=g) * (10.0 if step < 3 else 1e-3) for s in shapes]
        if step == 4:  # a scheduler changed the learning rate: the device-side value wins over the launch argument
            for grp in opt.param_groups:
                grp["lr"] = 1e-4
            lr_dev.fill_(1e-4)
        for p, gr in zip(ref_p, gs):
            p.grad = gr.clone()
        tn = torch.nn.utils.clip_grad_norm_(ref_p, 1.0)
        opt.step()
        dg = [x.cuda() * 4.0 for x in gs]  # as if SUM-all-reduced over 4 ranks: grad_scale = 1/4 undoes it
        nparts = ctypes.c_int(0)
        L.call("b2c_multi_sumsq", arr(dg), cnt, n, part, ctypes.byref(nparts))
        L.call("b2c_multi_adam", arr(dp), arr(dg), arr(dm), arr(dv), cnt, n, part, nparts.value, nrm, 0.25, 1.0, 2e-4,
               lr_dev, 0.5, 0.999, 1e-8, 0 if step == 2 else step,
               torch.tensor([2], dtype=torch.int64, device="cuda") if step == 2 else None)
        assert abs(float(nrm.sqrt()) - float(tn)) / float(tn) < 1e-5
        for a, b in zip(dp, ref_p):
            assert rel(a, b) < 1e-6


def test_fused_clip_adam_flat_buckets():
    """ops.FusedClipAdam re-homes parameters / gradients / moments in flat buckets: parameters keep their values and
    state-dict identity, zero_grad is one memset, a step equals torch's clip_grad_norm_ + Adam."""
    from b200cgan import ops
    g = torch.Generator().manual_seed(18)
    mods = [torch.nn.Conv2d(3, 8, 3), torch.nn.Linear(10, 3)]
    params = [p for m in mods for p in m.parameters()]
    ref = [torch.nn.Parameter(p.detach().clone()) for p in params]
    for m in mods:
        m.cuda()
    params = [p for m in mods for p in m.parameters()]
    opt = ops.FusedClipAdam(params, lr=2e-4, betas=(0.5, 0.999))
    ropt = torch.optim.Adam(ref, lr=2e-4, betas=(0.5, 0.999))
    for p, r in zip(params, ref):
        assert torch.equal(p.detach().cpu(), r.detach()) and p.grad is p._b2c_sink and float(p.grad.abs().sum()) == 0
    for step in range(3):
        opt.zero_grad()
        for p, r in zip(params, ref):
            gr = torch.randn(p.shape, generator=g) * 3
            p._b2c_sink.add_(gr.cuda())  # what the weight-gradient kernels do
            r.grad = gr.clone()
        tn = torch.nn.utils.clip_grad_norm_(ref, 1.0)
        ropt.step()
        gn = opt.step()
        assert abs(float(gn.sqrt()) - float(tn)) / float(tn) < 1e-5
        for p, r in zip(params, ref):
            assert rel(p, r) < 1e-6
    assert torch.equal(mods[1].state_dict()["weight"].cpu(), params[2].detach().cpu())
