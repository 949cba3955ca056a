"""Teacher-forced per-layer parity of the DISCRIMINATOR, first AND second order (north_star: per-layer outputs and gradients
<= 2e-2 in bf16).

The gradient penalty (cgan/losses.py:185-233) differentiates through D's first backward, so every block of D takes part in a
map  F: (a_in, g_out; W) -> (a_out, g_in)  with  g_in = J(a_in)^T g_out  (the block's own backward), and the critic's
backward pushes cotangents (k on a_out, h on g_in) through F: that is where Conv4s2Dgrad / INActBwd / LReluBwd / FirstDgrad /
HeadDgrad and their backward-of-backward compose.  Here one critic loss (real + fake + penalty) is run through the kernels
with tensor hooks that record, for every block, exactly the tensors it saw in the penalty branch - its input activation
a_in, the first-backward gradient g_out that entered it, and the second-backward cotangents k and h - and each block is
then re-run ALONE on those tensors and compared with the oracle block (cgan/models.py:236-258 restated in torch fp32 on
bf16 storage, differentiated with create_graph=True): a_out, g_in, and the three vector-Jacobian products d/d a_in,
d/d g_out, d/d weight_orig (through W / sigma(u, v) with the u, v of that forward instance) all within 2e-2."""
import pytest
import torch
import torch.nn.functional as F

import cgan_oracle as O
from test_models_gpu import build, cuda_batch, rel

pytestmark = pytest.mark.gpu
TOL = 2e-2


def _r(t):  # bf16 storage, straight-through (what O.emulate_bf16 does)
    return t + (t.to(torch.bfloat16).to(t.dtype) - t).detach()


def oracle_norm_block(a, w_orig, u, v, bias):
    """SN-conv -> InstanceNorm -> LeakyReLU (cgan/models.py:236-242) with the given power-iteration vectors."""
    wm = w_orig.reshape(w_orig.shape[0], -1)
    sigma = torch.dot(u, torch.mv(wm, v))
    z = F.conv2d(_r(a), _r(w_orig) / sigma, bias, 2, 1)  # bf16 master weight, 1/sigma on the fp32 accumulators
    return _r(F.leaky_relu(F.instance_norm(_r(z), eps=1e-5), 0.2))


def oracle_first_block(x, w_orig, u, v, bias):
    wm = w_orig.reshape(w_orig.shape[0], -1)
    sigma = torch.dot(u, torch.mv(wm, v))
    return _r(F.leaky_relu(F.conv2d(_r(x), _r(w_orig) / sigma, bias, 2, 1), 0.2))


def vjp2(fn, a_in, g_out, k, h, params):
    """(a_out, g_in, d a_in, d g_out, [d param]) of F with cotangents (k, h)."""
    a_out = fn(a_in)
    (g_in,) = torch.autograd.grad(a_out, a_in, g_out, create_graph=True)
    grads = torch.autograd.grad([a_out, g_in], [a_in, g_out] + params, [k, h], allow_unused=True)
    return a_out, g_in, grads[0], grads[1], list(grads[2:])


@pytest.mark.parametrize("B,S", [(2, 64), (2, 128)])
def test_discriminator_blocks_teacher_forced_second_order(B, S):
    from b200cgan import ops
    st = O.init_state(42)
    _, D = build(st)
    D.train()
    batch, _, alphas = O.make_batch(B, S, 0)
    cb = cuda_batch(batch)
    # ---- one penalty branch through the kernels, recording what every block saw
    interp = ops.Interp(cb["pred"], cb["gt"], cb["pred"], cb["refined"], alphas[0].cuda())
    ip = torch.empty_like(interp.a0).requires_grad_(True)
    io = torch.empty_like(interp.a1).requires_grad_(True)
    ctxs = D.weight_ctxs()
    acts = [D.first_block(ip, io, ctxs[0], interp)]
    for idx, wctx in zip((2, 5, 8), ctxs[1:]):
        acts.append(D.norm_block(idx, acts[-1], wctx))
    out = ops.HeadConv.apply(acts[-1], D.model[11].weight)
    rec = [dict() for _ in acts]  # per activation a_l: first-backward gradient g and second-backward cotangents k (on a_l), h (on g)
    for r, a in zip(rec, acts):
        def hook(g, r=r):
            if "g" not in r:       # first backward (create_graph): the gradient entering the block below through a_l
                r["g"] = g
                g.register_hook(lambda hh, r=r: r.__setitem__("h", hh.detach().clone()))
            else:                  # second backward: cotangent on the forward activation itself
                r["k"] = g.detach().clone()
        a.register_hook(hook)
    ones = torch.ones_like(out)
    with ops.backward_flags(weight_grads=False):
        g0, g1 = torch.autograd.grad(out, [ip, io], ones, create_graph=True)
    hx = []
    g0.register_hook(lambda t: hx.append(t.detach().clone()))
    g1.register_hook(lambda t: hx.append(t.detach().clone()))
    gp = ops.GradNormPenalty.apply(g0, g1)
    gp.backward()
    worst = {}
    # ---- blocks model.2 / 5 / 8: F(a_in, g_out) with cotangents (k on a_out, h on g_in)
    for li, idx in enumerate((2, 5, 8), start=1):
        conv, wctx = D.model[idx], ctxs[li]
        a_in_v, g_out_v, h_v = acts[li - 1].detach(), rec[li]["g"].detach(), rec[li - 1]["h"]
        k_v = rec[li].get("k")  # the last block's activation only feeds the head: no second-order cotangent reaches it
        if k_v is None:
            k_v = torch.zeros_like(acts[li])
        conv.weight_orig.grad = None
        a_in = a_in_v.clone().requires_grad_(True)
        g_out = g_out_v.clone().requires_grad_(True)
        w2 = ops.WeightCtx(conv.weight_orig, wctx.sig, wctx.u, wctx.v)
        got = vjp2(lambda a: D.norm_block(idx, a, w2), a_in, g_out, k_v, h_v, [conv.weight_orig])
        c = lambda t: t.detach().float().cpu()
        a_c, g_c = c(a_in_v), c(g_out_v).requires_grad_(True)
        w_c = st["D"][f"model.{idx}.weight_orig"].clone().requires_grad_(True)
        b_c = st["D"][f"model.{idx}.bias"].clone()
        if idx == 2:
            # model.2's data gradient carries the LeakyReLU backward of model.0-1 (fused epilogue), i.e. it is the gradient
            # with respect to the PRE-activation z0: the oracle block therefore starts from z0 = lrelu^-1(a_in)
            a_c = torch.where(a_c > 0, a_c, a_c / 0.2).requires_grad_(True)
            fn = lambda zz: oracle_norm_block(F.leaky_relu(zz, 0.2), w_c, c(wctx.u), c(wctx.v), b_c)
        else:
            a_c = a_c.requires_grad_(True)
            fn = lambda a: oracle_norm_block(a, w_c, c(wctx.u), c(wctx.v), b_c)
        ref = vjp2(fn, a_c, g_c, c(k_v), c(h_v), [w_c])
        errs = {"a_out": rel(got[0], ref[0]), "g_in": rel(got[1], ref[1]), "d_a_in": rel(got[2], ref[2]),
                "d_g_out": rel(got[3], ref[3]), "d_weight_orig": rel(got[4][0], ref[4][0])}
        worst[f"model.{idx}"] = errs
        for name, e in errs.items():
            assert e < TOL, (idx, name, e, errs)
    # ---- head model.11: a_out = conv(a4, W5); g_in = dgrad(g_out, W5)
    w5 = D.model[11].weight
    w5.grad = None
    a_in = acts[3].detach().clone().requires_grad_(True)
    g_out = ones.clone().requires_grad_(True)
    k_v = torch.zeros_like(out)  # the score map only feeds autograd.grad: no cotangent on it in the penalty branch
    got = vjp2(lambda a: ops.HeadConv.apply(a, w5), a_in, g_out, k_v, rec[3]["h"], [w5])
    c = lambda t: t.detach().float().cpu()
    a_c, g_c = c(acts[3]).requires_grad_(True), c(ones).requires_grad_(True)
    w_c = st["D"]["model.11.weight"].clone().requires_grad_(True)
    ref = vjp2(lambda a: F.conv2d(a, w_c, None, 1, 1), a_c, g_c, c(k_v), c(rec[3]["h"]), [w_c])
    errs = {"a_out": rel(got[0], ref[0]), "g_in": rel(got[1], ref[1]), "d_a_in": rel(got[2], ref[2]),
            "d_g_out": rel(got[3], ref[3]), "d_weight": rel(got[4][0], ref[4][0])}
    worst["model.11"] = errs
    for name, e in errs.items():
        assert e < TOL, ("head", name, e, errs)
    # ---- first block model.0 on the interpolates: g_in = (d/d ip, d/d io), cotangents (k on a_1, (h0, h1) on g_in)
    c0, wctx = D.model[0], ctxs[0]
    c0.weight_orig.grad = None
    c0.bias.grad = None
    al = alphas[0].cuda().reshape(-1, 1, 1, 1)
    xp = (al * cb["pred"] + (1 - al) * cb["pred"]).detach()
    xo = (al * cb["gt"] + (1 - al) * cb["refined"]).detach()
    x_in = [xp.clone().requires_grad_(True), xo.clone().requires_grad_(True)]
    g_out = rec[0]["g"].detach().clone().requires_grad_(True)
    w2 = ops.WeightCtx(c0.weight_orig, wctx.sig, wctx.u, wctx.v)
    # acts[0] is "premasked": its LeakyReLU backward lives in model.2's dgrad epilogue, so the gradient / cotangent that
    # reach this block (rec[0]["g"], rec[0]["k"]) are with respect to the pre-activation z0; the oracle block is the conv
    a_out = ops.FirstConv.apply(x_in[0], x_in[1], c0.weight_orig, c0.bias, w2, 0.2, True)
    gi = torch.autograd.grad(a_out, x_in, g_out, create_graph=True)
    grads = torch.autograd.grad([a_out, gi[0], gi[1]], [g_out, c0.weight_orig], [rec[0]["k"], hx[0], hx[1]], allow_unused=True)
    c = lambda t: t.detach().float().cpu()
    x_c = torch.cat([c(xp), c(xo)], 1).requires_grad_(True)
    g_c = c(rec[0]["g"]).requires_grad_(True)
    w_c = st["D"]["model.0.weight_orig"].clone().requires_grad_(True)
    b_c = st["D"]["model.0.bias"].clone()
    wm = w_c.reshape(64, -1)
    z0 = F.conv2d(_r(x_c), _r(w_c) / torch.dot(c(wctx.u), torch.mv(wm, c(wctx.v))), b_c, 2, 1)
    (gi_c,) = torch.autograd.grad(z0, x_c, g_c, create_graph=True)
    rg = torch.autograd.grad([z0, gi_c], [g_c, w_c], [c(rec[0]["k"]), torch.cat([c(hx[0]), c(hx[1])], 1)])
    errs = {"a_out": rel(a_out, _r(F.leaky_relu(z0, 0.2))), "g_in": rel(torch.cat([gi[0], gi[1]], 1), gi_c),
            "d_g_out": rel(grads[0], rg[0]), "d_weight_orig": rel(grads[1], rg[1])}
    worst["model.0"] = errs
    for name, e in errs.items():
        assert e < TOL, ("first", name, e, errs)
    print("teacher-forced D block errors (first and second order):",
          {n: {k: round(v, 5) for k, v in e.items()} for n, e in worst.items()})
