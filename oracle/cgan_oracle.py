"""CPU oracle for the cGAN patch-calibrator train step.  TEST INFRASTRUCTURE ONLY.

This is a functional, torch-CPU restatement of the reference's hot path.  It is NOT part of
the product: only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and only as the checker / timed CPU baseline.  The
product path (``b200cgan``) never imports it and fails loudly without its CUDA library.

Parity pin: the reference has no tests or golden vectors of its own, so this file
is pinned against outputs of the UNMODIFIED reference modules run in the build container
(``oracle/make_golden.py`` imports ``/root/reference/cgan/{models,losses}.py`` and writes
``tests/golden/*.json``; ``tests/test_oracle_golden.py`` replays them with this file).

Reference lines restated (paths relative to /root/reference):
  g_forward           cgan/models.py:54-83,89-141      (UNetDown / UNetUp / GeneratorUNet)
  simple_forward      cgan/models.py:147-216           (GeneratorSimpleRegressor, eval mode)
  sn_weight           torch/nn/utils/spectral_norm.py:92-114 (legacy hook attached at cgan/models.py:236-238)
  d_forward           cgan/models.py:222-258           (Discriminator)
  init_state          cgan/models.py:37-48 + cgan/cgan_train_enhanced.py:229-232,256-257
  smooth_clamp        cgan/losses.py:99-106
  apply_delta_to_bbox cgan/losses.py:108-150
  eiou_loss           cgan/losses.py:19-73
  iou_metric          cgan/losses.py:152-183
  gradient_penalty    cgan/losses.py:185-233
  clip_and_adam       torch clip_grad_norm_ + optim.Adam as called at cgan/cgan_train_enhanced.py:256-257,331-332,368-369
  train_step          cgan/cgan_train_enhanced.py:297-369 (loop body; get_refined_patch_batch replaced
                      by an injected ``refined`` tensor, dropout masks and GP alpha injected)
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

G_DOWN = [("down1", 3, 64, False, False), ("down2", 64, 128, True, False),
          ("down3", 128, 256, True, False), ("down4", 256, 512, True, True)]
G_UP = [("up1", 512, 256, True), ("up2", 512, 128, True), ("up3", 256, 64, False)]
D_BLOCKS = [(0, 6, 64, False), (2, 64, 128, True), (5, 128, 256, True), (8, 256, 512, True)]
D_BIAS_UNOBSERVABLE = ("model.2.bias", "model.5.bias", "model.8.bias")  # feed affine-less IN: true grad 0


# ----------------------------------------------------------------------------- init
def init_state(seed: int = 42, spectral_norm: bool = True, dtype=torch.float32) -> Dict[str, Dict[str, torch.Tensor]]:
    """Draw initial weights with the same RNG consumption order as the reference trainer
    (ctor of G, ctor of D, then .apply(weights_init_normal) on G then D)."""
    torch.manual_seed(seed)
    g: Dict[str, torch.Tensor] = {}
    holders = []
    for name, cin, cout, _, _ in G_DOWN:
        m = torch.nn.Conv2d(cin, cout, 4, 2, 1, bias=False); holders.append((name + ".model.0.weight", m))
    for name, cin, cout, _ in G_UP:
        m = torch.nn.ConvTranspose2d(cin, cout, 4, 2, 1, bias=False); holders.append((name + ".model.0.weight", m))
    m = torch.nn.ConvTranspose2d(128, 64, 4, 2, 1, bias=False); holders.append(("up4.0.weight", m))
    fc = torch.nn.Linear(64, 4)
    d_mods = []
    for idx, cin, cout, _ in D_BLOCKS:
        conv = torch.nn.Conv2d(cin, cout, 4, 2, 1)
        if spectral_norm:
            conv = torch.nn.utils.spectral_norm(conv)
        d_mods.append((idx, conv))
    head = torch.nn.Conv2d(512, 1, 4, 1, 1, bias=False)
    # weights_init_normal: every class whose name contains "Conv" gets N(0, .02) weights, zero bias
    for key, m in holders:
        torch.nn.init.normal_(m.weight.data, 0.0, 0.02)
        g[key] = m.weight.data.clone()
    g["fc_delta.1.weight"] = fc.weight.data.clone()
    g["fc_delta.1.bias"] = fc.bias.data.clone()
    d: Dict[str, torch.Tensor] = {}
    for idx, conv in d_mods:
        torch.nn.init.normal_(conv.weight.data, 0.0, 0.02)  # aliases weight_orig under legacy SN
        torch.nn.init.constant_(conv.bias.data, 0.0)
        if spectral_norm:
            d[f"model.{idx}.bias"] = conv.bias.data.clone()
            d[f"model.{idx}.weight_orig"] = conv.weight_orig.data.clone()
            d[f"model.{idx}.weight_u"] = conv.weight_u.data.clone()
            d[f"model.{idx}.weight_v"] = conv.weight_v.data.clone()
        else:
            d[f"model.{idx}.weight"] = conv.weight.data.clone()
            d[f"model.{idx}.bias"] = conv.bias.data.clone()
    torch.nn.init.normal_(head.weight.data, 0.0, 0.02)
    d["model.11.weight"] = head.weight.data.clone()
    g = {k: v.to(dtype) for k, v in g.items()}
    d = {k: v.to(dtype) for k, v in d.items()}
    return {"G": g, "D": d, "optG": new_adam_state(g_param_names(g), g), "optD": new_adam_state(d_param_names(d), d)}


def g_param_names(g) -> List[str]:
    return [k for k in g]


def d_param_names(d) -> List[str]:
    return [k for k in d if not (k.endswith("weight_u") or k.endswith("weight_v"))]


def new_adam_state(names, params):
    return {"step": 0, "m": {k: torch.zeros_like(params[k]) for k in names},
            "v": {k: torch.zeros_like(params[k]) for k in names}}


# ----------------------------------------------------------------------------- bf16 storage emulation
# The B200 path stores tensor-core weights, conv outputs and activations in bf16 (fp32 accumulate).  Rounding a
# pre-activation moves a few elements across the ReLU / LeakyReLU kink, and because the reference's losses see the
# trunk only through InstanceNorm -> ReLU -> mean (a nearly scale/shift-invariant statistic), those few flips are a
# 10-20 % relative-L2 change of G's weight gradients at random init - for ANY bf16 implementation.
# `emulate_bf16()` applies the same forward roundings (straight-through in backward) so that gradient parity can be
# checked against "the reference's arithmetic on bf16 storage"; the default (off) is the reference's fp32 arithmetic.
_EMULATE = {"on": False}


class emulate_bf16:
    def __enter__(self):
        self.prev = _EMULATE["on"]; _EMULATE["on"] = True

    def __exit__(self, *a):
        _EMULATE["on"] = self.prev


def _r(t):
    if not _EMULATE["on"]:
        return t
    return t + (t.to(torch.bfloat16).to(t.dtype) - t).detach()


# ----------------------------------------------------------------------------- generator
def _inorm(x):
    return F.instance_norm(x, eps=1e-5)


def g_down_block(w, h, norm, mask=None):
    """UNetDown (cgan/models.py:54-66): conv k4s2p1 -> [IN] -> LeakyReLU(0.2) -> [dropout mask x2]."""
    h = F.conv2d(_r(h), _r(w), None, 2, 1)  # every conv (first layer included) feeds bf16 operands to the tensor cores
    if norm:
        h = _inorm(_r(h))
    h = F.leaky_relu(h, 0.2)
    if mask is not None:
        h = h * mask * 2.0
    return _r(h)


def g_up_block(w, h, skip, mask=None):
    """UNetUp (cgan/models.py:68-83): convT k4s2p1 -> IN -> ReLU -> [dropout] -> cat((out, skip), 1)."""
    h = _r(F.conv_transpose2d(h, _r(w), None, 2, 1))
    h = F.relu(_inorm(h))
    if mask is not None:
        h = h * mask * 2.0
    return torch.cat((_r(h), skip), 1)


def g_tail(w_up4, fc_w, fc_b, h, delta_scale, taps=None):
    """up4 + avg-pool + Linear(64,4) + Tanh + scale (cgan/models.py:112-123,139-141)."""
    h = F.relu(_inorm(_r(F.conv_transpose2d(h, _r(w_up4), None, 2, 1))))
    if taps is not None:
        taps["up4"] = h
    pooled = h.mean(dim=(2, 3))
    return torch.tanh(F.linear(pooled, fc_w, fc_b)) * delta_scale


def g_forward(g, x, delta_scale: float = 0.3, train: bool = True,
              masks: Optional[Sequence[torch.Tensor]] = None, taps: Optional[dict] = None):
    """masks: three {0,1} tensors for the dropout sites (down4, up1, up2) when train=True."""
    skips = []
    h = x
    mi = 0
    for name, _, _, norm, drop in G_DOWN:
        m = None
        if drop and train:
            m = masks[mi]; mi += 1
        h = g_down_block(g[name + ".model.0.weight"], h, norm, m)
        if taps is not None:
            taps[name] = h
        skips.append(h)
    d1, d2, d3, d4 = skips
    h = d4
    for (name, _, _, drop), skip in zip(G_UP, (d3, d2, d1)):
        m = None
        if drop and train:
            m = masks[mi]; mi += 1
        h = g_up_block(g[name + ".model.0.weight"], h, skip, m)
        if taps is not None:
            taps[name] = h
    return g_tail(g["up4.0.weight"], g["fc_delta.1.weight"], g["fc_delta.1.bias"], h, delta_scale, taps)


SIMPLE_CONVS = [0, 3, 7, 10, 14, 17, 21, 24]  # indices of the Conv2d entries of GeneratorSimpleRegressor.features


def simple_forward(g, x, delta_scale: float = 0.3, masks=None):
    """GeneratorSimpleRegressor (cgan/models.py:162-216): [conv3x3 + IN + ReLU] x 2 + maxpool, four times; avg-pool ->
    Linear 512-256-64-4 (ReLU + Dropout(0.5) between) -> tanh -> * delta_scale.  g: its state dict (differentiable when its
    tensors require grad).  masks: None = eval mode, else the two keep-masks (B,256), (B,64) of the train-mode dropouts."""
    h = x
    for i, idx in enumerate(SIMPLE_CONVS):
        h = F.conv2d(_r(h), _r(g[f"features.{idx}.weight"]), g[f"features.{idx}.bias"], 1, 1)
        h = _r(F.relu(_inorm(_r(h))))
        if i % 2 == 1:
            h = F.max_pool2d(h, 2, 2)
    f = h.mean(dim=(2, 3))
    h1 = F.relu(F.linear(f, g["regressor.2.weight"], g["regressor.2.bias"]))
    if masks is not None:
        h1 = h1 * masks[0] * 2.0
    h2 = F.relu(F.linear(h1, g["regressor.5.weight"], g["regressor.5.bias"]))
    if masks is not None:
        h2 = h2 * masks[1] * 2.0
    return torch.tanh(F.linear(h2, g["regressor.8.weight"], g["regressor.8.bias"])) * delta_scale


def g_mask_shapes(B: int, S: int):
    return [(B, 512, S // 16, S // 16), (B, 256, S // 8, S // 8), (B, 128, S // 4, S // 4)]


# ----------------------------------------------------------------------------- discriminator
def sn_weight(d, idx: int, train: bool):
    """Legacy spectral-norm hook: one power iteration (train only, in place on the buffers), W/sigma."""
    w = d[f"model.{idx}.weight_orig"]
    u, v = d[f"model.{idx}.weight_u"], d[f"model.{idx}.weight_v"]
    wm = w.reshape(w.shape[0], -1)
    if train:
        with torch.no_grad():
            v.copy_(F.normalize(torch.mv(wm.t(), u), dim=0, eps=1e-12))
            u.copy_(F.normalize(torch.mv(wm, v), dim=0, eps=1e-12))
    uc, vc = u.clone(), v.clone()
    sigma = torch.dot(uc, torch.mv(wm, vc))
    # bf16-storage emulation (no-op otherwise): the tensor-core operand is the bf16 master weight, 1/sigma is applied to
    # the fp32 accumulators (the weights are packed once per optimizer step, sigma changes every forward)
    return _r(w) / sigma


def d_forward(d, pred, other, train: bool = True, taps: Optional[dict] = None):
    sn = "model.0.weight_orig" in d
    h = torch.cat([pred, other], 1)
    for idx, _, _, norm in D_BLOCKS:
        w = sn_weight(d, idx, train) if sn else _r(d[f"model.{idx}.weight"])
        h = F.conv2d(_r(h), w, d[f"model.{idx}.bias"], 2, 1)
        if norm:
            h = _inorm(_r(h))
        h = _r(F.leaky_relu(h, 0.2))
        if taps is not None:
            taps[f"model.{idx}"] = h
    return F.conv2d(h, d["model.11.weight"], None, 1, 1)


# ----------------------------------------------------------------------------- losses
def smooth_clamp(x, lo, hi, temperature=0.5):
    return lo + (hi - lo) * torch.sigmoid((x - (lo + hi) / 2) / temperature)


def apply_delta_to_bbox(bbox, delta, training=True):
    r = 1.5
    dc = smooth_clamp(delta, -r, r) if training else torch.clamp(delta, -r, r)
    cx = bbox[:, 0] + dc[:, 0] * bbox[:, 2]
    cy = bbox[:, 1] + dc[:, 1] * bbox[:, 3]
    w = bbox[:, 2] * torch.exp(torch.clamp(dc[:, 2], -1.0, 1.0))
    h = bbox[:, 3] * torch.exp(torch.clamp(dc[:, 3], -1.0, 1.0))
    if training:
        cx, cy = smooth_clamp(cx, 0.05, 0.95), smooth_clamp(cy, 0.05, 0.95)
        w, h = smooth_clamp(w, 0.02, 0.8), smooth_clamp(h, 0.02, 0.8)
    else:
        cx, cy = cx.clamp(0.05, 0.95), cy.clamp(0.05, 0.95)
        w, h = w.clamp(0.02, 0.8), h.clamp(0.02, 0.8)
    return torch.stack([cx, cy, w, h], -1)


def _corners(b):
    return b[:, 0] - b[:, 2] / 2, b[:, 1] - b[:, 3] / 2, b[:, 0] + b[:, 2] / 2, b[:, 1] + b[:, 3] / 2


def iou_metric(p, t, eps=1e-6):
    px1, py1, px2, py2 = _corners(p)
    tx1, ty1, tx2, ty2 = _corners(t)
    iw = (torch.min(px2, tx2) - torch.max(px1, tx1)).clamp(min=0)
    ih = (torch.min(py2, ty2) - torch.max(py1, ty1)).clamp(min=0)
    inter = iw * ih
    union = (px2 - px1) * (py2 - py1) + (tx2 - tx1) * (ty2 - ty1) - inter
    return inter / (union + eps)


def eiou_loss(p, t, eps=1e-6):
    px1, py1, px2, py2 = _corners(p)
    tx1, ty1, tx2, ty2 = _corners(t)
    iou = iou_metric(p, t, eps)
    ew = torch.max(px2, tx2) - torch.min(px1, tx1)
    eh = torch.max(py2, ty2) - torch.min(py1, ty1)
    cdist = (p[:, 0] - t[:, 0]) ** 2 + (p[:, 1] - t[:, 1]) ** 2
    e = iou - cdist / (ew ** 2 + eh ** 2 + eps) - (p[:, 2] - t[:, 2]) ** 2 / (ew ** 2 + eps) \
        - (p[:, 3] - t[:, 3]) ** 2 / (eh ** 2 + eps)
    return 1 - e.mean()


def gradient_penalty(d, real, fake, alpha, train=True):
    """alpha: (B,1,1,1) injected U[0,1) draw.  Returns the GP scalar with a graph (create_graph)."""
    B = real[0].shape[0]
    ip = (alpha * real[0] + (1 - alpha) * fake[0]).detach().requires_grad_(True)
    io = (alpha * real[1] + (1 - alpha) * fake[1]).detach().requires_grad_(True)
    out = d_forward(d, ip, io, train)
    gp_, go_ = torch.autograd.grad(out, [ip, io], torch.ones_like(out), create_graph=True, retain_graph=True)
    norm = torch.sqrt((gp_.reshape(B, -1) ** 2).sum(1) + (go_.reshape(B, -1) ** 2).sum(1) + 1e-12)
    return ((norm - 1) ** 2).mean()


# ----------------------------------------------------------------------------- optimizer
def clip_and_adam(params, grads, opt, names, lr=2e-4, beta1=0.5, beta2=0.999, eps=1e-8, max_norm=1.0):
    """clip_grad_norm_(max_norm) then one torch-default Adam step, in place.  Returns the pre-clip norm."""
    total = torch.sqrt(sum((grads[k].double() ** 2).sum() for k in names)).to(grads[names[0]].dtype)
    coef = min(float(max_norm / (total + 1e-6)), 1.0)
    opt["step"] += 1
    t = opt["step"]
    bc1, bc2 = 1 - beta1 ** t, 1 - beta2 ** t
    with torch.no_grad():
        for k in names:
            gk = grads[k] * coef
            opt["m"][k].mul_(beta1).add_(gk, alpha=1 - beta1)
            opt["v"][k].mul_(beta2).addcmul_(gk, gk, value=1 - beta2)
            denom = (opt["v"][k].sqrt() / math.sqrt(bc2)).add_(eps)
            params[k].addcdiv_(opt["m"][k], denom, value=-lr / bc1)
    return float(total)


# ----------------------------------------------------------------------------- the step
def critic_grads(d, pred, gt, refined, alpha, lambda_gp=1.0):
    """One critic loss + grads (no update).  Returns (scalars, grads dict)."""
    names = d_param_names(d)
    leaves = {k: d[k].detach().requires_grad_(True) for k in names}
    dd = dict(d); dd.update(leaves)
    real = d_forward(dd, pred, gt, True)
    fake = d_forward(dd, pred, refined, True)
    gp = gradient_penalty(dd, (pred, gt), (pred, refined), alpha, True)
    wd = real.mean() - fake.mean()
    loss = -wd + lambda_gp * gp
    gs = torch.autograd.grad(loss, [leaves[k] for k in names], allow_unused=True)
    grads = {k: (torch.zeros_like(d[k]) if gi is None else gi) for k, gi in zip(names, gs)}
    return {"d_loss": float(loss.detach()), "gp": float(gp.detach()), "wd": float(wd.detach())}, grads


def generator_grads(g, d, pred, refined, pred_box, delta_true, masks, delta_scale=0.3, lambda_iou=1.0):
    names = g_param_names(g)
    leaves = {k: g[k].detach().requires_grad_(True) for k in names}
    delta = g_forward(leaves, pred, delta_scale, True, masks)
    cal = apply_delta_to_bbox(pred_box, delta, True)
    gtb = apply_delta_to_bbox(pred_box, delta_true, True)
    liou = eiou_loss(cal, gtb)
    with torch.no_grad():  # D still runs in train mode: advances weight_u / weight_v
        lw = -d_forward(d, pred, refined, True).mean()
    loss = lambda_iou * liou + lw
    gs = torch.autograd.grad(lambda_iou * liou, [leaves[k] for k in names])
    return {"loss_G": float(loss), "loss_iou": float(liou), "loss_wgan": float(lw)}, dict(zip(names, gs)), delta.detach()


def train_step(state, batch, masks_per_gcall, alphas, *, n_critic=2, lambda_gp=1.0, lambda_iou=1.0,
               delta_scale=0.3, lr=2e-4, beta1=0.5, beta2=0.999):
    """One batch of the reference loop: n_critic critic updates (each with a train-mode,
    no-grad G forward that only consumes dropout draws), then one generator update.

    batch: dict(pred, gt, refined, pred_box, delta_true); masks_per_gcall: n_critic+1 mask triples;
    alphas: n_critic tensors (B,1,1,1).  State is updated in place.  Returns the logged scalars."""
    g, d = state["G"], state["D"]
    pred, gt, refined = batch["pred"], batch["gt"], batch["refined"]
    out = {"d_loss": 0.0, "gp": 0.0, "wd": 0.0}
    for c in range(n_critic):
        with torch.no_grad():
            g_forward(g, pred, delta_scale, True, masks_per_gcall[c])  # reference :311; result unused here
        sc, grads = critic_grads(d, pred, gt, refined, alphas[c], lambda_gp)
        out["d_gradnorm_%d" % c] = clip_and_adam(d, grads, state["optD"], d_param_names(d), lr, beta1, beta2)
        for k in ("d_loss", "gp", "wd"):
            out[k] += sc[k] / n_critic
    sc, grads, delta = generator_grads(g, d, pred, refined, batch["pred_box"], batch["delta_true"],
                                       masks_per_gcall[n_critic], delta_scale, lambda_iou)
    out["g_gradnorm"] = clip_and_adam(g, grads, state["optG"], g_param_names(g), lr, beta1, beta2)
    out.update(sc)
    out["delta"] = delta
    return out


# ----------------------------------------------------------------------------- synthetic data
def make_batch(B: int, S: int, step: int = 0, dtype=torch.float32):
    gen = torch.Generator("cpu").manual_seed(1234 + step)
    u = lambda *s: torch.rand(*s, generator=gen)
    pred, gt, refined = (u(B, 3, S, S) * 2 - 1 for _ in range(3))
    box = torch.cat([u(B, 2) * 0.6 + 0.2, u(B, 2) * 0.3 + 0.1], 1)
    dt = (torch.randn(B, 4, generator=gen) * torch.tensor([0.17, 0.23, 0.32, 0.32])).clamp(-1.5, 1.5)
    masks = [[(u(*s) < 0.5).float() for s in g_mask_shapes(B, S)] for _ in range(8)]
    alphas = [u(B, 1, 1, 1) for _ in range(8)]
    b = {"pred": pred, "gt": gt, "refined": refined, "pred_box": box, "delta_true": dt}
    b = {k: v.to(dtype) for k, v in b.items()}
    return b, [[m.to(dtype) for m in t] for t in masks], [a.to(dtype) for a in alphas]
