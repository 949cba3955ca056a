"""torch.autograd Functions over the libb2c.so kernels (the "PyTorch custom ops -> C-ABI -> sm_100a kernels" layer).

Activations are logical-NCHW / physical-NHWC (torch.channels_last) bf16 CUDA tensors; parameters and losses fp32.
The Functions are closed under differentiation where the reference needs it: compute_gradient_penalty
(cgan/losses.py:213-220) differentiates THROUGH the first backward of D, so conv-dgrad, LeakyReLU-backward and
InstanceNorm-backward are themselves Functions with a backward.  Everything G-side is
first-order only.  No torch conv / norm / matmul is called anywhere in this file.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

BF16 = torch.bfloat16
CL = torch.channels_last


def cl_empty(N, C, H, W, dtype=BF16, device=None):
    return torch.empty((N, C, H, W), dtype=dtype, device=device or torch.device("cuda"), memory_format=CL)


def as_nhwc(t):
    """-> (tensor, ld): t usable by the kernels as NHWC bf16 with pixel stride ld (copies only if it must)."""
    if t.dtype != BF16:
        t = t.to(BF16)
    N, C, H, W = t.shape
    s = t.stride()
    ld = s[3] if W > 1 else (s[2] if H > 1 else C)
    ok = (s[1] == 1 or C == 1) and (W == 1 or s[3] == ld) and (H == 1 or s[2] == W * ld) and \
         (N == 1 or s[0] == H * W * ld) and ld % 8 == 0 and ld >= C and t.data_ptr() % 16 == 0
    if not ok:
        t = t.contiguous(memory_format=CL)
        if t.stride(1) != 1:  # degenerate shapes where channels_last is ambiguous
            t = t.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)
        ld = C
    return t, ld


def dense_nhwc(t):
    t, ld = as_nhwc(t)
    if ld != t.shape[1]:
        t = t.contiguous(memory_format=CL)
    return t


class WeightCtx:
    """One forward instance of a conv weight: the fp32 master tensor, the spectral-norm (u, v, sigma) snapshot used by
    THAT forward (torch spectral_norm.py:97-112 advances u/v every train-mode forward) and lazily packed bf16 forms."""

    def __init__(self, weight, sig=None, u=None, v=None):
        self.weight, self.sig, self.u, self.v = weight, sig, u, v
        self._c = self._t = None
        self.want_stats = False  # ask the forward conv for the InstanceNorm statistics of its output ...
        self._stats = None       # ... which take_stats() hands to INAct / INActPoolFC

    def take_stats(self):
        zs, self._stats = self._stats, None
        return zs

    @property
    def inv_sigma(self):
        return None if self.sig is None else self.sig[1:]

    def pack_c(self):
        if self._c is None:
            d0, d1 = self.weight.shape[:2]
            self._c = torch.empty(d0, 16 * d1, dtype=BF16, device=self.weight.device)
            _lib.call("b2c_pack_w_c", self.weight.detach(), self.inv_sigma, self._c, d0, d1)
        return self._c

    def pack_t(self):
        if self._t is None:
            d0, d1 = self.weight.shape[:2]
            self._t = torch.empty(4, d1, 4 * d0, dtype=BF16, device=self.weight.device)
            _lib.call("b2c_pack_w_t", self.weight.detach(), self.inv_sigma, self._t, d0, d1)
        return self._t

    def finish_wgrad(self, dwe):
        """dL/dW_eff -> dL/dW_orig through W_eff = W / sigma(u, v)   (SURVEY.md App. B)."""
        if self.sig is None:
            return dwe
        R = self.weight.shape[0]
        K = self.weight.numel() // R
        dw = torch.empty_like(dwe)
        ws = torch.empty(592, dtype=torch.float32, device=dwe.device)
        _lib.call("b2c_sn_bwd", dwe, self.weight.detach(), self.u, self.v, self.sig, ws, dw, R, K, 0)
        return dw


# ------------------------------------------------------------------------------------------------ raw kernel helpers
def _stats_buf(N, C, device):
    return torch.empty(N, C, 2, dtype=torch.float32, device=device)


def _conv_raw(x, wctx, bias, want_stats=False):
    x = dense_nhwc(x)
    N, Cin, H, W = x.shape
    Cout = wctx.weight.shape[0]
    y = cl_empty(N, Cout, H // 2, W // 2, device=x.device)
    zs = _stats_buf(N, Cout, x.device) if want_stats else None
    _lib.call("b2c_conv4s2_bf16", x, Cin, wctx.pack_c(), bias, y, Cout, zs, N, H, W, Cin, Cout)
    if want_stats:
        wctx._stats = zs
    return y


def _convt_raw(x, wpack_t, Cout, wctx=None, act=None, slope=1.0):
    x, ld = as_nhwc(x)
    N, Cin, H, W = x.shape
    y = cl_empty(N, Cout, 2 * H, 2 * W, device=x.device)
    zs = _stats_buf(N, Cout, x.device) if (wctx is not None and wctx.want_stats) else None
    _lib.call("b2c_convt4s2_bf16", x, ld, wpack_t, None, y, Cout, zs, act, float(slope), N, H, W, Cin, Cout)
    if zs is not None:
        wctx._stats = zs
    return y


def _wgrad_raw(small, large):
    """-> fp32 [Cs][Cl][4][4] = sum small[n,sy,sx,a] large[n,2sy+ky-1,2sx+kx-1,b]."""
    small, lds = as_nhwc(small)
    large = dense_nhwc(large)
    N, Cs, Hs, Ws = small.shape
    Cl = large.shape[1]
    ws = torch.empty(16, Cs, Cl, dtype=torch.float32, device=small.device)
    dw = torch.empty(Cs, Cl, 4, 4, dtype=torch.float32, device=small.device)
    _lib.call("b2c_wgrad4s2_bf16", small, lds, large, Cl, ws, dw, 0, N, Hs, Ws, Cs, Cl, _lib.sm_count())
    return dw


# ------------------------------------------------------------------------------------------------ conv k4 s2 p1
class Conv4s2(torch.autograd.Function):
    """y = conv2d(x, W_eff, bias, stride 2, pad 1).  Replaces nn.Conv2d at cgan/models.py:57,236 (Cin >= 64)."""

    @staticmethod
    def forward(ctx, x, weight, bias, wctx, bias_feeds_norm, in_lrelu_slope=None):
        """in_lrelu_slope: x is the output of a LeakyReLU(slope) whose producer leaves its activation backward to this
        conv (FirstConv(premasked=True)): the returned input gradient is already multiplied by LeakyReLU'."""
        ctx.wctx, ctx.has_bias, ctx.bias_feeds_norm = wctx, bias is not None, bias_feeds_norm
        ctx.in_lrelu_slope = in_lrelu_slope
        ctx.save_for_backward(x)
        return _conv_raw(x, wctx, bias, wctx.want_stats)

    @staticmethod
    def backward(ctx, gy):
        (x,) = ctx.saved_tensors
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            if ctx.in_lrelu_slope is None:
                gx = Conv4s2Dgrad.apply(gy, ctx.wctx.weight, ctx.wctx, x.shape[2], x.shape[3])
            elif x.shape[1] == 64:  # fused into the dgrad epilogue (Cout = 64 kernel)
                gx = Conv4s2Dgrad.apply(gy, ctx.wctx.weight, ctx.wctx, x.shape[2], x.shape[3], x, ctx.in_lrelu_slope)
            else:
                gx = LReluBwd.apply(Conv4s2Dgrad.apply(gy, ctx.wctx.weight, ctx.wctx, x.shape[2], x.shape[3]), x,
                                    ctx.in_lrelu_slope)
        if ctx.needs_input_grad[1]:
            with torch.no_grad():
                gw = ctx.wctx.finish_wgrad(_wgrad_raw(gy.detach(), x))
        if ctx.has_bias and ctx.needs_input_grad[2]:
            with torch.no_grad():
                if ctx.bias_feeds_norm:  # an affine-less InstanceNorm follows: d/d bias == 0 exactly
                    gb = torch.zeros(gy.shape[1], dtype=torch.float32, device=gy.device)
                else:
                    gb = channel_sum(gy.detach())
        return gx, gw, gb, None, None, None


class Conv4s2Dgrad(torch.autograd.Function):
    """dx = dgrad(gy, W_eff) = convT(gy, W_eff).  Differentiable: d/d gy = conv(h, W_eff), d/d W_eff = wgrad(gy, h)."""

    @staticmethod
    def forward(ctx, gy, weight, wctx, H, W, act=None, slope=1.0):
        """act / slope: out = dgrad * LeakyReLU'(act) (the activation backward of the layer below, fused)."""
        ctx.wctx, ctx.slope, ctx.has_act = wctx, slope, act is not None
        Cin = wctx.weight.shape[1]
        if act is not None:
            act = dense_nhwc(act)
            ctx.save_for_backward(gy, act)
        else:
            ctx.save_for_backward(gy)
        return _convt_raw(gy, wctx.pack_t(), Cin, None, act, slope)

    @staticmethod
    def backward(ctx, h):
        gy = ctx.saved_tensors[0]
        if ctx.has_act:  # out = m * dgrad(gy): the cotangent reaches dgrad as m * h
            h = LReluBwd.apply(h, ctx.saved_tensors[1], ctx.slope)
        g_gy = gw = None
        if ctx.needs_input_grad[0]:
            g_gy = Conv4s2.apply(h, ctx.wctx.weight, None, ctx.wctx, False)
        if ctx.needs_input_grad[1]:
            with torch.no_grad():
                gw = ctx.wctx.finish_wgrad(_wgrad_raw(gy.detach(), h.detach()))
        return g_gy, gw, None, None, None, None, None


class ConvT4s2(torch.autograd.Function):
    """y = conv_transpose2d(x, W, stride 2, pad 1), W: (Cin, Cout, 4, 4).  Replaces cgan/models.py:72,113.  First order."""

    @staticmethod
    def forward(ctx, x, weight, wctx):
        ctx.wctx = wctx
        ctx.save_for_backward(x)
        return _convt_raw(x, wctx.pack_t(), weight.shape[1], wctx)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gy):
        (x,) = ctx.saved_tensors
        gx = gw = None
        if ctx.needs_input_grad[0]:
            gx = _conv_raw(gy, ctx.wctx, None)  # C-form pack of (Cin, Cout, 4, 4): output channels = Cin
        if ctx.needs_input_grad[1]:
            gw = _wgrad_raw(x, gy)
        return gx, gw, None


def channel_sum(g):
    """sum over (n, h, w) of an NHWC bf16 tensor -> fp32 [C] (bias gradient)."""
    g, ld = as_nhwc(g)
    N, C, H, W = g.shape
    zs = torch.empty(N, C, 2, dtype=torch.float32, device=g.device)
    _lib.call("b2c_in_stats", g, ld, zs, N, H * W, C)
    out = torch.zeros(C, dtype=torch.float32, device=g.device)
    # zs[:, :, 0] summed over n: one strided rowsum per channel is wasteful; transpose-free trick: rowsum over (N) rows
    tmp = zs[:, :, 0].t().contiguous()  # [C][N], tiny
    if N % 4 == 0:
        _lib.call("b2c_rowsum", tmp, N, C, 0, 1.0, out)
    else:
        pad = torch.zeros(C, (N + 3) // 4 * 4, dtype=torch.float32, device=g.device)
        pad[:, :N] = tmp
        _lib.call("b2c_rowsum", pad, pad.shape[1], C, 0, 1.0, out)
    return out


# ------------------------------------------------------------------------------------------------ first layer (Cin 3 / 6)
def _f32c(t):
    return None if t is None else t.detach().to(torch.float32).contiguous()


def _first_fprop_raw(a0, a1, wctx, bias, slope):
    a0, a1 = _f32c(a0), _f32c(a1)
    N, _, H, W = a0.shape
    y = cl_empty(N, 64, H // 2, W // 2, device=a0.device)
    _lib.call("b2c_first_fprop", a0, a1, None, None, None, wctx.weight.detach(), wctx.inv_sigma, bias, y, N, H, W, slope)
    return y


def _first_wgrad_raw(a0, a1, gz, wctx, want_bias):
    a0, a1 = _f32c(a0), _f32c(a1)
    gz = dense_nhwc(gz.detach())
    N, _, H, W = a0.shape
    dw = torch.zeros_like(wctx.weight, dtype=torch.float32)
    db = torch.zeros(64, dtype=torch.float32, device=dw.device) if want_bias else None
    _lib.call("b2c_first_wgrad", a0, a1, None, None, None, gz, dw, db, N, H, W)
    return dw, db


class FirstConv(torch.autograd.Function):
    """act = LeakyReLU_slope(conv2d(cat(a0, a1), W_eff, bias, 2, 1)) on NCHW fp32 inputs (cgan/models.py:103,245,257)."""

    @staticmethod
    def forward(ctx, a0, a1, weight, bias, wctx, slope, premasked=False):
        """premasked: the (only) consumer of the output is a Conv4s2(in_lrelu_slope=slope), which returns its input
        gradient already multiplied by LeakyReLU' - no separate activation-backward pass here."""
        y = _first_fprop_raw(a0, a1, wctx, bias, slope)
        ctx.premasked = premasked
        ctx.wctx, ctx.slope, ctx.has_a1, ctx.has_bias = wctx, slope, a1 is not None, bias is not None
        ctx.save_for_backward(a0, a1, y)
        return y

    @staticmethod
    def backward(ctx, g):
        a0, a1, y = ctx.saved_tensors
        gz = LReluBwd.apply(g, y, ctx.slope) if (ctx.slope != 1.0 and not ctx.premasked) else g
        g0 = g1 = gw = gb = None
        if ctx.needs_input_grad[0] or (ctx.has_a1 and ctx.needs_input_grad[1]):
            g0, g1 = FirstDgrad.apply(gz, ctx.wctx.weight, ctx.wctx, ctx.has_a1)
            if not ctx.has_a1:
                g1 = None
        if ctx.needs_input_grad[2] or (ctx.has_bias and ctx.needs_input_grad[3]):
            with torch.no_grad():
                dwe, gb = _first_wgrad_raw(a0, a1, gz, ctx.wctx, ctx.has_bias)
                gw = ctx.wctx.finish_wgrad(dwe)
        return g0, g1, gw, gb, None, None, None


class FirstDgrad(torch.autograd.Function):
    """(gx0, gx1) = dgrad of the first conv, fp32 NCHW.  Differentiable for the gradient penalty."""

    @staticmethod
    def forward(ctx, gz, weight, wctx, has_a1):
        gzd = dense_nhwc(gz)
        N, _, Ho, Wo = gzd.shape
        ctx.wctx, ctx.has_a1 = wctx, has_a1
        ctx.save_for_backward(gzd)
        g0 = torch.empty(N, 3, 2 * Ho, 2 * Wo, dtype=torch.float32, device=gzd.device)
        g1 = torch.empty_like(g0) if has_a1 else None
        ws = torch.empty(32 * 576, dtype=BF16, device=gzd.device)
        _lib.call("b2c_first_dgrad", gzd, wctx.weight.detach(), wctx.inv_sigma, g0, g1, ws, N, 2 * Ho, 2 * Wo)
        if g1 is None:
            g1 = g0.new_zeros(())
            ctx.mark_non_differentiable(g1)
        return g0, g1

    @staticmethod
    def backward(ctx, h0, h1):
        (gz,) = ctx.saved_tensors
        h1 = h1 if ctx.has_a1 else None
        g_gz = gw = None
        if ctx.needs_input_grad[0]:
            g_gz = FirstConv.apply(h0, h1, ctx.wctx.weight, None, ctx.wctx, 1.0)
        if ctx.needs_input_grad[1]:
            with torch.no_grad():
                dwe, _ = _first_wgrad_raw(h0, h1, gz, ctx.wctx, False)
                gw = ctx.wctx.finish_wgrad(dwe)
        return g_gz, gw, None, None


class LReluBwd(torch.autograd.Function):
    """out = g * (y > 0 ? 1 : slope): linear in g, so its own backward is the same op (LeakyReLU'' = 0 a.e.)."""

    @staticmethod
    def forward(ctx, g, y, slope):
        g, y = dense_nhwc(g), dense_nhwc(y)
        ctx.slope = slope
        ctx.save_for_backward(y)
        out = torch.empty_like(g)
        _lib.call("b2c_lrelu_bwd", g, y, out, g.numel(), slope)
        return out

    @staticmethod
    def backward(ctx, h):
        (y,) = ctx.saved_tensors
        return LReluBwd.apply(h, y, ctx.slope), None, None


# ------------------------------------------------------------------------------------------------ InstanceNorm + act (+dropout, +concat)
def _in_stats(z, ld):
    N, C, H, W = z.shape
    zs = torch.empty(N, C, 2, dtype=torch.float32, device=z.device)
    _lib.call("b2c_in_stats", z, ld, zs, N, H * W, C)
    return zs


class INAct(torch.autograd.Function):
    """dropout(act(InstanceNorm2d(z))) [cat skip].  cgan/models.py:59-62,73-77,82,114,241-242.  mask: dense u8 NHWC
    keep-mask or None.  With `skip`, the result is cat((out, skip), 1) written in place (no separate concat pass)."""

    @staticmethod
    def forward(ctx, z, slope, mask, skip, zs=None):
        z, ldz = as_nhwc(z)
        N, C, H, W = z.shape
        if zs is None:
            zs = _in_stats(z, ldz)
        Cs = 0 if skip is None else skip.shape[1]
        out = cl_empty(N, C + Cs, H, W, device=z.device)
        _lib.call("b2c_in_act_fwd", z, ldz, zs, mask, slope, out, C + Cs, None, 0, N, H * W, C)
        if skip is not None:
            s, lds = as_nhwc(skip)
            _lib.call("b2c_copy_strided", s, lds, out[:, C:], C + Cs, N * H * W, Cs)
        ctx.slope, ctx.mask, ctx.C, ctx.has_skip = slope, mask, C, skip is not None
        ctx.save_for_backward(z, zs)
        return out

    @staticmethod
    def backward(ctx, g):
        z, zs = ctx.saved_tensors
        gz = gskip = None
        if ctx.needs_input_grad[0]:
            gz = INActBwd.apply(g[:, :ctx.C] if ctx.has_skip else g, z, zs, ctx.mask, ctx.slope)
        if ctx.has_skip and ctx.needs_input_grad[3]:
            gskip = g[:, ctx.C:]
        return gz, None, None, gskip, None


class SkipTap(torch.autograd.Function):
    """(x, x) for a tensor with two consumers (U-Net skip, cgan/models.py:224-230): the two incoming gradients - one dense,
    one a channel slice of the concat gradient - are summed by one strided kernel instead of autograd's generic add."""

    @staticmethod
    def forward(ctx, x):
        return x.view_as(x), x.view_as(x)

    @staticmethod
    def backward(ctx, ga, gb):
        if ga is None or gb is None:
            return ga if gb is None else gb
        return AddStrided.apply(ga, gb)


class AddStrided(torch.autograd.Function):
    """a + b on NHWC bf16 tensors that may be channel slices; linear, so its backward passes g to both."""

    @staticmethod
    def forward(ctx, a, b):
        a, lda = as_nhwc(a)
        b, ldb = as_nhwc(b)
        N, C, H, W = a.shape
        out = cl_empty(N, C, H, W, device=a.device)
        _lib.call("b2c_add_strided", a, lda, b, ldb, out, C, N * H * W, C)
        return out

    @staticmethod
    def backward(ctx, g):
        return g, g


class INActBwd(torch.autograd.Function):
    """g_z = d/dz of dropout(act(IN(z))) applied to g.  Its backward is the InstanceNorm backward-of-backward."""

    @staticmethod
    def forward(ctx, g, z, zs, mask, slope):
        g, ldg = as_nhwc(g)
        z, ldz = as_nhwc(z)
        N, C, H, W = z.shape
        ws = torch.empty(N, C, 2, dtype=torch.float32, device=z.device)
        gz = cl_empty(N, C, H, W, device=z.device)
        _lib.call("b2c_in_act_bwd", g, ldg, None, z, ldz, zs, mask, slope, ws, gz, C, N, H * W, C)
        ctx.mask, ctx.slope = mask, slope
        ctx.save_for_backward(g, z, zs)
        return gz

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, h):
        g, z, zs = ctx.saved_tensors
        g, ldg = as_nhwc(g)
        z, ldz = as_nhwc(z)
        h, ldh = as_nhwc(h)
        N, C, H, W = z.shape
        ws = torch.empty(N, C, 5, dtype=torch.float32, device=z.device)
        gg, gx = cl_empty(N, C, H, W, device=z.device), cl_empty(N, C, H, W, device=z.device)
        _lib.call("b2c_in_act_bwd2", h, ldh, g, ldg, z, ldz, zs, ctx.mask, ctx.slope, ws, gg, C, gx, C, N, H * W, C)
        return gg, gx, None, None, None


class INActPoolFC(torch.autograd.Function):
    """delta = tanh(Linear(mean_hw(relu(IN(z))))) * scale: G's up4 tail + head (cgan/models.py:114,118-123,139-141)
    without ever materialising the activated (B,64,S,S) tensor."""

    @staticmethod
    def forward(ctx, z, fc_w, fc_b, slope, scale, zs=None):
        z, ldz = as_nhwc(z)
        N, C, H, W = z.shape
        if zs is None:
            zs = _in_stats(z, ldz)
        ps = torch.empty(N, C, dtype=torch.float32, device=z.device)
        _lib.call("b2c_in_act_pool_fwd", z, ldz, zs, slope, ps, N, H * W, C)
        delta = torch.empty(N, 4, dtype=torch.float32, device=z.device)
        th = torch.empty_like(delta)
        _lib.call("b2c_pool_fc_fwd", ps, fc_w.detach(), fc_b.detach(), scale, H * W, delta, th, N, C)
        ctx.slope, ctx.scale = slope, scale
        ctx.save_for_backward(z, zs, ps, th, fc_w)
        return delta

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gd):
        z, zs, ps, th, fc_w = ctx.saved_tensors
        z, ldz = as_nhwc(z)
        N, C, H, W = z.shape
        dW, db = torch.empty_like(fc_w), torch.empty(4, dtype=torch.float32, device=z.device)
        gvec = torch.empty(N, C, dtype=torch.float32, device=z.device)
        _lib.call("b2c_pool_fc_bwd", gd.contiguous(), th, ps, fc_w.detach(), ctx.scale, H * W, dW, db, gvec, N, C)
        ws = torch.empty(N, C, 2, dtype=torch.float32, device=z.device)
        gz = cl_empty(N, C, H, W, device=z.device)
        _lib.call("b2c_in_act_bwd", None, 0, gvec, z, ldz, zs, None, ctx.slope, ws, gz, C, N, H * W, C)
        return gz, dW, db, None, None, None


# ------------------------------------------------------------------------------------------------ PatchGAN head
class HeadConv(torch.autograd.Function):
    """score = conv2d(a, W5, None, stride 1, pad 1), Cout = 1, fp32 out (cgan/models.py:252)."""

    @staticmethod
    def forward(ctx, a, weight):
        a = dense_nhwc(a)
        N, C, h, _ = a.shape
        out = torch.empty(N, 1, h - 1, h - 1, dtype=torch.float32, device=a.device)
        _lib.call("b2c_head_fprop", a, weight.detach(), out, N, h, C)
        ctx.save_for_backward(a, weight)
        return out

    @staticmethod
    def backward(ctx, gy):
        a, weight = ctx.saved_tensors
        ga = gw = None
        if ctx.needs_input_grad[0]:
            ga = HeadDgrad.apply(gy, weight, a.shape[2])
        if ctx.needs_input_grad[1]:
            with torch.no_grad():
                gw = _head_wgrad_raw(a, gy.detach(), weight)
        return ga, gw


def _head_wgrad_raw(a, gy, weight):
    a = dense_nhwc(a)
    N, C, h, _ = a.shape
    dw = torch.zeros_like(weight, dtype=torch.float32)
    _lib.call("b2c_head_wgrad", a, gy.to(torch.float32).contiguous(), dw, N, h, C)
    return dw


class HeadDgrad(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gy, weight, h):
        gy = gy.to(torch.float32).contiguous()
        N, C = gy.shape[0], weight.shape[1]
        ga = cl_empty(N, C, h, h, device=gy.device)
        _lib.call("b2c_head_dgrad", gy, weight.detach(), ga, N, h, C)
        ctx.save_for_backward(gy, weight)
        return ga

    @staticmethod
    def backward(ctx, hh):
        gy, weight = ctx.saved_tensors
        g_gy = gw = None
        if ctx.needs_input_grad[0]:
            g_gy = HeadConv.apply(hh, weight)
        if ctx.needs_input_grad[1]:
            with torch.no_grad():
                gw = _head_wgrad_raw(hh.detach(), gy, weight)
        return g_gy, gw, None


# ------------------------------------------------------------------------------------------------ scalar losses
class MapMean(torch.autograd.Function):
    """mean over every element of an fp32 map (torch.mean at cgan/cgan_train_enhanced.py:327,362)."""

    @staticmethod
    def forward(ctx, x):
        x = x.contiguous()
        ctx.shape = x.shape
        out = torch.zeros(1, dtype=torch.float32, device=x.device)
        _lib.call("b2c_rowsum", x, x.numel(), 1, 0, 1.0 / x.numel(), out)
        return out.reshape(())

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        n = 1
        for s in ctx.shape:
            n *= s
        out = torch.empty(ctx.shape, dtype=torch.float32, device=g.device)
        _lib.call("b2c_fill_scaled", out, n, g.contiguous(), 1.0 / n)
        return out


class GradNormPenalty(torch.autograd.Function):
    """mean_b (sqrt(|g0_b|^2 + |g1_b|^2 + 1e-12) - 1)^2   (cgan/losses.py:223-231).  Backward emits the seed of the
    double backward, (2/B)(1 - 1/r_b) g, in one pass per tensor."""

    @staticmethod
    def forward(ctx, g0, g1):
        g0, g1 = g0.contiguous(), g1.contiguous()
        B = g0.shape[0]
        n = g0.numel() // B
        ss = torch.zeros(B, dtype=torch.float32, device=g0.device)
        _lib.call("b2c_rowsum", g0, n, B, 1, 1.0, ss)
        _lib.call("b2c_rowsum", g1, n, B, 1, 1.0, ss)
        gp = torch.empty(1, dtype=torch.float32, device=g0.device)
        coef = torch.empty(B, dtype=torch.float32, device=g0.device)
        _lib.call("b2c_gp_tail", ss, gp, coef, B)
        ctx.save_for_backward(g0, g1, coef)
        return gp.reshape(())

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gs):
        g0, g1, coef = ctx.saved_tensors
        B = g0.shape[0]
        n = g0.numel() // B
        gs = gs.contiguous()
        h0, h1 = torch.empty_like(g0), torch.empty_like(g1)
        _lib.call("b2c_scale_rows", g0, coef, gs, h0, n, B)
        _lib.call("b2c_scale_rows", g1, coef, gs, h1, n, B)
        return h0, h1


class ApplyDelta(torch.autograd.Function):
    """cgan/losses.py:108-150 (gradient w.r.t. delta only; boxes are data)."""

    @staticmethod
    def forward(ctx, bbox, delta, training):
        bbox, delta = bbox.to(torch.float32).contiguous(), delta.to(torch.float32).contiguous()
        out = torch.empty_like(delta)
        _lib.call("b2c_apply_delta", bbox, delta, int(training), out, None, None, delta.shape[0])
        ctx.training = training
        ctx.save_for_backward(bbox, delta)
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        bbox, delta = ctx.saved_tensors
        gd = torch.empty_like(delta)
        _lib.call("b2c_apply_delta", bbox, delta, int(ctx.training), None, g.contiguous(), gd, delta.shape[0])
        return None, gd, None


class EIoU(torch.autograd.Function):
    """1 - mean(EIoU(pred, target))   (cgan/losses.py:19-73; gradient w.r.t. pred only)."""

    @staticmethod
    def forward(ctx, pred, target, eps):
        pred, target = pred.to(torch.float32).contiguous(), target.to(torch.float32).contiguous()
        loss = torch.empty(1, dtype=torch.float32, device=pred.device)
        _lib.call("b2c_eiou", pred, target, eps, loss, None, None, None, pred.shape[0])
        ctx.eps = eps
        ctx.save_for_backward(pred, target)
        return loss.reshape(())

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        pred, target = ctx.saved_tensors
        gp = torch.empty_like(pred)
        _lib.call("b2c_eiou", pred, target, ctx.eps, None, g.contiguous(), gp, None, pred.shape[0])
        return gp, None, None


def iou_rows(pred, target, eps=1e-6):
    pred, target = pred.to(torch.float32).contiguous(), target.to(torch.float32).contiguous()
    out = torch.empty(pred.shape[0], dtype=torch.float32, device=pred.device)
    _lib.call("b2c_eiou", pred, target, eps, None, None, None, out, pred.shape[0])
    return out


# ------------------------------------------------------------------------------------------------ optimizer
class FusedClipAdam:
    """clip_grad_norm_(max_norm) + Adam in two multi-tensor launches (cgan/cgan_train_enhanced.py:256-257,331-332,368-369).
    Torch-default Adam semantics (eps 1e-8, no weight decay, bias correction), one shared step counter."""

    def __init__(self, params, lr=2e-4, betas=(0.5, 0.999), eps=1e-8, max_norm=1.0):
        self.params = [p for p in params]
        assert 1 <= len(self.params) <= 16
        self.lr, self.betas, self.eps, self.max_norm = lr, betas, eps, max_norm
        self.m = [torch.zeros_like(p, dtype=torch.float32) for p in self.params]
        self.v = [torch.zeros_like(p, dtype=torch.float32) for p in self.params]
        self.step_count = 0
        self.step_base = None  # int64 device scalar added to step_count in the kernel (CUDA-graph replays)
        self.gnorm_sq = torch.zeros(1, dtype=torch.float32, device=self.params[0].device)
        n = len(self.params)
        self._n = (ctypes.c_longlong * n)(*[p.numel() for p in self.params])
        self._arr = lambda ts: (ctypes.c_void_p * n)(*[t.data_ptr() for t in ts])

    def zero_grad(self):
        for p in self.params:
            p.grad = None

    def step(self):
        """Returns the device scalar holding the squared pre-clip gradient norm (no host sync)."""
        n = len(self.params)
        grads = []
        for p in self.params:
            g = p.grad if p.grad is not None else torch.zeros_like(p)
            grads.append(g.to(torch.float32).contiguous())
        self.step_count += 1
        _lib.call("b2c_multi_sumsq", self._arr(grads), self._n, n, self.gnorm_sq)
        _lib.call("b2c_multi_adam", self._arr([p.data for p in self.params]), self._arr(grads), self._arr(self.m),
                  self._arr(self.v), self._n, n, self.gnorm_sq, self.max_norm, self.lr, self.betas[0], self.betas[1],
                  self.eps, self.step_count, self.step_base)
        return self.gnorm_sq
