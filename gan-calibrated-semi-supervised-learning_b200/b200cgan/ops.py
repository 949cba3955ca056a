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


# ------------------------------------------------------------------------------------------------ backward switches
class _Flags:
    """Which parts of a backward pass are live.  torch.autograd.Function.needs_input_grad is static, so the pass
    autograd.grad(inputs=[interpolates], create_graph=True) of the gradient penalty (cgan/losses.py:213-220) would also
    compute every weight gradient (and throw it away), and d_loss.backward() would compute the gradient of the
    interpolates once more.  compute_gradient_penalty / CGANTrainer switch those dead branches off."""
    weight_grads = True   # False inside the inputs-only first backward of the gradient penalty
    first_dgrad = True    # False inside CGANTrainer's d_loss.backward(): nobody reads d loss / d patches


FLAGS = _Flags()


class backward_flags:
    def __init__(self, **kw):
        self.kw = kw

    def __enter__(self):
        self.prev = {k: getattr(FLAGS, k) for k in self.kw}
        for k, v in self.kw.items():
            setattr(FLAGS, k, v)

    def __exit__(self, *a):
        for k, v in self.prev.items():
            setattr(FLAGS, k, v)


# ------------------------------------------------------------------------------------------------ gradient sinks
# A parameter with a sink (installed by FusedClipAdam: a view into the network's flat fp32 gradient bucket, also visible as
# param.grad) receives its weight gradients IN PLACE from the kernels: every wgrad-type kernel accumulates into the sink,
# backward() returns None for that parameter, and autograd launches no add kernels (a D weight gets 3-4 contributions per
# critic step: real, fake and the two terms of the penalty).  Parameters without a sink get ordinary returned gradients,
# so the modules also work under a plain torch loop (torch.optim.Adam + clip_grad_norm_).
def sink_of(param):
    return getattr(param, "_b2c_sink", None)


_SINK_LISTENER = None  # callable(param) invoked after every kernel that accumulated into param's sink (train_step.GradBuckets)


def set_sink_listener(fn):
    global _SINK_LISTENER
    _SINK_LISTENER = fn


def _sink_written(*params):
    if _SINK_LISTENER is not None:
        for p in params:
            if p is not None:
                _SINK_LISTENER(p)


_SN_PENDING = {}  # id(weight_orig) -> (param, [(u, v, sig, part, nparts)]): rank-1 terms owed to the sink (flush_sn_pending)


def _sn_defer(wctx, part, nparts):
    _SN_PENDING.setdefault(id(wctx.weight), (wctx.weight, []))[1].append((wctx.u, wctx.v, wctx.sig, part, nparts))


def flush_sn_pending(params=None):
    """Spectral-norm backward, second half: the weight-gradient kernels put sum_i dWe_i / sigma_i into the sinks; this ONE
    launch subtracts sum_i (<dWe_i, W> / sigma_i^2) u_i v_i^T for every layer and forward instance.
    params: only these parameters (a gradient bucket about to be all-reduced); default: everything pending."""
    if not _SN_PENDING:
        return
    if params is None:
        items = list(_SN_PENDING.values())
        _SN_PENDING.clear()
    else:
        items = [_SN_PENDING.pop(id(p)) for p in params if id(p) in _SN_PENDING]
        if not items:
            return
    nl = len(items)
    ninst = [len(lst) for _, lst in items]
    tot = sum(ninst)
    P, I = ctypes.c_void_p, ctypes.c_int
    grad = (P * nl)(*[sink_of(p).data_ptr() for p, _ in items])
    R = (I * nl)(*[p.shape[0] for p, _ in items])
    K = (I * nl)(*[p.numel() // p.shape[0] for p, _ in items])
    flat = [inst for _, lst in items for inst in lst]
    _lib.call("b2c_sn_rank1", grad, R, K, (I * nl)(*ninst), nl, (P * tot)(*[f[0].data_ptr() for f in flat]),
              (P * tot)(*[f[1].data_ptr() for f in flat]), (P * tot)(*[f[2].data_ptr() for f in flat]),
              (P * tot)(*[f[3].data_ptr() for f in flat]), (I * tot)(*[f[4] for f in flat]))


# ------------------------------------------------------------------------------------------------ packed weights
_PACKS = {}  # id(param) -> {"key": ..., "c": tensor | None, "t": tensor | None}; buffers are persistent (CUDA-graph safe)


def _pack_entry(param):
    key = (param._version, getattr(param, "_b2c_epoch", 0), param.data_ptr())
    e = _PACKS.get(id(param))
    if e is None:
        e = _PACKS[id(param)] = {"key": key, "c": None, "t": None, "c_ok": False, "t_ok": False, "ref": param}
    elif e["key"] != key or e["ref"] is not param:
        e.update(key=key, c_ok=False, t_ok=False, k3_ok=False, ref=param)
    return e


def invalidate_packs():
    """Forget which bf16 packs are current (buffers stay): the next use of every weight re-packs it.  CGANTrainer calls
    this before capturing a CUDA graph so that the capture contains every pack the replays need."""
    for e in _PACKS.values():
        e["c_ok"] = e["t_ok"] = e["k3_ok"] = False


def pack_all(params):
    """Re-pack, in ONE launch, every stale bf16 pack of these parameters whose buffer already exists (i.e. that has been
    used in that layout before).  FusedClipAdam calls it right after its update: the 26 lazy per-weight pack launches of a
    step become 3, and a captured step starts with current packs."""
    jobs = []
    for p in params:
        e = _PACKS.get(id(p))
        if e is None or e["ref"] is not p or (e["c"] is None and e["t"] is None):
            continue
        e = _pack_entry(p)
        d0, d1 = p.shape[:2]
        for kind, name in ((0, "c"), (1, "t")):
            if e[name] is not None and not e[name + "_ok"]:
                jobs.append((p.data_ptr(), e[name].data_ptr(), d0, d1, kind))
                e[name + "_ok"] = True
    if not jobs:
        return
    n = len(jobs)
    P, I = ctypes.c_void_p, ctypes.c_int
    _lib.call("b2c_pack_w_multi", (P * n)(*[j[0] for j in jobs]), (P * n)(*[j[1] for j in jobs]), (I * n)(*[j[2] for j in jobs]),
              (I * n)(*[j[3] for j in jobs]), (I * n)(*[j[4] for j in jobs]), n)


def packed_c(param):
    """bf16 [D0][16*D1] K-major operand of the UN-SCALED master weight; re-packed only after the weight changed
    (torch version counter, or FusedClipAdam's epoch stamp) - once per optimizer step instead of once per forward."""
    e = _pack_entry(param)
    if not e["c_ok"]:
        d0, d1 = param.shape[:2]
        if e["c"] is None:
            e["c"] = torch.empty(d0, 16 * d1, dtype=BF16, device=param.device)
        _lib.call("b2c_pack_w_c", param.detach(), None, e["c"], d0, d1)
        e["c_ok"] = True
    return e["c"]


def packed_t(param):
    e = _pack_entry(param)
    if not e["t_ok"]:
        d0, d1 = param.shape[:2]
        if e["t"] is None:
            e["t"] = torch.empty(4, d1, 4 * d0, dtype=BF16, device=param.device)
        _lib.call("b2c_pack_w_t", param.detach(), None, e["t"], d0, d1)
        e["t_ok"] = True
    return e["t"]


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
    """One forward instance of a conv weight: the fp32 master parameter and the spectral-norm (u, v, sigma) snapshot used
    by THAT forward (torch spectral_norm.py:97-112 advances u/v every train-mode forward).  The tensor-core operands are
    the cached packs of the un-scaled weight; 1/sigma rides along as the `scale` of the GEMM epilogues."""

    def __init__(self, weight, sig=None, u=None, v=None):
        self.weight, self.sig, self.u, self.v = weight, sig, u, v
        self.want_stats = False  # ask the forward conv for the InstanceNorm statistics of its output ...
        self._stats = None       # ... which take_stats() hands to INAct / INActPoolFC

    def take_stats(self):
        zs, self._stats = self._stats, None
        self.want_stats = False  # later uses of this context (double backward) are plain convs
        return zs

    @property
    def inv_sigma(self):
        return None if self.sig is None else self.sig[1:]

    def pack_c(self):
        return packed_c(self.weight)

    def pack_t(self):
        return packed_t(self.weight)

    def finish_wgrad(self, dwe):
        """dL/dW_eff -> dL/dW_orig through W_eff = W / sigma(u, v)   (SURVEY.md App. B); sink-less path."""
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
    """InstanceNorm statistics (sum z, sum z^2) per (image, channel): 64-bit fixed point, value = int64 * 2^-30."""
    return torch.empty(N, C, 2, dtype=torch.int64, device=device)


def _conv_raw(x, wctx, bias, want_stats=False):
    x = dense_nhwc(x)
    N, Cin, H, W = x.shape
    Cout = wctx.weight.shape[0]
    y = cl_empty(N, Cout, H // 2, W // 2, device=x.device)
    zs = _stats_buf(N, Cout, x.device) if want_stats else None
    _lib.call("b2c_conv4s2_bf16", x, Cin, wctx.pack_c(), wctx.inv_sigma, bias, y, Cout, zs, N, H, W, Cin, Cout)
    if want_stats:
        wctx._stats = zs
    return y


# Sign bits of the LeakyReLU outputs of the first layers, keyed by the activation's data pointer: FirstConv(premasked=True)
# leaves them (b2c_first_fprop_bits), the Cout = 64 data gradient of the conv above reads them instead of the activation tensor
# (8 instead of 128 bytes per pixel in its epilogue).  Only FirstConv outputs are ever looked up, and each FirstConv forward
# overwrites the entry of the pointer it returns, so a recycled pointer cannot carry stale bits.
_SIGN_BITS = {}


def _sign_bits_of(act):
    e = _SIGN_BITS.get(act.data_ptr())
    if e is not None and e.numel() * 64 == act.numel():
        return e
    return None


def _convt_raw(x, wpack_t, Cout, wctx=None, act=None, slope=1.0, scale=None):
    x, ld = as_nhwc(x)
    N, Cin, H, W = x.shape
    y = cl_empty(N, Cout, 2 * H, 2 * W, device=x.device)
    zs = _stats_buf(N, Cout, x.device) if (wctx is not None and wctx.want_stats) else None
    bits = _sign_bits_of(act) if (act is not None and Cout == 64 and zs is None) else None
    if bits is not None:
        _lib.call("b2c_convt4s2_signbits_bf16", x, ld, wpack_t, scale, y, Cout, bits, float(slope), N, H, W, Cin, Cout)
        return y
    _lib.call("b2c_convt4s2_bf16", x, ld, wpack_t, scale, None, y, Cout, zs, act, float(slope), N, H, W, Cin, Cout)
    if zs is not None:
        wctx._stats = zs
    return y


_WGRAD_DET = __import__("os").environ.get("B2C_WGRAD_DET", "0") == "1"


def _wgrad(wctx, small, large):
    """Weight gradient sum small[n,sy,sx,a] large[n,2sy+ky-1,2sx+kx-1,b] -> [Cs][Cl][4][4] of wctx.weight.  With a sink:
    accumulated in place (times 1/sigma, rank-1 term deferred) and None is returned; else returned as a tensor."""
    small, lds = as_nhwc(small)
    large = dense_nhwc(large)
    N, Cs, Hs, Ws = small.shape
    Cl = large.shape[1]
    dev = small.device
    # one [16][Cs][Cl] slab: split-K partials are reduced with fp32 vector atomics.  B2C_WGRAD_DET=1 asks for one slab per
    # split instead (fixed-order reduction, bit-reproducible, measured 1.2-2.6x slower at the step's shapes: 9-74 splits)
    nws = _lib.query("b2c_wgrad4s2_ws_floats", N, Hs, Ws, Cs, Cl, _lib.sm_count()) if _WGRAD_DET else 16 * Cs * Cl
    ws = torch.empty(nws, dtype=torch.float32, device=dev)
    sink = sink_of(wctx.weight)
    if sink is None:
        dw = torch.empty(Cs, Cl, 4, 4, dtype=torch.float32, device=dev)
        _lib.call("b2c_wgrad4s2_bf16", small, lds, large, Cl, ws, nws, dw, 0, None, None, None, None, N, Hs, Ws, Cs, Cl,
                  _lib.sm_count())
        return wctx.finish_wgrad(dw)
    if wctx.sig is None:
        _lib.call("b2c_wgrad4s2_bf16", small, lds, large, Cl, ws, nws, sink, 1, None, None, None, None, N, Hs, Ws, Cs, Cl,
                  _lib.sm_count())
    else:
        part, nparts = torch.empty(1024, dtype=torch.float32, device=dev), ctypes.c_int(0)
        _lib.call("b2c_wgrad4s2_bf16", small, lds, large, Cl, ws, nws, sink, 1, wctx.inv_sigma, wctx.weight.detach(), part,
                  ctypes.byref(nparts), N, Hs, Ws, Cs, Cl, _lib.sm_count())
        _sn_defer(wctx, part, nparts.value)
    _sink_written(wctx.weight)
    return None


# ------------------------------------------------------------------------------------------------ conv k4 s2 p1
class Conv4s2(torch.autograd.Function):
    """y = conv2d(x, W_eff, bias, stride 2, pad 1).  Replaces nn.Conv2d at cgan/models.py:57,236 (Cin >= 64)."""

    @staticmethod
    def forward(ctx, x, weight, bias, wctx, bias_feeds_norm, in_lrelu_slope=None):
        """in_lrelu_slope: x is the output of a LeakyReLU(slope) whose producer leaves its activation backward to this
        conv (FirstConv(premasked=True)): the returned input gradient is already multiplied by LeakyReLU'."""
        ctx.wctx, ctx.has_bias, ctx.bias_feeds_norm = wctx, bias is not None, bias_feeds_norm
        ctx.bias_param = bias if isinstance(bias, torch.nn.Parameter) else None
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
        if ctx.needs_input_grad[1] and FLAGS.weight_grads:
            with torch.no_grad():
                gw = _wgrad(ctx.wctx, gy.detach(), x)
        if ctx.has_bias and ctx.needs_input_grad[2] and FLAGS.weight_grads:
            with torch.no_grad():
                bsink = sink_of(ctx.bias_param) if ctx.bias_param is not None else None
                if ctx.bias_feeds_norm:  # an affine-less InstanceNorm follows: d/d bias == 0 exactly
                    gb = None if bsink is not None else torch.zeros(gy.shape[1], dtype=torch.float32, device=gy.device)
                else:
                    gb = channel_sum(gy.detach())
                    if bsink is not None:
                        bsink.add_(gb)
                        gb = None
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
        return _convt_raw(gy, wctx.pack_t(), Cin, None, act, slope, wctx.inv_sigma)

    @staticmethod
    def backward(ctx, h):
        gy = ctx.saved_tensors[0]
        if ctx.has_act:  # out = m * dgrad(gy): the cotangent reaches dgrad as m * h
            h = LReluBwd.apply(h, ctx.saved_tensors[1], ctx.slope)
        g_gy = gw = None
        if ctx.needs_input_grad[0]:
            g_gy = Conv4s2.apply(h, ctx.wctx.weight, None, ctx.wctx, False)
        if ctx.needs_input_grad[1] and FLAGS.weight_grads:
            with torch.no_grad():
                gw = _wgrad(ctx.wctx, gy.detach(), h.detach())
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
        if ctx.needs_input_grad[1] and FLAGS.weight_grads:
            gw = _wgrad(ctx.wctx, x, gy)
        return gx, gw, None


def channel_sum(g):
    """sum over (n, h, w) of an NHWC bf16 tensor -> fp32 [C] (bias gradient)."""
    g, ld = as_nhwc(g)
    N, C, H, W = g.shape
    zs = _stats_buf(N, C, g.device)
    _lib.call("b2c_in_stats", g, ld, zs, N, H * W, C)
    out = torch.zeros(C, dtype=torch.float32, device=g.device)
    # zs[:, :, 0] summed over n: one strided rowsum per channel is wasteful; transpose-free trick: rowsum over (N) rows
    tmp = (zs[:, :, 0].to(torch.float64) * 2.0 ** -30).to(torch.float32).t().contiguous()  # [C][N], tiny
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


class Interp:
    """GP interpolation x = alpha * a + (1 - alpha) * b (cgan/losses.py:203-204) fused into the first-layer gathers:
    a0/a1 = real (pred, other), b0/b1 = fake, alpha: fp32 (N,) device tensor."""

    def __init__(self, a0, a1, b0, b1, alpha):
        self.a0, self.a1, self.b0, self.b1 = _f32c(a0), _f32c(a1), _f32c(b0), _f32c(b1)
        self.alpha = alpha.detach().to(torch.float32).reshape(-1).contiguous()


def _first_fprop_raw(a0, a1, wctx, bias, slope, interp=None, want_bits=False):
    if interp is not None:
        a0, a1, b0, b1, al = interp.a0, interp.a1, interp.b0, interp.b1, interp.alpha
    else:
        a0, a1, b0, b1, al = _f32c(a0), _f32c(a1), None, None, None
    N, _, H, W = a0.shape
    y = cl_empty(N, 64, H // 2, W // 2, device=a0.device)
    if want_bits:
        bits = torch.empty(N * (H // 2) * (W // 2), dtype=torch.int64, device=a0.device)
        _lib.call("b2c_first_fprop_bits", a0, a1, b0, b1, al, wctx.weight.detach(), wctx.inv_sigma, bias, y, N, H, W, slope, bits)
        if len(_SIGN_BITS) > 64:
            _SIGN_BITS.clear()
        _SIGN_BITS[y.data_ptr()] = bits
        return y
    _SIGN_BITS.pop(y.data_ptr(), None)
    _lib.call("b2c_first_fprop", a0, a1, b0, b1, al, wctx.weight.detach(), wctx.inv_sigma, bias, y, N, H, W, slope)
    return y


def _first_wgrad(a0, a1, gz, wctx, bias_param, interp=None):
    """-> (gw, gb); None where the gradient went into the parameter's sink."""
    if interp is not None:
        a0, a1, b0, b1, al = interp.a0, interp.a1, interp.b0, interp.b1, interp.alpha
    else:
        a0, a1, b0, b1, al = _f32c(a0), _f32c(a1), None, None, None
    gz = dense_nhwc(gz.detach())
    N, _, H, W = a0.shape
    dev = gz.device
    wsink = sink_of(wctx.weight)
    bsink = sink_of(bias_param) if bias_param is not None else None
    if wsink is None:
        dw = torch.empty_like(wctx.weight, dtype=torch.float32)
        _lib.call("b2c_zero", dw, dw.numel() * 4)
        db = None
        if bias_param is not None and bsink is None:
            db = torch.empty(64, dtype=torch.float32, device=dev)
            _lib.call("b2c_zero", db, 256)
        _lib.call("b2c_first_wgrad", a0, a1, b0, b1, al, gz, dw, db if bsink is None else bsink, None, None, None, None,
                  N, H, W)
        return wctx.finish_wgrad(dw), db
    db = None
    if bias_param is not None and bsink is None:
        db = torch.empty(64, dtype=torch.float32, device=dev)
        _lib.call("b2c_zero", db, 256)
    dbp = bsink if bsink is not None else db
    if wctx.sig is None:
        _lib.call("b2c_first_wgrad", a0, a1, b0, b1, al, gz, wsink, dbp, None, None, None, None, N, H, W)
    else:
        part, nparts = torch.empty(1024, dtype=torch.float32, device=dev), ctypes.c_int(0)
        _lib.call("b2c_first_wgrad", a0, a1, b0, b1, al, gz, wsink, dbp, wctx.inv_sigma, wctx.weight.detach(), part,
                  ctypes.byref(nparts), N, H, W)
        _sn_defer(wctx, part, nparts.value)
    _sink_written(wctx.weight, bias_param if bsink is not None else None)
    return None, db


class FirstConv(torch.autograd.Function):
    """act = LeakyReLU_slope(conv2d(cat(a0, a1), W_eff, bias, 2, 1)) on NCHW fp32 inputs (cgan/models.py:103,245,257)."""

    @staticmethod
    def forward(ctx, a0, a1, weight, bias, wctx, slope, premasked=False, interp=None):
        """premasked: the (only) consumer of the output is a Conv4s2(in_lrelu_slope=slope), which returns its input
        gradient already multiplied by LeakyReLU' - no separate activation-backward pass here.
        interp: the layer input is alpha * real + (1 - alpha) * fake, formed inside the gather; a0 / a1 are then only
        autograd PLACEHOLDERS of the interpolates (uninitialised, never read) that receive d out / d interpolate."""
        y = _first_fprop_raw(a0, a1, wctx, bias, slope, interp, want_bits=premasked)
        ctx.premasked, ctx.interp = premasked, interp
        ctx.wctx, ctx.slope, ctx.has_a1, ctx.has_bias = wctx, slope, a1 is not None, bias is not None
        ctx.bias_param = bias if isinstance(bias, torch.nn.Parameter) else None
        if interp is None:
            ctx.save_for_backward(a0, a1, y)
        else:
            ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, g):
        if ctx.interp is None:
            a0, a1, y = ctx.saved_tensors
        else:
            a0 = a1 = None
            (y,) = ctx.saved_tensors
        gz = LReluBwd.apply(g, y, ctx.slope) if (ctx.slope != 1.0 and not ctx.premasked) else g
        g0 = g1 = gw = gb = None
        if (ctx.needs_input_grad[0] or (ctx.has_a1 and ctx.needs_input_grad[1])) and FLAGS.first_dgrad:
            g0, g1 = FirstDgrad.apply(gz, ctx.wctx.weight, ctx.wctx, ctx.has_a1)
            if not ctx.has_a1:
                g1 = None
        if (ctx.needs_input_grad[2] or (ctx.has_bias and ctx.needs_input_grad[3])) and FLAGS.weight_grads:
            with torch.no_grad():
                gw, gb = _first_wgrad(a0, a1, gz, ctx.wctx, ctx.bias_param if ctx.has_bias else None, ctx.interp)
        return g0, g1, gw, gb, None, None, None, None


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
            g1 = g0.new_empty(())
            ctx.mark_non_differentiable(g1)
        return g0, g1

    @staticmethod
    def backward(ctx, h0, h1):
        (gz,) = ctx.saved_tensors
        h1 = h1 if ctx.has_a1 else None
        g_gz = gw = None
        if ctx.needs_input_grad[0]:
            g_gz = FirstConv.apply(h0, h1, ctx.wctx.weight, None, ctx.wctx, 1.0)
        if ctx.needs_input_grad[1] and FLAGS.weight_grads:
            with torch.no_grad():
                gw, _ = _first_wgrad(h0, h1, gz, ctx.wctx, None)
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
    zs = _stats_buf(N, C, z.device)
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
        _lib.call("b2c_in_act_bwd", g, ldg, None, None, z, ldz, zs, mask, slope, ws, gz, C, N, H * W, C)
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
        ps = torch.empty(2, N, C, dtype=torch.int64, device=z.device)  # planar fixed-point (sum act(x_hat), sum act'(x_hat))
        _lib.call("b2c_in_act_pool_fwd", z, ldz, zs, slope, ps, N, H * W, C)
        delta = torch.empty(N, 4, dtype=torch.float32, device=z.device)
        th = torch.empty_like(delta)
        _lib.call("b2c_pool_fc_fwd", ps, fc_w.detach(), fc_b.detach(), scale, H * W, delta, th, N, C)
        ctx.slope, ctx.scale, ctx.fc_b = slope, scale, fc_b
        ctx.save_for_backward(z, zs, ps, th, fc_w)
        return delta

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gd):
        z, zs, ps, th, fc_w = ctx.saved_tensors
        z, ldz = as_nhwc(z)
        N, C, H, W = z.shape
        wsink, bsink = sink_of(fc_w), sink_of(ctx.fc_b)
        dW = wsink if wsink is not None else torch.empty_like(fc_w)
        db = bsink if bsink is not None else torch.empty(4, dtype=torch.float32, device=z.device)
        gvec = torch.empty(N, C, dtype=torch.float32, device=z.device)
        _lib.call("b2c_pool_fc_bwd", gd.contiguous(), th, ps, fc_w.detach(), ctx.scale, H * W, dW, db, gvec, N, C)
        ws = torch.empty(N, C, 2, dtype=torch.float32, device=z.device)
        gz = cl_empty(N, C, H, W, device=z.device)
        # the plane sums of this backward follow from the forward's (sum act, sum act'): ONE apply pass over z
        _lib.call("b2c_in_act_bwd", None, 0, gvec, ps, z, ldz, zs, None, ctx.slope, ws, gz, C, N, H * W, C)
        _sink_written(fc_w if wsink is not None else None, ctx.fc_b if bsink is not None else None)
        return gz, (None if wsink is not None else dW), (None if bsink is not None else db), None, None, None


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
        if ctx.needs_input_grad[1] and FLAGS.weight_grads:
            with torch.no_grad():
                gw = _head_wgrad_raw(a, gy.detach(), weight)
        return ga, gw


def _head_wgrad_raw(a, gy, weight):
    a = dense_nhwc(a)
    N, C, h, _ = a.shape
    sink = sink_of(weight)
    if sink is not None:  # accumulated in place
        _lib.call("b2c_head_wgrad", a, gy.to(torch.float32).contiguous(), sink, N, h, C)
        _sink_written(weight)
        return None
    dw = torch.empty_like(weight, dtype=torch.float32)
    _lib.call("b2c_zero", dw, dw.numel() * 4)
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
        if ctx.needs_input_grad[1] and FLAGS.weight_grads:
            with torch.no_grad():
                gw = _head_wgrad_raw(hh.detach(), gy, weight)
        return g_gy, gw, None


# ------------------------------------------------------------------------------------------------ scalar losses
class MapMean(torch.autograd.Function):
    """mean over every element of an fp32 map (torch.mean at cgan/cgan_train_enhanced.py:327,362)."""

    @staticmethod
    def forward(ctx, x, scale=1.0):
        x = x.contiguous()
        ctx.shape, ctx.scale = x.shape, scale
        out = torch.empty(1, dtype=torch.float32, device=x.device)
        _lib.call("b2c_map_mean", x, x.numel(), scale / x.numel(), out)
        return out.reshape(())

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        n = 1
        for s in ctx.shape:
            n *= s
        out = torch.empty(ctx.shape, dtype=torch.float32, device=g.device)
        _lib.call("b2c_fill_scaled", out, n, g.contiguous(), ctx.scale / n)
        return out, None


class CriticLoss(torch.autograd.Function):
    """d_loss = -(mean(real) - mean(fake)) + lambda_gp * gp in one launch (cgan/cgan_train_enhanced.py:327-328).
    Returns a 3-vector (d_loss, wd, gp); differentiate element 0."""

    @staticmethod
    def forward(ctx, real, fake, gp, lambda_gp):
        real, fake = real.contiguous(), fake.contiguous()
        out = torch.empty(3, dtype=torch.float32, device=real.device)
        _lib.call("b2c_critic_loss", real, real.numel(), fake, fake.numel(), gp.reshape(1), float(lambda_gp), out)
        ctx.shapes, ctx.lam = (real.shape, fake.shape), float(lambda_gp)
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        g0 = g[0:1].contiguous()
        gr = torch.empty(ctx.shapes[0], dtype=torch.float32, device=g.device)
        gf = torch.empty(ctx.shapes[1], dtype=torch.float32, device=g.device)
        _lib.call("b2c_critic_loss_bwd", g0, gr, gr.numel(), gf, gf.numel())
        ggp = g0.reshape(()) if ctx.lam == 1.0 else g0.reshape(()) * ctx.lam
        return gr, gf, ggp, None


class GradNormPenalty(torch.autograd.Function):
    """mean_b (sqrt(|g0_b|^2 + |g1_b|^2 + 1e-12) - 1)^2   (cgan/losses.py:223-231).  Backward emits the seed of the
    double backward, (2/B)(1 - 1/r_b) g, in one pass per tensor.  Reductions are two-stage, fixed order."""

    @staticmethod
    def forward(ctx, g0, g1):
        g0, g1 = g0.contiguous(), g1.contiguous()
        B = g0.shape[0]
        n = g0.numel() // B
        part, splits = torch.empty(32 * B, dtype=torch.float32, device=g0.device), ctypes.c_int(0)
        _lib.call("b2c_gp_sumsq", g0, g1, n, B, part, ctypes.byref(splits))
        gp = torch.empty(1, dtype=torch.float32, device=g0.device)
        coef = torch.empty(B, dtype=torch.float32, device=g0.device)
        _lib.call("b2c_gp_tail", part, splits.value, gp, coef, B)
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


# ------------------------------------------------------------------------------------------------ simple regressor (forward)
def packed_k3(param):
    """bf16 [Cout][9*Cin] operand of a 3x3 conv weight, cached like packed_c / packed_t."""
    e = _pack_entry(param)
    if not e.get("k3_ok") or e.get("k3") is None:
        co, ci = param.shape[:2]
        if e.get("k3") is None:
            e["k3"] = torch.empty(co, 9 * ci, dtype=BF16, device=param.device)
        _lib.call("b2c_pack_w_k3", param.detach(), e["k3"], co, ci, 0)
        e["k3_ok"] = True
    return e["k3"]


def packed_k3_flip(param):
    """bf16 [Cin][9*Cout] data-gradient operand of a 3x3 conv weight (taps flipped, channels transposed), cached."""
    e = _pack_entry(param)
    if not e.get("k3f_ok") or e.get("k3f") is None:
        co, ci = param.shape[:2]
        if e.get("k3f") is None:
            e["k3f"] = torch.empty(ci, 9 * co, dtype=BF16, device=param.device)
        _lib.call("b2c_pack_w_k3", param.detach(), e["k3f"], co, ci, 1)
        e["k3f_ok"] = True
    return e["k3f"]


def _simple_forward(net, x, save):
    """GeneratorSimpleRegressor.forward (cgan/models.py:213-216) on the kernels: (B,3,S,S) fp32 -> (B,4).  save: keep what
    the backward needs (per conv: its input, pre-norm output, statistics, post-activation map; head activations, masks)."""
    from . import rng
    x = x.detach().to(torch.float32).contiguous()
    N, _, H, W = x.shape
    dev = x.device
    convs = [m for m in net.features if m.__class__.__name__ == "Conv2dK3S1"]
    a, saved = None, []
    for i, conv in enumerate(convs):
        cout = conv.out_channels
        z = cl_empty(N, cout, H, W, device=dev)
        zs = _stats_buf(N, cout, dev)
        if i == 0:
            _lib.call("b2c_conv3x3_first", x, conv.weight.detach(), conv.bias.detach(), z, N, H, W)
            _lib.call("b2c_in_stats", z, cout, zs, N, H * W, cout)
        else:
            _lib.call("b2c_conv3s1_bf16", a, conv.in_channels, packed_k3(conv.weight), None, conv.bias.detach(), z, cout, zs,
                      N, H, W, conv.in_channels, cout)
        inp = a
        a = cl_empty(N, cout, H, W, device=dev)
        _lib.call("b2c_in_act_fwd", z, cout, zs, None, 0.0, a, cout, None, 0, N, H * W, cout)
        if save:
            saved.append((inp, z, zs, a, H, W))
        if i % 2 == 1:
            p = cl_empty(N, cout, H // 2, W // 2, device=dev)
            _lib.call("b2c_maxpool2_fwd", a, p, N, H, W, cout)
            a, H, W = p, H // 2, W // 2
    l1, l2, l3 = net.regressor[2], net.regressor[5], net.regressor[8]
    m1 = m2 = None
    if net.training:  # Dropout(0.5) after each hidden ReLU
        m1, m2 = rng.next_mask(N * 256, dev), rng.next_mask(N * 64, dev)
    delta = torch.empty(N, 4, dtype=torch.float32, device=dev)
    acts = torch.empty(N, 836, dtype=torch.float32, device=dev) if save else None
    _lib.call("b2c_simple_head_fwd", a, H * W, l1.weight.detach(), l1.bias.detach(), l2.weight.detach(), l2.bias.detach(),
              l3.weight.detach(), l3.bias.detach(), m1, m2, float(net.delta_scale), delta, acts, N)
    return delta, (x, saved, acts, m1, m2, H * W)


@torch.no_grad()
def simple_regressor_forward(net, x):
    return _simple_forward(net, x, False)[0]


class SimpleRegressorFn(torch.autograd.Function):
    """The whole GeneratorSimpleRegressor as ONE differentiable op (first order: the generator is never under the gradient
    penalty).  Parameter order: the 8 conv (weight, bias) pairs, then the 3 Linear (weight, bias) pairs.  Conv biases feed an
    InstanceNorm2d, which removes per-channel constants: their gradient is zero in exact arithmetic (the reference's autograd
    returns rounding noise of the order 1e-9 there) and is returned as exact zeros."""

    @staticmethod
    def forward(ctx, net, x, *params):
        delta, state = _simple_forward(net, x, True)
        ctx.net, ctx.state = net, state
        return delta

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gdelta):
        net = ctx.net
        x, saved, acts, m1, m2, HWp = ctx.state
        ctx.state = None
        N, dev = x.shape[0], x.device
        convs = [m for m in net.features if m.__class__.__name__ == "Conv2dK3S1"]
        l1, l2, l3 = net.regressor[2], net.regressor[5], net.regressor[8]
        sm = _lib.sm_count()

        def target(p):  # (buffer the kernel writes, accumulate flag, value autograd should see)
            sink = sink_of(p)
            if sink is not None:
                return sink, 1, None
            buf = torch.empty_like(p, dtype=torch.float32)
            return buf, 0, buf

        lin = [target(q) for q in (l1.weight, l1.bias, l2.weight, l2.bias, l3.weight, l3.bias)]
        gv = torch.empty(N, 836, dtype=torch.float32, device=dev)
        hp = int(round(HWp ** 0.5))
        g = cl_empty(N, 512, hp, hp, device=dev)
        acc = lin[0][1]
        assert all(t[1] == acc for t in lin), "GeneratorSimpleRegressor: all or none of the head parameters may have a gradient sink"
        _lib.call("b2c_simple_head_bwd", gdelta.contiguous().float(), acts, l1.weight.detach(), l2.weight.detach(),
                  l3.weight.detach(), m1, m2, float(net.delta_scale), HWp, gv, g, lin[0][0], lin[1][0], lin[2][0], lin[3][0],
                  lin[4][0], lin[5][0], acc, N)
        _sink_written(*[q for q, t in zip((l1.weight, l1.bias, l2.weight, l2.bias, l3.weight, l3.bias), lin) if t[2] is None])
        conv_grads = [None] * 16
        for i in reversed(range(8)):
            conv = convs[i]
            inp, z, zs, a, H, W = saved[i]
            cout, cin = conv.out_channels, conv.in_channels
            if i % 2 == 1:  # MaxPool2d(2, 2) follows this block
                ga = cl_empty(N, cout, H, W, device=dev)
                _lib.call("b2c_maxpool2_bwd", a, g, ga, N, H, W, cout)
                g = ga
            ws2 = torch.empty(N, cout, 2, dtype=torch.float32, device=dev)
            gz = cl_empty(N, cout, H, W, device=dev)
            _lib.call("b2c_in_act_bwd", g, cout, None, None, z, cout, zs, None, 0.0, ws2, gz, cout, N, H * W, cout)
            wbuf, wacc, wret = target(conv.weight)
            if i == 0:
                if not wacc:
                    wbuf.zero_()
                _lib.call("b2c_conv3x3_first_wgrad", x, gz, wbuf, N, H, W)
            else:
                nws = 9 * cout * cin
                ws = torch.empty(nws, dtype=torch.float32, device=dev)
                _lib.call("b2c_wgrad3s1_bf16", gz, inp, ws, nws, wbuf, wacc, N, H, W, cout, cin, sm)
                g = cl_empty(N, cin, H, W, device=dev)
                _lib.call("b2c_conv3s1_bf16", gz, cout, packed_k3_flip(conv.weight), None, None, g, cin, None, N, H, W, cout, cin)
            conv_grads[2 * i] = wret
            bsink = sink_of(conv.bias)
            conv_grads[2 * i + 1] = None if bsink is not None else torch.zeros_like(conv.bias)
            _sink_written(conv.weight if wret is None else None, conv.bias if bsink is not None else None)
        return (None, None, *conv_grads, *[t[2] for t in lin])


def fill_(t, value):
    """t[...] = value for a contiguous fp32 tensor (b2c_fill_scaled instead of an ATen fill kernel)."""
    _lib.call("b2c_fill_scaled", t, t.numel(), None, float(value))
    return t


def iou_rows(pred, target, eps=1e-6):
    pred, target = pred.to(torch.float32).contiguous(), target.to(torch.float32).contiguous()
    out = torch.empty(pred.shape[0], dtype=torch.float32, device=pred.device)
    _lib.call("b2c_eiou", pred, target, eps, None, None, None, out, pred.shape[0])
    return out


# ------------------------------------------------------------------------------------------------ optimizer
class FusedClipAdam:
    """clip_grad_norm_(max_norm) + Adam in two launches over FLAT buckets (cgan/cgan_train_enhanced.py:256-257,331-332,
    368-369).  Torch-default Adam semantics (eps 1e-8, no weight decay, bias correction), one shared step counter.

    The parameters are re-homed as views of one contiguous fp32 buffer, their gradients as views of a second one (the
    gradient SINK the weight-gradient kernels accumulate into, also visible as param.grad), moments likewise: zero_grad is
    one memset, the norm one two-stage reduction, the update one 128-bit vectorised kernel, and a data-parallel all-reduce
    runs on the bucket itself (no gather / scatter copies).  `lr` lives on the device, so a scheduler can change it while
    the step is replayed from a captured CUDA graph."""

    def __init__(self, params, lr=2e-4, betas=(0.5, 0.999), eps=1e-8, max_norm=1.0):
        self.params = [p for p in params]
        assert self.params and all(p.is_cuda and p.dtype == torch.float32 for p in self.params)
        dev = self.params[0].device
        self.betas, self.eps, self.max_norm = betas, eps, max_norm
        offs, total = [], 0
        for p in self.params:
            offs.append(total)
            total += (p.numel() + 3) // 4 * 4  # 16-byte aligned slices
        self.total = total
        self.flat_p = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_m = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_v = torch.zeros(total, dtype=torch.float32, device=dev)
        self.slices = []
        for p, o in zip(self.params, offs):
            n = p.numel()
            self.flat_p[o:o + n].copy_(p.data.reshape(-1))
            p.data = self.flat_p[o:o + n].view_as(p)
            p._b2c_sink = self.flat_g[o:o + n].view_as(p)
            p.grad = p._b2c_sink
            self.slices.append((o, n))
        self.step_count = 0
        self.step_base = None  # int64 device scalar added to step_count in the kernel (CUDA-graph replays)
        self.grad_scale = 1.0  # data parallel: 1 / world_size (the all-reduce SUMs)
        self.gnorm_sq = torch.zeros(1, dtype=torch.float32, device=dev)  # |clipped-from gradient|^2 of the last step
        self._part = torch.empty(1024, dtype=torch.float32, device=dev)
        self.lr_dev = torch.full((1,), float(lr), dtype=torch.float32, device=dev)
        self._lr = float(lr)
        self._n = (ctypes.c_longlong * 1)(total)
        self._arr = lambda t: (ctypes.c_void_p * 1)(t.data_ptr())

    @property
    def lr(self):
        return self._lr

    @lr.setter
    def lr(self, value):  # e.g. ReduceLROnPlateau (cgan_train_enhanced.py:427-428); takes effect in captured graphs too
        self._lr = float(value)
        self.lr_dev.fill_(self._lr)

    def m_of(self, i):
        o, n = self.slices[i]
        return self.flat_m[o:o + n].view_as(self.params[i])

    def v_of(self, i):
        o, n = self.slices[i]
        return self.flat_v[o:o + n].view_as(self.params[i])

    def zero_grad(self):
        _lib.call("b2c_zero", self.flat_g, self.total * 4)
        for p in self.params:
            if p.grad is not p._b2c_sink:
                p.grad = p._b2c_sink

    def step(self):
        """Returns the device scalar holding the squared pre-clip gradient norm (no host sync)."""
        for p in self.params:  # a gradient that arrived through autograd (plain tensor) instead of the sink
            if p.grad is not None and p.grad.data_ptr() != p._b2c_sink.data_ptr():
                p._b2c_sink.copy_(p.grad)
                p.grad = p._b2c_sink
        self.step_count += 1
        nparts = ctypes.c_int(0)
        _lib.call("b2c_multi_sumsq", self._arr(self.flat_g), self._n, 1, self._part, ctypes.byref(nparts))
        _lib.call("b2c_multi_adam", self._arr(self.flat_p), self._arr(self.flat_g), self._arr(self.flat_m),
                  self._arr(self.flat_v), self._n, 1, self._part, nparts.value, self.gnorm_sq, float(self.grad_scale),
                  float(self.max_norm), self._lr, self.lr_dev, self.betas[0], self.betas[1], self.eps, self.step_count,
                  self.step_base)
        for p in self.params:
            p._b2c_epoch = getattr(p, "_b2c_epoch", 0) + 1  # cached bf16 packs of this weight are stale now
        pack_all(self.params)
        return self.gnorm_sq
