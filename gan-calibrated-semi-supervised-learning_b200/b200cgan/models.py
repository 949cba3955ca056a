"""Drop-in model classes for the reference's cgan/models.py, running on the libb2c.so kernels.

Same class names, constructor arguments, forward signatures, state-dict keys and `.apply(weights_init_normal)`
behaviour as /root/reference/cgan/models.py:37-141,222-258:
  G: down{1..4}.model.0.weight, up{1..3}.model.0.weight, up4.0.weight, fc_delta.1.{weight,bias}
  D: model.{0,2,5,8}.{bias,weight_orig,weight_u,weight_v}, model.11.weight   (spectral_norm=False: .weight/.bias)
Constructing G then D under the same torch seed draws the same initial parameters as the reference (same tensor
shapes initialised in the same order).  The nn.InstanceNorm2d / LeakyReLU / ReLU / Dropout entries of each
nn.Sequential are kept as parameter-less markers (indices and printing match); the arithmetic runs fused in ops.py.
Inputs are NCHW fp32 CUDA patches whose size is a multiple of 16; activations are channels_last bf16.
"""
from __future__ import annotations

import math
from pathlib import Path

import torch
import torch.nn as nn
import torch.nn.functional as F
import yaml

from . import _lib, ops, rng

CONFIG_PATH = Path(__file__).parent / "config.yaml"


def _default(key):
    with open(CONFIG_PATH, "r", encoding="utf-8") as f:
        return yaml.safe_load(f)[key]


def weights_init_normal(m: nn.Module) -> None:
    """Conv* weights ~ N(0, 0.02), biases 0; norm layers with affine weights ~ N(1, 0.02) (cgan/models.py:37-48)."""
    classname = m.__class__.__name__
    if classname.find("Conv") != -1:
        nn.init.normal_(m.weight.data, 0.0, 0.02)
        if m.bias is not None:
            nn.init.constant_(m.bias.data, 0.0)
    elif classname.find("BatchNorm") != -1 or classname.find("InstanceNorm") != -1:
        if m.weight is not None:
            nn.init.normal_(m.weight.data, 1.0, 0.02)
        if m.bias is not None:
            nn.init.constant_(m.bias.data, 0.0)


def _torch_conv_init(weight, bias, fan_in):
    nn.init.kaiming_uniform_(weight, a=math.sqrt(5))
    if bias is not None:
        bound = 1 / math.sqrt(fan_in) if fan_in > 0 else 0
        nn.init.uniform_(bias, -bound, bound)


class Conv2dK4S2(nn.Module):
    """nn.Conv2d(cin, cout, 4, stride=2, padding=1) on the tcgen05 implicit-GEMM kernel (first layer: CUDA cores)."""

    def __init__(self, in_ch, out_ch, bias=True):
        super().__init__()
        self.in_channels, self.out_channels = in_ch, out_ch
        self.weight = nn.Parameter(torch.empty(out_ch, in_ch, 4, 4))
        self.bias = nn.Parameter(torch.empty(out_ch)) if bias else None
        _torch_conv_init(self.weight, self.bias, in_ch * 16)

    def weight_ctx(self):
        return ops.WeightCtx(self.weight)

    def param(self):
        return self.weight

    def extra_repr(self):
        return f"{self.in_channels}, {self.out_channels}, kernel_size=(4, 4), stride=(2, 2), padding=(1, 1), bias={self.bias is not None}"


class SpectralNormConv2dK4S2(nn.Module):
    """nn.utils.spectral_norm(nn.Conv2d(cin, cout, 4, 2, 1)): legacy hook semantics (torch spectral_norm.py:92-114):
    one power iteration per TRAIN-mode forward, in place on weight_u / weight_v; W_eff = weight_orig / sigma."""

    def __init__(self, in_ch, out_ch):
        super().__init__()
        self.in_channels, self.out_channels = in_ch, out_ch
        w = torch.empty(out_ch, in_ch, 4, 4)
        self.bias = nn.Parameter(torch.empty(out_ch))
        _torch_conv_init(w, self.bias, in_ch * 16)
        u = F.normalize(w.new_empty(out_ch).normal_(0, 1), dim=0, eps=1e-12)
        v = F.normalize(w.new_empty(in_ch * 16).normal_(0, 1), dim=0, eps=1e-12)
        self.weight_orig = nn.Parameter(w)
        self.register_buffer("weight_u", u)
        self.register_buffer("weight_v", v)

    @property
    def weight(self):  # what weights_init_normal touches: shares storage with weight_orig
        return self.weight_orig.detach()

    def param(self):
        return self.weight_orig

    def weight_ctx(self):
        return sn_weight_ctxs([self])[0]

    def extra_repr(self):
        return f"{self.in_channels}, {self.out_channels}, kernel_size=(4, 4), stride=(2, 2), padding=(1, 1), spectral_norm"


def sn_weight_ctxs(convs):
    """One power iteration (train mode) + sigma for ALL spectral-norm convs of a Discriminator forward in three launches
    (b2c_sn_power_iter_multi), with the per-forward-instance (sigma, u, v) snapshots the backward of THAT forward needs
    written by the same kernels into one scratch allocation.  -> [ops.WeightCtx] in the order of `convs`."""
    import ctypes
    n = len(convs)
    dev = convs[0].weight_orig.device
    shapes = [(c.weight_orig.shape[0], c.weight_orig.numel() // c.weight_orig.shape[0]) for c in convs]
    a4 = lambda v: (v + 3) // 4 * 4
    sizes = [(4, a4(R), a4(K), a4(K + (K + 31) // 32 + R)) for R, K in shapes]  # sig, u_snap, v_snap, ws
    buf = torch.empty(sum(sum(t) for t in sizes), dtype=torch.float32, device=dev)
    views, off = [], 0
    for (R, K), t in zip(shapes, sizes):
        sig = buf[off:off + 2]; off += t[0]
        us = buf[off:off + R]; off += t[1]
        vs = buf[off:off + K]; off += t[2]
        ws = buf[off:off + t[3]]; off += t[3]
        views.append((sig, us, vs, ws))
    P, I = ctypes.c_void_p, ctypes.c_int
    arr = lambda ts: (P * n)(*[t.data_ptr() for t in ts])
    _lib.call("b2c_sn_power_iter_multi", arr([c.weight_orig for c in convs]), arr([c.weight_u for c in convs]),
              arr([c.weight_v for c in convs]), arr([v[0] for v in views]), arr([v[3] for v in views]),
              arr([v[1] for v in views]), arr([v[2] for v in views]), (I * n)(*[r for r, _ in shapes]),
              (I * n)(*[k for _, k in shapes]), n, int(convs[0].training))
    return [ops.WeightCtx(c.weight_orig, v[0], v[1], v[2]) for c, v in zip(convs, views)]


class ConvTranspose2dK4S2(nn.Module):
    """nn.ConvTranspose2d(cin, cout, 4, stride=2, padding=1, bias=False) as four sub-pixel tcgen05 GEMMs."""

    def __init__(self, in_ch, out_ch):
        super().__init__()
        self.in_channels, self.out_channels = in_ch, out_ch
        self.weight = nn.Parameter(torch.empty(in_ch, out_ch, 4, 4))
        self.bias = None
        _torch_conv_init(self.weight, None, out_ch * 16)

    def extra_repr(self):
        return f"{self.in_channels}, {self.out_channels}, kernel_size=(4, 4), stride=(2, 2), padding=(1, 1), bias=False"


class HeadConv2d(nn.Module):
    """nn.Conv2d(512, 1, 4, stride=1, padding=1, bias=False): PatchGAN score head (bandwidth kernel)."""

    def __init__(self, in_ch):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(1, in_ch, 4, 4))
        self.bias = None
        _torch_conv_init(self.weight, None, in_ch * 16)


class DropoutSite(nn.Dropout):
    """nn.Dropout(p=0.5) marker that also owns the keep-mask source: `injected` (dense u8 NHWC tensor) for parity
    runs, else a counter-based mask from rng.next_mask (identical for 1-GPU and N-GPU runs of the same samples)."""

    injected = None

    def mask(self, N, H, W, C, device, training):
        if not training:
            return None
        if self.injected is not None:
            return self.injected
        return rng.next_mask(N * H * W * C, device)


def _conv_block(x, conv, slope, drop, training, skip=None, norm=True):
    """conv -> [InstanceNorm] -> LeakyReLU(slope) -> [Dropout]; first layers (Cin < 64) take NCHW fp32 input."""
    if conv.in_channels < 64:
        assert not norm and drop is None
        return ops.FirstConv.apply(x, None, conv.param(), conv.bias, conv.weight_ctx(), slope)
    if not norm:
        raise NotImplementedError("un-normalised conv blocks with Cin >= 64 do not occur in the reference models")
    wctx = conv.weight_ctx()
    wctx.want_stats = True  # the InstanceNorm statistics come out of the GEMM epilogue
    z = ops.Conv4s2.apply(x, conv.param(), conv.bias, wctx, norm)
    N, C, H, W = z.shape
    mask = drop.mask(N, H, W, C, z.device, training) if drop is not None else None
    return ops.INAct.apply(z, slope, mask, skip, wctx.take_stats())


class UNetDown(nn.Module):
    def __init__(self, in_size: int, out_size: int, normalize: bool = True, dropout: float | None = None):
        super().__init__()
        layers: list[nn.Module] = [Conv2dK4S2(in_size, out_size, bias=False)]
        if normalize:
            layers.append(nn.InstanceNorm2d(out_size))
        layers.append(nn.LeakyReLU(0.2, inplace=True))
        self._drop = None
        if dropout is not None and dropout > 0:
            assert dropout == 0.5, "only Dropout(0.5) is implemented (the reference uses no other value)"
            layers.append(DropoutSite(dropout))
        self.model = nn.Sequential(*layers)
        self._norm = normalize
        self._has_drop = dropout is not None and dropout > 0

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        drop = self.model[-1] if self._has_drop else None
        return _conv_block(x, self.model[0], 0.2, drop, self.training, norm=self._norm)


class UNetUp(nn.Module):
    def __init__(self, in_size: int, out_size: int, dropout: float | None = None):
        super().__init__()
        layers = [ConvTranspose2dK4S2(in_size, out_size), nn.InstanceNorm2d(out_size), nn.ReLU(inplace=True)]
        self._has_drop = dropout is not None and dropout > 0
        if self._has_drop:
            assert dropout == 0.5
            layers.append(DropoutSite(dropout))
        self.model = nn.Sequential(*layers)

    def forward(self, x: torch.Tensor, skip_input: torch.Tensor) -> torch.Tensor:
        conv = self.model[0]
        wctx = ops.WeightCtx(conv.weight)
        wctx.want_stats = True
        z = ops.ConvT4s2.apply(x, conv.weight, wctx)
        N, C, H, W = z.shape
        mask = self.model[-1].mask(N, H, W, C, z.device, self.training) if self._has_drop else None
        return ops.INAct.apply(z, 0.0, mask, skip_input, wctx.take_stats())  # cat((x, skip), 1) written in place


class GeneratorUNet(nn.Module):
    """4-down / 4-up U-Net trunk -> global average pool -> Linear(64,4) -> Tanh -> * delta_scale   (B,3,S,S) -> (B,4)."""

    def __init__(self, delta_scale: float = None):
        super().__init__()
        if delta_scale is None:
            delta_scale = _default("delta_scale")
        self.delta_scale = float(delta_scale)
        self.down1 = UNetDown(3, 64, normalize=False)
        self.down2 = UNetDown(64, 128)
        self.down3 = UNetDown(128, 256)
        self.down4 = UNetDown(256, 512, dropout=0.5)
        self.up1 = UNetUp(512, 256, dropout=0.5)
        self.up2 = UNetUp(512, 128, dropout=0.5)
        self.up3 = UNetUp(256, 64)
        self.up4 = nn.Sequential(ConvTranspose2dK4S2(128, 64), nn.InstanceNorm2d(64), nn.ReLU(inplace=True))
        self.avg_pool = nn.AdaptiveAvgPool2d(1)
        self.fc_delta = nn.Sequential(nn.Flatten(), nn.Linear(64, 4), nn.Tanh())

    def dropout_sites(self):
        return [self.down4.model[-1], self.up1.model[-1], self.up2.model[-1]]

    @staticmethod
    def dropout_numels(B, S):
        """Elements drawn per train-mode forward at the three sites (down4, up1, up2), in order."""
        return [B * 512 * (S // 16) ** 2, B * 256 * (S // 8) ** 2, B * 128 * (S // 4) ** 2]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        # each skip tensor has two consumers; SkipTap sums their gradients with one strided kernel
        d1, s1 = ops.SkipTap.apply(self.down1(x))
        d2, s2 = ops.SkipTap.apply(self.down2(d1))
        d3, s3 = ops.SkipTap.apply(self.down3(d2))
        d4 = self.down4(d3)
        u1 = self.up1(d4, s3)
        u2 = self.up2(u1, s2)
        u3 = self.up3(u2, s1)
        conv = self.up4[0]
        wctx = ops.WeightCtx(conv.weight)
        wctx.want_stats = True
        z = ops.ConvT4s2.apply(u3, conv.weight, wctx)
        fc = self.fc_delta[1]
        return ops.INActPoolFC.apply(z, fc.weight, fc.bias, 0.0, self.delta_scale, wctx.take_stats())


class Conv2dK3S1(nn.Module):
    """nn.Conv2d(cin, cout, 3, padding=1) with bias: the 9-tap tcgen05 GEMM (first layer, Cin = 3: CUDA cores)."""

    def __init__(self, in_ch, out_ch):
        super().__init__()
        self.in_channels, self.out_channels = in_ch, out_ch
        self.weight = nn.Parameter(torch.empty(out_ch, in_ch, 3, 3))
        self.bias = nn.Parameter(torch.empty(out_ch))
        _torch_conv_init(self.weight, self.bias, in_ch * 9)

    def extra_repr(self):
        return f"{self.in_channels}, {self.out_channels}, kernel_size=(3, 3), stride=(1, 1), padding=(1, 1)"


class GeneratorSimpleRegressor(nn.Module):
    """The reference's alternative generator (cgan/models.py:147-216; generator_type: "simple", cgan/config.yaml:21):
    4 blocks of [Conv3x3 + InstanceNorm + ReLU] x 2 + MaxPool2d(2), then AvgPool -> Linear 512-256-64-4 -> Tanh -> * delta_scale.
    Same constructor, state-dict keys (features.{0,3,7,10,14,17,21,24}.{weight,bias}, regressor.{2,5,8}.{weight,bias};
    4 833 412 parameters) and forward signature.  Forward and (first-order) backward run on the B200 kernels: the whole
    network is one autograd Function (ops.SimpleRegressorFn: 9-tap tcgen05 convs for fprop / dgrad, the split-K tcgen05
    weight-gradient kernel in its 3x3 mode, MaxPool / InstanceNorm+ReLU / head backward kernels); train mode draws the two
    Dropout(0.5) masks of the head from b200cgan.rng.  It trains under CGANTrainer like the U-Net (generator_type: simple)."""

    CHANNELS = [(3, 64), (64, 64), (64, 128), (128, 128), (128, 256), (256, 256), (256, 512), (512, 512)]

    def __init__(self, delta_scale: float = None):
        super().__init__()
        if delta_scale is None:
            delta_scale = _default("delta_scale")
        self.delta_scale = float(delta_scale)
        layers = []
        for i, (cin, cout) in enumerate(self.CHANNELS):
            layers += [Conv2dK3S1(cin, cout), nn.InstanceNorm2d(cout), nn.ReLU(inplace=True)]
            if i % 2 == 1:
                layers.append(nn.MaxPool2d(2, 2))
        self.features = nn.Sequential(*layers)
        self.regressor = nn.Sequential(nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(512, 256), nn.ReLU(inplace=True),
                                       nn.Dropout(0.5), nn.Linear(256, 64), nn.ReLU(inplace=True), nn.Dropout(0.5),
                                       nn.Linear(64, 4), nn.Tanh())

    def dropout_sites(self):
        return []

    @staticmethod
    def dropout_numels(B, S):
        return [B * 256, B * 64]

    def ordered_parameters(self):
        convs = [m for m in self.features if isinstance(m, Conv2dK3S1)]
        lins = [self.regressor[2], self.regressor[5], self.regressor[8]]
        return [p for m in convs + lins for p in (m.weight, m.bias)]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            return ops.SimpleRegressorFn.apply(self, x, *self.ordered_parameters())
        return ops.simple_regressor_forward(self, x)


class Discriminator(nn.Module):
    """Spectral-norm PatchGAN critic on (pred_patch, other_patch) pairs: 2 x (B,3,S,S) -> (B,1,S/16-1,S/16-1), no sigmoid."""

    def __init__(self, spectral_norm: bool = None):
        super().__init__()
        if spectral_norm is None:
            spectral_norm = _default("spectral_norm")

        def conv_block(in_ch: int, out_ch: int, norm: bool = True):
            conv = SpectralNormConv2dK4S2(in_ch, out_ch) if spectral_norm else Conv2dK4S2(in_ch, out_ch, bias=True)
            layers = [conv]
            if norm:
                layers.append(nn.InstanceNorm2d(out_ch))
            layers.append(nn.LeakyReLU(0.2, inplace=True))
            return layers

        self.model = nn.Sequential(*[
            *conv_block(6, 64, norm=False),
            *conv_block(64, 128),
            *conv_block(128, 256),
            *conv_block(256, 512),
            HeadConv2d(512),
        ])

    def weight_ctxs(self):
        """Per-forward weight contexts of model.{0,2,5,8}: with spectral norm, all four power iterations of this forward
        in three launches (torch runs the hook of every wrapped conv at the top of its forward, spectral_norm.py:92-114)."""
        convs = [self.model[i] for i in (0, 2, 5, 8)]
        if isinstance(convs[0], SpectralNormConv2dK4S2):
            return sn_weight_ctxs(convs)
        return [c.weight_ctx() for c in convs]

    def first_block(self, pred_patch, other_patch, wctx, _interp=None):
        """model.0-1: SN-conv(6 -> 64) + bias + LeakyReLU on cat(pred, other) (cat fused).  The LeakyReLU backward of this
        layer is fused into model.2's data gradient (premasked / in_lrelu_slope)."""
        c0 = self.model[0]
        return ops.FirstConv.apply(pred_patch, other_patch, c0.param(), c0.bias, wctx, 0.2, True, _interp)

    def norm_block(self, idx, a, wctx):
        """model.{2,5,8} + InstanceNorm + LeakyReLU."""
        conv = self.model[idx]
        wctx.want_stats = True  # InstanceNorm statistics from the GEMM epilogue
        z = ops.Conv4s2.apply(a, conv.param(), conv.bias, wctx, True, 0.2 if idx == 2 else None)
        return ops.INAct.apply(z, 0.2, None, None, wctx.take_stats())

    def forward(self, pred_patch: torch.Tensor, other_patch: torch.Tensor, _interp=None) -> torch.Tensor:
        """_interp (ops.Interp, internal): the gradient penalty's interpolates (cgan/losses.py:203-204) are formed inside
        the first-layer gather; pred_patch / other_patch are then autograd placeholders for them."""
        ctxs = self.weight_ctxs()
        a = self.first_block(pred_patch, other_patch, ctxs[0], _interp)
        for idx, wctx in zip((2, 5, 8), ctxs[1:]):
            a = self.norm_block(idx, a, wctx)
        return ops.HeadConv.apply(a, self.model[11].weight)
