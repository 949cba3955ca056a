"""Drop-in loss functions for the reference's cgan/losses.py (same names, arguments and return values), computed by
the fused kernels in csrc/loss.cu through ops.py."""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops


class EIoULoss(nn.Module):
    """1 - mean(IoU - rho^2(c)/c^2 - rho^2(w)/c_w^2 - rho^2(h)/c_h^2)   (cgan/losses.py:10-73)."""

    def __init__(self, eps=1e-6):
        super().__init__()
        self.eps = eps

    def forward(self, pred_boxes, target_boxes):
        return ops.EIoU.apply(pred_boxes, target_boxes, self.eps)


class HybridLoss(nn.Module):
    """Pure-EIoU regression loss; returns (lambda_iou * L, L)   (cgan/losses.py:75-97)."""

    def __init__(self, lambda_iou=1.0, **kwargs):
        super().__init__()
        self.lambda_iou = lambda_iou
        self.iou_loss = EIoULoss()
        print(f"Using Pure EIoU Loss (lambda_iou={lambda_iou})")

    def forward(self, pred_deltas, target_deltas, pred_boxes, target_boxes):
        iou_loss = self.iou_loss(pred_boxes, target_boxes)
        return self.lambda_iou * iou_loss, iou_loss


def smooth_clamp(x, min_val, max_val, temperature=0.5):
    """Sigmoid squashing into (min_val, max_val)   (cgan/losses.py:99-106).  Stand-alone helper (torch elementwise);
    the train step uses the fused apply_delta kernel instead."""
    center = (min_val + max_val) / 2
    return min_val + (max_val - min_val) * torch.sigmoid((x - center) / temperature)


def apply_delta_to_bbox(bbox, delta, training=True):
    """(N,4) boxes + (N,4) deltas -> calibrated boxes   (cgan/losses.py:108-150)."""
    return ops.ApplyDelta.apply(bbox, delta, bool(training))


def iou_metric(pred_boxes, target_boxes, eps=1e-6):
    """Row-wise IoU   (cgan/losses.py:152-183)."""
    return ops.iou_rows(pred_boxes, target_boxes, eps)


def compute_gradient_penalty(discriminator, real_samples, fake_samples, device, alpha=None):
    """WGAN-GP penalty on per-sample interpolates of (pred, other) pairs   (cgan/losses.py:185-233).
    The double backward runs through ops.Conv4s2Dgrad / INActBwd / FirstDgrad / HeadDgrad.

    alpha (optional, (B,1,1,1) or (B,)): the interpolation coefficients; default: B draws from the counter-based stream
    (b200cgan.rng, keyed by global sample index), where the reference calls torch.rand (cgan/losses.py:199).
    For a b200cgan Discriminator the interpolates are never materialised: alpha * real + (1 - alpha) * fake is formed
    inside the first conv's gather (and again inside its weight-gradient kernel), and the tensors differentiated
    against are uninitialised placeholders of the interpolates' shape.  During the inputs-only first backward the
    weight-gradient branches are switched off (they would be computed and thrown away)."""
    from . import rng
    from .models import Discriminator
    batch_size = real_samples[0].size(0)
    dev = real_samples[0].device
    if alpha is None:
        alpha = rng.next_uniform(batch_size, dev)
    alpha = alpha.to(dev)
    if isinstance(discriminator, Discriminator):
        interp = ops.Interp(real_samples[0], real_samples[1], fake_samples[0], fake_samples[1], alpha)
        ip = torch.empty_like(interp.a0).requires_grad_(True)  # placeholders: only their autograd identity is used
        io = torch.empty_like(interp.a1).requires_grad_(True)
        d_interpolated = discriminator(ip, io, _interp=interp)
    else:
        a = alpha.reshape(-1, 1, 1, 1)
        ip = torch.lerp(fake_samples[0], real_samples[0], a).detach().requires_grad_(True)
        io = torch.lerp(fake_samples[1], real_samples[1], a).detach().requires_grad_(True)
        d_interpolated = discriminator(ip, io)
    ones = torch.empty_like(d_interpolated)
    ops.fill_(ones, 1.0)
    with ops.backward_flags(weight_grads=False):
        g0, g1 = torch.autograd.grad(outputs=d_interpolated, inputs=[ip, io], grad_outputs=ones,
                                     create_graph=True, retain_graph=True, only_inputs=True)
    return ops.GradNormPenalty.apply(g0, g1)
