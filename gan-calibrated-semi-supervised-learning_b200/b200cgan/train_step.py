"""The cGAN train step: n_critic critic updates (WGAN-GP) then one generator update (EIoU), restating the loop body
of the reference trainer (cgan/cgan_train_enhanced.py:297-369) over the B200 modules.

Differences from the reference loop, all outside the arithmetic:
  * `refined_patch` is an input tensor (the reference re-crops it on the CPU with PIL from G's delta, :311-315,358-360,
    non-differentiably; SURVEY.md §2 #5 keeps that file I/O out of the hot path);
  * no .item() host syncs: logged scalars stay on the device until `StepOutput.to_host()` is called;
  * the generator step skips D's backward (the reference computes it and zeroes it next step: dead work), but still
    runs D's forward in train mode, which advances the spectral-norm u/v buffers exactly like the reference (:361);
  * clip_grad_norm_ + Adam are one fused multi-tensor update per network;
  * optional data-parallel gradient averaging (torch.distributed all-reduce) between backward and the update.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

from . import losses, ops


def allreduce_mean_(grads, group=None, world_size=None):
    """In-place average of a list of gradient tensors over the ranks of `group` with ONE collective on a flat bucket
    (every loss is a batch mean and every per-sample quantity is rank-local, so the average of the
    per-shard gradients equals the single-GPU gradient on the concatenated batch).  Backend-agnostic (NCCL on the
    B200s, gloo in the CPU tests)."""
    import torch.distributed as dist
    if world_size is None:
        world_size = dist.get_world_size(group)
    if world_size <= 1 or not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    flat.div_(world_size)
    off = 0
    for g in grads:
        g.copy_(flat[off:off + g.numel()].view_as(g))
        off += g.numel()


@dataclass
class StepOutput:
    scalars: Dict[str, torch.Tensor] = field(default_factory=dict)
    delta: Optional[torch.Tensor] = None

    def to_host(self):
        return {k: float(v) for k, v in self.scalars.items()}


class CGANTrainer:
    def __init__(self, netG, netD, *, lr=2e-4, beta1=0.5, beta2=0.999, n_critic=2, lambda_gp=1.0, lambda_iou=1.0,
                 process_group=None, world_size=1):
        self.netG, self.netD = netG, netD
        self.n_critic, self.lambda_gp, self.lambda_iou = n_critic, lambda_gp, lambda_iou
        self.optG = ops.FusedClipAdam(netG.parameters(), lr=lr, betas=(beta1, beta2))
        self.optD = ops.FusedClipAdam(netD.parameters(), lr=lr, betas=(beta1, beta2))
        self.pg, self.world_size = process_group, world_size
        self.alpha_source = None  # test hook: callable(critic_step) -> (B,1,1,1) tensor replacing torch.rand

    # ---- data parallel: average gradients over ranks (per-sample losses are batch means) ----
    def _allreduce(self, params):
        if self.world_size > 1:
            allreduce_mean_([p.grad for p in params if p.grad is not None], self.pg, self.world_size)

    def critic_step(self, pred, gt, refined, c=0):
        netD, netG = self.netD, self.netG
        self.optD.zero_grad()
        real_validity = netD(pred, gt)
        with torch.no_grad():
            netG(pred)  # reference :311: train-mode forward; its delta only feeds the CPU re-crop we take as input
        fake_validity = netD(pred, refined)
        if self.alpha_source is not None:
            orig = torch.rand
            torch.rand = lambda *a, **k: self.alpha_source(c)
            try:
                gp = losses.compute_gradient_penalty(netD, (pred, gt), (pred, refined), pred.device)
            finally:
                torch.rand = orig
        else:
            gp = losses.compute_gradient_penalty(netD, (pred, gt), (pred, refined), pred.device)
        wd = ops.MapMean.apply(real_validity) - ops.MapMean.apply(fake_validity)
        d_loss = -wd + self.lambda_gp * gp
        d_loss.backward()
        self._allreduce(self.optD.params)
        gn = self.optD.step()
        return d_loss.detach(), gp.detach(), wd.detach(), gn.clone()

    def generator_step(self, pred, refined, pred_box, delta_true):
        netD, netG = self.netD, self.netG
        self.optG.zero_grad()
        delta_pred = netG(pred)
        calibrated = losses.apply_delta_to_bbox(pred_box, delta_pred, training=True)
        with torch.no_grad():
            gt_boxes = losses.apply_delta_to_bbox(pred_box, delta_true, training=True)
        loss_iou = ops.EIoU.apply(calibrated, gt_boxes, 1e-6)
        with torch.no_grad():
            loss_wgan = -ops.MapMean.apply(netD(pred, refined))
        loss_reg = self.lambda_iou * loss_iou
        loss_reg.backward()
        self._allreduce(self.optG.params)
        gn = self.optG.step()
        return (loss_reg.detach() + loss_wgan), loss_iou.detach(), loss_wgan, gn.clone(), delta_pred.detach()

    def step(self, pred, gt, refined, pred_box, delta_true) -> StepOutput:
        self.netG.train(); self.netD.train()
        out = StepOutput()
        acc = None
        for c in range(self.n_critic):
            r = self.critic_step(pred, gt, refined, c)
            out.scalars["d_gradnorm_%d" % c] = r[3].sqrt().reshape(())
            acc = list(r[:3]) if acc is None else [a + b for a, b in zip(acc, r[:3])]
        for k, v in zip(("d_loss", "gp", "wd"), acc):
            out.scalars[k] = v / self.n_critic
        lg, liou, lw, gn, delta = self.generator_step(pred, refined, pred_box, delta_true)
        out.scalars.update({"loss_G": lg, "loss_iou": liou, "loss_wgan": lw, "g_gradnorm": gn.sqrt().reshape(())})
        out.delta = delta
        return out
