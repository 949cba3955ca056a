"""The cGAN train step: n_critic critic updates (WGAN-GP) then one generator update (EIoU), restating the loop body
of the reference trainer (cgan/cgan_train_enhanced.py:297-369) over the B200 modules.

Differences from the reference loop, all outside the arithmetic:
  * `refined_patch` is an input tensor (the reference re-crops it on the CPU with PIL from G's delta, :311-315,358-360,
    non-differentiably; SURVEY.md §2 #5 keeps that file I/O out of the hot path);
  * no .item() host syncs: logged scalars stay on the device until `StepOutput.to_host()` is called;
  * the generator step skips D's backward (the reference computes it and zeroes it next step: dead work), but still
    runs D's forward in train mode, which advances the spectral-norm u/v buffers exactly like the reference (:361);
  * clip_grad_norm_ + Adam are one fused multi-tensor update per network;
  * optional data-parallel gradient averaging (torch.distributed all-reduce) between backward and the update;
  * optional CUDA-graph mode (`use_cuda_graph=True`): after two eager steps the whole step (about 480 kernel launches
    plus the autograd bookkeeping around them) is captured once and replayed; the dropout stream position and the Adam
    step counts live in a device-side counter block that is advanced between replays, so every replay draws fresh
    masks / interpolation coefficients and applies the right bias corrections.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

from . import losses, ops, rng


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


class HostPrefetcher:
    """Double-buffered pinned-host -> device staging on a side stream, the data-loader half of the step: while step i
    runs, the H2D copy of batch i+1 is already in flight.  submit(host tensors) enqueues a copy, take() returns the
    device tensors of the oldest submitted batch (the current stream waits for its copy), release() marks them
    reusable once the consumer has been enqueued."""

    def __init__(self, example, device):
        self.stream = torch.cuda.Stream(device)
        self.bufs = [[torch.empty_like(t, device=device) for t in example] for _ in range(2)]
        self.ready = [torch.cuda.Event() for _ in range(2)]
        self.free = [torch.cuda.Event() for _ in range(2)]
        self.n_sub = self.n_take = 0

    def submit(self, host_tensors):
        b = self.n_sub % 2
        self.stream.wait_event(self.free[b])
        with torch.cuda.stream(self.stream):
            for d, h in zip(self.bufs[b], host_tensors):
                d.copy_(h, non_blocking=True)
            self.ready[b].record(self.stream)
        self.n_sub += 1

    def take(self):
        b = self.n_take % 2
        torch.cuda.current_stream().wait_event(self.ready[b])
        self.n_take += 1
        return self.bufs[b]

    def release(self):
        self.free[(self.n_take - 1) % 2].record(torch.cuda.current_stream())


@dataclass
class StepOutput:
    scalars: Dict[str, torch.Tensor] = field(default_factory=dict)
    delta: Optional[torch.Tensor] = None

    def to_host(self):
        return {k: float(v) for k, v in self.scalars.items()}


class CGANTrainer:
    def __init__(self, netG, netD, *, lr=2e-4, beta1=0.5, beta2=0.999, n_critic=2, lambda_gp=1.0, lambda_iou=1.0,
                 process_group=None, world_size=1, use_cuda_graph=False):
        self.netG, self.netD = netG, netD
        self.n_critic, self.lambda_gp, self.lambda_iou = n_critic, lambda_gp, lambda_iou
        self.optG = ops.FusedClipAdam(netG.parameters(), lr=lr, betas=(beta1, beta2))
        self.optD = ops.FusedClipAdam(netD.parameters(), lr=lr, betas=(beta1, beta2))
        self.pg, self.world_size = process_group, world_size
        self.alpha_source = None  # test hook: callable(critic_step) -> (B,1,1,1) tensor replacing torch.rand
        self.use_cuda_graph = use_cuda_graph
        self._graph = None        # captured step (torch.cuda.CUDAGraph)
        self._eager_steps = 0
        self._static_in = self._static_out = self._ctr = self._ctr_inc = None

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
        """One train step.  Device tensors in, device scalars out (StepOutput.to_host() copies them to the host)."""
        args = (pred, gt, refined, pred_box, delta_true)
        if not self.use_cuda_graph or self.alpha_source is not None or \
                any(s.injected is not None for s in self.netG.dropout_sites()):
            return self._step_eager(*args)
        if self._graph is None:
            if self._eager_steps < 2:  # warm-up: function attributes, allocator, NCCL communicators
                return self._step_eager(*args)
            self._capture(args)
        elif any(a.shape != s.shape for a, s in zip(args, self._static_in)):
            raise ValueError("CUDA-graph mode was captured for inputs of shape %s" % (tuple(self._static_in[0].shape),))
        for s, a in zip(self._static_in, args):
            if s.data_ptr() != a.data_ptr():
                s.copy_(a, non_blocking=True)
        self._graph.replay()
        self._ctr.add_(self._ctr_inc)  # dropout stream position, D / G Adam step counts
        self.optD.host_steps += self.n_critic
        self.optG.host_steps += 1
        return self._static_out

    def static_inputs(self):
        """The captured step's input buffers (pred, gt, refined, pred_box, delta_true), or None before capture: a
        caller may write the next batch straight into them (e.g. the H2D copy) and pass them to step()."""
        return self._static_in

    def disable_cuda_graph(self):
        """Back to eager stepping (e.g. to bracket individual kernels with events): the device counters are folded back
        into the host-side optimizer step counts and dropout stream position (one host sync)."""
        if self._graph is not None:
            pos, dsteps, gsteps = (int(v) for v in self._ctr.tolist())
            rng.set_device_base(None)
            rng.manual_seed(rng.get_state()["seed"], pos)
            self.optD.step_base = self.optG.step_base = None
            self.optD.step_count, self.optG.step_count = dsteps, gsteps
            self._graph = None
        self.use_cuda_graph = False

    def _capture(self, args):
        dev = args[0].device
        self._static_in = [torch.empty_like(a) for a in args]
        for s, a in zip(self._static_in, args):
            s.copy_(a)
        # device counter block: [dropout stream position, optD steps so far, optG steps so far]
        st = rng.get_state()
        self._ctr = torch.tensor([st["offset"], self.optD.step_count, self.optG.step_count], dtype=torch.int64, device=dev)
        self.optD.host_steps, self.optG.host_steps = self.optD.step_count, self.optG.step_count
        rng.manual_seed(st["seed"], 0)
        rng.set_device_base(self._ctr[0:1])
        self.optD.step_base, self.optG.step_base = self._ctr[1:2], self._ctr[2:3]
        self.optD.step_count = self.optG.step_count = 0
        torch.cuda.synchronize()
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._static_out = self._step_eager(*self._static_in)
        # host-side ordinals recorded in the graph are relative to the start of a step; freeze them
        per_step = rng.get_state()["offset"]
        self._ctr_inc = torch.tensor([per_step, self.n_critic, 1], dtype=torch.int64, device=dev)
        rng.manual_seed(st["seed"], 0)
        self.optD.step_count = self.optG.step_count = 0

    def _step_eager(self, pred, gt, refined, pred_box, delta_true) -> StepOutput:
        self._eager_steps += 1
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
