"""The cGAN train step: n_critic critic updates (WGAN-GP) then one generator update (EIoU), restating the loop body
of the reference trainer (cgan/cgan_train_enhanced.py:297-369) over the B200 modules.

Differences from the reference loop, all outside the arithmetic:
  * `refined_patch` is either an input tensor (every BASELINE config: synthetic paired patches) or, with `refine_fn`,
    recomputed on the device from G's current delta in every critic step and in the generator step, as the reference
    does on the CPU with PIL (:311-315,358-360; b200cgan.refine is the CUDA crop + letterbox + bicubic kernel).  Without
    `refine_fn` the critic steps' no-grad generator forward (:311) has no consumer and is not executed; its dropout
    draws are skipped in the random stream so that the masks of the later forwards do not depend on that choice;
  * no .item() host syncs: logged scalars stay on the device until `StepOutput.to_host()` makes ONE copy;
  * the generator step skips D's backward (the reference computes it and zeroes it next step: dead work), but still
    runs D's forward in train mode, which advances the spectral-norm u/v buffers exactly like the reference (:361);
  * parameters, gradients and Adam moments of each network live in flat fp32 buckets (ops.FusedClipAdam): the weight-
    gradient kernels accumulate straight into the gradient bucket, zero_grad is one memset, clip_grad_norm_ + Adam are two
    launches per network, and the optional data-parallel all-reduce (torch.distributed) runs on the bucket itself with
    the 1/world_size folded into the Adam kernel;
  * optional CUDA-graph mode (`use_cuda_graph=True`): after two eager steps the whole step (about 480 kernel launches
    plus the autograd bookkeeping around them) is captured once and replayed; the dropout stream position and the Adam
    step counts live in a device-side counter block that is advanced between replays, so every replay draws fresh
    masks / interpolation coefficients and applies the right bias corrections.
"""
from __future__ import annotations

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


class GradBuckets:
    """Gradient all-reduce overlapped with backward (north_star).  The flat gradient bucket of a network is cut
    into contiguous slices in REVERSE parameter order - the order in which backward completes them (D: model.11 -> model.8 ->
    model.5 -> model.2 / model.0; G: fc -> up4 -> ... -> down1).  The weight-gradient kernels report every write into a
    parameter's sink (ops.set_sink_listener); the number of writes each parameter receives per backward is learned in the
    first step, and from then on a slice is SUM-all-reduced on a side stream the moment its last write has been enqueued
    (after that slice's spectral-norm rank-1 terms were flushed), while backward keeps running on the main stream.  finish()
    reduces whatever is left and joins the streams.  The 1/world factor is applied inside the Adam kernel."""

    def __init__(self, opt, group, min_bytes=None):
        import os
        if min_bytes is None:
            min_bytes = int(os.environ.get("B2C_DP_BUCKET_MB", "4")) << 20
        self.overlap = os.environ.get("B2C_DP_OVERLAP", "1") != "0"
        self.opt, self.group = opt, group
        self.index = {id(p): i for i, p in enumerate(opt.params)}
        self.expected = None            # writes per parameter and backward (learned)
        self.counts = [0] * len(opt.params)
        self.stream = torch.cuda.Stream(opt.flat_g.device)
        self.slices = []                # (lo, hi, [param indices]) in completion order
        idxs, lo_hi = [], None
        for i in reversed(range(len(opt.params))):
            o, n = opt.slices[i]
            lo_hi = (o, lo_hi[1]) if lo_hi else (o, o + (n + 3) // 4 * 4)
            idxs.append(i)
            if (lo_hi[1] - lo_hi[0]) * 4 >= min_bytes or i == 0:
                self.slices.append((lo_hi[0], lo_hi[1], idxs))
                idxs, lo_hi = [], None
        self.done = [False] * len(self.slices)
        self.learning = True

    def begin(self):
        self.counts = [0] * len(self.counts)
        self.done = [False] * len(self.slices)

    def on_write(self, param):
        i = self.index.get(id(param))
        if i is None:
            return
        self.counts[i] += 1
        if self.expected is None or not self.overlap:
            return
        for k, (lo, hi, idxs) in enumerate(self.slices):
            if self.done[k]:
                continue
            if all(self.counts[j] >= self.expected[j] for j in idxs):
                self._reduce(k)
            else:
                break  # slices complete in order

    def _reduce(self, k):
        import torch.distributed as dist
        lo, hi, idxs = self.slices[k]
        ops.flush_sn_pending([self.opt.params[j] for j in idxs])
        main = torch.cuda.current_stream()
        self.stream.wait_stream(main)
        with torch.cuda.stream(self.stream):
            dist.all_reduce(self.opt.flat_g[lo:hi], op=dist.ReduceOp.SUM, group=self.group)
        self.done[k] = True

    def finish(self):
        if self.expected is None:
            self.expected = list(self.counts)  # learned: from the next backward on, slices go out as they complete
        if not self.overlap:  # B2C_DP_OVERLAP=0: one all-reduce of the whole bucket after backward, on the main stream
            import torch.distributed as dist
            ops.flush_sn_pending(list(self.opt.params))
            dist.all_reduce(self.opt.flat_g, op=dist.ReduceOp.SUM, group=self.group)
            return
        for k in range(len(self.slices)):
            if not self.done[k]:
                self._reduce(k)
        torch.cuda.current_stream().wait_stream(self.stream)


class StepOutput:
    """Logged scalars of one step, kept on the device.  to_host() concatenates them and makes ONE device->host copy."""

    def __init__(self, n_critic, lambda_iou):
        self.n_critic, self.lambda_iou = n_critic, lambda_iou
        self.critic = []      # per critic step: 3-vector (d_loss, wd, gp)
        self.gnorm_sq = None  # (n_critic + 1,): squared pre-clip gradient norms of the D updates, then of the G update
        self.loss_iou = self.loss_wgan = None
        self.delta: Optional[torch.Tensor] = None

    def _decode(self, v):
        c = self.n_critic
        out = {"d_loss": sum(v[3 * i] for i in range(c)) / c, "wd": sum(v[3 * i + 1] for i in range(c)) / c,
               "gp": sum(v[3 * i + 2] for i in range(c)) / c}
        for i in range(c):
            out["d_gradnorm_%d" % i] = v[3 * c + i] ** 0.5
        out["g_gradnorm"] = v[4 * c] ** 0.5
        out["loss_iou"], out["loss_wgan"] = v[4 * c + 1], v[4 * c + 2]
        out["loss_G"] = self.lambda_iou * out["loss_iou"] + out["loss_wgan"]
        return out

    def _flat(self):
        return torch.cat([t.reshape(-1) for t in (*self.critic, self.gnorm_sq, self.loss_iou, self.loss_wgan)])

    def to_host(self):
        return self._decode(self._flat().tolist())

    @property
    def scalars(self) -> Dict[str, torch.Tensor]:
        """The same quantities as device tensors (no host sync)."""
        return self._decode(self._flat())


def masks_to_device(masks, device="cuda"):
    """{0,1} NCHW dropout masks (one per site) -> the dense u8 NHWC buffers DropoutSite.injected expects."""
    return [m.permute(0, 2, 3, 1).contiguous().to(torch.uint8).to(device).reshape(-1) for m in masks]


class CGANTrainer:
    def __init__(self, netG, netD, *, lr=2e-4, beta1=0.5, beta2=0.999, n_critic=2, lambda_gp=1.0, lambda_iou=1.0,
                 process_group=None, world_size=1, rank=None, use_cuda_graph=False, refine_fn=None):
        self.netG, self.netD = netG, netD
        self.n_critic, self.lambda_gp, self.lambda_iou = n_critic, lambda_gp, lambda_iou
        self.optG = ops.FusedClipAdam(netG.parameters(), lr=lr, betas=(beta1, beta2))
        self.optD = ops.FusedClipAdam(netD.parameters(), lr=lr, betas=(beta1, beta2))
        self.pg, self.world_size = process_group, world_size
        self.optG.grad_scale = self.optD.grad_scale = 1.0 / max(world_size, 1)
        self.refine_fn = refine_fn  # callable(delta (B,4)) -> refined patches (B,3,S,S); None: `refined` is an input
        # parity hooks: callable(critic_step) -> (B,) / (B,1,1,1) alpha, callable(g_call) -> 3 dense u8 NHWC masks, where
        # g_call counts the reference's generator forwards of a batch (critic step c -> c, generator step -> n_critic)
        self.alpha_source = None
        self.mask_source = None
        self.use_cuda_graph = use_cuda_graph
        self._graph = None        # captured step (torch.cuda.CUDAGraph)
        self._eager_steps = 0
        self._static_in = self._static_out = self._ctr = self._ctr_inc = None
        self._static_alpha = self._static_masks = None  # device buffers the captured graph reads the injected draws from
        dev = self.optD.flat_p.device
        self._seed3(dev); self._seed1(dev)  # backward seeds: created now, not inside a graph capture
        self._buckets = {}
        if world_size > 1:
            self._buckets = {id(self.optD): GradBuckets(self.optD, process_group), id(self.optG): GradBuckets(self.optG, process_group)}
            if rank is not None:
                rng.set_shard(rank, world_size)  # draws keyed by GLOBAL sample index: 1 x (N*B) == N x B

    # ---- data parallel: SUM the flat gradient buckets over ranks, slice by slice as backward completes them (GradBuckets);
    # ---- the 1/world is applied inside the Adam kernel ----
    def _allreduce(self, opt):
        if self.world_size > 1:
            self._buckets[id(opt)].finish()

    def _begin_backward(self, opt):
        if self.world_size > 1:
            b = self._buckets[id(opt)]
            b.begin()
            ops.set_sink_listener(b.on_write)

    def _seed3(self, dev):  # d/d (d_loss, wd, gp) = (1, 0, 0): persistent, so backward() launches no fill kernel
        if getattr(self, "_s3", None) is None or self._s3.device != dev:
            self._s3 = torch.tensor([1.0, 0.0, 0.0], dtype=torch.float32, device=dev)
        return self._s3

    def _seed1(self, dev):
        if getattr(self, "_s1", None) is None or self._s1.device != dev:
            self._s1 = torch.tensor(float(self.lambda_iou), dtype=torch.float32, device=dev)
        return self._s1

    # ---- random draws -------------------------------------------------------------------------------------------------
    def _alpha(self, c, B, dev):
        if self._static_alpha is not None:
            return self._static_alpha[c]
        if self.alpha_source is not None:
            return self.alpha_source(c).to(dev, torch.float32).reshape(-1)
        return None  # compute_gradient_penalty draws from the counter-based stream

    def _set_masks(self, gcall):
        sites = self.netG.dropout_sites()
        if self._static_masks is not None:
            ms = self._static_masks[gcall]
        elif self.mask_source is not None:
            ms = self.mask_source(gcall)
        else:
            ms = [None] * len(sites)
        for site, m in zip(sites, ms):
            site.injected = m

    def _skip_generator_draws(self, pred):
        """The reference's critic-step generator forward (:311) is not executed when nobody consumes its delta; keep the
        random stream where it would be."""
        if self.mask_source is None and self._static_masks is None:
            for n in self.netG.dropout_numels(pred.shape[0], pred.shape[2]):
                rng.skip(n)

    # ---- the two halves of a step ---------------------------------------------------------------------------------------
    def critic_step(self, pred, gt, refined, c=0, gnorm_out=None):
        netD, netG = self.netD, self.netG
        self.optD.zero_grad()
        real_validity = netD(pred, gt)
        if self.refine_fn is not None:
            self._set_masks(c)
            with torch.no_grad():
                refined = self.refine_fn(netG(pred))  # reference :311-315: fake patch from G's CURRENT delta
        else:
            self._skip_generator_draws(pred)
        fake_validity = netD(pred, refined)
        gp = losses.compute_gradient_penalty(netD, (pred, gt), (pred, refined), pred.device,
                                             alpha=self._alpha(c, pred.shape[0], pred.device))
        cl = ops.CriticLoss.apply(real_validity, fake_validity, gp, self.lambda_gp)  # (d_loss, wd, gp)
        self._begin_backward(self.optD)
        with ops.backward_flags(first_dgrad=False):  # nobody reads d loss / d interpolates
            torch.autograd.backward(cl, self._seed3(cl.device))  # d_loss = cl[0]
        ops.set_sink_listener(None)
        ops.flush_sn_pending()
        self._allreduce(self.optD)
        if gnorm_out is not None:
            self.optD.gnorm_sq = gnorm_out
        self.optD.step()
        return cl.detach()

    def generator_step(self, pred, refined, pred_box, delta_true, gnorm_out=None):
        netD, netG = self.netD, self.netG
        self.optG.zero_grad()
        self._set_masks(self.n_critic)
        delta_pred = netG(pred)
        calibrated = losses.apply_delta_to_bbox(pred_box, delta_pred, training=True)
        with torch.no_grad():
            gt_boxes = losses.apply_delta_to_bbox(pred_box, delta_true, training=True)
        loss_iou = ops.EIoU.apply(calibrated, gt_boxes, 1e-6)
        with torch.no_grad():
            if self.refine_fn is not None:
                refined = self.refine_fn(delta_pred.detach())  # reference :358-360
            loss_wgan = ops.MapMean.apply(netD(pred, refined), -1.0)  # D still runs in train mode (advances u / v)
        self._begin_backward(self.optG)
        torch.autograd.backward(loss_iou, self._seed1(loss_iou.device))  # loss_G's live part: lambda_iou * loss_iou
        ops.set_sink_listener(None)
        self._allreduce(self.optG)
        if gnorm_out is not None:
            self.optG.gnorm_sq = gnorm_out
        self.optG.step()
        return loss_iou.detach(), loss_wgan, delta_pred.detach()

    # ---- public entry point ----------------------------------------------------------------------------------------------
    def step(self, pred, gt, refined, pred_box, delta_true) -> StepOutput:
        """One train step.  Device tensors in, device scalars out (StepOutput.to_host() copies them to the host).
        `refined` may be None when a refine_fn is installed."""
        args = (pred, gt, refined, pred_box, delta_true)
        if not self.use_cuda_graph:
            return self._step_eager(*args)
        if self._graph is None:
            if self._eager_steps < 2:  # warm-up: function attributes, allocator, NCCL communicators
                return self._step_eager(*args)
            self._capture(args)
        elif any(a is not None and a.shape != s.shape for a, s in zip(args, self._static_in)):
            raise ValueError("CUDA-graph mode was captured for inputs of shape %s" % (tuple(self._static_in[0].shape),))
        for s, a in zip(self._static_in, args):
            if a is not None and s.data_ptr() != a.data_ptr():
                s.copy_(a, non_blocking=True)
        self._load_injected()
        # weights changed through torch since the last step (load_state_dict, manual edits): the captured step only re-packs
        # after its own optimizer updates, so refresh the stale bf16 packs here (no launch when nothing changed)
        ops.pack_all(self.optD.params)
        ops.pack_all(self.optG.params)
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
            self._static_alpha = self._static_masks = None
        self.use_cuda_graph = False

    def _load_injected(self):
        """Parity runs in graph mode: this step's injected alpha / masks go into the buffers the captured kernels read."""
        if self._static_alpha is not None:
            for c, buf in enumerate(self._static_alpha):
                buf.copy_(self.alpha_source(c).to(buf.device, torch.float32).reshape(-1), non_blocking=True)
        if self._static_masks is not None:
            for gcall, bufs in self._static_masks.items():
                for buf, m in zip(bufs, self.mask_source(gcall)):
                    buf.copy_(m, non_blocking=True)

    def _capture(self, args):
        dev = args[0].device
        self._static_in = [None if a is None else torch.empty_like(a) for a in args]
        for s, a in zip(self._static_in, args):
            if a is not None:
                s.copy_(a)
        B = args[0].shape[0]
        if self.alpha_source is not None:
            self._static_alpha = [torch.empty(B, dtype=torch.float32, device=dev) for _ in range(self.n_critic)]
        if self.mask_source is not None:
            live = ([*range(self.n_critic)] if self.refine_fn is not None else []) + [self.n_critic]
            self._static_masks = {g: [torch.empty_like(m) for m in self.mask_source(g)] for g in live}
        self._load_injected()
        # device counter block: [dropout stream position, optD steps so far, optG steps so far]
        st = rng.get_state()
        self._ctr = torch.tensor([st["offset"], self.optD.step_count, self.optG.step_count], dtype=torch.int64, device=dev)
        self.optD.host_steps, self.optG.host_steps = self.optD.step_count, self.optG.step_count
        rng.manual_seed(st["seed"], 0)
        rng.set_device_base(self._ctr[0:1])
        self.optD.step_base, self.optG.step_base = self._ctr[1:2], self._ctr[2:3]
        self.optD.step_count = self.optG.step_count = 0
        # every optimizer step re-packs its network's weights itself (ops.pack_all): the capture then contains exactly the
        # packs the replays need, provided the packs are current when it starts
        ops.pack_all(self.optD.params)
        ops.pack_all(self.optG.params)
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
        out = StepOutput(self.n_critic, self.lambda_iou)
        out.gnorm_sq = torch.empty(self.n_critic + 1, dtype=torch.float32, device=pred.device)
        for c in range(self.n_critic):
            out.critic.append(self.critic_step(pred, gt, refined, c, out.gnorm_sq[c:c + 1]))
        out.loss_iou, out.loss_wgan, out.delta = self.generator_step(pred, refined, pred_box, delta_true,
                                                                     out.gnorm_sq[self.n_critic:])
        return out
