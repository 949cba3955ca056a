"""Epoch-level driver around CGANTrainer: validation, LR-on-plateau, best-checkpoint, early stopping, NaN guard - the parts
of the reference trainer that surround its hot loop (cgan/cgan_train_enhanced.py:260-261,273-288,395-420,427-428,474-508),
restated over the B200 modules.  Dataset / wandb / argparse glue stays out.

  validate()            :395-420  generator in eval mode, hard-clamp apply_delta_to_bbox(training=False), iou_metric;
                                  the per-batch sums stay on the device, ONE host read per epoch (the reference does two
                                  .item() syncs per batch)
  ReduceLROnPlateau     :260-261,427-428  torch.optim.lr_scheduler.ReduceLROnPlateau(mode='max', factor=0.5, patience=5)
                                  restated for ops.FusedClipAdam, whose `lr` lives on the device: a reduction takes effect
                                  inside an already captured CUDA graph
  save_best_checkpoint  :481-489  the same dict ('generator', 'discriminator', 'epoch', 'delta_iou', 'config') that
                                  cgan/inference.py:104-147 (and b200cgan.inference.resolve_checkpoint) reads
  fit()                 :273-508  the epoch loop: train batches through CGANTrainer.step, validation, schedulers, history,
                                  NaN/Inf guard (:474-478), early stopping on delta-IoU (:481-503), training_history.json
"""
from __future__ import annotations

import json
import math
from pathlib import Path
from typing import Dict, Iterable, Optional, Sequence

import torch

from . import losses, ops


@torch.no_grad()
def validate(netG, val_batches: Iterable[Sequence[torch.Tensor]]) -> Dict[str, float]:
    """val_batches yields (pred_patch, delta_true, pred_box) device tensors.  -> mean IoU before / after calibration and
    their difference (cgan/cgan_train_enhanced.py:395-420)."""
    was_training = netG.training
    netG.eval()
    acc = None  # device [sum iou_before, sum iou_after]
    n = 0
    for pred_patch, delta_true, pred_box in val_batches:
        delta_pred = netG(pred_patch)
        calibrated = losses.apply_delta_to_bbox(pred_box, delta_pred, training=False)
        gt_boxes = losses.apply_delta_to_bbox(pred_box, delta_true, training=False)
        both = torch.stack([losses.iou_metric(pred_box, gt_boxes).sum(), losses.iou_metric(calibrated, gt_boxes).sum()])
        acc = both if acc is None else acc + both
        n += pred_patch.shape[0]
    if was_training:
        netG.train()
    before, after = (acc / max(1, n)).tolist() if acc is not None else (0.0, 0.0)
    return {"mean_iou_before": before, "mean_iou_after": after, "delta_iou": after - before}


class ReduceLROnPlateau:
    """torch.optim.lr_scheduler.ReduceLROnPlateau restated (same defaults: threshold 1e-4 'rel', cooldown 0, min_lr 0,
    eps 1e-8) for an optimizer exposing a settable `lr` (ops.FusedClipAdam).  The reference builds two of them with
    mode='max', factor=0.5, patience=5 and steps both with the validation delta-IoU."""

    def __init__(self, optimizer, mode="max", factor=0.5, patience=5, threshold=1e-4, threshold_mode="rel", cooldown=0,
                 min_lr=0.0, eps=1e-8):
        if factor >= 1.0:
            raise ValueError("Factor should be < 1.0.")
        self.optimizer, self.mode, self.factor, self.patience = optimizer, mode, factor, patience
        self.threshold, self.threshold_mode, self.cooldown, self.min_lr, self.eps = threshold, threshold_mode, cooldown, min_lr, eps
        self.best = -math.inf if mode == "max" else math.inf
        self.num_bad_epochs, self.cooldown_counter, self.last_epoch = 0, 0, 0

    def is_better(self, a, best):
        if self.mode == "min" and self.threshold_mode == "rel":
            return a < best * (1.0 - self.threshold)
        if self.mode == "min":
            return a < best - self.threshold
        if self.threshold_mode == "rel":
            return a > best * (self.threshold + 1.0)
        return a > best + self.threshold

    def step(self, metric):
        current = float(metric)
        self.last_epoch += 1
        if self.is_better(current, self.best):
            self.best, self.num_bad_epochs = current, 0
        else:
            self.num_bad_epochs += 1
        if self.cooldown_counter > 0:
            self.cooldown_counter -= 1
            self.num_bad_epochs = 0
        if self.num_bad_epochs > self.patience:
            old = self.optimizer.lr
            new = max(old * self.factor, self.min_lr)
            if old - new > self.eps:
                self.optimizer.lr = new
            self.cooldown_counter, self.num_bad_epochs = self.cooldown, 0


def save_best_checkpoint(path, netG, netD, epoch: int, delta_iou: float, config: Optional[dict]) -> None:
    """The reference's G_best.pth dict (cgan/cgan_train_enhanced.py:483-489); tensors are saved as plain CPU copies so the
    file loads anywhere (the live parameters are views of the optimizer's flat buckets)."""
    cpu = lambda sd: {k: v.detach().cpu().clone() for k, v in sd.items()}
    torch.save({"generator": cpu(netG.state_dict()), "discriminator": cpu(netD.state_dict()), "epoch": epoch,
                "delta_iou": delta_iou, "config": config or {}}, str(path))


def fit(trainer, train_batches, val_batches, *, n_epochs: int, save_dir, config: Optional[dict] = None, patience: int = 10,
        min_delta: float = 1e-4, log=print):
    """Epoch loop of cgan/cgan_train_enhanced.py:273-508 over a CGANTrainer.  train_batches / val_batches are callables
    returning an iterable of device-tensor tuples per epoch: (pred, gt, refined, pred_box, delta_true) and
    (pred, delta_true, pred_box).  Returns the training history (also written to save_dir/training_history.json)."""
    out_root = Path(save_dir)
    out_root.mkdir(parents=True, exist_ok=True)
    ckpt_best = out_root / "G_best.pth"
    sched_G = ReduceLROnPlateau(trainer.optG, mode="max", factor=0.5, patience=5)
    sched_D = ReduceLROnPlateau(trainer.optD, mode="max", factor=0.5, patience=5)
    best_val_iou, epochs_no_improve, history = -1.0, 0, []
    names = {"loss_G": "loss_G", "loss_D": "d_loss", "loss_iou": "loss_iou", "loss_wgan": "loss_wgan", "loss_gp": "gp",
             "wasserstein_distance": "wd"}
    for epoch in range(1, n_epochs + 1):
        trainer.netG.train(); trainer.netD.train()
        sums, nb = None, 0
        for batch in train_batches():
            out = trainer.step(*batch)
            flat = out._flat()  # device vector of this step's scalars; summed on the device, read once per epoch
            sums = flat.clone() if sums is None else sums + flat
            nb += 1
        stats = {k: 0.0 for k in names}
        if nb:
            dec = out._decode((sums / nb).tolist())
            stats = {k: dec[v] for k, v in names.items()}
        val = validate(trainer.netG, val_batches())
        delta_iou = val["delta_iou"]
        sched_G.step(delta_iou)
        sched_D.step(delta_iou)
        history.append({"epoch": epoch, **val, **stats, "lr_G": trainer.optG.lr, "lr_D": trainer.optD.lr})
        log(f"[Epoch {epoch}/{n_epochs}] G: {stats['loss_G']:.3f} D: {stats['loss_D']:.3f} EIoU: {stats['loss_iou']:.3f} "
            f"WGAN: {stats['loss_wgan']:.3f} GP: {stats['loss_gp']:.3f} WD: {stats['wasserstein_distance']:.3f} "
            f"dIoU: {delta_iou:.4f} Before: {val['mean_iou_before']:.4f} After: {val['mean_iou_after']:.4f}")
        if not (math.isfinite(stats["loss_G"]) and math.isfinite(stats["loss_D"])):
            log("Warning: NaN or Inf detected in losses! Stopping training to prevent further instability.")
            break
        if delta_iou > best_val_iou + min_delta:
            best_val_iou, epochs_no_improve = delta_iou, 0
            save_best_checkpoint(ckpt_best, trainer.netG, trainer.netD, epoch, delta_iou, config)
            log(f"New best model saved! Delta IoU = {best_val_iou:.4f}")
        else:
            epochs_no_improve += 1
            if epochs_no_improve >= patience:
                log("Early stopping triggered.")
                break
    with open(out_root / "training_history.json", "w") as f:
        json.dump(history, f, indent=2)
    return history
