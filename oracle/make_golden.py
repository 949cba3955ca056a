"""Generate tests/golden/*.json from the UNMODIFIED reference modules (build container only).

Imports /root/reference/cgan/{models,losses}.py, builds the nets exactly as the trainer does
(cgan/cgan_train_enhanced.py:229-232,253-257), and drives them with a restatement of the loop body
(:297-369) in which only the file-I/O helper ``get_refined_patch_batch`` is replaced by an injected
tensor, the three nn.Dropout modules of G by injected masks and ``torch.rand`` inside
``compute_gradient_penalty`` by an injected alpha.  The outputs pin
``oracle/cgan_oracle.py``; /root/reference does not exist on the GPU box, so only the JSON travels.

Run:  PYTHONDONTWRITEBYTECODE=1 python oracle/make_golden.py
"""
from __future__ import annotations

import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "oracle"))
sys.path.insert(0, "/root/reference/cgan")
sys.dont_write_bytecode = True

import models as ref_models  # noqa: E402  (the reference)
import losses as ref_losses  # noqa: E402
import cgan_oracle as O  # noqa: E402


class InjectedDropout(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.mask = None

    def forward(self, x):
        return x * self.mask * 2.0 if self.training else x


def build_reference(seed=42):
    torch.manual_seed(seed)
    G = ref_models.GeneratorUNet(delta_scale=0.3)
    D = ref_models.Discriminator(spectral_norm=True)
    G.apply(ref_models.weights_init_normal)
    D.apply(ref_models.weights_init_normal)
    drops = []
    for blk in (G.down4, G.up1, G.up2):
        blk.model[3] = InjectedDropout(); drops.append(blk.model[3])
    return G, D, drops


def tstats(t):
    t = t.detach().double()
    return [float(t.sum()), float(t.abs().sum()), float((t * t).sum())]


def run(B, S, steps, n_critic=2, lambda_gp=1.0):
    G, D, drops = build_reference()
    G.train(); D.train()
    crit = ref_losses.HybridLoss(lambda_iou=1.0)
    optG = torch.optim.Adam(G.parameters(), lr=2e-4, betas=(0.5, 0.999))
    optD = torch.optim.Adam(D.parameters(), lr=2e-4, betas=(0.5, 0.999))
    rec = {"B": B, "S": S, "n_critic": n_critic, "lambda_gp": lambda_gp, "steps": [],
           "init": {"G": {k: tstats(v) for k, v in G.state_dict().items()},
                    "D": {k: tstats(v) for k, v in D.state_dict().items()}},
           "n_params": {"G": sum(p.numel() for p in G.parameters()), "D": sum(p.numel() for p in D.parameters())}}
    for step in range(steps):
        batch, masks, alphas = O.make_batch(B, S, step)
        pred, gt, refined = batch["pred"], batch["gt"], batch["refined"]
        srec = {}
        if step == 0:
            # forward-only known answers on the initial weights (eval G; train D on a deep copy)
            import copy
            G.eval()
            with torch.no_grad():
                srec["g_eval_delta"] = G(pred).tolist()
            G.train()
            D2 = copy.deepcopy(D)
            with torch.no_grad():
                o = D2(pred, gt)
            srec["d_out_shape"] = list(o.shape)
            srec["d_out_stats"] = tstats(o)
            srec["d_out_first"] = o.flatten()[:8].tolist()
        dl = gpl = wdl = 0.0
        for c in range(n_critic):
            optD.zero_grad()
            real_v = D(pred, gt)
            with torch.no_grad():
                for dm, m in zip(drops, masks[c]):
                    dm.mask = m
                G(pred)
            fake_v = D(pred, refined)
            orig_rand = torch.rand
            torch.rand = lambda *a, **k: alphas[c].clone()
            try:
                gp = ref_losses.compute_gradient_penalty(D, (pred, gt), (pred, refined), torch.device("cpu"))
            finally:
                torch.rand = orig_rand
            wd = real_v.mean() - fake_v.mean()
            d_loss = -wd + lambda_gp * gp
            d_loss.backward()
            gn = torch.nn.utils.clip_grad_norm_(D.parameters(), max_norm=1.0)
            if step == 0 and c == 0:
                srec["d_grad_stats_clipped"] = {k: tstats(p.grad) for k, p in D.named_parameters()}
            optD.step()
            srec["d_gradnorm_%d" % c] = float(gn)
            dl += d_loss.item() / n_critic; gpl += gp.item() / n_critic; wdl += wd.item() / n_critic
        optG.zero_grad()
        for dm, m in zip(drops, masks[n_critic]):
            dm.mask = m
        delta = G(pred)
        cal = ref_losses.apply_delta_to_bbox(batch["pred_box"], delta, training=True)
        gtb = ref_losses.apply_delta_to_bbox(batch["pred_box"], batch["delta_true"], training=True)
        loss_reg, loss_iou = crit(delta, batch["delta_true"], cal, gtb)
        fv = D(pred, refined)
        loss_w = -fv.mean()
        loss_G = loss_reg + loss_w
        loss_G.backward()
        gn = torch.nn.utils.clip_grad_norm_(G.parameters(), max_norm=1.0)
        if step == 0:
            srec["g_grad_stats_clipped"] = {k: tstats(p.grad) for k, p in G.named_parameters()}
        optG.step()
        srec.update({"d_loss": dl, "gp": gpl, "wd": wdl, "loss_G": loss_G.item(), "loss_iou": loss_iou.item(),
                     "loss_wgan": loss_w.item(), "g_gradnorm": float(gn), "delta": delta.detach().tolist()})
        rec["steps"].append(srec)
    rec["final"] = {"G": {k: tstats(v) for k, v in G.state_dict().items()},
                    "D": {k: tstats(v) for k, v in D.state_dict().items()}}
    return rec


def loss_vectors():
    """Known answers for the (B,4) loss helpers, straight from cgan/losses.py."""
    gen = torch.Generator().manual_seed(7)
    box = torch.cat([torch.rand(16, 2, generator=gen) * 0.6 + 0.2, torch.rand(16, 2, generator=gen) * 0.3 + 0.1], 1)
    delta = torch.randn(16, 4, generator=gen) * 0.6
    dt = torch.randn(16, 4, generator=gen) * 0.3
    out = {"box": box.tolist(), "delta": delta.tolist(), "delta_true": dt.tolist()}
    for tr in (True, False):
        a = ref_losses.apply_delta_to_bbox(box, delta, training=tr)
        b = ref_losses.apply_delta_to_bbox(box, dt, training=tr)
        out["apply_%s" % tr] = a.tolist()
        out["eiou_%s" % tr] = float(ref_losses.EIoULoss()(a, b))
        out["iou_%s" % tr] = ref_losses.iou_metric(a, b).tolist()
    d = delta.clone().requires_grad_(True)
    l = ref_losses.EIoULoss()(ref_losses.apply_delta_to_bbox(box, d, True), ref_losses.apply_delta_to_bbox(box, dt, True))
    l.backward()
    out["eiou_grad_delta"] = d.grad.tolist()
    out["smooth_clamp"] = ref_losses.smooth_clamp(torch.linspace(-3, 3, 13), -1.5, 1.5).tolist()
    return out


if __name__ == "__main__":
    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    torch.set_num_threads(8)
    gold = ROOT / "tests" / "golden"
    gold.mkdir(parents=True, exist_ok=True)
    (gold / "losses.json").write_text(json.dumps(loss_vectors()))
    for B, S, steps in ((2, 64, 3), (2, 128, 2)):
        (gold / f"step_B{B}_S{S}.json").write_text(json.dumps(run(B, S, steps)))
        print("wrote", B, S)
