"""The reference's own loop body (cgan/cgan_train_enhanced.py:297-369), line for line, over the drop-in modules: stock
`torch.optim.Adam`, `torch.nn.utils.clip_grad_norm_`, `.item()`, `compute_gradient_penalty(netD, real, fake, device)`
with the reference's four arguments.  This is INTEGRATION.md section 1's claim ("nothing else in the trainer changes"):
parameters receive ordinary `.grad` tensors (no gradient sinks, no fused optimizer, no CGANTrainer), the penalty runs
`torch.autograd.grad(create_graph=True)` through the kernels' double backward.  Checked against the CPU oracle on the
same dropout masks / alpha, and against `CGANTrainer` on the same batch."""
import pytest
import torch

import cgan_oracle as O
from test_models_gpu import KEYS9, attach_sources, build, cuda_batch, inject_masks, rel

pytestmark = pytest.mark.gpu


class Args:  # cgan/config.yaml defaults the reference's argparse would deliver
    n_critic, lambda_gp, lambda_iou, lr, beta1, beta2, img_size = 2, 1.0, 1.0, 2e-4, 0.5, 0.999, 128


def reference_loop_body(netG, netD, optimizer_G, optimizer_D, criterion_hybrid, args, batch, device, get_refined_patch_batch):
    """cgan_train_enhanced.py:297-369 with its own variable names; returns the epoch_stats increments."""
    from b200cgan.losses import apply_delta_to_bbox, compute_gradient_penalty
    pred_patch, gt_patch, delta_true, pred_box, img_path = batch
    d_loss_epoch = gp_loss_epoch = wd_epoch = 0.0
    gnorms = []
    for critic_step in range(args.n_critic):
        optimizer_D.zero_grad()
        real_validity = netD(pred_patch, gt_patch)
        with torch.no_grad():
            delta_pred_detached = netG(pred_patch)
            refined_patch = get_refined_patch_batch(img_path, pred_box, delta_pred_detached, args.img_size, device,
                                                    fallback_patches=pred_patch)
        fake_validity = netD(pred_patch, refined_patch)
        gradient_penalty = compute_gradient_penalty(netD, (pred_patch, gt_patch), (pred_patch, refined_patch), device)
        wasserstein_distance = torch.mean(real_validity) - torch.mean(fake_validity)
        d_loss = -wasserstein_distance + args.lambda_gp * gradient_penalty
        d_loss.backward()
        gnorms.append(float(torch.nn.utils.clip_grad_norm_(netD.parameters(), max_norm=1.0)))
        optimizer_D.step()
        d_loss_epoch += d_loss.item()
        gp_loss_epoch += gradient_penalty.item()
        wd_epoch += wasserstein_distance.item()
    optimizer_G.zero_grad()
    delta_pred = netG(pred_patch)
    calibrated_boxes = apply_delta_to_bbox(pred_box, delta_pred, training=True)
    gt_boxes = apply_delta_to_bbox(pred_box, delta_true, training=True)
    loss_G_reg, loss_iou = criterion_hybrid(delta_pred, delta_true, calibrated_boxes, gt_boxes)
    refined_patch_for_G = get_refined_patch_batch(img_path, pred_box, delta_pred, args.img_size, device,
                                                  fallback_patches=pred_patch)
    fake_validity_for_G = netD(pred_patch, refined_patch_for_G)
    loss_WGAN_G = -torch.mean(fake_validity_for_G)
    loss_G = loss_G_reg + loss_WGAN_G
    loss_G.backward()
    g_gradnorm = float(torch.nn.utils.clip_grad_norm_(netG.parameters(), max_norm=1.0))
    optimizer_G.step()
    return {"d_loss": d_loss_epoch / args.n_critic, "gp": gp_loss_epoch / args.n_critic, "wd": wd_epoch / args.n_critic,
            "loss_G": loss_G.item(), "loss_iou": loss_iou.item(), "loss_wgan": loss_WGAN_G.item(), "g_gradnorm": g_gradnorm,
            "d_gradnorm_0": gnorms[0], "d_gradnorm_1": gnorms[1], "delta": delta_pred.detach().float().cpu()}


@pytest.mark.parametrize("B,S", [(2, 64), (4, 128)])
def test_reference_loop_body_runs_unchanged_on_dropin_modules(B, S, monkeypatch):
    from b200cgan import rng
    from b200cgan.losses import HybridLoss
    from b200cgan.train_step import CGANTrainer
    args = Args()
    st = O.init_state(42)
    netG, netD = build(st)
    netG.train(); netD.train()
    optimizer_G = torch.optim.Adam(netG.parameters(), lr=args.lr, betas=(args.beta1, args.beta2))
    optimizer_D = torch.optim.Adam(netD.parameters(), lr=args.lr, betas=(args.beta1, args.beta2))
    criterion_hybrid = HybridLoss(lambda_iou=args.lambda_iou)

    steps = 2
    # a trainer on a second copy of the same weights steps the same batches through the fused path
    G2, D2 = build(st)
    tr = CGANTrainer(G2, D2, lr=args.lr, beta1=args.beta1, beta2=args.beta2, n_critic=args.n_critic, lambda_gp=args.lambda_gp,
                     lambda_iou=args.lambda_iou)
    for step in range(steps):
        batch, masks, alphas = O.make_batch(B, S, step)
        cb = cuda_batch(batch)
        # the reference's random draws, replaced by the oracle's: dropout masks per generator forward, alpha per penalty
        g_calls = {"n": 0}
        orig_forward = netG.forward

        def g_forward(x, _orig=orig_forward):
            inject_masks(netG, masks[g_calls["n"]])
            g_calls["n"] += 1
            return _orig(x)
        monkeypatch.setattr(netG, "forward", g_forward)
        a_calls = {"n": 0}

        def next_uniform(n, device):
            a = alphas[a_calls["n"]].reshape(-1).to(device)
            a_calls["n"] += 1
            return a
        monkeypatch.setattr(rng, "next_uniform", next_uniform)
        # the reference re-crops `refined` from the image files with PIL; the synthetic batch carries it (tests/golden do the same)
        refined = lambda img_path, pred_box, delta, img_size, device, fallback_patches=None: cb["refined"]

        got = reference_loop_body(netG, netD, optimizer_G, optimizer_D, criterion_hybrid, args,
                                  (cb["pred"], cb["gt"], cb["delta_true"], cb["pred_box"], None), "cuda", refined)
        monkeypatch.undo()
        assert g_calls["n"] == args.n_critic + 1 and a_calls["n"] == args.n_critic
        ref = O.train_step(st, batch, masks, alphas, n_critic=args.n_critic, lambda_gp=args.lambda_gp)
        for k in KEYS9:
            assert abs(got[k] - ref[k]) <= 3e-2 * abs(ref[k]) + 5e-3, (step, k, got[k], ref[k])
        assert rel(got["delta"], ref["delta"]) < 5e-2
        # ordinary autograd state: every parameter got a dense .grad of its own shape
        for p in list(netG.parameters()) + list(netD.parameters()):
            assert p.grad is not None and p.grad.shape == p.shape and p.grad.dtype == torch.float32

        attach_sources(tr, masks, alphas)
        fused = tr.step(cb["pred"], cb["gt"], cb["refined"], cb["pred_box"], cb["delta_true"]).to_host()
        for k in KEYS9:
            assert abs(got[k] - fused[k]) <= 2e-2 * abs(fused[k]) + 5e-3, (step, k, got[k], fused[k])

    # post-step weights: stock Adam + clip_grad_norm_ against the oracle's, and against the fused clip+Adam kernel
    for k, p in list(netG.state_dict().items()) + list(netD.state_dict().items()):
        if k in O.D_BIAS_UNOBSERVABLE:
            continue
        r = st["G"][k] if k in st["G"] else st["D"][k]
        assert rel(p, r) < 2e-2, k
    for (k, p), q in zip(list(netG.state_dict().items()) + list(netD.state_dict().items()),
                         list(G2.state_dict().values()) + list(D2.state_dict().values())):
        if k in O.D_BIAS_UNOBSERVABLE:
            continue
        assert rel(p, q) < 2e-2, k
