"""The train step at the other BASELINE patch sizes / batch shapes, one step each against the CPU oracle on the same
weights, inputs, dropout masks and GP alpha: configs[3] (512x512 patches, activation-heavy), the reference's own
default img_size 128 with an odd batch, batch 1 (every per-sample statistic degenerates to one image), and patch sizes
that are not powers of two (96, 160, 192: the reference CLI takes any multiple of 16, cgan_train_enhanced.py:147)."""
import pytest
import torch

import cgan_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,S", [(2, 512), (5, 128), (1, 256), (2, 96), (3, 160), (2, 192), (2, 48), (2, 80)])
def test_one_step_vs_oracle(B, S):
    from b200cgan.models import Discriminator, GeneratorUNet
    from b200cgan.train_step import CGANTrainer
    st = O.init_state(42)
    G, D = GeneratorUNet(0.3), Discriminator(True)
    G.load_state_dict(st["G"]); D.load_state_dict(st["D"])
    G, D = G.cuda(), D.cuda()
    tr = CGANTrainer(G, D, n_critic=2, lambda_gp=1.0)
    batch, masks, alphas = O.make_batch(B, S, 0)
    from test_models_gpu import attach_sources
    attach_sources(tr, masks, alphas)
    out = tr.step(*[batch[k].cuda() for k in ("pred", "gt", "refined", "pred_box", "delta_true")])
    got = out.to_host()
    ref = O.train_step(st, batch, masks, alphas, n_critic=2, lambda_gp=1.0)
    assert tuple(out.delta.shape) == (B, 4)
    for k in ("d_loss", "gp", "wd", "loss_iou", "loss_G", "loss_wgan", "g_gradnorm", "d_gradnorm_0", "d_gradnorm_1"):
        assert abs(got[k] - ref[k]) <= 3e-2 * abs(ref[k]) + 5e-3, (k, got[k], ref[k])
    rel = float((out.delta.float().cpu() - ref["delta"]).norm() / (ref["delta"].norm() + 1e-30))
    assert rel < 5e-2, rel
