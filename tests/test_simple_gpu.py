"""GeneratorSimpleRegressor on the kernels: the 9-tap tcgen05 conv (b2c_conv3s1_bf16) and MaxPool2d kernels
against torch, and the whole eval forward against the oracle (pinned to the reference module by tests/test_simple_host.py),
at two patch sizes; the inference entry point accepts generator_type "simple" (cgan/inference.py:131-135)."""
import pytest
import torch
import torch.nn.functional as F

import cgan_oracle as O
from test_kernels_gpu import nchw, nhwc, q, rel
from test_simple_host import build_cpu

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,H,W,Cin,Cout", [(2, 16, 16, 64, 64), (3, 32, 32, 64, 128), (2, 8, 8, 256, 512), (1, 64, 64, 128, 128),
                                            (5, 4, 4, 512, 512), (2, 128, 128, 64, 64)])
def test_conv3s1_and_maxpool(N, H, W, Cin, Cout):
    from b200cgan import _lib as L
    g = torch.Generator().manual_seed(31)
    x = q(torch.randn(N, Cin, H, W, generator=g))
    w = torch.randn(Cout, Cin, 3, 3, generator=g) * 0.05
    b = torch.randn(Cout, generator=g)
    wp = torch.empty(Cout, 9 * Cin, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_pack_w_k3", w.cuda(), wp, Cout, Cin, 0)
    y = torch.empty(N, H, W, Cout, dtype=torch.bfloat16, device="cuda")
    zsi = torch.empty(N, Cout, 2, dtype=torch.int64, device="cuda")
    L.call("b2c_conv3s1_bf16", nhwc(x), Cin, wp, None, b.cuda(), y, Cout, zsi, N, H, W, Cin, Cout)
    ref = F.conv2d(x, q(w), b, 1, 1)
    assert rel(nchw(y), ref) < 1e-2
    yf = y.float()
    assert rel(zsi[:, :, 1].double() * 2.0 ** -30, (yf * yf).sum((1, 2))) < 1e-4
    # data-gradient form: conv with the flipped, transposed weights == conv_transpose of the same weights
    wf = torch.empty(Cin, 9 * Cout, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_pack_w_k3", w.cuda(), wf, Cout, Cin, 1)
    gy = q(torch.randn(N, Cout, H, W, generator=g))
    dx = torch.empty(N, H, W, Cin, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_conv3s1_bf16", nhwc(gy), Cout, wf, None, None, dx, Cin, None, N, H, W, Cout, Cin)
    assert rel(nchw(dx), torch.nn.grad.conv2d_input((N, Cin, H, W), q(w), gy, 1, 1)) < 1e-2
    # MaxPool2d(2,2) forward / backward
    a = q(torch.randn(N, Cout, H, W, generator=g)).requires_grad_(True)
    p = torch.empty(N, H // 2, W // 2, Cout, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_maxpool2_fwd", nhwc(a.detach()), p, N, H, W, Cout)
    pr = F.max_pool2d(a, 2, 2)
    assert torch.equal(nchw(p), pr.detach())
    gp = q(torch.randn(N, Cout, H // 2, W // 2, generator=g))
    pr.backward(gp)
    ga = torch.empty(N, H, W, Cout, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_maxpool2_bwd", nhwc(a.detach()), nhwc(gp), ga, N, H, W, Cout)
    assert torch.equal(nchw(ga), a.grad)


@pytest.mark.parametrize("B,S", [(3, 64), (2, 128)])
def test_simple_regressor_forward_vs_oracle(B, S):
    net = build_cpu()
    sd = {k: v.clone().float() for k, v in net.state_dict().items()}
    net = net.cuda().eval()
    x = torch.rand(B, 3, S, S, generator=torch.Generator().manual_seed(4)) * 2 - 1
    with torch.no_grad():
        got = net(x.cuda())
        ref = O.simple_forward(sd, x, 0.3)
        with O.emulate_bf16():
            ref16 = O.simple_forward(sd, x, 0.3)
    assert tuple(got.shape) == (B, 4)
    assert rel(got, ref) < 2e-2 and rel(got, ref16) < 2e-2, (rel(got, ref), rel(got, ref16))
    assert net(x.cuda()).requires_grad  # autograd enabled: the differentiable op (ops.SimpleRegressorFn)
    net.train()
    with torch.no_grad():
        d1, d2 = net(x.cuda()), net(x.cuda())
    assert not torch.equal(d1, d2)  # train mode: the head's Dropout(0.5) masks come from the counter-based stream


def test_inference_cli_accepts_simple_generator(tmp_path):
    import numpy as np
    from PIL import Image
    from b200cgan import inference as I
    net = build_cpu()
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    rng = np.random.default_rng(6)
    Image.fromarray(rng.integers(0, 256, size=(100, 140, 3), dtype=np.uint8), "RGB").save(tmp_path / "im.png")
    rows = [[1, 0.5, 0.5, 0.3, 0.4], [0, 0.3, 0.6, 0.2, 0.2, 0.77]]
    I.save_yolo_txt(tmp_path / "p.txt", rows)
    torch.save({"generator": sd, "config": {"delta_scale": 0.3, "generator_type": "simple"}}, tmp_path / "G_best.pth")
    I.main(["--weights", str(tmp_path / "G_best.pth"), "--image", str(tmp_path / "im.png"), "--pred_txt", str(tmp_path / "p.txt"),
            "--out_txt", str(tmp_path / "o.txt"), "--img_size", "64"])
    got = I.load_yolo_txt(tmp_path / "o.txt")
    img = Image.open(tmp_path / "im.png").convert("RGB")
    for g, r in zip(got, rows):
        patch = I.patch_to_tensor(I.letterbox(I.crop_patch(img, r[1:5]), 64)).unsqueeze(0)
        with torch.no_grad():
            d = O.simple_forward({k: v.float() for k, v in sd.items()}, patch, 0.3)[0]
        want = I.apply_delta_to_bbox_inference(r[1:5], d)
        assert np.allclose(g[1:5], want, atol=3e-3), (g, want)


@pytest.mark.parametrize("N,H,W,Cin,Cout", [(2, 16, 16, 64, 64), (3, 32, 32, 64, 128), (2, 8, 8, 256, 512), (5, 4, 4, 512, 512),
                                            (2, 64, 64, 128, 128)])
def test_wgrad3s1(N, H, W, Cin, Cout):
    """3x3 / stride-1 mode of the split-K tcgen05 weight-gradient kernel (Cout = 64: the zero-filled upper half of the M tile)."""
    from b200cgan import _lib as L
    g = torch.Generator().manual_seed(33)
    a = q(torch.randn(N, Cin, H, W, generator=g))
    gz = q(torch.randn(N, Cout, H, W, generator=g))
    ref = torch.nn.grad.conv2d_weight(a.double(), (Cout, Cin, 3, 3), gz.double(), 1, 1)
    for det in (False, True):
        nws = L.query("b2c_wgrad3s1_ws_floats", N, H, W, Cout, Cin, 148) if det else 9 * Cout * Cin
        ws = torch.empty(nws, dtype=torch.float32, device="cuda")
        dw = torch.full((Cout, Cin, 3, 3), 7.0, device="cuda")
        L.call("b2c_wgrad3s1_bf16", nhwc(gz), nhwc(a), ws, nws, dw, 0, N, H, W, Cout, Cin, 148)
        assert rel(dw, ref) < 1e-4, det
        L.call("b2c_wgrad3s1_bf16", nhwc(gz), nhwc(a), ws, nws, dw, 1, N, H, W, Cout, Cin, 148)
        assert rel(dw, 2 * ref) < 1e-4, det


@pytest.mark.parametrize("N,S", [(2, 32), (3, 64), (1, 256)])
def test_first_conv3x3_wgrad_and_head_bwd(N, S):
    from b200cgan import _lib as L
    g = torch.Generator().manual_seed(34)
    x = torch.randn(N, 3, S, S, generator=g)
    gz = q(torch.randn(N, 64, S, S, generator=g))
    dw = torch.zeros(64, 3, 3, 3, device="cuda")
    L.call("b2c_conv3x3_first_wgrad", x.cuda(), nhwc(gz), dw, N, S, S)
    assert rel(dw, torch.nn.grad.conv2d_weight(x.double(), (64, 3, 3, 3), gz.double(), 1, 1)) < 1e-4
    # regression head: forward with saved activations, then backward, against autograd
    hw = 4
    a = q(torch.randn(N, 512, 2, 2, generator=g)).requires_grad_(True)
    Ws = [torch.randn(256, 512, generator=g) * 0.05, torch.randn(256, generator=g) * 0.1, torch.randn(64, 256, generator=g) * 0.08,
          torch.randn(64, generator=g) * 0.1, torch.randn(4, 64, generator=g) * 0.2, torch.randn(4, generator=g) * 0.1]
    Ws = [w.requires_grad_(True) for w in Ws]
    m1, m2 = (torch.rand(N, 256, generator=g) < 0.5), (torch.rand(N, 64, generator=g) < 0.5)
    f = a.mean((2, 3))
    h1 = F.relu(F.linear(f, Ws[0], Ws[1])) * m1 * 2.0
    h2 = F.relu(F.linear(h1, Ws[2], Ws[3])) * m2 * 2.0
    ref = torch.tanh(F.linear(h2, Ws[4], Ws[5])) * 0.3
    gd = torch.randn(N, 4, generator=g)
    ref.backward(gd)
    c = [w.detach().cuda() for w in Ws]
    delta = torch.empty(N, 4, device="cuda")
    acts = torch.empty(N, 836, device="cuda")
    m1d, m2d = m1.to(torch.uint8).cuda().reshape(-1), m2.to(torch.uint8).cuda().reshape(-1)
    L.call("b2c_simple_head_fwd", nhwc(a.detach()), hw, c[0], c[1], c[2], c[3], c[4], c[5], m1d, m2d, 0.3, delta, acts, N)
    assert rel(delta, ref) < 1e-5
    gv = torch.empty(N, 836, device="cuda")
    ga = torch.empty(N, 2, 2, 512, dtype=torch.bfloat16, device="cuda")
    d = [torch.full_like(w, 3.0) for w in c]
    L.call("b2c_simple_head_bwd", gd.cuda(), acts, c[0], c[2], c[4], m1d, m2d, 0.3, hw, gv, ga, d[0], d[1], d[2], d[3], d[4], d[5], 0, N)
    for got, w in zip(d, Ws):
        assert rel(got, w.grad) < 1e-4
    assert rel(nchw(ga), a.grad) < 5e-3
    L.call("b2c_simple_head_bwd", gd.cuda(), acts, c[0], c[2], c[4], m1d, m2d, 0.3, hw, gv, ga, d[0], d[1], d[2], d[3], d[4], d[5], 1, N)
    assert rel(d[0], 2 * Ws[0].grad) < 1e-4


@pytest.mark.parametrize("B,S,train", [(3, 64, False), (2, 128, True)])
def test_simple_regressor_training_gradients_vs_oracle(B, S, train, monkeypatch):
    """Forward + backward of the whole network on the kernels (ops.SimpleRegressorFn) against autograd through the CPU oracle
    on bf16-rounded storage: delta and every weight gradient; train mode with the oracle's dropout masks."""
    from b200cgan import rng
    net = build_cpu()
    sd = {k: v.clone().float().requires_grad_(True) for k, v in net.state_dict().items()}
    net = net.cuda()
    net.train(train)
    gen = torch.Generator().manual_seed(5)
    x = torch.rand(B, 3, S, S, generator=gen) * 2 - 1
    r = torch.randn(B, 4, generator=gen)
    masks = None
    if train:
        masks = [(torch.rand(B, 256, generator=gen) < 0.5).float(), (torch.rand(B, 64, generator=gen) < 0.5).float()]
        queue = [m.to(torch.uint8).reshape(-1).cuda() for m in masks]
        monkeypatch.setattr(rng, "next_mask", lambda n, dev: queue.pop(0))
    delta = net(x.cuda())
    assert delta.requires_grad
    (delta * r.cuda()).sum().backward()
    with O.emulate_bf16():
        ref = O.simple_forward(sd, x, 0.3, masks)
    (ref * r).sum().backward()
    assert rel(delta, ref) < 2e-2
    worst = 0.0
    for k, p in net.named_parameters():
        if k.startswith("features") and k.endswith("bias"):
            assert float(p.grad.abs().max()) == 0.0  # feeds an InstanceNorm: exactly zero here, rounding noise in the oracle
            continue
        # end to end through 8 x (InstanceNorm -> ReLU) and 4 max-pools two bf16 pipelines flip a fraction of the kinks / argmaxes
        # (the same loose bound as the U-Net's end-to-end gradients; the kernels are held to 1e-4 .. 1e-2 above)
        assert rel(p.grad, sd[k].grad) < 0.35, (k, rel(p.grad, sd[k].grad))
        worst = max(worst, rel(p.grad, sd[k].grad))
    print("worst end-to-end gradient deviation", worst)


def test_simple_regressor_trains_under_cgan_trainer():
    """generator_type: simple through CGANTrainer (fused clip + Adam on gradient sinks): two steps run, losses are finite and
    the generator's weights move."""
    from b200cgan.models import Discriminator
    from b200cgan.train_step import CGANTrainer
    torch.manual_seed(0)
    G = build_cpu().cuda()
    D = Discriminator(True).cuda()
    w0 = G.features[3].weight.detach().clone()
    tr = CGANTrainer(G, D, n_critic=2, lambda_gp=1.0)
    batch, _, _ = O.make_batch(2, 64, 0)
    cb = {k: v.cuda() for k, v in batch.items()}
    for _ in range(2):
        out = tr.step(cb["pred"], cb["gt"], cb["refined"], cb["pred_box"], cb["delta_true"]).to_host()
    assert all(torch.isfinite(torch.tensor(v)) for v in out.values())
    assert float((G.features[3].weight.detach() - w0).abs().max()) > 0
