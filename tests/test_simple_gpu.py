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
    with pytest.raises(NotImplementedError):
        net(x.cuda())  # autograd enabled: no backward is built for this model
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
