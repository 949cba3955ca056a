"""b2c_in_act_bwd_static: the allocation-free variant of b2c_in_act_bwd (scratch in a file-scope device array of the
library, address resolved with cudaGetSymbolAddress).  It has no counterpart in the reference; it exists for callers that
cannot allocate (CUDA graph capture, serving loops).  Checked through the C-ABI against the existing entry point on the same
inputs, on two calls in a row (the second with different data and a different shape: the shared array keeps the previous
call's contents and must be cleared by the call itself), and for the size limit.  Collected last (file name), runs alone too."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_PLANES = 131072  # include/b2c.h: N * C <= 131072


def _nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()


def _rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _case(L, N, C, H, slope, drop, seed):
    g = torch.Generator().manual_seed(seed)
    HW = H * H
    z = (torch.randn(N, C, H, H, generator=g) * 1.5 + 0.3).to(torch.bfloat16).float()
    ga = torch.randn(N, C, H, H, generator=g).to(torch.bfloat16).float()
    mk = (torch.rand(N, H, H, C, generator=g) < 0.5).to(torch.uint8).cuda() if drop else None
    zd, gd = _nhwc(z), _nhwc(ga)
    zs = torch.empty(N, C, 2, dtype=torch.int64, device="cuda")
    L.call("b2c_in_stats", zd, C, zs, N, HW, C)
    ws = torch.empty(N, C, 2, dtype=torch.float32, device="cuda")
    ref = torch.empty(N, H, H, C, dtype=torch.bfloat16, device="cuda")
    L.call("b2c_in_act_bwd", gd, C, None, None, zd, C, zs, mk, slope, ws, ref, C, N, HW, C)
    got = torch.full((N, H, H, C), float("nan"), dtype=torch.bfloat16, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    L.call("b2c_in_act_bwd_static", gd, C, zd, C, zs, mk, slope, got, C, N, HW, C)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before  # nothing went through torch's allocator
    assert bool(torch.isfinite(got.float()).all())
    # same arithmetic, different summation order of the plane sums (cluster / fused kernels vs two-pass with fp32 atomics):
    # the tolerance of tests/test_kernels2_gpu.py for this kernel's bf16 output
    assert _rel(got.float(), ref.float()) < 1e-2, (N, C, H, slope, drop)


def test_static_scratch_variant_matches_existing_entry_point_twice():
    from b200cgan import _lib as L
    # first call: big planes (the existing entry point takes its cluster kernel here), large plane sums left in the array
    _case(L, 4, 128, 64, 0.2, False, 1)
    # second call in a row: other data, other shape, dropout mask; then a third with small planes
    _case(L, 3, 256, 16, 0.0, True, 2)
    _case(L, 2, 512, 4, 0.2, False, 3)


def test_static_scratch_variant_rejects_sizes_beyond_the_limit():
    from b200cgan import _lib as L
    C, H = 512, 2
    N = MAX_PLANES // C + 1   # N * C = 131584 > 131072
    z = torch.zeros(N, H, H, C, dtype=torch.bfloat16, device="cuda")
    zs = torch.zeros(N, C, 2, dtype=torch.int64, device="cuda")
    out = torch.empty_like(z)
    with pytest.raises(L.B2CError, match="exceed the static scratch"):
        L.call("b2c_in_act_bwd_static", z, C, z, C, zs, None, 0.2, out, C, N, H * H, C)
    # exactly at the limit is accepted (a constant plane: zero gradient out)
    N = MAX_PLANES // C
    z = torch.ones(N, H, H, C, dtype=torch.bfloat16, device="cuda")
    zs = torch.zeros(N, C, 2, dtype=torch.int64, device="cuda")
    L.call("b2c_in_stats", z, C, zs, N, H * H, C)
    out = torch.full_like(z, float("nan"))
    L.call("b2c_in_act_bwd_static", z, C, z, C, zs, None, 0.2, out, C, N, H * H, C)
    torch.cuda.synchronize()
    assert float(out.float().abs().max()) < 1e-3
