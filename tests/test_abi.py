"""CPU-only checks of the drop-in boundary: libb2c.so loads, exports every symbol include/b2c.h declares, and the
Python binding refuses to run without a CUDA tensor (there is no CPU fallback)."""
import ctypes

import pytest
import torch

from b200cgan import _lib


def test_header_parses_and_library_exports_every_symbol():
    protos = _lib.parse_header()
    assert len(protos) >= 33 and "b2c_conv4s2_bf16" in protos and "b2c_multi_adam" in protos
    assert _lib.LIBPATH.exists(), "build with __graft_entry__.build()"
    lib = ctypes.CDLL(str(_lib.LIBPATH))
    missing = [n for n in protos if not hasattr(lib, n)]
    assert not missing, missing
    lib.b2c_abi_version.restype = ctypes.c_int
    assert lib.b2c_abi_version() == 2


def test_binding_has_no_cpu_fallback():
    x = torch.zeros(8)
    with pytest.raises(_lib.B2CError):
        _lib.call("b2c_fill_scaled", x, 8, x, 1.0)


def test_drop_in_surface_names_and_state_dict_keys():
    """Same class names / state-dict keys as cgan/models.py; construction needs no GPU."""
    from b200cgan import losses, models
    G, D = models.GeneratorUNet(0.3), models.Discriminator(True)
    assert sum(p.numel() for p in G.parameters()) == 6294788
    assert sum(p.numel() for p in D.parameters()) == 2767808
    gk = set(G.state_dict())
    assert gk == {f"down{i}.model.0.weight" for i in range(1, 5)} | {f"up{i}.model.0.weight" for i in range(1, 4)} | \
        {"up4.0.weight", "fc_delta.1.weight", "fc_delta.1.bias"}
    dk = set(D.state_dict())
    assert dk == {f"model.{i}.{s}" for i in (0, 2, 5, 8) for s in ("bias", "weight_orig", "weight_u", "weight_v")} | \
        {"model.11.weight"}
    assert set(models.Discriminator(False).state_dict()) == \
        {f"model.{i}.{s}" for i in (0, 2, 5, 8) for s in ("bias", "weight")} | {"model.11.weight"}
    G.apply(models.weights_init_normal); D.apply(models.weights_init_normal)
    assert float(D.model[0].bias.abs().max()) == 0.0 and abs(float(D.model[2].weight_orig.std()) - 0.02) < 2e-3
    for name in ("EIoULoss", "HybridLoss", "smooth_clamp", "apply_delta_to_bbox", "iou_metric", "compute_gradient_penalty"):
        assert hasattr(losses, name)


def test_environment_switches_are_documented():
    """Every B2C_* variable the library or the package reads appears in DESIGN.md's switches table, and vice versa."""
    import re
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    pkg = root / "gan-calibrated-semi-supervised-learning_b200"
    read = set()
    for f in list((pkg / "csrc").glob("*.cu")) + list((pkg / "csrc").glob("*.cuh")) + list((pkg / "b200cgan").glob("*.py")):
        text = f.read_text()
        read |= set(re.findall(r'(?:getenv|env_int|environ\.get)\(\s*"(B2C_[A-Z0-9_]+)"', text))
    design = (root / "DESIGN.md").read_text()
    table = design[design.index("### Switches"):design.index("## 4. Oracle and parity")]
    documented = set(re.findall(r"`(B2C_[A-Z0-9_]+)", table))
    assert read - documented == set(), sorted(read - documented)
    assert documented - read == set(), sorted(documented - read)
