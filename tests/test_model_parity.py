"""CPU oracle parity: our UNet / Loss vs the reference's own code (/root/reference), same state_dict."""
import importlib.util
import os
import sys
import types

import pytest
import torch

from distributedpytorch_b200.models import UNet, count_parameters
from distributedpytorch_b200.utils.loss import Loss

REF = "/root/reference"


def _load_reference():
    if not os.path.isdir(REF):
        pytest.skip("reference tree not mounted")
    # matplotlib is imported by the reference's utils/utils.py but is not installed here: stub it
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        mpl.pyplot = types.ModuleType("matplotlib.pyplot")
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = mpl.pyplot

    def load(name, path, package=None):
        spec = importlib.util.spec_from_file_location(name, path, submodule_search_locations=None)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
        return mod

    pkg = types.ModuleType("refmodel")
    pkg.__path__ = [os.path.join(REF, "model")]
    sys.modules["refmodel"] = pkg
    load("refmodel.unet_parts", os.path.join(REF, "model", "unet_parts.py"))
    ref_unet = load("refmodel.unet_model", os.path.join(REF, "model", "unet_model.py"))
    ref_utils = load("refutils_utils", os.path.join(REF, "utils", "utils.py"))
    return ref_unet.UNet, ref_utils.Loss


def test_state_dict_keys_shapes_and_param_count():
    RefUNet, _ = _load_reference()
    ours, ref = UNet(), RefUNet()
    sd_o, sd_r = ours.state_dict(), ref.state_dict()
    assert list(sd_o.keys()) == list(sd_r.keys())
    for k in sd_r:
        assert sd_o[k].shape == sd_r[k].shape, k
    assert count_parameters(ours) == 7_760_097


def test_forward_matches_reference_including_crop():
    RefUNet, _ = _load_reference()
    torch.manual_seed(0)
    ours, ref = UNet(), RefUNet()
    ours.load_state_dict(ref.state_dict())
    for shape in [(2, 3, 64, 96), (1, 3, 70, 100)]:  # the second exercises the CenterCrop of the skips
        x = torch.rand(*shape)
        with torch.no_grad():
            a, b = ours(x), ref(x)
        assert a.shape == b.shape
        assert torch.allclose(a, b, atol=1e-6), (a - b).abs().max()


def test_loss_matches_reference():
    _, RefLoss = _load_reference()
    torch.manual_seed(1)
    p = torch.rand(2, 1, 16, 24).clamp(1e-4, 1 - 1e-4)
    t = (torch.rand(2, 1, 16, 24) > 0.5).float()
    for dw in (1, 0, 0.5):
        assert torch.allclose(Loss(dw)(p, t), RefLoss(dw)(p.clone(), t), atol=1e-6)


def test_milesial_variant_parameter_count_and_shapes():
    m = UNet(variant="milesial", n_classes=2)
    assert count_parameters(m) == 31_037_698
    y = m(torch.rand(1, 3, 32, 48))
    assert y.shape == (1, 2, 32, 48)
    mb = UNet(variant="milesial", n_classes=2, bilinear=True)
    assert mb(torch.rand(1, 3, 32, 48)).shape == (1, 2, 32, 48)
