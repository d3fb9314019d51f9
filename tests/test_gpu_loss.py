"""Fused L1 + SSIM loss kernels (csrc/loss.cu through the C ABI and the utils.loss_utils drop-in) against the
reference-made fixtures and the CPU oracle.  fp32 with a different (separable) summation order than cuDNN's/MKL's
2-D depthwise convolution: values within 2e-6 absolute, gradients within 1e-4 relative of the largest entry."""
import os

import numpy as np
import pytest
import torch

from harness import add_paths

add_paths()
from oracle import loss_reference as lr  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda"
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "loss_l1_ssim.npz"))


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-30))


@pytest.mark.parametrize("name", ["a", "b", "c", "d"])
def test_fused_loss_matches_reference_fixture(name):
    from utils import loss_utils as lu
    img = torch.from_numpy(GOLD[f"{name}_image"]).to(DEV).requires_grad_(True)
    gt = torch.from_numpy(GOLD[f"{name}_gt"]).to(DEV)
    lam = float(GOLD[f"{name}_lam"])
    loss, ll1 = lu.fused_l1_ssim_loss(img, gt, lam)
    loss.backward()
    assert abs(ll1.item() - float(GOLD[f"{name}_l1"])) <= 2e-6
    assert abs(loss.item() - float(GOLD[f"{name}_loss"])) <= 2e-6
    assert abs(lu.ssim(img.detach(), gt).item() - float(GOLD[f"{name}_ssim"])) <= 2e-6
    assert rel(img.grad.cpu(), torch.from_numpy(GOLD[f"{name}_grad"])) <= 1e-4
    assert not ll1.requires_grad


@pytest.mark.parametrize("name", ["m1", "m2"])
def test_masked_loss_and_depth_l1_match_reference_fixture(name):
    """fused_l1_ssim_loss(mask=alpha_mask) and masked_l1_loss against the reference-made fixtures
    (train_single.py:116-125 with `image *= alpha_mask`, and the depth loss of :133-139)."""
    from utils import loss_utils as lu
    img = torch.from_numpy(GOLD[f"{name}_image"]).to(DEV).requires_grad_(True)
    gt = torch.from_numpy(GOLD[f"{name}_gt"]).to(DEV)
    mask = torch.from_numpy(GOLD[f"{name}_mask"]).to(DEV)
    loss, ll1 = lu.fused_l1_ssim_loss(img, gt, float(GOLD[f"{name}_lam"]), mask=mask)
    loss.backward()
    assert abs(ll1.item() - float(GOLD[f"{name}_l1"])) <= 2e-6
    assert abs(loss.item() - float(GOLD[f"{name}_loss"])) <= 2e-6
    assert rel(img.grad.cpu(), torch.from_numpy(GOLD[f"{name}_grad"])) <= 1e-4
    inv = torch.from_numpy(GOLD[f"{name}_inv"]).to(DEV).requires_grad_(True)
    mono = torch.from_numpy(GOLD[f"{name}_mono"]).to(DEV)
    dmask = torch.from_numpy(GOLD[f"{name}_dmask"]).to(DEV)
    ld = lu.masked_l1_loss(inv, mono, dmask)
    (3.0 * ld).backward()
    assert abs(ld.item() - float(GOLD[f"{name}_depth_l1"])) <= 1e-7
    assert rel(inv.grad.cpu(), 3.0 * torch.from_numpy(GOLD[f"{name}_depth_grad"])) <= 1e-6
    # no mask = plain mean absolute difference
    assert abs(lu.masked_l1_loss(inv.detach(), mono).item() - torch.abs(inv.detach() - mono).mean().item()) <= 1e-7


def test_masked_loss_full_size_against_oracle():
    """1080p: masked photometric loss and depth loss against the CPU oracle on the same seeded inputs."""
    from utils import loss_utils as lu
    g = torch.Generator().manual_seed(21)
    gt = torch.rand(3, 1080, 1920, generator=g)
    img = (gt + 0.1 * torch.randn(3, 1080, 1920, generator=g)).clamp(0, 1)
    mask = (torch.rand(1, 1080, 1920, generator=g) > 0.2).float()
    a = img.clone().requires_grad_(True)
    want, _ = lr.l1_ssim_loss(a, gt, 0.2, mask=mask)
    want.backward()
    b = img.to(DEV).requires_grad_(True)
    got, _ = lu.fused_l1_ssim_loss(b, gt.to(DEV), 0.2, mask=mask.to(DEV))
    got.backward()
    assert abs(got.item() - want.item()) <= 2e-6
    assert rel(b.grad.cpu(), a.grad) <= 1e-4
    assert float(b.grad.cpu()[:, mask[0] == 0].abs().max()) == 0.0    # masked-out pixels receive nothing


@pytest.mark.parametrize("shape", [(3, 1080, 1920), (3, 137, 211), (1, 16, 16), (2, 3, 40, 56), (3, 5, 300)])
def test_ssim_matches_cpu_oracle(shape):
    from utils import loss_utils as lu
    g = torch.Generator().manual_seed(sum(shape))
    gt = torch.rand(shape, generator=g)
    img = (gt + 0.2 * torch.randn(shape, generator=g)).clamp(0, 1)
    a = img.clone().requires_grad_(True)
    b4 = gt if gt.dim() == 4 else gt.unsqueeze(0)
    want_map = lr.ssim_map(a if a.dim() == 4 else a.unsqueeze(0), b4)
    want = want_map.mean()
    up = torch.tensor(0.37)
    (want * up).backward()
    x = img.to(DEV).requires_grad_(True)
    got = lu.ssim(x, gt.to(DEV))
    (got * up.to(DEV)).backward()
    assert abs(got.item() - want.item()) <= 2e-6
    assert rel(x.grad.cpu(), a.grad) <= 1e-4
    m = lu.ssim_map(img.to(DEV), gt.to(DEV)).cpu()
    assert float((m.reshape(want_map.shape) - want_map.detach()).abs().max()) <= 2e-5
    if len(shape) == 4:   # size_average=False: one value per batch entry (loss_utils.py:62-63)
        per = lu.ssim(img.to(DEV), gt.to(DEV), size_average=False).cpu()
        assert torch.allclose(per, want_map.detach().mean(1).mean(1).mean(1), atol=2e-6)


def test_identical_images_and_argument_checks():
    from utils import loss_utils as lu
    x = torch.rand(3, 64, 64, device=DEV)
    assert abs(lu.ssim(x, x.clone()).item() - 1.0) <= 1e-6
    loss, ll1 = lu.fused_l1_ssim_loss(x, x.clone(), 0.2)
    assert abs(loss.item()) <= 1e-6 and ll1.item() == 0.0
    # what the kernels do not take goes through the reference's torch formula (CPU tensors, fp64, other windows)
    assert abs(lu.ssim(x.cpu(), x.cpu()).item() - 1.0) <= 1e-6
    assert abs(lu.ssim(x.double(), x.double()).item() - 1.0) <= 1e-9
    assert abs(lu.ssim(x, x.clone(), window_size=7).item() - 1.0) <= 1e-6
    with pytest.raises(RuntimeError):
        lu.ssim(x, x[:, :32])
    # the training losses have no such path
    with pytest.raises(ValueError):
        lu.fused_l1_ssim_loss(x.cpu(), x.cpu(), 0.2)
    with pytest.raises(TypeError):
        lu.fused_l1_ssim_loss(x.double(), x.double(), 0.2)
    # no gradient requested: no derivative maps are produced, the value is the same
    y = (x + 0.1).clamp(0, 1)
    with torch.no_grad():
        v0 = lu.ssim(y, x)
    v1 = lu.ssim(y.clone().requires_grad_(True), x)
    assert v0.item() == pytest.approx(v1.item(), abs=1e-7)


def test_ssim_gradient_of_second_image_and_other_windows():
    """The drop-in ``ssim`` differentiates BOTH images like the reference's torch version (the second through a
    swapped pass, SSIM being symmetric); window sizes other than 11 take the torch formula."""
    from utils import loss_utils as lu
    g = torch.Generator().manual_seed(31)
    a0 = torch.rand(3, 90, 130, generator=g).to(DEV)
    b0 = (a0 + 0.1 * torch.randn(3, 90, 130, generator=g).to(DEV)).clamp(0, 1)
    a, b = a0.clone().requires_grad_(True), b0.clone().requires_grad_(True)
    s = lu.ssim(a, b)
    s.backward()
    ar, br = a0.clone().requires_grad_(True), b0.clone().requires_grad_(True)
    sr = lu._ssim_torch(ar, br, 11, True)
    sr.backward()
    assert abs(float(s) - float(sr)) <= 2e-6
    for got, want in ((a.grad, ar.grad), (b.grad, br.grad)):
        assert got is not None
        assert float((got - want).abs().max()) <= 1e-4 * float(want.abs().max())
    assert abs(float(lu.ssim(a0, b0, 7)) - float(lu._ssim_torch(a0, b0, 7, True))) == 0.0


def test_exposure_clamp_matches_torch_ops():
    """The tail of render(): matmul(image.permute(1,2,0), E[:3,:3]).permute(2,0,1) + E[:3,3,None,None], clamp(0,1)
    (gaussian_renderer/__init__.py:115-118) fused; values and all gradients against the torch ops."""
    from gaussian_renderer import exposure_clamp
    g = torch.Generator().manual_seed(9)
    for H, W in ((37, 53), (1080, 1920)):
        img = (torch.rand(3, H, W, generator=g) * 1.6 - 0.3).to(DEV)          # values outside [0, 1] on both sides
        img[0, 0, 0], img[1, 0, 1] = 0.0, 1.0                                   # the bounds themselves pass gradient
        E = (torch.eye(3, 4) + 0.1 * torch.randn(3, 4, generator=g)).to(DEV)
        up = torch.randn(3, H, W, generator=g).to(DEV)
        for use_e in (True, False):
            a, ea = img.clone().requires_grad_(True), E.clone().requires_grad_(True)
            b, eb = img.clone().requires_grad_(True), E.clone().requires_grad_(True)
            ours = exposure_clamp(a, ea if use_e else None)
            pre = b
            if use_e:
                pre = torch.matmul(pre.permute(1, 2, 0), eb[:3, :3]).permute(2, 0, 1) + eb[:3, 3, None, None]
            ref = pre.clamp(0, 1)
            assert float((ours - ref).abs().max()) <= 5e-7
            (ours * up).sum().backward()
            (ref * up).sum().backward()
            # a pixel whose unclamped value sits within rounding of a bound may gate differently on the two sides
            away = ((pre.detach().abs() > 1e-6) & ((pre.detach() - 1).abs() > 1e-6)).all(dim=0)
            if not use_e:
                away = torch.ones_like(away)   # no arithmetic in front of the clamp: the gates are identical
            assert float(away.float().mean()) > 0.999
            assert float(((a.grad - b.grad).abs() * away).max()) <= 2e-6 * float(b.grad.abs().max())
            if use_e:
                assert float((ea.grad - eb.grad).abs().max()) <= 1e-4 * float(eb.grad.abs().max())
