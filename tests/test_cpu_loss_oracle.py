"""The CPU restatement of the photometric loss (oracle/loss_reference.py) against fixtures produced by the
reference's own utils/loss_utils.py (tests/golden/make_golden_loss.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import loss_reference as lr

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "loss_l1_ssim.npz"))


@pytest.mark.parametrize("name", ["a", "b", "c", "d"])
def test_loss_oracle_matches_reference_fixture(name):
    img = torch.from_numpy(GOLD[f"{name}_image"]).requires_grad_(True)
    gt = torch.from_numpy(GOLD[f"{name}_gt"])
    lam = float(GOLD[f"{name}_lam"])
    loss, ll1 = lr.l1_ssim_loss(img, gt, lam)
    loss.backward()
    assert abs(ll1.item() - float(GOLD[f"{name}_l1"])) <= 1e-7
    assert abs(lr.ssim(img.detach(), gt).item() - float(GOLD[f"{name}_ssim"])) <= 1e-6
    assert abs(loss.item() - float(GOLD[f"{name}_loss"])) <= 1e-6
    np.testing.assert_allclose(img.grad.numpy(), GOLD[f"{name}_grad"], rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("name", ["m1", "m2"])
def test_masked_loss_and_depth_l1_oracle_match_reference_fixture(name):
    """Alpha-masked photometric loss (train_single.py:116-125) and the depth loss (:133-139)."""
    img = torch.from_numpy(GOLD[f"{name}_image"]).requires_grad_(True)
    gt, mask = torch.from_numpy(GOLD[f"{name}_gt"]), torch.from_numpy(GOLD[f"{name}_mask"])
    loss, ll1 = lr.l1_ssim_loss(img, gt, float(GOLD[f"{name}_lam"]), mask=mask)
    loss.backward()
    assert abs(ll1.item() - float(GOLD[f"{name}_l1"])) <= 1e-7
    assert abs(loss.item() - float(GOLD[f"{name}_loss"])) <= 1e-6
    np.testing.assert_allclose(img.grad.numpy(), GOLD[f"{name}_grad"], rtol=1e-5, atol=1e-9)
    inv = torch.from_numpy(GOLD[f"{name}_inv"]).requires_grad_(True)
    ld = lr.depth_l1(inv, torch.from_numpy(GOLD[f"{name}_mono"]), torch.from_numpy(GOLD[f"{name}_dmask"]))
    ld.backward()
    assert abs(ld.item() - float(GOLD[f"{name}_depth_l1"])) <= 1e-7
    np.testing.assert_array_equal(inv.grad.numpy(), GOLD[f"{name}_depth_grad"])


def test_window_is_the_reference_window():
    w = lr.gaussian(11, 1.5)
    assert abs(float(w.sum()) - 1.0) < 1e-6 and torch.equal(w, w.flip(0)) and float(w[5]) == float(w.max())
    assert lr.create_window(11, 3).shape == (3, 1, 11, 11)


def test_torch_formula_paths_of_the_drop_in_match_reference_fixture():
    """utils.loss_utils on inputs the kernels do not take (CPU tensors, window 7, per-batch means) and the
    evaluation-only ssim_masked, against outputs of the reference's own module
    (tests/golden/make_golden_ssim_variants.py)."""
    from utils import loss_utils as lu
    G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "loss_ssim_variants.npz"))
    a, b, mask = (torch.from_numpy(G[k]) for k in ("a", "b", "mask"))
    assert abs(lu.ssim(a, b, 7).item() - float(G["ssim_w7"])) <= 1e-7
    np.testing.assert_allclose(lu.ssim(a, b, 11, False).numpy(), G["ssim_w11_batch"], rtol=0, atol=1e-7)
    assert abs(lu.ssim_masked(a, b, mask).item() - float(G["masked_mean"])) <= 1e-7
    np.testing.assert_allclose(lu.ssim_masked(a, b, mask, size_average=False).numpy(), G["masked_map"], rtol=0, atol=1e-7)
    assert abs(lu.ssim_masked(a, b, mask[:, 0]).item() - float(G["masked_mean_3d"])) <= 1e-7
    assert lu.create_window(11, 3).shape == (3, 1, 11, 11)
    # gradients flow to both images on the torch path, as in the reference
    x, y = a.clone().requires_grad_(True), b.clone().requires_grad_(True)
    lu.ssim(x, y).backward()
    assert x.grad is not None and y.grad is not None
