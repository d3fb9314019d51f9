"""Config 2 (BASELINE.json: single-chunk training step, train_single.py:111-141) value parity: plain ``render()`` on
a reference-style GaussianModel of 1 M Gaussians with the skybox rows at the FRONT of the store, depth output,
learnable exposure, then the training loss  0.8 L1 + 0.2 (1 - SSIM) on ``image * alpha_mask``  plus the masked
inverse-depth L1 -- ours (raw-store kernels + fused loss) against the reference rasterizer under a restatement of
the reference's render glue (harness/refglue.py ref_render) and the reference's torch loss formulas."""
import pytest
import torch

from harness import refglue, scenes, workload
from tests import util
from tests.util import bits

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _step(arm, sc, cam, sky, gt, alpha_mask, mono, depth_mask, depth_w):
    model = workload.SingleModel(sc, sky, camera_names=(cam.image_name,))
    with torch.no_grad():   # a non-trivial exposure, so that its gradient and the clamp matter
        model._exposure[0, :3, :3] += torch.tensor([[0.05, -0.02, 0.0], [0.01, 0.03, -0.01], [0.0, 0.02, -0.04]],
                                                   device=DEV)
        model._exposure[0, :3, 3] += torch.tensor([0.02, -0.01, 0.03], device=DEV)
    bg = torch.rand(3, generator=torch.Generator().manual_seed(3)).to(DEV)
    if arm == "ours":
        from utils.loss_utils import fused_l1_ssim_loss, masked_l1_loss
        out = workload.ours_render(cam, model, bg, use_trained_exp=True)
        photo, _ = fused_l1_ssim_loss(out["render"], gt, 0.2, mask=alpha_mask)
        ld = masked_l1_loss(out["depth"], mono, depth_mask)
    else:
        from oracle import loss_reference as lr
        out = workload.ref_render(cam, model, bg, use_trained_exp=True)
        photo, _ = lr.l1_ssim_loss(out["render"], gt, 0.2, mask=alpha_mask)
        ld = lr.depth_l1(out["depth"], mono, depth_mask)
    loss = photo + depth_w * ld
    loss.backward()
    torch.cuda.synchronize()
    return model, out, loss.detach()


@pytest.mark.parametrize("n,sky,w,h", [(1_000_000, 100_000, 1920, 1080), (60_000, 2_000, 800, 600)])
def test_train_single_step_values_match_reference(n, sky, w, h, monkeypatch):
    util.need_ref()
    import diff_gaussian_rasterization as dgr
    sc = workload.single_chunk_scene(n, sky, 6, DEV)
    cam = scenes.street_camera(w, h, seed=4, device=DEV, yaw_deg=8.0, x=-15.0)
    g = torch.Generator().manual_seed(17)
    gt = torch.rand(3, h, w, generator=g).to(DEV)
    alpha_mask = (torch.rand(1, h, w, generator=g) > 0.1).float().to(DEV)
    mono = torch.rand(1, h, w, generator=g).to(DEV)
    depth_mask = (torch.rand(1, h, w, generator=g) > 0.3).float().to(DEV)
    calls = []
    orig = dgr._RasterizeGaussiansRaw.forward
    monkeypatch.setattr(dgr._RasterizeGaussiansRaw, "forward",
                        staticmethod(lambda *a, **k: (calls.append(1), orig(*a, **k))[1]))
    m_o, out_o, loss_o = _step("ours", sc, cam, sky, gt, alpha_mask, mono, depth_mask, 0.7)
    assert calls, "render() did not take the raw-store path"
    # two more runs of ours: both arms reduce with float atomics (tests/util.py assert_grad_close, ours_again)
    again = [_step("ours", sc, cam, sky, gt, alpha_mask, mono, depth_mask, 0.7) for _ in range(2)]
    for m2, out2, _ in again:
        assert torch.equal(bits(out2["render"]), bits(out_o["render"]))
    refs = [_step("ref", sc, cam, sky, gt, alpha_mask, mono, depth_mask, 0.7) for _ in range(3)]
    m_r, out_r, loss_r = refs[0]
    assert set(out_o) == set(out_r) == {"render", "depth", "viewspace_points", "visibility_filter", "radii"}
    assert torch.equal(bits(out_o["render"]), bits(out_r["render"])), "render() image differs"
    assert torch.equal(bits(out_o["depth"]), bits(out_r["depth"]))
    assert out_o["visibility_filter"].dtype == torch.int64
    assert torch.equal(out_o["visibility_filter"], out_r["visibility_filter"])
    assert torch.equal(out_o["radii"], out_r["radii"])
    # some skybox rows (the first `sky` rows of the store) are on screen
    assert int((out_o["visibility_filter"] < sky).sum()) > 0
    assert abs(float(loss_o) - float(loss_r)) <= 2e-6 * abs(float(loss_r)), (float(loss_o), float(loss_r))
    names = ["xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation"]
    for i, name in enumerate(names):
        util.assert_grad_close(m_o.parameters()[i].grad, refs[0][0].parameters()[i].grad,
                               refs[1][0].parameters()[i].grad, name, more_refs=(refs[2][0].parameters()[i].grad,),
                               ours_again=tuple(x[0].parameters()[i].grad for x in again))
    util.assert_grad_close(m_o._exposure.grad, refs[0][0]._exposure.grad, refs[1][0]._exposure.grad, "exposure",
                           more_refs=(refs[2][0]._exposure.grad,))
    util.assert_grad_close(out_o["viewspace_points"].grad, refs[0][1]["viewspace_points"].grad,
                           refs[1][1]["viewspace_points"].grad, "viewspace_points",
                           more_refs=(refs[2][1]["viewspace_points"].grad,))
