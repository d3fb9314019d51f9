"""Gradient parity at the size the benchmark runs (BASELINE.json config 4): render_post forward + backward on a
~5 M-slot cut (+100 k skybox) of the 16.9 M-row level-23 street hierarchy at 1920x1080, one camera looking
along the street (vanishing-point tiles with 20 k+ list entries: the backward's work items, the snapshots at
the list cuts) and one looking across it, against the unmodified reference CUDA extensions under the
reference's Python glue.  All six raw-parameter gradients and viewspace_points.grad."""
import pytest
import torch

from harness import refglue, workload
from tests import util
from tests.util import bits

pytestmark = pytest.mark.gpu
DEV = "cuda"
SKY = 100_000
W, H = 1920, 1080


@pytest.fixture(scope="module")
def scene():
    util.need_ref()
    attrs = workload.build_scene(23, SKY, 0, DEV, free_corridor=True)
    cams = workload.cameras(W, H, 2, DEV)   # yaw -5 deg (along the street) and 90 + 6 deg (across)
    yield attrs, cams
    del attrs
    torch.cuda.empty_cache()


def _touched_rows(cut, n):
    """Store rows a render of this cut can add into: children, parents, skybox."""
    K = cut.count
    t = torch.zeros(n, dtype=torch.bool, device=DEV)
    t[cut.render_indices[:K].long()] = True
    par = cut.parent_indices[:K]
    t[par[par >= 0].long()] = True
    t[n - SKY:] = True
    return t


@pytest.mark.parametrize("cam_index", [0, 1])
def test_render_post_gradients_at_benchmark_size(scene, cam_index):
    import gaussian_hierarchy
    attrs, cams = scene
    cam = cams[cam_index]
    n = attrs["means3D"].size(0)
    bg = torch.zeros(3, device=DEV)
    g = torch.Generator().manual_seed(1234 + cam_index)
    dcol = (torch.randn(3, H, W, generator=g) / (W * H)).to(DEV)
    ddep = (torch.randn(1, H, W, generator=g) / (W * H)).to(DEV)

    probe = workload.HierModel(attrs, SKY)
    limit = workload.find_limit(gaussian_hierarchy._C, probe, cams[0], 5_000_000, DEV)
    del probe

    def run(fn, gh):
        model = workload.HierModel(attrs, SKY)
        cut = workload.Cut(n, DEV)
        cut.compute(gh, model, cam, limit)
        out = fn(cam, model, bg, cut, do_depth=True)
        ((out["render"] * dcol).sum() + (out["depth"] * ddep).sum()).backward()
        torch.cuda.synchronize()
        return model, cut, out

    m_o, cut_o, out_o = run(workload.ours_render_post, gaussian_hierarchy._C)
    K = cut_o.count
    assert 4_000_000 < K < 6_000_000
    touched = _touched_rows(cut_o, n)
    rows = touched.nonzero().flatten()
    names = ["xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation"]
    # keep only the rows a render can touch (and proof that nothing else was written) -- the dense tensors of
    # four runs would be 100 GB
    ours = {}
    for name, p in zip(names, m_o.parameters()):
        assert not bool(p.grad[~touched].any()), f"{name}: gradient outside the rows of the cut"
        ours[name] = p.grad[rows].clone()
    ours["viewspace_points"] = out_o["viewspace_points"].grad[:K + SKY].clone()
    assert float(out_o["viewspace_points"].grad[K + SKY:].abs().sum()) == 0.0
    img_o, dep_o = out_o["render"].clone(), out_o["depth"].clone()
    vis_o, radii_o = out_o["visibility_filter"].clone(), out_o["radii"].clone()
    del m_o, out_o
    torch.cuda.empty_cache()
    # a second run of ours: both arms reduce with float atomics (tests/util.py assert_grad_close, ours_again)
    m_o2, _, out_o2 = run(workload.ours_render_post, gaussian_hierarchy._C)
    assert torch.equal(bits(out_o2["render"]), bits(img_o))
    ours2 = {name: p.grad[rows].clone() for name, p in zip(names, m_o2.parameters())}
    ours2["viewspace_points"] = out_o2["viewspace_points"].grad[:K + SKY].clone()
    del m_o2, out_o2
    torch.cuda.empty_cache()

    refs = []
    for rep in range(3):
        m_r, cut_r, out_r = run(workload.ref_render_post, refglue.ref_gh())
        if rep == 0:
            assert cut_r.count == K
            assert torch.equal(cut_r.render_indices[:K], cut_o.render_indices[:K])
            assert torch.equal(bits(out_r["render"]), bits(img_o)), "image differs at benchmark size"
            assert torch.equal(bits(out_r["depth"]), bits(dep_o))
            assert torch.equal(out_r["visibility_filter"], vis_o)
            assert torch.equal(out_r["radii"], radii_o)
        d = {name: p.grad[rows].clone() for name, p in zip(names, m_r.parameters())}
        d["viewspace_points"] = out_r["viewspace_points"].grad[:K + SKY].clone()
        refs.append(d)
        del m_r, out_r
        torch.cuda.empty_cache()

    n_grad_rows = int((ours["opacity"].flatten() != 0).sum())
    assert n_grad_rows > 50_000, f"only {n_grad_rows} rows received a gradient"
    for name in names + ["viewspace_points"]:
        util.assert_grad_close(ours[name], refs[0][name], refs[1][name], name, more_refs=(refs[2][name],),
                               ours_again=(ours2[name],))
        for r in refs:
            del r[name]
        del ours[name], ours2[name]
