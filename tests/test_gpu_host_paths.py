"""The two host paths over the same C ABI -- csrc/torch_binding.cpp (lib/_h3dgs_torch.so, the default) and the ctypes
binding in diff_gaussian_rasterization/_C.py -- must be interchangeable: identical forward bits (same kernels, same
launch order), gradients equal up to the float-atomic reduction order, in the dense, pre-zeroed and compact modes."""
import pytest
import torch

from harness import scenes, workload
from tests.util import bits

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _run(host, monkeypatch, raw, sparse_mode):
    import diff_gaussian_rasterization as dgr
    import gaussian_hierarchy
    from gaussian_renderer import render_post
    monkeypatch.setattr(dgr._C, "HOST", host)
    assert (dgr._C.ext() is None) == (host == "ctypes")
    sky = 500
    attrs = workload.build_scene(15, sky, 9, DEV)
    cam = scenes.street_camera(800, 600, seed=3, device=DEV, yaw_deg=5.0, x=-12.0)
    bg = torch.tensor([0.1, 0.0, 0.2], device=DEV)
    n = attrs["means3D"].size(0)
    g = torch.Generator().manual_seed(4)
    dcol = (torch.randn(3, 600, 800, generator=g) / 480000).to(DEV)
    ddep = (torch.randn(1, 600, 800, generator=g) / 480000).to(DEV)
    model = workload.HierModel(attrs, sky) if raw else workload.ActivatedModel(attrs, sky)
    limit = workload.find_limit(gaussian_hierarchy._C, model, cam, 12000, DEV)
    cut = workload.Cut(n, DEV)
    cut.compute(gaussian_hierarchy._C, model, cam, limit)
    K = cut.count
    sg = dgr.SparseGradients() if sparse_mode else None
    out = render_post(cam, model, scenes.Pipe, bg, render_indices=cut.render_indices[:K],
                      parent_indices=cut.parent_indices[:K], interpolation_weights=cut.interpolation_weights,
                      num_node_kids=cut.num_siblings, do_depth=True, sparse_grads=sg)
    ((out["render"] * dcol).sum() + (out["depth"] * ddep).sum()).backward()
    torch.cuda.synchronize()
    if sparse_mode:
        names = list(sg.grads)
        grads = {k: sg.dense(k) for k in names}
    else:
        grads = {str(i): p.grad for i, p in enumerate(model.parameters())}
    grads["viewspace_points"] = out["viewspace_points"].grad
    return out, grads


@pytest.mark.parametrize("raw,sparse_mode", [(True, False), (False, False), (True, True)])
def test_extension_and_ctypes_host_paths_agree(monkeypatch, raw, sparse_mode):
    import diff_gaussian_rasterization as dgr
    monkeypatch.setattr(dgr._C, "HOST", "ext")
    dgr._C.ext()   # raises when lib/_h3dgs_torch.so is missing: the default host path must be the one that runs
    out_e, g_e = _run("ext", monkeypatch, raw, sparse_mode)
    out_c, g_c = _run("ctypes", monkeypatch, raw, sparse_mode)
    for k in ("render", "depth"):
        assert torch.equal(bits(out_e[k]), bits(out_c[k])), k
    assert torch.equal(out_e["radii"], out_c["radii"])
    assert torch.equal(out_e["visibility_filter"], out_c["visibility_filter"])
    assert set(g_e) == set(g_c)
    for k in g_e:
        a, b = g_e[k], g_c[k]
        assert a.shape == b.shape, k
        den = float(b.abs().max())
        assert den > 0
        assert float((a - b).abs().max()) <= 3e-3 * den, k


def test_extension_reports_errors_like_the_ctypes_binding(monkeypatch):
    import diff_gaussian_rasterization as dgr
    monkeypatch.setattr(dgr._C, "HOST", "ext")
    z = torch.zeros(4, 3, device=DEV)
    eye = torch.eye(4, device=DEV)
    args = lambda m: (torch.zeros(3, device=DEV), None, None, None, None, m, None, torch.ones(4, 1, device=DEV),
                      torch.ones(4, 3, device=DEV), torch.ones(4, 4, device=DEV), 1.0, None, eye, eye, 1.0, 1.0, 64, 64,
                      torch.zeros(4, 1, 3, device=DEV), 0, torch.zeros(3, device=DEV), False, False, False)
    with pytest.raises(RuntimeError, match="num_points, 3"):
        dgr._C.rasterize_gaussians(*args(torch.zeros(4, 2, device=DEV)))
    with pytest.raises(TypeError, match="float32"):
        dgr._C.rasterize_gaussians(*args(z.double()))
    out = dgr._C.rasterize_gaussians(*args(z))
    assert len(out) == 7 and out[1].shape == (3, 64, 64)


class _GuardedTorch:
    """Stands in for the `torch` module inside diff_gaussian_rasterization._C: every 1-D uint8 CUDA buffer (the
    opaque geometry / binning / image / backward workspaces, sized by the library's own *_bytes queries) is carved
    out of a larger allocation with 4 KiB of 0xA5 in front and behind."""
    GUARD = 4096

    def __init__(self):
        self.blocks = []

    def __getattr__(self, name):
        return getattr(torch, name)

    def empty(self, *args, **kw):
        shape = args[0] if args else kw.get("size")
        if kw.get("dtype") is torch.uint8 and isinstance(shape, tuple) and len(shape) == 1 and \
                torch.device(kw.get("device", "cpu")).type == "cuda" and len(args) <= 1:
            n = int(shape[0])
            pad = (-n) % 256
            whole = torch.full((n + pad + 2 * self.GUARD,), 0xA5, dtype=torch.uint8, device=kw["device"])
            self.blocks.append((whole, n))
            return whole[self.GUARD:self.GUARD + n]
        return torch.empty(*args, **kw)

    def check(self):
        assert self.blocks, "no workspace went through the guarded allocator"
        for whole, n in self.blocks:
            G = self.GUARD
            assert bool((whole[:G] == 0xA5).all()), f"a kernel wrote in front of a {n}-byte workspace"
            assert bool((whole[G + n:] == 0xA5).all()), f"a kernel wrote behind a {n}-byte workspace"


@pytest.mark.parametrize("w,h,levels,target", [(800, 600, 15, 12000), (1920, 1080, 17, 60000), (1040, 72, 14, 6000)])
def test_kernels_stay_inside_their_workspaces(monkeypatch, w, h, levels, target):
    """compute-sanitizer is not available on the GPU pool: the library's workspace carving (common.cuh GeomWs /
    BinWs / ImageWs / BwdWs, sized by counts that partly live on the device: strip records, expansion chunks) is
    checked with guard bands instead.  ctypes host path, forward + backward of a cut with parents, at three image
    shapes (1920 wide: two 64-tile strips per tile row, the second partly filled; 1040 x 72: 65 tile columns, a
    one-tile second strip)."""
    import diff_gaussian_rasterization as dgr
    import gaussian_hierarchy
    from gaussian_renderer import render_post
    monkeypatch.setattr(dgr._C, "HOST", "ctypes")
    guarded = _GuardedTorch()
    monkeypatch.setattr(dgr._C, "torch", guarded)
    sky = 300
    attrs = workload.build_scene(levels, sky, 5, DEV)
    cam = scenes.street_camera(w, h, seed=2, device=DEV, yaw_deg=3.0, x=-8.0)
    bg = torch.tensor([0.1, 0.0, 0.2], device=DEV)
    n = attrs["means3D"].size(0)
    model = workload.HierModel(attrs, sky)
    limit = workload.find_limit(gaussian_hierarchy._C, model, cam, target, DEV)
    cut = workload.Cut(n, DEV)
    cut.compute(gaussian_hierarchy._C, model, cam, limit)
    K = cut.count
    out = render_post(cam, model, scenes.Pipe, bg, render_indices=cut.render_indices[:K],
                      parent_indices=cut.parent_indices[:K], interpolation_weights=cut.interpolation_weights,
                      num_node_kids=cut.num_siblings, do_depth=True)
    (out["render"].sum() + out["depth"].sum()).backward()
    torch.cuda.synchronize()
    assert int(out["radii"].numel()) > 0
    guarded.check()
    assert len(guarded.blocks) >= 3, [b[1] for b in guarded.blocks]
