"""GPU parity of the CUDA path (through the drop-in Python API -> C ABI) against
  (1) the committed golden fixtures tests/golden/*.npz -- outputs of the unmodified reference -- and
  (2) the CPU oracle (oracle/) on seeded inputs the oracle finishes in seconds.
Neither needs oracle/_ref or /root/reference at run time.

Bars: radii, tile counts, sort keys, sorted instance list, tile ranges, n_contrib, cut index arrays and
interpolation weights bit-exact; images bit-exact against the golden (same device exp); images within 2e-6
and gradients within 1e-4 relative (north_star) against the CPU oracle / golden gradients."""
import os

import numpy as np
import pytest
import torch

from harness import refglue, scenes, workload
from oracle import cpu_reference as cr
from tests.util import bits

pytestmark = pytest.mark.gpu
DEV = "cuda"
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REL_TOL = 1e-4
GRADS = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
         "dL_drotations"]


def _t(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    return t if dtype is None else t.to(dtype)


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-30))


def _golden_args(g):
    e = torch.Tensor([])
    has_sh = "shs" in g.files
    interp = "ts" in g.files
    ei = torch.empty(0, dtype=torch.int32, device=DEV)
    ef = torch.empty(0, dtype=torch.float32, device=DEV)
    ts = _t(g["ts"]) if interp else ef
    kids = _t(g["kids"]) if interp else ei
    common = dict(bg=_t(g["bg"]), ts=ts, kids=kids, means3D=_t(g["means3D"]),
                  colors=e if has_sh else _t(g["colors_precomp"]), opac=_t(g["opacities"]), scales=_t(g["scales"]),
                  rots=_t(g["rotations"]), mod=float(g["scale_modifier"]), view=_t(g["viewmatrix"]),
                  proj=_t(g["projmatrix"]), tx=float(g["tanfovx"]), ty=float(g["tanfovy"]), H=int(g["H"]),
                  W=int(g["W"]), sh=_t(g["shs"]) if has_sh else e, deg=int(g["degree"]), campos=_t(g["campos"]),
                  depth=bool(g["depth"]), ei=ei)
    return common


@pytest.mark.parametrize("name", ["raster_plain", "raster_interp", "raster_deg1_nodepth", "raster_colors_precomp"])
def test_rasterizer_against_reference_golden(name):
    import diff_gaussian_rasterization as dgr
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    a = _golden_args(g)
    e = torch.Tensor([])
    n, W, H = int(g["n"]), a["W"], a["H"]
    fargs = (a["bg"], a["ei"], a["ei"], a["ts"], a["kids"], a["means3D"], a["colors"], a["opac"], a["scales"],
             a["rots"], a["mod"], e, a["view"], a["proj"], a["tx"], a["ty"], H, W, a["sh"], a["deg"], a["campos"],
             False, False, a["depth"])
    R, color, radii, geom, binning, img, invd = dgr._C.rasterize_gaussians(*fargs)
    torch.cuda.synchronize()
    assert R == int(g["num_rendered"])
    assert torch.equal(radii.cpu(), torch.from_numpy(g["radii"]))
    dv = dgr._C.debug_view(n, W, H, R, geom, binning, img)
    vis = _t(g["radii"]) > 0
    assert torch.equal(dv["tiles_touched"][vis].cpu().int(), torch.from_numpy(g["ref_tiles_touched"])[vis.cpu()])
    assert torch.equal(bits(dv["depths"][vis]).cpu(), bits(torch.from_numpy(g["ref_depths"])[vis.cpu()]))
    assert torch.equal(dv["point_list_keys"].cpu().view(torch.int64), torch.from_numpy(g["ref_point_list_keys"]))
    assert torch.equal(dv["point_list"].cpu().view(torch.int32), torch.from_numpy(g["ref_point_list"]))
    assert torch.equal(dv["ranges"].cpu().view(torch.int32).flatten(), torch.from_numpy(g["ref_ranges"]))
    assert torch.equal(dv["n_contrib"].cpu().view(torch.int32), torch.from_numpy(g["ref_n_contrib"]))
    assert torch.equal(bits(dv["final_T"]).cpu(), bits(torch.from_numpy(g["ref_final_T"])))
    assert torch.equal(bits(color).cpu(), bits(torch.from_numpy(g["color"]))), "image not bit-identical to golden"
    assert invd.shape == tuple(g["invdepth"].shape)
    if a["depth"]:
        assert torch.equal(bits(invd).cpu(), bits(torch.from_numpy(g["invdepth"])))
    bargs = (a["bg"], a["ei"], a["ei"], a["ts"], a["kids"], a["means3D"], radii, a["colors"], a["opac"],
             a["scales"], a["rots"], a["mod"], e, a["view"], a["proj"], a["tx"], a["ty"], _t(g["dL_dcolor"]),
             _t(g["dL_dinvdepth"]) if a["depth"] else torch.empty(0, device=DEV), a["sh"], a["deg"], a["campos"],
             geom, R, binning, img, False)
    grads = dgr._C.rasterize_gaussians_backward(*bargs)
    torch.cuda.synchronize()
    checked = 0
    for k, ours in zip(GRADS, grads):
        ref, ref2 = g[k], g[k + "_run2"]
        if ref.size == 0 or not np.any(ref) or ours.numel() == 0:
            # gradients of inputs that were not given come back empty here (the reference materialises the
            # intermediate dL_dcolors / dL_dcov3D even then; Python never reads them, __init__.py:132-142)
            continue
        den = np.abs(ref).max()
        noise = np.abs(ref2 - ref).max() / den
        err = np.abs(ours.cpu().numpy().reshape(ref.shape) - ref).max() / den
        assert err <= max(REL_TOL, 3 * noise), f"{k}: {err:.3e} (reference run-to-run {noise:.3e})"
        checked += 1
    assert checked >= 5


def test_cut_against_reference_golden():
    import gaussian_hierarchy
    C = gaussian_hierarchy._C
    g = np.load(os.path.join(GOLDEN, "cut_l9.npz"))
    nodes, boxes = _t(g["nodes"]), _t(g["boxes"])
    N = nodes.size(0)
    for ci, vp in enumerate(g["viewpoints"]):
        for li, limit in enumerate(g["limits"]):
            key = f"c{ci}_l{li}_"
            ri, pi, ni = (torch.zeros(N, dtype=torch.int32, device=DEV) for _ in range(3))
            ts = torch.zeros(N, device=DEV)
            kids = torch.zeros(N, dtype=torch.int32, device=DEV)
            K = C.expand_to_size(nodes, boxes, float(limit), _t(vp), torch.zeros(3), ri, pi, ni)
            assert K == int(g[key + "count"])
            C.get_interpolation_weights(ni[:K].contiguous(), float(limit), nodes, boxes, torch.from_numpy(vp),
                                        torch.zeros(3), ts, kids)
            torch.cuda.synchronize()
            assert torch.equal(ri[:K].cpu(), torch.from_numpy(g[key + "render_indices"]))
            assert torch.equal(pi[:K].cpu(), torch.from_numpy(g[key + "parent_indices"]))
            assert torch.equal(ni[:K].cpu(), torch.from_numpy(g[key + "nodes_for_render_indices"]))
            assert torch.equal(kids[:K].cpu(), torch.from_numpy(g[key + "kids"]))
            assert torch.equal(bits(ts[:K]).cpu(), bits(torch.from_numpy(g[key + "ts"])))
    for k in g.files:
        if k.startswith("target_"):
            assert torch.equal(C.expand_to_target(nodes.cpu(), int(k[7:])), torch.from_numpy(g[k]))


def test_render_post_against_reference_golden():
    """Fused cut -> render_post forward/backward against the golden produced by the reference rasterizer under
    its Python glue."""
    g = np.load(os.path.join(GOLDEN, "render_post_l8.npz"))
    sky, W, H = int(g["sky"]), int(g["W"]), int(g["H"])
    attrs = {k: _t(g[k]) for k in ("means3D", "shs", "opacities", "scales", "rotations", "nodes", "boxes")}
    import math
    from types import SimpleNamespace
    import gaussian_hierarchy
    cam = SimpleNamespace(image_width=W, image_height=H, FoVx=2 * math.atan(float(g["tanfovx"])),
                          FoVy=2 * math.atan(float(g["tanfovy"])), world_view_transform=_t(g["viewmatrix"]),
                          full_proj_transform=_t(g["projmatrix"]), camera_center=_t(g["campos"]), image_name="golden")
    model = workload.ActivatedModel(attrs, sky)
    cut = workload.Cut(attrs["means3D"].size(0), DEV)
    cut.compute(gaussian_hierarchy._C, model, cam, float(g["limit"]))
    K = cut.count
    assert K == int(g["K"])
    assert torch.equal(cut.render_indices[:K].cpu(), torch.from_numpy(g["render_indices"]))
    assert torch.equal(bits(cut.interpolation_weights[:K]).cpu(), bits(torch.from_numpy(g["ts"])))
    out = workload.ours_render_post(cam, model, _t(g["bg"]), cut, do_depth=True)
    ((out["render"] * _t(g["dL_dcolor"])).sum() + (out["depth"] * _t(g["dL_dinvdepth"])).sum()).backward()
    torch.cuda.synchronize()
    assert torch.equal(bits(out["render"]).cpu(), bits(torch.from_numpy(g["render"])))
    assert torch.equal(bits(out["depth"]).cpu(), bits(torch.from_numpy(g["depth"])))
    assert torch.equal(out["visibility_filter"].cpu(), torch.from_numpy(g["visibility_filter"]))
    assert torch.equal(out["radii"].cpu(), torch.from_numpy(g["radii_visible"]))
    for nm, p in zip(["g_xyz", "g_features", "g_opacity", "g_scaling", "g_rotation"], model.parameters()):
        rows, ref, ref2 = g[nm + "_rows"], g[nm], g[nm + "_run2"]
        den = np.abs(ref).max()
        noise = np.abs(ref2 - ref).max() / den
        ours = p.grad.cpu().numpy()
        err = np.abs(ours[rows] - ref).max() / den
        assert err <= max(REL_TOL, 3 * noise), f"{nm}: {err:.3e} (noise {noise:.3e})"
        mask = np.ones(ours.shape[0], bool)
        mask[rows] = False
        assert not ours[mask].any()


@pytest.mark.parametrize("n,w,h,interp,seed", [(6000, 320, 200, False, 1), (9000, 400, 240, True, 2)])
def test_rasterizer_against_cpu_oracle(n, w, h, interp, seed):
    """Seeded inputs, CUDA path vs oracle: integers exact, image 2e-6, gradients 1e-4 relative."""
    import diff_gaussian_rasterization as dgr
    sc = scenes.street_gaussians(n, seed=seed, device=DEV, extent=(25.0, 12.0, 10.0))
    cam = scenes.street_camera(w, h, seed=seed, device=DEV, yaw_deg=7.0, x=-22.0)
    bg = torch.tensor([0.2, 0.1, 0.3], device=DEV)
    ts = kids = None
    if interp:
        gen = torch.Generator().manual_seed(seed)
        ts = torch.rand(n, generator=gen).to(DEV)
        kids = torch.randint(1, 5, (n,), generator=gen, dtype=torch.int32).to(DEV)
    rs = refglue.make_settings(dgr.GaussianRasterizationSettings, cam, bg, interpolation_weights=ts,
                               num_node_kids=kids)
    gen = torch.Generator().manual_seed(seed + 10)
    dcol = (torch.randn(3, h, w, generator=gen) / (w * h)).to(DEV)
    ddep = (torch.randn(1, h, w, generator=gen) / (w * h)).to(DEV)
    p = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    m2d = torch.zeros_like(p["means3D"], requires_grad=True)
    color, radii, depth = dgr.GaussianRasterizer(rs)(means3D=p["means3D"], means2D=m2d, opacities=p["opacities"],
                                                     shs=p["shs"], scales=p["scales"], rotations=p["rotations"])
    ((color * dcol).sum() + (depth * ddep).sum()).backward()
    torch.cuda.synchronize()
    c = lambda t: t.detach().cpu().numpy()  # noqa: E731
    o = cr.rasterize_forward(c(sc["means3D"]), c(sc["opacities"]), c(sc["scales"]), c(sc["rotations"]),
                             c(cam.world_view_transform).ravel(), c(cam.full_proj_transform).ravel(),
                             c(cam.camera_center), c(bg), w, h, rs.tanfovx, rs.tanfovy, shs=c(sc["shs"]),
                             ts=None if ts is None else c(ts), kids=None if kids is None else c(kids))
    assert np.array_equal(c(radii), o["radii"])
    # Without interpolation weights the CUDA path and the oracle differ only by device vs host expf.
    # With them the forward uses the hardware __powf (MUFU lg2/ex2, forward.cu:543), which the host
    # cannot reproduce: a pixel whose alpha lands within an ulp of the 1/255 or T < 1e-4 threshold may
    # keep or drop one contribution (<= alpha*T*c ~ 4e-3).  Such pixels must stay rare and small.
    dcolor = np.abs(c(color) - o["color"]).max(axis=0)
    ddepth = np.abs(c(depth) - o["invdepth"])[0] / max(1.0, np.abs(o["invdepth"]).max())
    if not interp:
        assert dcolor.max() <= 2e-6 and ddepth.max() <= 2e-6
    else:
        assert (dcolor > 2e-6).mean() <= 1e-3 and dcolor.max() <= 5e-3, (float((dcolor > 2e-6).mean()), float(dcolor.max()))
        assert (ddepth > 2e-6).mean() <= 1e-3 and ddepth.max() <= 5e-3
    gr = cr.rasterize_backward(o, c(dcol), c(ddep))
    for key, ours in (("dL_dmeans3D", p["means3D"].grad), ("dL_dsh", p["shs"].grad),
                      ("dL_dopacity", p["opacities"].grad), ("dL_dscales", p["scales"].grad),
                      ("dL_drotations", p["rotations"].grad), ("dL_dmeans2D", m2d.grad)):
        ref = gr[key]
        err = np.abs(c(ours).reshape(ref.shape) - ref).max() / (np.abs(ref).max() + 1e-30)
        assert err <= REL_TOL, f"{key}: {err:.3e}"
