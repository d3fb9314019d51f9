"""GPU parity of the rasterizer against the unmodified reference CUDA build (oracle/_ref, sm_100a):
sort keys, instance lists, tile ranges, radii, n_contrib and the forward images bit for bit; gradients
within the north-star tolerance.  All calls go through the drop-in Python API and hence the C ABI."""
import pytest
import torch

from harness import refglue, scenes
from tests import util
from tests.util import bits, relerr

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _settings(cam, bg, do_depth=True, ts=None, kids=None, degree=3):
    import diff_gaussian_rasterization as dgr
    return refglue.make_settings(dgr.GaussianRasterizationSettings, cam, bg, sh_degree=degree,
                                 do_depth=do_depth, interpolation_weights=ts, num_node_kids=kids)


def _compare_forward(sc, cam, rs, args, n):
    import diff_gaussian_rasterization as dgr
    W, H = cam.image_width, cam.image_height
    R, color, radii, geom, binning, img, invd = dgr._C.rasterize_gaussians(*args)
    Rr, color_r, radii_r, geom_r, binning_r, img_r, invd_r = refglue.ref_hr().rasterize_gaussians(*args)
    torch.cuda.synchronize()
    assert R == Rr
    assert torch.equal(radii, radii_r)
    dv = dgr._C.debug_view(n, W, H, R, geom, binning, img)
    rv = refglue.parse_ref_buffers(geom_r, binning_r, img_r, n, Rr, W, H)
    vis = radii_r > 0
    assert torch.equal(dv["tiles_touched"], rv["tiles_touched"])
    assert torch.equal(bits(dv["depths"][vis]), bits(rv["depths"][vis]))
    rec = dv["records"].view(n, 16)[vis]
    assert torch.equal(bits(rec[:, 0:2]), bits(rv["means2D"][vis]))
    assert torch.equal(bits(rec[:, 4:8]), bits(rv["conic_opacity"][vis]))
    if R > 0:
        assert torch.equal(dv["point_list_keys"], rv["point_list_keys"]), "sort keys differ"
        assert torch.equal(dv["point_list"], rv["point_list"]), "sorted instance list differs"
        assert torch.equal(dv["ranges"], rv["ranges"]), "tile ranges differ"
        assert torch.equal(dv["n_contrib"], rv["n_contrib"])
        assert torch.equal(bits(dv["final_T"]), bits(rv["final_T"]))
        # size-independent properties of the binning
        keys = dv["point_list_keys"]
        assert bool((keys[1:] >= keys[:-1]).all()), "keys not sorted"
        assert int(dv["tiles_touched"].sum()) == R
        rg = dv["ranges"].view(-1, 2).long()
        assert int((rg[:, 1] - rg[:, 0]).sum()) == R, "tile ranges do not partition the list"
    assert relerr(color, color_r) <= util.REL_TOL
    assert torch.equal(bits(color), bits(color_r)), "image not bit-identical to the reference"
    if invd.numel():
        assert torch.equal(bits(invd), bits(invd_r))
    else:
        assert invd.shape == invd_r.shape
    return R


@pytest.mark.parametrize("n,w,h,yaw,interp", [
    (20000, 640, 480, 5.0, False),
    (100000, 1024, 768, 5.0, False),      # BASELINE.json config 1 size
    (100000, 1024, 768, 95.0, True),
    (30000, 1000, 700, 182.0, True),      # image size not a multiple of the tile size
])
def test_forward_bit_exact(n, w, h, yaw, interp):
    util.need_ref()
    sc, cam, ts, kids = util.make_scene(n, w, h, seed=n % 7, dev=DEV, yaw=yaw, interp=interp)
    bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
    rs = _settings(cam, bg, ts=ts, kids=kids)
    R = _compare_forward(sc, cam, rs, util.forward_args(rs, sc), n)
    assert R > 0


def _run_backward(fn, sc, dcol, ddep, extra=()):
    names = ("means3D", "shs", "opacities", "scales", "rotations") + tuple(extra)
    p = {k: sc[k].clone().requires_grad_(True) for k in names}
    m2d = torch.zeros_like(p["means3D"], requires_grad=True)
    color, radii, depth = fn(p, m2d)
    loss = (color * dcol).sum()
    if depth.numel():
        loss = loss + (depth * ddep).sum()
    loss.backward()
    g = {k: v.grad for k, v in p.items()}
    g["means2D"] = m2d.grad
    return color.detach(), radii, g


@pytest.mark.parametrize("n,w,h,interp,depth", [
    (20000, 640, 480, False, True),
    (100000, 1024, 768, False, True),
    (100000, 1024, 768, True, True),
    (50000, 800, 600, True, False),
])
def test_backward_matches_reference(n, w, h, interp, depth):
    util.need_ref()
    import diff_gaussian_rasterization as dgr
    sc, cam, ts, kids = util.make_scene(n, w, h, seed=3, dev=DEV, interp=interp)
    bg = torch.tensor([0.3, 0.1, 0.2], device=DEV)
    rs = _settings(cam, bg, do_depth=depth, ts=ts, kids=kids)
    g = torch.Generator().manual_seed(7)
    dcol = (torch.randn(3, h, w, generator=g) / (w * h)).to(DEV)
    ddep = (torch.randn(1, h, w, generator=g) / (w * h)).to(DEV)

    def ours(p, m2d):
        return dgr.GaussianRasterizer(rs)(means3D=p["means3D"], means2D=m2d, opacities=p["opacities"],
                                          shs=p["shs"], scales=p["scales"], rotations=p["rotations"])

    def ref(p, m2d):
        return refglue.ref_rasterize(rs, p["means3D"], m2d, p["opacities"], shs=p["shs"], scales=p["scales"],
                                     rotations=p["rotations"])

    co, ro, go = _run_backward(ours, sc, dcol, ddep)
    cr, rr, gr = _run_backward(ref, sc, dcol, ddep)
    more = [_run_backward(ref, sc, dcol, ddep)[2] for _ in range(4)]
    gr2 = more.pop()
    torch.cuda.synchronize()
    assert torch.equal(bits(co), bits(cr))
    for k in go:
        util.assert_grad_close(go[k], gr[k], gr2[k], k, more_refs=[m[k] for m in more])


@pytest.mark.parametrize("interp,depth", [(False, True), (True, True), (True, False)])
def test_backward_long_lists_cut_into_work_items(interp, depth):
    """Lists longer than the backward's work-item length (1024 entries, csrc/common.cuh BLEND_SEG): faint, wide
    Gaussians on a small image, so that pixels keep blending thousands of entries deep and the deeper work
    items start from the forward's snapshots at the cuts.  Gradients against the reference build."""
    util.need_ref()
    import diff_gaussian_rasterization as dgr
    n, w, h = 120000, 320, 240
    sc, cam, ts, kids = util.make_scene(n, w, h, seed=11, dev=DEV, interp=interp)
    sc["opacities"] = (sc["opacities"] * 0.02).contiguous()
    sc["scales"] = (sc["scales"] * 2.0).contiguous()
    bg = torch.tensor([0.3, 0.1, 0.2], device=DEV)
    rs = _settings(cam, bg, do_depth=depth, ts=ts, kids=kids)
    args = util.forward_args(rs, sc)
    R, color, radii, geom, binning, img, invd = dgr._C.rasterize_gaussians(*args)
    dv = dgr._C.debug_view(n, w, h, R, geom, binning, img)
    deep = int(dv["tile_last"].max())
    assert deep > 3 * 1024, f"scene too shallow for this test: deepest last contributor {deep}"
    assert int((dv["tile_last"] > 1024).sum()) >= 10
    g = torch.Generator().manual_seed(7)
    dcol = (torch.randn(3, h, w, generator=g) / (w * h)).to(DEV)
    ddep = (torch.randn(1, h, w, generator=g) / (w * h)).to(DEV)

    def ours(p, m2d):
        return dgr.GaussianRasterizer(rs)(means3D=p["means3D"], means2D=m2d, opacities=p["opacities"],
                                          shs=p["shs"], scales=p["scales"], rotations=p["rotations"])

    def ref(p, m2d):
        return refglue.ref_rasterize(rs, p["means3D"], m2d, p["opacities"], shs=p["shs"], scales=p["scales"],
                                     rotations=p["rotations"])

    co, ro, go = _run_backward(ours, sc, dcol, ddep)
    cr, rr, gr = _run_backward(ref, sc, dcol, ddep)
    more = [_run_backward(ref, sc, dcol, ddep)[2] for _ in range(4)]
    gr2 = more.pop()
    torch.cuda.synchronize()
    assert torch.equal(bits(co), bits(cr))
    for k in go:
        util.assert_grad_close(go[k], gr[k], gr2[k], k, more_refs=[m[k] for m in more])


@pytest.mark.parametrize("n", [1024, 1025, 2048, 2049, 4100])
def test_backward_work_item_boundaries(n):
    """List lengths at and around the cuts of the backward work items (1024 entries): n faint Gaussians in four
    clusters inside ONE tile, each pixel blending only its own cluster's quarter, so the last contributors lie
    at the very end of an n-entry list and every work item of the tile has live pixels.  Against the reference."""
    util.need_ref()
    import diff_gaussian_rasterization as dgr
    w = h = 64
    cam = scenes.street_camera(w, h, seed=1, device=DEV)
    g = torch.Generator().manual_seed(n)
    centre = torch.tensor([[38.0, 38.0], [42.0, 38.0], [38.0, 42.0], [42.0, 42.0]])[torch.arange(n) % 4]
    pix = centre + (torch.rand(n, 2, generator=g) - 0.5) * 0.6
    z = 8.0 + 2.0 * torch.rand(n, generator=g)
    xv = z * ((pix[:, 0] + 0.5) / (w / 2) - 1.0)           # fovx = fovy = 90 degrees: ndc = x / z
    yv = z * ((pix[:, 1] + 0.5) / (h / 2) - 1.0)
    V = cam.world_view_transform.cpu()
    pos = cam.camera_center.cpu() + xv[:, None] * V[:3, 0] + yv[:, None] * V[:3, 1] + z[:, None] * V[:3, 2]
    rot = torch.zeros(n, 4)
    rot[:, 0] = 1.0
    sc = {"means3D": pos.float().to(DEV).contiguous(),
          # ~1.6 px on screen; mildly anisotropic: with equal scales the rotation gradient is analytically zero and
          # both arms return rounding noise for it (relative error of order 1 between two runs of the reference)
          "scales": ((0.05 * z)[:, None] * torch.tensor([1.0, 0.85, 1.15])).float().to(DEV).contiguous(),
          "rotations": rot.to(DEV), "opacities": torch.full((n, 1), 0.0062, device=DEV),
          "shs": (torch.randn(n, 16, 3, generator=g) * 0.2).to(DEV).contiguous()}
    bg = torch.tensor([0.3, 0.1, 0.2], device=DEV)
    rs = _settings(cam, bg, do_depth=True)
    R, color, radii, geom, binning, img, invd = dgr._C.rasterize_gaussians(*util.forward_args(rs, sc))
    dv = dgr._C.debug_view(n, w, h, R, geom, binning, img)
    rg = dv["ranges"].view(-1, 2).long()
    assert int((rg[:, 1] - rg[:, 0]).max()) == n, "all Gaussians must share one tile"
    assert int(dv["tile_last"].max()) > n - 40, f"deepest last contributor {int(dv['tile_last'].max())} of {n}"
    gen = torch.Generator().manual_seed(7)
    dcol = torch.randn(3, h, w, generator=gen).to(DEV)
    ddep = torch.randn(1, h, w, generator=gen).to(DEV)

    def ours(p, m2d):
        return dgr.GaussianRasterizer(rs)(means3D=p["means3D"], means2D=m2d, opacities=p["opacities"],
                                          shs=p["shs"], scales=p["scales"], rotations=p["rotations"])

    def ref(p, m2d):
        return refglue.ref_rasterize(rs, p["means3D"], m2d, p["opacities"], shs=p["shs"], scales=p["scales"],
                                     rotations=p["rotations"])

    co, ro, go = _run_backward(ours, sc, dcol, ddep)
    cr, rr, gr = _run_backward(ref, sc, dcol, ddep)
    more = [_run_backward(ref, sc, dcol, ddep)[2] for _ in range(4)]
    gr2 = more.pop()
    torch.cuda.synchronize()
    assert torch.equal(bits(co), bits(cr))
    for k in go:
        util.assert_grad_close(go[k], gr[k], gr2[k], k, more_refs=[m[k] for m in more])


def test_precomputed_colors_and_low_sh_degree():
    util.need_ref()
    import diff_gaussian_rasterization as dgr
    n, w, h = 30000, 640, 480
    sc, cam, _, _ = util.make_scene(n, w, h, seed=1, dev=DEV)
    bg = torch.zeros(3, device=DEV)
    # precomputed colours
    rs = _settings(cam, bg)
    colors = torch.rand(n, 3, device=DEV)
    args = util.forward_args(rs, sc, colors=colors)
    _compare_forward(sc, cam, rs, args, n)
    # active SH degree below the stored one
    for deg in (0, 1, 2):
        rs = _settings(cam, bg, degree=deg)
        _compare_forward(sc, cam, rs, util.forward_args(rs, sc, degree=deg), n)

    # gradients of precomputed colours
    g = torch.Generator().manual_seed(2)
    dcol = (torch.randn(3, h, w, generator=g) / (w * h)).to(DEV)
    rs = _settings(cam, bg, do_depth=False)

    def run(fn):
        c = colors.clone().requires_grad_(True)
        m = sc["means3D"].clone().requires_grad_(True)
        m2d = torch.zeros_like(m, requires_grad=True)
        color, _, _ = fn(m, m2d, c)
        (color * dcol).sum().backward()
        return c.grad, m.grad

    ours = run(lambda m, m2d, c: dgr.GaussianRasterizer(rs)(means3D=m, means2D=m2d, opacities=sc["opacities"],
                                                             colors_precomp=c, scales=sc["scales"],
                                                             rotations=sc["rotations"]))
    ref = run(lambda m, m2d, c: refglue.ref_rasterize(rs, m, m2d, sc["opacities"], colors_precomp=c,
                                                      scales=sc["scales"], rotations=sc["rotations"]))
    ref2 = run(lambda m, m2d, c: refglue.ref_rasterize(rs, m, m2d, sc["opacities"], colors_precomp=c,
                                                       scales=sc["scales"], rotations=sc["rotations"]))
    util.assert_grad_close(ours[0], ref[0], ref2[0], "colors_precomp")
    util.assert_grad_close(ours[1], ref[1], ref2[1], "means3D")


def test_edge_cases():
    util.need_ref()
    import diff_gaussian_rasterization as dgr
    w, h = 320, 240
    cam = scenes.street_camera(w, h, seed=0, device=DEV, yaw_deg=0.0, x=0.0)
    bg = torch.tensor([0.5, 0.25, 0.125], device=DEV)
    rs = _settings(cam, bg)
    # everything behind the camera: no instances; the reference returns a zero image in that case
    sc = scenes.street_gaussians(1000, seed=0, device=DEV)
    sc["means3D"][:, 0] = -50.0 - sc["means3D"][:, 0].abs()
    R = _compare_forward(sc, cam, rs, util.forward_args(rs, sc), 1000)
    assert R == 0
    # a single Gaussian covering exactly one tile: the reference's range quirk leaves it unrendered
    one = {k: v[:1].clone() for k, v in scenes.street_gaussians(4, seed=1, device=DEV).items()}
    one["means3D"][0] = torch.tensor([5.0, 0.3, 2.0], device=DEV)
    one["scales"][0] = 0.002
    _compare_forward(one, cam, rs, util.forward_args(rs, one), 1)
    # P == 0
    empty = {k: v[:0].clone() for k, v in one.items()}
    R, color, radii, *_ = dgr._C.rasterize_gaussians(*util.forward_args(rs, empty))
    assert R == 0 and radii.numel() == 0 and float(color.abs().sum()) == 0.0
    # bad shape raises like AT_ERROR does
    with pytest.raises(RuntimeError):
        bad = dict(one); bad["means3D"] = one["means3D"].view(-1)
        dgr._C.rasterize_gaussians(*util.forward_args(rs, bad))
    # wrapper argument checks (HR/diff_gaussian_rasterization/__init__.py:186-190)
    with pytest.raises(Exception, match="excatly one"):
        dgr.GaussianRasterizer(rs)(means3D=one["means3D"], means2D=one["means3D"], opacities=one["opacities"],
                                   scales=one["scales"], rotations=one["rotations"])
    with pytest.raises(Exception, match="exactly one"):
        dgr.GaussianRasterizer(rs)(means3D=one["means3D"], means2D=one["means3D"], opacities=one["opacities"],
                                   shs=one["shs"], scales=one["scales"])


def test_mark_visible():
    import diff_gaussian_rasterization as dgr
    sc, cam, _, _ = util.make_scene(50000, 640, 480, seed=4, dev=DEV)
    rs = _settings(cam, torch.zeros(3, device=DEV))
    vis = dgr.GaussianRasterizer(rs).markVisible(sc["means3D"])
    p = torch.cat([sc["means3D"], torch.ones(50000, 1, device=DEV)], dim=1) @ cam.world_view_transform
    expect = p[:, 2] > 0.2
    # fp32 rounding at the plane itself may differ between the matmul and the kernel's FMA chain
    undecided = (p[:, 2] - 0.2).abs() < 1e-5
    assert vis.dtype == torch.bool
    assert bool(((vis == expect) | undecided).all())


def test_staging_by_bulk_copy_engine_matches(monkeypatch, tmp_path):
    """H3DGS_STAGE=tma (cp.async.bulk + mbarrier) must give the same bits as the LDGSTS path; run in a
    subprocess because the choice is latched at first use.  Both processes load the SAME saved inputs: torch's
    CPU exp() that makes the scene's scales is not bit-reproducible from process to process on every host
    (3 of 120 processes differed by an ulp in a block of rows on the GPU box, harness/diag_stage_modes.py)."""
    import os
    import subprocess
    import sys
    sc0, _, ts0, kids0 = util.make_scene(60000, 800, 600, seed=2, dev="cpu", interp=True)
    scene_file = str(tmp_path / "scene.pt")
    torch.save({"sc": sc0, "ts": ts0, "kids": kids0}, scene_file)
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from harness import add_paths; add_paths()\n"
        "import torch, hashlib\n"
        "from tests import util\n"
        "from tests.test_gpu_rasterizer import _settings\n"
        "import diff_gaussian_rasterization as dgr\n"
        "_, cam, _, _ = util.make_scene(16, 800, 600, seed=2, dev='cuda', interp=True)\n"
        "blob = torch.load(%r)\n"
        "sc = {k: v.cuda() for k, v in blob['sc'].items()}; ts = blob['ts'].cuda(); kids = blob['kids'].cuda()\n"
        "rs = _settings(cam, torch.zeros(3, device='cuda'), ts=ts, kids=kids)\n"
        "p = {k: sc[k].clone().requires_grad_(True) for k in ('means3D','shs','opacities','scales','rotations')}\n"
        "m2d = torch.zeros_like(p['means3D'], requires_grad=True)\n"
        "c, r, d = dgr.GaussianRasterizer(rs)(means3D=p['means3D'], means2D=m2d, opacities=p['opacities'], shs=p['shs'], scales=p['scales'], rotations=p['rotations'])\n"
        "(c.sum() + d.sum()).backward()\n"
        "torch.cuda.synchronize()\n"
        "h = hashlib.sha256(c.detach().cpu().numpy().tobytes() + d.detach().cpu().numpy().tobytes()).hexdigest()\n"
        "print('HASH', h, float(p['means3D'].grad.abs().sum()))\n"
        % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), scene_file))
    outs = []
    for mode in ("ldgsts", "tma"):
        env = dict(os.environ, H3DGS_STAGE=mode)
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        line = [l for l in r.stdout.splitlines() if l.startswith("HASH")][0].split()
        outs.append((line[1], float(line[2])))
    assert outs[0][0] == outs[1][0], "image differs between staging modes"
    assert abs(outs[0][1] - outs[1][1]) <= 1e-4 * abs(outs[0][1])


def test_long_thin_splats_bit_exact():
    """Splats thousands of pixels long and sub-pixel thin: the rounding of a dx^2 + 2 b dx dy + c dy^2 grows with
    the cancelling terms, so the blend kernels' per-block cull (blend_common.cuh block_hit) scales its margin
    with them.  n_contrib, final_T and the image must stay bit-identical to the reference's per-pixel test."""
    util.need_ref()
    n, w, h = 6000, 1280, 720
    sc, cam, _, _ = util.make_scene(n, w, h, seed=5, dev=DEV)
    g = torch.Generator().manual_seed(9)
    k = 3000
    # major sigma 4..20 m at 3..60 m distance (thousands of pixels), minor sigma 0.5 mm (the 0.3 px^2 dilation
    # then sets the width), faint enough that pixels blend hundreds of them
    sc["scales"][:k] = torch.stack([4.0 + 16.0 * torch.rand(k, generator=g), torch.full((k,), 5e-4),
                                    torch.full((k,), 5e-4)], dim=1).to(DEV)
    sc["opacities"][:k] = (0.02 + 0.2 * torch.rand(k, 1, generator=g)).to(DEV)
    bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
    rs = _settings(cam, bg)
    R = _compare_forward(sc, cam, rs, util.forward_args(rs, sc), n)
    assert R > 0


def test_more_than_65535_tiles():
    """4112 x 4096 pixels = 257 x 256 = 65 792 tiles: tile ids no longer fit 16 bits (the reference's key takes 17
    tile bits; here 5 strips per tile row, 1280 strips, two sort passes).  Keys, list, ranges and image against the
    reference build, bit for bit."""
    util.need_ref()
    import diff_gaussian_rasterization as dgr
    n, w, h = 40000, 4112, 4096
    sc, cam, _, _ = util.make_scene(n, w, h, seed=8, dev=DEV)
    bg = torch.tensor([0.2, 0.1, 0.05], device=DEV)
    rs = _settings(cam, bg)
    args = util.forward_args(rs, sc)
    R = _compare_forward(sc, cam, rs, args, n)
    assert R > 0
    _, _, _, geom, binning, img, _ = dgr._C.rasterize_gaussians(*args)
    dv = dgr._C.debug_view(n, w, h, R, geom, binning, img)
    assert int((dv["point_list_keys"] >> 32).max()) > 65535, "no instance landed in a tile above 65535"


@pytest.mark.parametrize("mode", [1, 2, 3])
def test_wide_strip_keys_and_two_sort_passes(mode):
    """The binning sort keys one record per (slot, tile row, 64-tile strip) by strip id: u16 keys and one 8-bit pass
    up to 256 strips.  u32 keys (more than 65 536 strips) and a second pass only appear for images no test can
    afford; h3dgs_debug_set(2, mode) forces them (bit 0: u32 keys, bit 1: two passes).  Instance list, ranges, keys
    and image must not change by a bit."""
    import diff_gaussian_rasterization as dgr
    import h3dgs_native as nat
    n, w, h = 30000, 1280, 720
    sc, cam, _, _ = util.make_scene(n, w, h, seed=12, dev=DEV)
    bg = torch.tensor([0.05, 0.1, 0.2], device=DEV)
    rs = _settings(cam, bg)
    args = util.forward_args(rs, sc)

    def run():
        R, color, radii, geom, binning, img, invd = dgr._C.rasterize_gaussians(*args)
        dv = dgr._C.debug_view(n, w, h, R, geom, binning, img)
        torch.cuda.synchronize()
        return R, color.clone(), {k: dv[k].clone() for k in ("point_list", "point_list_keys", "ranges", "n_contrib")}

    R0, c0, d0 = run()
    assert R0 > 0
    prev = nat.lib().h3dgs_debug_set(2, mode)
    try:
        R1, c1, d1 = run()
    finally:
        nat.lib().h3dgs_debug_set(2, prev)
    assert R1 == R0 and torch.equal(bits(c0), bits(c1))
    for k in d0:
        assert torch.equal(d0[k], d1[k]), k
