"""Pins the CPU oracle (oracle/h3dgs_oracle.c, oracle/cpu_reference.py) against the golden vectors under
tests/golden/ -- outputs of the UNMODIFIED reference (its CUDA extensions run on a B200, its Python optimizer
run on CPU; see tests/golden/make_golden.py).  Runs without a GPU.

Bars: everything the reference computes before its first exp() -- per-Gaussian projection (depths, pixel
centres, cov3D, conic, AA-compensated opacity, SH colours), radii, tile counts, sort keys, sorted instance
list, tile ranges, cut index arrays and interpolation weights -- is BIT-EXACT.  Blended images, final_T and
the gradients go through exp/pow, whose host and device implementations differ in the last ulp: images
within 1e-6 absolute, n_contrib exact, gradients within 1e-4 relative (north_star tolerance) of the
reference, whose own run-to-run spread (float atomics) is stored beside each gradient.
"""
import os

import numpy as np
import pytest

from oracle import cpu_reference as cr

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GRAD_NAMES = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
              "dL_drotations"]
REL_TOL = 1e-4


def _load(name):
    path = os.path.join(GOLDEN, name + ".npz")
    if not os.path.exists(path):
        pytest.skip(f"{name}.npz not generated yet")
    return np.load(path)


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.int32)


def _forward(g):
    kw = dict(shs=g["shs"]) if "shs" in g else dict(colors_precomp=g["colors_precomp"])
    if "ts" in g:
        kw.update(ts=g["ts"], kids=g["kids"])
    return cr.rasterize_forward(g["means3D"], g["opacities"], g["scales"], g["rotations"], g["viewmatrix"],
                                g["projmatrix"], g["campos"], g["bg"], int(g["W"]), int(g["H"]),
                                float(g["tanfovx"]), float(g["tanfovy"]), degree=int(g["degree"]),
                                scale_modifier=float(g["scale_modifier"]), do_depth=bool(g["depth"]), **kw)


CASES = ["raster_plain", "raster_interp", "raster_deg1_nodepth", "raster_colors_precomp"]


@pytest.mark.parametrize("name", CASES)
def test_preprocess_and_binning_bit_exact(name):
    g = _load(name)
    o = _forward(g)
    vis = g["radii"] > 0
    assert vis.sum() > 100, "fixture must have visible Gaussians"
    assert np.array_equal(o["radii"], g["radii"])
    assert np.array_equal(o["tiles_touched"][vis].astype(np.int32), g["ref_tiles_touched"][vis])
    assert not o["tiles_touched"][~vis].any()
    for k in ("depths", "means2D", "cov3D", "conic_opacity"):
        assert np.array_equal(_bits(o[k][vis]), _bits(g["ref_" + k][vis])), k
    if "shs" in g:
        assert np.array_equal(_bits(o["rgb"][vis]), _bits(g["ref_rgb"][vis]))
        assert np.array_equal(o["clamped"][vis].astype(bool), g["ref_clamped"].reshape(-1, 3)[vis].astype(bool))
    assert o["num_rendered"] == int(g["num_rendered"]) > 0
    assert np.array_equal(o["point_list_keys"].view(np.int64), g["ref_point_list_keys"])
    assert np.array_equal(o["point_list"].view(np.int32), g["ref_point_list"])
    assert np.array_equal(o["ranges"].view(np.int32).ravel(), g["ref_ranges"])
    # size-independent properties of the binning
    keys = o["point_list_keys"]
    assert np.all(keys[1:] >= keys[:-1])
    r = o["ranges"].astype(np.int64)
    assert (r[:, 1] - r[:, 0]).sum() == o["num_rendered"]


@pytest.mark.parametrize("name", CASES)
def test_blend_forward_matches_reference(name):
    g = _load(name)
    o = _forward(g)
    assert np.array_equal(o["n_contrib"].view(np.int32), g["ref_n_contrib"])
    assert np.abs(o["final_T"] - g["ref_final_T"]).max() <= 1e-6
    assert np.abs(o["color"] - g["color"]).max() <= 1e-6
    if bool(g["depth"]):
        assert o["invdepth"].shape == g["invdepth"].shape
        assert np.abs(o["invdepth"] - g["invdepth"]).max() <= 1e-6 * max(1.0, np.abs(g["invdepth"]).max())
    else:
        assert g["invdepth"].shape[0] == 0 and o["invdepth"].shape[0] == 0


@pytest.mark.parametrize("name", CASES)
def test_backward_matches_reference(name):
    g = _load(name)
    o = _forward(g)
    gr = cr.rasterize_backward(o, g["dL_dcolor"], g["dL_dinvdepth"] if bool(g["depth"]) else None)
    checked = 0
    for k in GRAD_NAMES:
        ref, ref2 = g[k], g[k + "_run2"]
        if ref.size == 0:
            continue
        den = np.abs(ref).max()
        if den == 0:   # e.g. dL_dcov3D / dL_dsh when the corresponding input was not given
            assert not np.any(gr[k]) or k in ("dL_dcolors",), k
            continue
        noise = np.abs(ref2 - ref).max() / den
        err = np.abs(gr[k].reshape(ref.shape) - ref).max() / den
        assert err <= max(REL_TOL, 3 * noise), f"{k}: {err:.3e} (reference run-to-run {noise:.3e})"
        checked += 1
    assert checked >= 6


def test_adam_matches_reference_ouradam():
    """scene/OurAdam.py run as-is on CPU (make_golden.py cpu): 4 steps, sparse / dense (`relevant` empty)."""
    g = _load("adam_ouradam")
    widths, lrs, rows, steps = g["widths"], g["lrs"], int(g["rows"]), int(g["steps"])
    P = [g[f"p{i}_init"].copy() for i in range(len(widths))]
    M = [np.zeros_like(p) for p in P]
    V = [np.zeros_like(p) for p in P]
    for s in range(steps):
        rel = g[f"s{s}_relevant"]
        for i in range(len(widths)):
            before = P[i].copy()
            cr.sparse_adam(P[i], g[f"s{s}_g{i}"], M[i], V[i], rel, float(lrs[i]), step=s + 1)
            for got, want, nm in ((P[i], g[f"s{s}_p{i}"], "param"), (M[i], g[f"s{s}_m{i}"], "exp_avg"),
                                  (V[i], g[f"s{s}_v{i}"], "exp_avg_sq")):
                err = np.abs(got - want).max() / (np.abs(want).max() + 1e-30)
                assert err <= 1e-6, f"step {s} tensor {i} {nm}: {err:.3e}"
            if len(rel):
                untouched = np.setdiff1d(np.arange(rows), rel)
                assert np.array_equal(P[i][untouched], before[untouched]), "rows outside `relevant` must not move"


def test_cut_matches_reference():
    g = _load("cut_l9")
    nodes, boxes = g["nodes"], g["boxes"]
    nonempty = 0
    for ci, vp in enumerate(g["viewpoints"]):
        for li, limit in enumerate(g["limits"]):
            key = f"c{ci}_l{li}_"
            K, ri, pi, ni = cr.expand_to_size(nodes, boxes, float(limit), vp)
            assert K == int(g[key + "count"]), key
            assert np.array_equal(ri, g[key + "render_indices"])
            assert np.array_equal(pi, g[key + "parent_indices"])
            assert np.array_equal(ni, g[key + "nodes_for_render_indices"])
            if K:
                ts, kids = cr.get_interpolation_weights(ni, float(limit), nodes, boxes, vp)
                assert np.array_equal(_bits(ts), _bits(g[key + "ts"])), key
                assert np.array_equal(kids, g[key + "kids"])
                nonempty += 1
    assert nonempty >= 10
    for k in g.files:
        if k.startswith("target_"):
            assert np.array_equal(cr.expand_to_target(nodes, int(k[7:])), g[k]), k


def test_render_post_glue_matches_reference():
    """Hierarchy cut -> gather + lerp glue -> rasterizer fwd/bwd -> scatter to store rows, against the
    reference rasterizer under a restatement of its Python glue (harness/refglue.py)."""
    g = _load("render_post_l8")
    K, sky, W, H = int(g["K"]), int(g["sky"]), int(g["W"]), int(g["H"])
    cutK, ri, pi, ni = cr.expand_to_size(g["nodes"], g["boxes"], float(g["limit"]), g["campos"])
    assert cutK == K
    assert np.array_equal(ri, g["render_indices"]) and np.array_equal(pi, g["parent_indices"])
    ts, kids = cr.get_interpolation_weights(ni, float(g["limit"]), g["nodes"], g["boxes"], g["campos"])
    assert np.array_equal(_bits(ts), _bits(g["ts"]))
    attrs, w, sign = cr.render_post_gather(g["means3D"], g["shs"], g["opacities"], g["scales"], g["rotations"],
                                           ri, pi, ts, sky)
    kids_full = np.concatenate([kids, np.ones(sky, np.int32)])
    assert np.array_equal(kids_full, g["kids"])
    o = cr.rasterize_forward(attrs["means3D"], attrs["opacities"], attrs["scales"], attrs["rotations"],
                             g["viewmatrix"], g["projmatrix"], g["campos"], g["bg"], W, H, float(g["tanfovx"]),
                             float(g["tanfovy"]), shs=attrs["shs"], ts=w, kids=kids_full, do_depth=True)
    vis = o["radii"] > 0
    assert np.array_equal(vis, g["visibility_filter"])
    assert np.array_equal(o["radii"][vis], g["radii_visible"])
    assert np.abs(np.clip(o["color"], 0, 1) - g["render"]).max() <= 1e-6
    assert np.abs(o["invdepth"] - g["depth"]).max() <= 1e-6 * max(1.0, np.abs(g["depth"]).max())
    # the glue clamps the image to [0,1] before the loss: gradient passes where 0 <= c <= 1
    dcol = g["dL_dcolor"] * ((o["color"] >= 0) & (o["color"] <= 1))
    gr = cr.rasterize_backward(o, dcol, g["dL_dinvdepth"])
    store = cr.render_post_scatter(dict(means3D=gr["dL_dmeans3D"], shs=gr["dL_dsh"], opacities=gr["dL_dopacity"],
                                        scales=gr["dL_dscales"], rotations=gr["dL_drotations"]),
                                   g["means3D"].shape[0], ri, pi, ts, sign, sky)
    for nm, key in (("g_xyz", "means3D"), ("g_features", "shs"), ("g_opacity", "opacities"),
                    ("g_scaling", "scales"), ("g_rotation", "rotations")):
        rows, ref, ref2 = g[nm + "_rows"], g[nm], g[nm + "_run2"]
        den = np.abs(ref).max()
        noise = np.abs(ref2 - ref).max() / den
        err = np.abs(store[key][rows] - ref).max() / den
        assert err <= max(REL_TOL, 3 * noise), f"{nm}: {err:.3e} (reference run-to-run {noise:.3e})"
        others = np.setdiff1d(np.arange(g["means3D"].shape[0]), rows)
        assert not store[key][others].any(), f"{nm}: gradient on rows the reference leaves at zero"
    den = np.abs(g["g_viewspace"]).max()
    assert np.abs(gr["dL_dmeans2D"] - g["g_viewspace"]).max() / den <= REL_TOL


def test_edge_cases_empty_and_all_culled():
    cam = dict(viewmatrix=np.eye(4, dtype=np.float32).ravel(), projmatrix=np.eye(4, dtype=np.float32).ravel(),
               campos=np.zeros(3, np.float32), bg=np.array([0.2, 0.4, 0.6], np.float32), W=40, H=24, tanfovx=1.0,
               tanfovy=0.6)
    z = lambda *s: np.zeros(s, np.float32)  # noqa: E731
    o = cr.rasterize_forward(z(0, 3), z(0, 1), z(0, 3), z(0, 4), shs=z(0, 16, 3), **cam)
    assert o["num_rendered"] == 0 and not o["color"].any()      # reference returns the zero-filled image (R == 0)
    # everything behind the near plane (z <= 0.2): culled, radii stay 0
    means = np.array([[0, 0, -1.0], [0, 0, 0.1]], np.float32)
    rot = np.tile(np.array([1, 0, 0, 0], np.float32), (2, 1))
    o = cr.rasterize_forward(means, np.full((2, 1), 0.5, np.float32), np.full((2, 3), 0.1, np.float32), rot,
                             shs=z(2, 16, 3), **cam)
    assert not o["radii"].any() and o["num_rendered"] == 0
    gr = cr.rasterize_backward(o, z(3, 24, 40), z(1, 24, 40))
    assert all(not v.any() for v in gr.values())
