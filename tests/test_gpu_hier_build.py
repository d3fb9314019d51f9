"""GPU hierarchy build (csrc/hier_build.cu) against
  * the committed golden fixture written by the reference Creator's own classes (tests/golden/hier_build_601.npz),
  * the numpy oracle (oracle/hier_build.py) on other seeded inputs, including 1, 2 and 3 Gaussians and coordinate ties,
  * the reference tool run live when oracle/_ref/ref_gh_tool travelled to the box,
and, at a size the oracle cannot reach, the structural properties of a Creator hierarchy (pre-order layout, boxes of
parents contain their children, every input Gaussian is exactly one leaf row).
Integer outputs and boxes must be identical; merged attributes within 1e-4 (written in tests/hier_build_util.py)."""
import os

import numpy as np
import pytest
import torch

from harness import ROOT, add_paths, ref_creator
from oracle import hier_build as hb
from tests import hier_build_util as u

add_paths()

pytestmark = pytest.mark.gpu


def _gpu_build(g, align=True):
    import gaussian_hierarchy._C as gh
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in g.items()}
    out = gh.build_hierarchy(t["pos"], t["shs"], t["opacity"], t["scale"], t["rot"], align=align, return_debug=True)
    names = ("positions", "shs", "opacities", "log_scales", "rotations", "nodes", "boxes", "covariances", "leaf_order")
    res = {k: v.cpu().numpy() for k, v in zip(names, out)}
    res["shs"] = res["shs"].reshape(len(res["shs"]), 48)
    res["opacities"] = res["opacities"].reshape(-1)
    res["boxes"] = res["boxes"].reshape(len(res["boxes"]), 8)
    return res


def _fixture():
    z = np.load(os.path.join(ROOT, "tests", "golden", "hier_build_601.npz"))
    g = {k[3:]: z[k] for k in z.files if k.startswith("in_")}
    aligned = {k[4:]: z[k] for k in z.files if k.startswith("out_")}
    merged = dict(aligned, log_scales=z["merged_log_scales"], rotations=z["merged_rotations"])
    return g, aligned, merged


@pytest.mark.parametrize("align", [False, True])
def test_matches_the_reference_creator_fixture(align):
    g, aligned, merged = _fixture()
    got = _gpu_build(g, align)
    rep = u.compare_builds(got, aligned if align else merged, align)
    print("hier_build vs reference fixture, align =", align, rep)
    assert rep["covariance"] < 5e-5


@pytest.mark.parametrize("n,seed", [(1, 5), (2, 3), (3, 4), (5, 9), (64, 6), (257, 7), (1000, 8), (4099, 10)])
def test_matches_the_oracle(n, seed):
    g = u.synthetic_gaussians(n, seed)
    for align in (False, True):
        got = _gpu_build(g, align)
        want = hb.build(**g, align=align)
        rep = u.compare_builds(got, want, align)
        assert np.array_equal(got["leaf_order"], want["leaf_order"]), "leaves hold different input Gaussians"
        # the moment-matched covariances the merges accumulated (float32, same operation order as the reference)
        den = np.abs(want["covariances"]).max(axis=1)
        cerr = (np.abs(got["covariances"] - want["covariances"]).max(axis=1) / den).max()
        assert cerr < 1e-4, cerr
        print(f"hier_build vs oracle n={n} align={align}", rep, "moments", float(cerr))


def test_ties_at_the_median_go_by_gaussian_index():
    """Equal coordinates (a quantised street scan): std::nth_element may order them either way; the oracle and the
    CUDA build both order ties by Gaussian index."""
    g = u.synthetic_gaussians(513, 21)
    g["pos"] = (np.round(g["pos"] * 2.0) / 2.0).astype(np.float32)   # many duplicates on every axis
    g["pos"][::7, 0] = -0.0
    got = _gpu_build(g, True)
    want = hb.build(**g, align=True)
    assert np.array_equal(got["nodes"], want["nodes"])
    assert np.array_equal(got["leaf_order"], want["leaf_order"])
    assert np.array_equal(got["boxes"].view(np.uint32), want["boxes"].view(np.uint32))


@pytest.mark.skipif(not ref_creator.available(), reason="oracle/_ref/ref_gh_tool not on this box")
@pytest.mark.parametrize("n,seed", [(2, 3), (300, 12), (20000, 13)])
def test_matches_the_reference_tool_live(n, seed):
    g = u.synthetic_gaussians(n, seed)
    ref, secs = ref_creator.run_build(**g, align=True)
    got = _gpu_build(g, True)
    rep = u.compare_builds(got, ref, True)
    print(f"hier_build vs reference Creator live n={n}", rep, "reference seconds", secs)


def test_structure_at_a_million_gaussians():
    n = 1_000_003
    g = u.synthetic_gaussians(n, 31)
    got = _gpu_build(g, True)
    nodes, boxes = got["nodes"], got["boxes"]
    nn = 2 * n - 1
    assert nodes.shape == (nn, 7)
    depth, parent, start, cl, cm, sc, cc = nodes.T
    assert parent[0] == -1 and depth[0] == int(np.ceil(np.log2(n)))
    assert np.array_equal(np.sort(start), np.arange(nn)), "one Gaussian row per node"
    leaves = cl == 1
    assert leaves.sum() == n and (depth[leaves] == 0).all() and (cc[leaves] == 0).all() and (cm[leaves] == 0).all()
    inner = ~leaves
    assert (cc[inner] == 2).all() and (cm[inner] == 1).all()
    k = np.flatnonzero(inner)
    c0 = sc[k]
    assert (parent[c0] == k).all() and (parent[c0 + 1] == k).all()
    assert (depth[k] == np.maximum(depth[c0], depth[c0 + 1]) + 1).all()
    assert (start[c0] == start[k] + 1).all(), "pre-order: the first child follows its parent"
    # boxes of parents contain their children; w = largest extent
    for c in (c0, c0 + 1):
        assert (boxes[k, 0:3] <= boxes[c, 0:3]).all() and (boxes[k, 4:7] >= boxes[c, 4:7]).all()
    ext = (boxes[k, 4:7] - boxes[k, 0:3]).max(axis=1)
    assert np.array_equal(boxes[k, 3], ext) and np.array_equal(boxes[k, 7], ext)
    lo = got["leaf_order"]
    assert np.array_equal(np.sort(lo[lo >= 0]), np.arange(n)), "every input Gaussian is exactly one leaf"
    rows = start[leaves]
    assert np.array_equal(got["positions"][rows], g["pos"][lo[rows]])
    # a leaf's box is its own bounds (position +- 3 max scale)
    r = (np.float32(3.0) * g["scale"].max(axis=1)).astype(np.float32)[lo[rows]]
    assert np.array_equal(boxes[leaves, 0:3], (g["pos"][lo[rows]] - r[:, None]).astype(np.float32))
    for name in ("positions", "shs", "opacities", "log_scales", "rotations"):
        assert np.isfinite(got[name]).all(), name
    # merged Gaussians lie inside their node's box and the root blends everything: weights sum to one
    assert (got["positions"][start[k]] >= boxes[k, 0:3] - 1e-3).all() and (got["positions"][start[k]] <= boxes[k, 4:7] + 1e-3).all()
