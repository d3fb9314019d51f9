"""Incremental cut (h3dgs_cut_step / gaussian_hierarchy._C.IncrementalCut) against the reference's own
Switching::changeToSizeStep compiled from its source (oracle/_ref/ref_switch_tool) on a camera path: after every step
the node list, the split-driven render / parent / node-of-render arrays and the counts must be identical; and, with
or without the reference tool, iterating the step converges to the cut expand_to_size computes in one go."""
import os
import subprocess
import tempfile

import numpy as np
import pytest
import torch

from harness import ROOT, add_paths, scenes

add_paths()

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOOL = os.path.join(ROOT, "oracle", "_ref", "ref_switch_tool")


def _path(views):
    g = torch.Generator().manual_seed(4)
    cams = []
    for i in range(views):
        x = -60.0 + 120.0 * i / max(views - 1, 1)
        cams.append([x, float(torch.rand(1, generator=g)) * 4 - 2, 1.7, 0.02 + 0.03 * float(torch.rand(1, generator=g))])
    return cams


def _run_reference(nodes, boxes, cams, steps, max_nodes):
    with tempfile.TemporaryDirectory() as d:
        a, b = os.path.join(d, "in.blob"), os.path.join(d, "out.blob")
        with open(a, "wb") as f:
            f.write(np.array([nodes.shape[0], max_nodes, len(cams), steps], np.int32).tobytes())
            f.write(np.ascontiguousarray(nodes, np.int32).tobytes())
            f.write(np.ascontiguousarray(boxes, np.float32).reshape(-1).tobytes())
            f.write(np.array(cams, np.float32).tobytes())
        subprocess.run([TOOL, a, b], check=True)
        raw = np.fromfile(b, np.int32)
    out, off = [], 0
    for _ in range(len(cams) * steps):
        ok, n, r, need = raw[off:off + 4]
        off += 4
        rec = {"ok": int(ok), "N": int(n), "R": int(r), "need": int(need)}
        if ok:
            rec["nodes"] = raw[off:off + n]; off += n
            rec["render"] = raw[off:off + r]; off += r
            rec["parent"] = raw[off:off + r]; off += r
            rec["nodes_of"] = raw[off:off + r]; off += r
        rec["expand"] = raw[off:off + need]; off += need
        out.append(rec)
    assert off == raw.size
    return out


@pytest.mark.skipif(not os.path.exists(TOOL), reason="oracle/_ref/ref_switch_tool not on this box")
@pytest.mark.parametrize("levels,max_frac", [(12, 1.0), (15, 1.0), (14, 0.05)])
def test_every_step_matches_the_reference_changeToSizeStep(levels, max_frac):
    import gaussian_hierarchy._C as gh
    attrs = scenes.build_hierarchy(levels, seed=3, device="cpu")
    nodes, boxes = attrs["nodes"].numpy(), attrs["boxes"].numpy()
    if max_frac == 1.0:
        nodes = nodes.copy()
        nodes[::37, 5] = np.where(nodes[::37, 6] > 0, -1, nodes[::37, 5])   # some nodes' children "not resident"
    max_nodes = max(8, int(nodes.shape[0] * max_frac))     # 5 %: the list outgrows its capacity on the way down
    cams, steps = _path(5), 6
    ref = _run_reference(nodes, boxes, cams, steps, max_nodes)
    cut = gh.IncrementalCut(torch.from_numpy(nodes).to(DEV), torch.from_numpy(boxes).to(DEV), max_nodes=max_nodes)
    k = 0
    saw_fail = saw_need = False
    for cam in cams:
        vp = torch.tensor(cam[:3], device=DEV)
        for _ in range(steps):
            ok, n, r = cut.step(cam[3], vp)
            want = ref[k]
            k += 1
            assert int(ok) == want["ok"] and cut.need_expansion == want["need"]
            saw_fail |= not ok
            saw_need |= want["need"] > 0
            got_expand = np.sort(cut.nodes_to_expand[:cut.need_expansion].cpu().numpy())
            assert np.array_equal(got_expand, np.sort(want["expand"]))
            if not ok:
                continue
            assert (n, r) == (want["N"], want["R"])
            assert np.array_equal(cut.current_nodes.cpu().numpy(), want["nodes"])
            assert np.array_equal(cut.render_indices[:r].cpu().numpy(), want["render"])
            assert np.array_equal(cut.parent_indices[:r].cpu().numpy(), want["parent"])
            assert np.array_equal(cut.nodes_of_render_indices[:r].cpu().numpy(), want["nodes_of"])
    if max_frac < 1.0:
        assert saw_fail, "the capacity case never overflowed"
    else:
        assert saw_need


def test_iterating_the_step_converges_to_expand_to_size():
    import gaussian_hierarchy._C as gh
    attrs = scenes.build_hierarchy(14, seed=5, device=DEV)
    nodes, boxes = attrs["nodes"], attrs["boxes"]
    n = nodes.size(0)
    vp = torch.tensor([3.0, -1.0, 1.8], device=DEV)
    target = 0.03
    cut = gh.IncrementalCut(nodes, boxes)
    prev = None
    for _ in range(40):
        ok, cnt, r = cut.step(target, vp)
        assert ok
        state = (cnt, r)
        if state == prev:
            break
        prev = state
    ri = torch.zeros(n, dtype=torch.int32, device=DEV)
    pi, ni = torch.zeros_like(ri), torch.zeros_like(ri)
    K = gh.expand_to_size(nodes, boxes, target, vp, torch.zeros(3), ri, pi, ni)
    assert r == K
    got = torch.stack((cut.render_indices[:r], cut.parent_indices[:r]), 1)
    want = torch.stack((ri[:K], pi[:K]), 1)
    # same set of (gaussian, parent gaussian) pairs; the incremental list keeps its own order
    assert torch.equal(got[got[:, 0].argsort()], want[want[:, 0].argsort()])
