"""CPU-only checks: the C-ABI library loads and exports every symbol include/h3dgs_b200.h declares, the
Python mirror keeps the reference's interface, host-side (.hier I/O, DFS cut) results match the
reference build and an independent restatement.  No compute kernels are launched here."""
import os
import re

import numpy as np
import pytest
import torch

from harness import ROOT, refglue, scenes


def _declared():
    text = open(os.path.join(ROOT, "include", "h3dgs_b200.h")).read()
    return sorted(set(re.findall(r"H3DGS_API[^;(]*?\b(h3dgs_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import ctypes
    import h3dgs_native as nat
    names = _declared()
    assert len(names) >= 20
    handle = ctypes.CDLL(nat.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), f"{n} declared in the header but not exported"
        assert n in nat.SIGNATURES, f"{n} has no ctypes prototype"
    assert sorted(nat.SIGNATURES) == names
    assert nat.lib().h3dgs_abi_version() == nat.ABI_VERSION == 8


def test_workspace_queries_are_sane():
    import h3dgs_native as nat
    L = nat.lib()
    assert L.h3dgs_geom_bytes(0) > 0
    assert L.h3dgs_geom_bytes(2_000_000) > L.h3dgs_geom_bytes(1_000_000) >= 1_000_000 * (64 + 4 + 24)
    assert L.h3dgs_image_bytes(1920, 1080) >= 1920 * 1080 * 24
    assert L.h3dgs_binning_bytes(1_000_000, 1920, 1080) >= 20 * 1_000_000
    # the depth sort of the visible slots is part of the geometry workspace (sized by P, before R is known)
    assert L.h3dgs_geom_bytes(1_000_000) >= 1_000_000 * (64 + 4 + 24 + 20)
    assert L.h3dgs_backward_workspace_bytes(1000) >= 48 * 1000
    assert L.h3dgs_expand_workspace_bytes(1000) >= 8 * 1000


def test_usage_errors_are_reported_not_crashed():
    import ctypes
    import h3dgs_native as nat
    L = nat.lib()
    rc = L.h3dgs_forward_begin(None, None, None, None, 0, None, None, 0, None)
    assert rc == -1 and b"null" in L.h3dgs_last_error()
    cnt = ctypes.c_int32(7)
    assert L.h3dgs_expand_to_size(-5, 1.0, None, None, None, None, None, None, None, 0, ctypes.byref(cnt), None) == -1
    assert L.h3dgs_sparse_adam(None, 0, None, 0, 0, None) == -1
    # ABI v5 additions: argument checks come before any device work
    assert L.h3dgs_zero_fill(None, 3, 0, None) == -1 and L.h3dgs_zero_fill(None, 0, 0, None) == 0
    assert L.h3dgs_zero_fill(None, 0, 9, None) == -1
    span = (nat.Span * 1)(nat.Span(None, -4))
    assert L.h3dgs_zero_fill(span, 1, 0, None) == -1
    assert L.h3dgs_pack_rows(None, 2, None, 5, None, None) == -1
    rt = (nat.RowTensor * 1)()
    assert L.h3dgs_unpack_rows(rt, 1, None, 0, None, 1.0, None) == 0      # no rows: nothing to do
    assert L.h3dgs_unpack_rows(rt, 1, None, 3, None, 1.0, None) == -1     # rows / payload missing
    w = (ctypes.c_float * 11)(*([1.0 / 11] * 11))
    assert L.h3dgs_ssim_forward(3, 0, 16, None, None, None, w, None, None, None, None, None, None) == 0   # empty image
    assert L.h3dgs_ssim_forward(3, 8, 8, None, None, None, w, None, None, None, None, None, None) == -1
    assert L.h3dgs_ssim_backward(3, 8, 8, None, None, None, w, None, None, None, None, None, 1.0, 0.0, None, None) == -1
    assert L.h3dgs_masked_l1_forward(0, None, None, None, None, None) == 0      # empty
    assert L.h3dgs_masked_l1_forward(-1, None, None, None, None, None) == -1
    assert L.h3dgs_masked_l1_forward(16, None, None, None, None, None) == -1
    assert L.h3dgs_masked_l1_backward(16, None, None, None, None, 1.0, None, None) == -1
    assert L.h3dgs_blend_stats(None) == 0
    with pytest.raises(nat.H3dgsError):
        nat.check(-1, "x")


def test_settings_fields_match_reference_order():
    import diff_gaussian_rasterization as dgr
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
        "projmatrix", "sh_degree", "campos", "prefiltered", "debug", "render_indices", "parent_indices",
        "interpolation_weights", "num_node_kids", "do_depth")
    for name in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert callable(getattr(dgr._C, name))
    import gaussian_hierarchy
    for name in ("load_hierarchy", "write_hierarchy", "expand_to_target", "expand_to_size",
                 "get_interpolation_weights"):
        assert callable(getattr(gaussian_hierarchy._C, name))
    import gaussian_renderer
    for name in ("render", "render_post", "render_coarse"):
        assert callable(getattr(gaussian_renderer, name))


def test_rasterizer_argument_checks_raise_like_the_reference():
    import diff_gaussian_rasterization as dgr
    r = dgr.GaussianRasterizer(None)
    x = torch.zeros(2, 3)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(means3D=x, means2D=x, opacities=x[:, :1])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(means3D=x, means2D=x, opacities=x[:, :1], shs=x, colors_precomp=x)
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(means3D=x, means2D=x, opacities=x[:, :1], shs=x)
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(means3D=x, means2D=x, opacities=x[:, :1], shs=x, scales=x, rotations=x, cov3D_precomp=x)


def test_our_adam_constructor_validation_and_state_layout():
    from scene.OurAdam import Adam
    p = torch.nn.Parameter(torch.zeros(4, 3))
    with pytest.raises(ValueError, match="Invalid learning rate"):
        Adam([p], lr=-1.0)
    with pytest.raises(ValueError, match="Invalid epsilon"):
        Adam([p], eps=-1.0)
    with pytest.raises(ValueError, match="beta parameter at index 0"):
        Adam([p], betas=(1.0, 0.9))
    opt = Adam([{"params": [p], "lr": 0.1, "name": "xyz"}], lr=0.0, eps=1e-15)
    assert opt.param_groups[0]["eps"] == 1e-15 and opt.param_groups[0]["betas"] == (0.9, 0.999)
    p.grad = torch.ones_like(p)
    with pytest.raises(RuntimeError, match="CUDA"):   # no CPU fallback
        opt.step(torch.tensor([0, 1]))


def _iterative_expand(nodes, target):
    # independent restatement of GH/traversal.cpp:15-38
    out, stack = [], [0]
    while stack:
        i = stack.pop()
        depth, parent, start, leafs, merged, sc, cc = [int(v) for v in nodes[i]]
        out.extend(range(start, start + leafs))
        if depth <= target:
            out.extend(range(start + leafs, start + leafs + merged))
        else:
            stack.extend(range(sc + cc - 1, sc - 1, -1))
    return np.array(out, dtype=np.int32)


@pytest.mark.parametrize("levels", [1, 5, 9])
def test_expand_to_target(levels):
    import gaussian_hierarchy
    attrs = scenes.build_hierarchy(levels, seed=levels, device="cpu")
    nodes = attrs["nodes"]
    for target in (0, 1, 2, levels, levels + 3):
        got = gaussian_hierarchy._C.expand_to_target(nodes, target)
        assert got.dtype == torch.int32
        assert np.array_equal(got.numpy(), _iterative_expand(nodes.numpy(), target))
        if refglue.ref_gh() is not None:
            assert torch.equal(got, refglue.ref_gh().expand_to_target(nodes, target))
    # leaf cut of a DFS-ordered hierarchy: ascending Gaussian ids, one per leaf
    leaves = gaussian_hierarchy._C.expand_to_target(nodes, 0)
    assert leaves.numel() == 1 << levels
    assert bool((leaves[1:] > leaves[:-1]).all())


def _hier_tensors(levels, seed):
    a = scenes.build_hierarchy(levels, seed=seed, device="cpu")
    return (a["means3D"], a["shs"], a["opacities"], a["log_scales"], a["rotations"], a["nodes"], a["boxes"])


def test_hier_write_read_roundtrip_and_half_precision(tmp_path):
    import gaussian_hierarchy
    C = gaussian_hierarchy._C
    pos, shs, op, sc, rot, nodes, boxes = _hier_tensors(7, 3)
    path = str(tmp_path / "a.hier")
    C.write_hierarchy(path, pos, shs, op, sc, rot, nodes, boxes)          # compressed, like the reference binding
    got = C.load_hierarchy(path)
    assert [tuple(t.shape) for t in got] == [tuple(t.shape) for t in (pos, shs, op, sc, rot, nodes, boxes)]
    assert torch.equal(got[0], pos)                                        # positions stay fp32
    f16 = lambda t: torch.from_numpy(t.numpy().astype(np.float16).astype(np.float32))
    assert torch.equal(got[1], f16(shs)) and torch.equal(got[2], f16(op))
    assert torch.equal(got[3], f16(sc)) and torch.equal(got[4], f16(rot))
    assert torch.equal(got[5], nodes) and torch.equal(got[6], f16(boxes))
    # fp32 container round-trips exactly
    path2 = str(tmp_path / "b.hier")
    C.write_hierarchy(path2, pos, shs, op, sc, rot, nodes, boxes, False)
    for a, b in zip(C.load_hierarchy(path2), (pos, shs, op, sc, rot, nodes, boxes)):
        assert torch.equal(a, b)
    expect = 4 + pos.size(0) * (12 + 16 + 12 + 4 + 192) + 4 + nodes.size(0) * (28 + 32)
    assert os.path.getsize(path2) == expect
    assert os.path.getsize(path) == 4 + pos.size(0) * (12 + 8 + 6 + 2 + 96) + 4 + nodes.size(0) * (20 + 16)


def test_hier_special_values_and_errors(tmp_path):
    import gaussian_hierarchy
    import h3dgs_native as nat
    C = gaussian_hierarchy._C
    pos, shs, op, sc, rot, nodes, boxes = _hier_tensors(3, 1)
    # rounding edge cases of the fp16 conversion: ties, subnormals, overflow, signed zero, inf
    vals = torch.tensor([0.0, -0.0, 1.0, 1.0009765625, 1.00048828125, 1.00146484375, 65504.0, 65519.9, 65520.0,
                         1e6, 6.1e-5, 5.96e-8, 2.98e-8, 2.9802322e-8, 3.0e-8, -1e-10, float("inf"), -float("inf")])
    sc = sc.clone()
    sc.view(-1)[:vals.numel()] = vals
    path = str(tmp_path / "c.hier")
    C.write_hierarchy(path, pos, shs, op, sc, rot, nodes, boxes)
    got = C.load_hierarchy(path)[3]
    want = torch.from_numpy(sc.numpy().astype(np.float16).astype(np.float32))
    assert torch.equal(got.view(torch.int32), want.view(torch.int32))
    with pytest.raises(RuntimeError, match="File not found"):
        C.load_hierarchy(str(tmp_path / "missing.hier"))
    bad = nodes.clone(); bad[0, 0] = 40000
    with pytest.raises(nat.H3dgsError, match="Would lose information"):
        C.write_hierarchy(str(tmp_path / "d.hier"), pos, shs, op, sc, rot, bad, boxes)


def _numpy_hier_bytes(pos, shs, op, sc, rot, nodes, boxes):
    """Independent construction of the compressed container (GH/hierarchy_writer.cpp:59-118)."""
    P, N = pos.size(0), nodes.size(0)
    h = lambda t: t.numpy().astype(np.float16).tobytes()
    hn = np.zeros(N, dtype=np.dtype([("parent", "<i4"), ("start", "<i4"), ("start_children", "<i4"),
                                     ("dccc", "<i2", (4,))]))
    nd = nodes.numpy()
    hn["parent"], hn["start"], hn["start_children"] = nd[:, 1], nd[:, 2], nd[:, 5]
    hn["dccc"] = np.stack([nd[:, 0], nd[:, 6], nd[:, 3], nd[:, 4]], axis=1).astype(np.int16)
    return (np.int32(-P).tobytes() + pos.numpy().tobytes() + h(rot) + h(sc) + h(op) + h(shs) +
            np.int32(N).tobytes() + hn.tobytes() + h(boxes))


def test_hier_files_interchange_with_reference_build(tmp_path):
    import gaussian_hierarchy
    C = gaussian_hierarchy._C
    t = _hier_tensors(6, 9)
    ours_path = str(tmp_path / "ours.hier")
    C.write_hierarchy(ours_path, *t)
    assert open(ours_path, "rb").read() == _numpy_hier_bytes(*t), "compressed .hier bytes differ from the format"
    ref = refglue.ref_gh()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    # the reference loader reads our file and agrees with our loader on every tensor
    for a, b in zip(C.load_hierarchy(ours_path), ref.load_hierarchy(ours_path)):
        assert torch.equal(a, b)
    # (the reference's write_hierarchy binding segfaults on CPU tensors in this environment, so the
    #  byte-level comparison against the reference writer runs on the GPU box with CUDA tensors)


def test_hier_files_match_reference_writer_and_loader_fixture(tmp_path):
    """Host-side .hier I/O against files written / read by the reference's own C++ classes (tests/hier_fixture.py)."""
    from tests import hier_fixture
    hier_fixture.check(tmp_path, "cpu")


def test_torch_extension_loads_and_matches_the_abi():
    """lib/_h3dgs_torch.so (csrc/torch_binding.cpp) is the default host path: it must be built, import without a
    GPU, resolve libh3dgs_b200.so through its $ORIGIN rpath and agree on the ABI version."""
    import diff_gaussian_rasterization as dgr
    import h3dgs_native as nat
    E = dgr._C.ext()
    assert E is not None, "torch extension not built (python street-sparse-3dgs_b200/build.py)"
    assert E.abi_version() == nat.ABI_VERSION
    for name in ("rasterize_forward", "rasterize_backward", "row_map", "Prezero"):
        assert hasattr(E, name)


def test_overlay_installer_replaces_three_files_and_restores_them(tmp_path):
    """Python puts the script's directory first, so the reference's own gaussian_renderer /
    scene / utils win over PYTHONPATH; the three pure-Python drop-ins are overlaid into the checkout instead."""
    import importlib.util
    import os
    pkg = os.path.join(ROOT, "street-sparse-3dgs_b200")
    spec = importlib.util.spec_from_file_location("h3dgs_install", os.path.join(pkg, "install_into_reference.py"))
    inst = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(inst)
    for rel in inst.FILES + ("scene/__init__.py", "utils/general_utils.py"):
        p = tmp_path / rel
        p.parent.mkdir(parents=True, exist_ok=True)
        p.write_text("# reference " + rel + "\n")
    assert inst.main([str(tmp_path)]) == 0
    for rel in inst.FILES:
        text = (tmp_path / rel).read_text()
        assert text.startswith(inst.MARK) and open(os.path.join(pkg, rel)).read() in text
        assert (tmp_path / (rel + ".reference_orig")).read_text() == "# reference " + rel + "\n"
    assert (tmp_path / "scene/__init__.py").read_text() == "# reference scene/__init__.py\n"   # untouched
    assert inst.main([str(tmp_path), "--undo"]) == 0
    for rel in inst.FILES:
        assert (tmp_path / rel).read_text() == "# reference " + rel + "\n"
    # what the reference imports from utils.loss_utils across its scripts (train_*.py, render_hierarchy_final.py:5,154)
    src = open(os.path.join(pkg, "utils", "loss_utils.py")).read()
    for name in ("def l1_loss", "def l2_loss", "def ssim(", "def ssim_masked", "def create_window", "def gaussian("):
        assert name in src, name
