#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE.

TEST INFRASTRUCTURE.  The reference has no tests, golden images or known-answer vectors for this path
(SURVEY.md section 4), so the oracle is pinned against outputs of the reference itself:

  * GPU part  (``python tests/golden/make_golden.py gpu``, needs a CUDA device and oracle/_ref/*.so --
    the reference's CUDA extensions compiled from /root/reference by oracle/build_ref.py): runs
    ``rasterize_gaussians`` / ``rasterize_gaussians_backward`` (HR/rasterize_points.cu:35-242) and
    ``expand_to_size`` / ``get_interpolation_weights`` (GH/torch/torch_interface.cpp:77-120) on small
    seeded scenes and stores inputs, every intermediate buffer the reference keeps (decoded with the
    layouts of HR/cuda_rasterizer/rasterizer_impl.cu:159-199) and all outputs.
  * CPU part  (``python tests/golden/make_golden.py cpu``, needs /root/reference): imports the
    reference's scene/OurAdam.py and the compiled gaussian_hierarchy extension's CPU entry points
    (expand_to_target, write_hierarchy/load_hierarchy) and stores their outputs.

Outputs are written to $H3DGS_GOLDEN_OUT (default: this directory) as compressed .npz files.
The fixtures are committed; this script documents exactly how they were made.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from harness import add_paths  # noqa: E402

add_paths()
from harness import refglue, scenes  # noqa: E402

OUT = os.environ.get("H3DGS_GOLDEN_OUT", HERE)


def _np(t):
    return t.detach().cpu().contiguous().numpy().copy()   # copy: CPU tensors share memory with numpy


def _save(name, d):
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **d)
    print(f"[golden] {path}: {os.path.getsize(path) / 1024:.0f} KiB, {len(d)} arrays")


def raster_case(name, n, w, h, seed, interp=False, depth=True, degree=3, colors_precomp=False, yaw=5.0,
                extent=(14.0, 12.0, 8.0), bg=(0.1, 0.2, 0.3), scale_modifier=1.0):
    dev = "cuda"
    sc = scenes.street_gaussians(n, seed=seed, device=dev, extent=extent)
    cam = scenes.street_camera(w, h, seed=seed, device=dev, yaw_deg=yaw, x=-extent[0] * 0.9)
    bgt = torch.tensor(bg, device=dev)
    ts = kids = None
    if interp:
        g = torch.Generator().manual_seed(seed + 5)
        ts = torch.rand(n, generator=g)
        ts[::3] = 1.0
        ts = ts.to(dev)
        kids = torch.randint(1, 6, (n,), generator=g, dtype=torch.int32).to(dev)
    rs = refglue.make_settings(refglue.RefSettings, cam, bgt, sh_degree=degree, do_depth=depth,
                               interpolation_weights=ts, num_node_kids=kids, scale_modifier=scale_modifier)
    g = torch.Generator().manual_seed(seed + 7)
    dcol = (torch.randn(3, h, w, generator=g) / (w * h)).to(dev)
    ddep = (torch.randn(1, h, w, generator=g) / (w * h)).to(dev)
    colors = None
    if colors_precomp:
        colors = torch.rand(n, 3, generator=g).to(dev)

    C = refglue.ref_hr()
    e = torch.Tensor([])
    fargs = (rs.bg, rs.render_indices, rs.parent_indices, rs.interpolation_weights, rs.num_node_kids,
             sc["means3D"], e if colors is None else colors, sc["opacities"], sc["scales"], sc["rotations"],
             rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
             rs.image_width, e if colors is not None else sc["shs"], rs.sh_degree, rs.campos, False, False,
             rs.do_depth)
    R, color, radii, geom, binning, img, invdepth = C.rasterize_gaussians(*fargs)
    torch.cuda.synchronize()
    bufs = refglue.parse_ref_buffers(geom, binning, img, n, R, w, h)
    bargs = (rs.bg, rs.render_indices, rs.parent_indices, rs.interpolation_weights, rs.num_node_kids,
             sc["means3D"], radii, e if colors is None else colors, sc["opacities"], sc["scales"],
             sc["rotations"], rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
             dcol, ddep if depth else torch.empty(0, device=dev), e if colors is not None else sc["shs"],
             rs.sh_degree, rs.campos, geom, R, binning, img, False)
    # the reference's gradients are not deterministic (float atomics): store two runs so that the
    # tests know the spread
    g1 = C.rasterize_gaussians_backward(*bargs)
    g2 = C.rasterize_gaussians_backward(*bargs)
    torch.cuda.synchronize()
    names = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations"]
    d = dict(
        n=np.int64(n), W=np.int64(w), H=np.int64(h), degree=np.int64(degree), depth=np.int64(depth),
        interp=np.int64(interp), scale_modifier=np.float32(scale_modifier),
        tanfovx=np.float64(rs.tanfovx), tanfovy=np.float64(rs.tanfovy),
        means3D=_np(sc["means3D"]), shs=_np(sc["shs"]), opacities=_np(sc["opacities"]),
        scales=_np(sc["scales"]), rotations=_np(sc["rotations"]),
        viewmatrix=_np(rs.viewmatrix), projmatrix=_np(rs.projmatrix), campos=_np(rs.campos), bg=_np(bgt),
        dL_dcolor=_np(dcol), dL_dinvdepth=_np(ddep),
        num_rendered=np.int64(R), color=_np(color), radii=_np(radii), invdepth=_np(invdepth))
    if ts is not None:
        d["ts"] = _np(ts)
        d["kids"] = _np(kids)
    if colors is not None:
        d["colors_precomp"] = _np(colors)
        del d["shs"]
    for k in ("depths", "clamped", "means2D", "cov3D", "conic_opacity", "rgb", "tiles_touched", "point_list",
              "point_list_keys", "ranges", "final_T", "n_contrib"):
        if k in bufs:
            d["ref_" + k] = _np(bufs[k])
    for k, a, b in zip(names, g1, g2):
        d[k] = _np(a)
        d[k + "_run2"] = _np(b)
    _save(name, d)


def cut_case(name, levels, seed, limits, w=640, h=480):
    dev = "cuda"
    attrs = scenes.build_hierarchy(levels, seed=seed, device=dev, extent=(30.0, 12.0, 10.0))
    gh = refglue.ref_gh()
    nodes, boxes = attrs["nodes"], attrs["boxes"]
    N = nodes.size(0)
    d = dict(nodes=_np(nodes), boxes=_np(boxes), limits=np.asarray(limits, dtype=np.float32))
    cams = [scenes.street_camera(w, h, seed=seed + i, device=dev, x=x, yaw_deg=10.0 * i, height_m=hm)
            for i, (x, hm) in enumerate([(-25.0, 1.8), (0.0, 5.0), (80.0, 2.0)])]
    d["viewpoints"] = np.stack([_np(c.camera_center) for c in cams])
    for ci, cam in enumerate(cams):
        for li, limit in enumerate(limits):
            ri = torch.zeros(N, dtype=torch.int32, device=dev)
            pi = torch.zeros(N, dtype=torch.int32, device=dev)
            ni = torch.zeros(N, dtype=torch.int32, device=dev)
            tsb = torch.zeros(N, dtype=torch.float32, device=dev)
            kb = torch.zeros(N, dtype=torch.int32, device=dev)
            K = gh.expand_to_size(nodes, boxes, float(limit), cam.camera_center, torch.zeros(3), ri, pi, ni)
            if K > 0:   # the reference launches a 0-block grid for an empty cut (invalid configuration)
                gh.get_interpolation_weights(ni[:K].contiguous(), float(limit), nodes, boxes,
                                             cam.camera_center.cpu(), torch.zeros(3), tsb, kb)
            torch.cuda.synchronize()
            key = f"c{ci}_l{li}_"
            d[key + "count"] = np.int64(K)
            d[key + "render_indices"] = _np(ri[:K])
            d[key + "parent_indices"] = _np(pi[:K])
            d[key + "nodes_for_render_indices"] = _np(ni[:K])
            d[key + "ts"] = _np(tsb[:K])
            d[key + "kids"] = _np(kb[:K])
    nodes_cpu = nodes.cpu()
    for target in (0, 1, 2, 4, levels, levels + 3):
        d[f"target_{target}"] = _np(gh.expand_to_target(nodes_cpu, target))
    _save(name, d)


def render_post_case(name, levels, sky, w, h, target, seed):
    """The Python glue of gaussian_renderer/__init__.py:138-304 (restated in harness/refglue.py on the
    reference rasterizer) on a hierarchy cut with skybox rows: image + parameter gradients."""
    import types
    dev = "cuda"
    from harness import workload
    attrs = workload.build_scene(levels, sky, seed, dev)
    # shrink: workload.build_scene uses the 100 m street; fine for a fixture as long as it is small
    cam = scenes.street_camera(w, h, seed=seed, device=dev, yaw_deg=4.0, x=-10.0)
    bg = torch.tensor([0.05, 0.1, 0.15], device=dev)
    gh = refglue.ref_gh()
    model = workload.ActivatedModel(attrs, sky)
    limit = workload.find_limit(gh, model, cam, target, dev)
    n = attrs["means3D"].size(0)
    g = torch.Generator().manual_seed(11)
    dcol = (torch.randn(3, h, w, generator=g) / (w * h)).to(dev)
    ddep = (torch.randn(1, h, w, generator=g) / (w * h)).to(dev)
    runs = []
    for _ in range(2):
        m = workload.ActivatedModel(attrs, sky)
        cut = workload.Cut(n, dev)
        cut.compute(gh, m, cam, limit)
        out = workload.ref_render_post(cam, m, bg, cut, do_depth=True)
        ((out["render"] * dcol).sum() + (out["depth"] * ddep).sum()).backward()
        torch.cuda.synchronize()
        runs.append((m, cut, out))
    m, cut, out = runs[0]
    K = cut.count
    d = dict(W=np.int64(w), H=np.int64(h), sky=np.int64(sky), limit=np.float32(limit), K=np.int64(K),
             tanfovx=np.float64(np.tan(cam.FoVx * 0.5)), tanfovy=np.float64(np.tan(cam.FoVy * 0.5)),
             means3D=_np(attrs["means3D"]), shs=_np(attrs["shs"]), opacities=_np(attrs["opacities"]),
             scales=_np(attrs["scales"]), rotations=_np(attrs["rotations"]), nodes=_np(attrs["nodes"]),
             boxes=_np(attrs["boxes"]), viewmatrix=_np(cam.world_view_transform),
             projmatrix=_np(cam.full_proj_transform), campos=_np(cam.camera_center), bg=_np(bg),
             dL_dcolor=_np(dcol), dL_dinvdepth=_np(ddep),
             render_indices=_np(cut.render_indices[:K]), parent_indices=_np(cut.parent_indices[:K]),
             ts=_np(cut.interpolation_weights[:K]), kids=_np(cut.num_siblings[:K + sky]),
             render=_np(out["render"]), depth=_np(out["depth"]), radii_visible=_np(out["radii"]),
             visibility_filter=_np(out["visibility_filter"]))
    for nm, p, p2 in zip(["g_xyz", "g_features", "g_opacity", "g_scaling", "g_rotation"], m.parameters(),
                         runs[1][0].parameters()):
        # store sparse: most rows of the store receive no gradient
        gsum = p.grad.reshape(p.grad.size(0), -1).abs().sum(1) + p2.grad.reshape(p.grad.size(0), -1).abs().sum(1)
        rows = gsum.nonzero().flatten()
        d[nm + "_rows"] = _np(rows)
        d[nm] = _np(p.grad[rows])
        d[nm + "_run2"] = _np(p2.grad[rows])
    d["g_viewspace"] = _np(out["viewspace_points"].grad[:K + sky])
    _save(name, d)


def gpu():
    assert torch.cuda.is_available()
    assert refglue.ref_hr() is not None and refglue.ref_gh() is not None, "oracle/_ref not built"
    raster_case("raster_plain", n=1200, w=160, h=96, seed=3)
    raster_case("raster_interp", n=1200, w=144, h=112, seed=4, interp=True, yaw=-8.0)
    raster_case("raster_deg1_nodepth", n=600, w=100, h=70, seed=5, depth=False, degree=1, bg=(0.0, 0.0, 0.0),
                scale_modifier=0.8)
    raster_case("raster_colors_precomp", n=500, w=96, h=64, seed=6, colors_precomp=True)
    cut_case("cut_l9", levels=9, seed=2, limits=[1e-9, 0.002, 0.02, 0.2, 5.0, 1e6])
    render_post_case("render_post_l8", levels=8, sky=40, w=128, h=80, target=300, seed=8)


def cpu():
    """Golden vectors from reference code that runs without a GPU."""
    import importlib.util
    ref = os.environ.get("H3DGS_REFERENCE_ROOT", "/root/reference")
    spec = importlib.util.spec_from_file_location("ref_ouradam", os.path.join(ref, "scene", "OurAdam.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    # torch >= 2.5 renamed the Optimizer hook the reference calls first (OurAdam.py:113); give the old
    # name back (a no-op outside CUDA-graph capture) instead of touching the reference
    if not hasattr(mod.Adam, "_cuda_graph_capture_health_check"):
        mod.Adam._cuda_graph_capture_health_check = lambda self: None
    g = torch.Generator().manual_seed(21)
    rows = 257
    widths = [3, 3, 45, 1, 3, 4]
    lrs = [1.6e-4, 0.0025, 0.0025 / 20.0, 0.05, 0.005, 0.001]
    params = [torch.randn(rows, wd, generator=g).requires_grad_(True) for wd in widths]
    d = dict(widths=np.asarray(widths), lrs=np.asarray(lrs, dtype=np.float64), rows=np.int64(rows))
    for i, p in enumerate(params):
        d[f"p{i}_init"] = _np(p)
    opt = mod.Adam([{"params": [p], "lr": lr} for p, lr in zip(params, lrs)], lr=0.0, eps=1e-15)
    steps = 4
    d["steps"] = np.int64(steps)
    for s in range(steps):
        if s == 2:
            rel = torch.empty(0, dtype=torch.long)     # empty `relevant` => dense update (OurAdam.py:214-228)
        else:
            rel = torch.randperm(rows, generator=g)[: 40 + 30 * s].sort().values
        d[f"s{s}_relevant"] = _np(rel)
        for i, p in enumerate(params):
            p.grad = torch.randn(rows, widths[i], generator=g) * (10.0 ** float(torch.randint(-6, 1, (1,), generator=g)))
            d[f"s{s}_g{i}"] = _np(p.grad)
        opt.step(rel)
        for i, p in enumerate(params):
            st = opt.state[p]
            d[f"s{s}_p{i}"] = _np(p)
            d[f"s{s}_m{i}"] = _np(st["exp_avg"])
            d[f"s{s}_v{i}"] = _np(st["exp_avg_sq"])
    _save("adam_ouradam", d)



def hier_case():
    """.hier written by the reference's writer (GH/hierarchy_writer.cpp:27-118, compressed fp16 form) and
    read back by its loader (GH/hierarchy_loader.cpp:26-128).  Runs with CUDA tensors like the reference's
    own caller (scene/gaussian_model.py:437-445); the binding crashes on CPU tensors in this image."""
    gh = refglue.ref_gh()
    attrs = scenes.build_hierarchy(5, seed=9, device="cuda", extent=(10.0, 5.0, 5.0))
    path = os.path.join(OUT, "ref_written_l5.hier")
    gh.write_hierarchy(path, attrs["means3D"], attrs["shs"], attrs["opacities"], attrs["log_scales"],
                       attrs["rotations"], attrs["nodes"], attrs["boxes"])
    pos, shs, alpha, scale, rot, nodes, boxes = gh.load_hierarchy(path)
    _save("hier_roundtrip_l5", dict(
        in_means3D=_np(attrs["means3D"]), in_shs=_np(attrs["shs"]), in_opacities=_np(attrs["opacities"]),
        in_log_scales=_np(attrs["log_scales"]), in_rotations=_np(attrs["rotations"]),
        in_nodes=_np(attrs["nodes"]), in_boxes=_np(attrs["boxes"]),
        pos=_np(pos), shs=_np(shs), alpha=_np(alpha), scale=_np(scale), rot=_np(rot), nodes=_np(nodes),
        boxes=_np(boxes), file_bytes=np.frombuffer(open(path, "rb").read(), dtype=np.uint8)))
    os.unlink(path)


if __name__ == "__main__":
    which = sys.argv[1:] or ["gpu"]
    if "gpu" in which:
        gpu()
    if "cpu" in which:
        cpu()
    if "hier" in which:
        hier_case()
