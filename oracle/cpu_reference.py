"""TEST INFRASTRUCTURE ONLY -- numpy/ctypes front end of the CPU oracle (oracle/h3dgs_oracle.c).

Restates, on the host, the path the reference executes on the GPU:

  rasterize_forward / rasterize_backward   HR/rasterize_points.cu:35-242 -> HR/cuda_rasterizer/rasterizer_impl.cu:203-510
  render_post / render_post_backward       gaussian_renderer/__init__.py:138-304 (Python gather + lerp glue, interp_python=True)
  expand_to_size / get_interpolation_weights / expand_to_target
                                           GH/torch/torch_interface.cpp:69-120
  sparse_adam                              scene/OurAdam.py:105-175,249-337

Parity status: pinned against tests/golden/*.npz (outputs of the unmodified reference, see
tests/golden/make_golden.py) by tests/test_cpu_oracle.py.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os

import numpy as np

from . import build_oracle

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = build_oracle.LIB if os.path.exists(build_oracle.LIB) else build_oracle.build()
        _lib = C.CDLL(path)
        _lib.orc_bin.restype = C.c_int64
        _lib.orc_expand_to_target.restype = C.c_int64
        _lib.orc_expand_to_size.restype = C.c_int
        _lib.orc_higher_msb.restype = C.c_uint32
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.int32)


def rasterize_forward(means3D, opacities, scales, rotations, viewmatrix, projmatrix, campos, bg, W, H, tanfovx,
                      tanfovy, shs=None, colors_precomp=None, degree=3, scale_modifier=1.0, ts=None, kids=None,
                      do_depth=True, cov3D_precomp=None):
    """RasterizeGaussiansCUDA (HR/rasterize_points.cu:35-136).  Returns a dict with every buffer the
    reference keeps (names of HR/cuda_rasterizer/rasterizer_impl.h:21-62) plus color / radii / invdepth."""
    L = lib()
    means3D, opacities = _f32(means3D), _f32(opacities)
    scales, rotations, shs, colors_precomp = _f32(scales), _f32(rotations), _f32(shs), _f32(colors_precomp)
    cov3D_precomp = _f32(cov3D_precomp)
    viewmatrix, projmatrix, campos, bg = _f32(viewmatrix), _f32(projmatrix), _f32(campos), _f32(bg)
    ts, kids = _f32(ts), _i32(kids)
    P = means3D.shape[0]
    M = 0 if shs is None else shs.shape[1]
    o = dict(radii=np.zeros(P, np.int32), means2D=np.zeros((P, 2), np.float32), depths=np.zeros(P, np.float32),
             cov3D=np.zeros((P, 6), np.float32), rgb=np.zeros((P, 3), np.float32),
             conic_opacity=np.zeros((P, 4), np.float32), tiles_touched=np.zeros(P, np.uint32),
             clamped=np.zeros((P, 3), np.uint8), rects=np.zeros((P, 2), np.int32))
    L.orc_preprocess_fwd(C.c_int(P), C.c_int(degree), C.c_int(M), _p(means3D), _p(scales),
                         C.c_float(scale_modifier), _p(rotations), _p(opacities), _p(shs), _p(colors_precomp),
                         _p(cov3D_precomp), _p(viewmatrix), _p(projmatrix), _p(campos), C.c_int(W), C.c_int(H),
                         C.c_float(tanfovx), C.c_float(tanfovy), _p(o["radii"]), _p(o["means2D"]),
                         _p(o["depths"]), _p(o["cov3D"]), _p(o["rgb"]), _p(o["conic_opacity"]),
                         _p(o["tiles_touched"]), _p(o["clamped"]), _p(o["rects"]))
    if colors_precomp is not None:
        o["rgb"] = colors_precomp
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    R = int(o["tiles_touched"].sum(dtype=np.int64))
    o["num_rendered"] = R
    o["point_list_keys"] = np.zeros(max(R, 1), np.uint64)
    o["point_list"] = np.zeros(max(R, 1), np.uint32)
    o["ranges"] = np.zeros((tiles, 2), np.uint32)
    L.orc_bin(C.c_int(P), C.c_int(W), C.c_int(H), _p(o["radii"]), _p(o["means2D"]), _p(o["depths"]),
              _p(o["rects"]), _p(o["tiles_touched"]), _p(o["point_list_keys"]), _p(o["point_list"]),
              _p(o["ranges"]))
    o["point_list_keys"] = o["point_list_keys"][:R]
    o["point_list"] = o["point_list"][:R]
    o["color"] = np.zeros((3, H, W), np.float32)
    o["final_T"] = np.zeros(H * W, np.float32)
    o["n_contrib"] = np.zeros(H * W, np.uint32)
    o["invdepth"] = np.zeros((1, H, W), np.float32) if do_depth else np.zeros((0, H, W), np.float32)
    if R > 0:
        L.orc_blend_fwd(C.c_int(W), C.c_int(H), C.c_int(P), _p(o["ranges"]), _p(o["point_list"]),
                        _p(o["means2D"]), _p(o["rgb"]), _p(o["conic_opacity"]), _p(o["depths"]), _p(ts),
                        _p(kids), _p(bg), _p(o["color"]), _p(o["final_T"]), _p(o["n_contrib"]),
                        _p(o["invdepth"]) if do_depth else None)
    o["_in"] = dict(means3D=means3D, opacities=opacities, scales=scales, rotations=rotations, shs=shs,
                    colors_precomp=colors_precomp, viewmatrix=viewmatrix, projmatrix=projmatrix, campos=campos,
                    bg=bg, W=W, H=H, tanfovx=tanfovx, tanfovy=tanfovy, degree=degree,
                    scale_modifier=scale_modifier, ts=ts, kids=kids, M=M)
    return o


def rasterize_backward(fwd, dL_dcolor, dL_dinvdepth=None):
    """RasterizeGaussiansBackwardCUDA (HR/rasterize_points.cu:138-242): the 8 gradient tensors in the
    reference's return order, as a dict."""
    L = lib()
    i = fwd["_in"]
    P, W, H, M = i["means3D"].shape[0], i["W"], i["H"], i["M"]
    dL_dcolor = _f32(dL_dcolor)
    dL_dinvdepth = _f32(dL_dinvdepth) if dL_dinvdepth is not None and np.size(dL_dinvdepth) else None
    g = dict(dL_dmeans2D=np.zeros((P, 3), np.float32), dL_dconic=np.zeros((P, 2, 2), np.float32),
             dL_dopacity=np.zeros((P, 1), np.float32), dL_dcolors=np.zeros((P, 3), np.float32),
             dL_dinvdepths=np.zeros((P, 1), np.float32), dL_dmeans3D=np.zeros((P, 3), np.float32),
             dL_dcov3D=np.zeros((P, 6), np.float32), dL_dsh=np.zeros((P, M, 3), np.float32),
             dL_dscales=np.zeros((P, 3), np.float32), dL_drotations=np.zeros((P, 4), np.float32))
    if fwd["num_rendered"] == 0:
        return g
    L.orc_blend_bwd(C.c_int(W), C.c_int(H), C.c_int(P), _p(fwd["ranges"]), _p(fwd["point_list"]),
                    _p(fwd["means2D"]), _p(fwd["rgb"]), _p(fwd["conic_opacity"]), _p(fwd["depths"]), _p(i["ts"]),
                    _p(i["kids"]), _p(i["bg"]), _p(fwd["final_T"]), _p(fwd["n_contrib"]), _p(dL_dcolor),
                    _p(dL_dinvdepth), _p(g["dL_dmeans2D"]), _p(g["dL_dconic"]), _p(g["dL_dopacity"]),
                    _p(g["dL_dcolors"]), _p(g["dL_dinvdepths"]))
    L.orc_preprocess_bwd(C.c_int(P), C.c_int(i["degree"]), C.c_int(M), _p(i["means3D"]), _p(fwd["radii"]),
                         _p(i["shs"]), _p(fwd["clamped"]), _p(i["opacities"]), _p(i["scales"]),
                         _p(i["rotations"]), C.c_float(i["scale_modifier"]), _p(fwd["cov3D"]),
                         _p(i["viewmatrix"]), _p(i["projmatrix"]), C.c_int(W), C.c_int(H),
                         C.c_float(i["tanfovx"]), C.c_float(i["tanfovy"]), _p(i["campos"]),
                         _p(g["dL_dmeans2D"]), _p(g["dL_dconic"]),
                         _p(g["dL_dinvdepths"]) if dL_dinvdepth is not None else None, _p(g["dL_dopacity"]),
                         _p(g["dL_dmeans3D"]), _p(g["dL_dcolors"]), _p(g["dL_dcov3D"]), _p(g["dL_dsh"]),
                         _p(g["dL_dscales"]), _p(g["dL_drotations"]))
    return g


# ---------------------------------------------------------------------------------------------------
# render_post glue (gaussian_renderer/__init__.py:202-235): gather child + parent rows, lerp, append skybox
# ---------------------------------------------------------------------------------------------------
def render_post_gather(means3D, shs, opacities, scales, rotations, render_indices, parent_indices, ts, skybox):
    """fp32 lerp exactly as the torch expression `interps * child + interps_inv * parent` evaluates it
    (three separately rounded ops); parent == -1 indexes the last row with weight 0 (python negative index)."""
    ri = np.asarray(render_indices, dtype=np.int64)
    pi = np.asarray(parent_indices, dtype=np.int64)
    K = ri.shape[0]
    n = means3D.shape[0]
    t = np.asarray(ts[:K], dtype=np.float32)
    omt = (np.float32(1) - t).astype(np.float32)

    def lerp(a, sign=None):
        a = np.asarray(a, dtype=np.float32)
        tt = t.reshape((K,) + (1,) * (a.ndim - 1))
        oo = omt.reshape((K,) + (1,) * (a.ndim - 1))
        par = a[pi]
        if sign is not None:
            par = par * sign.reshape(K, 1)
        return (tt * a[ri]).astype(np.float32) + (oo * par).astype(np.float32)

    rot = np.asarray(rotations, dtype=np.float32)
    dots = np.einsum("ij,ij->i", rot[ri].astype(np.float64), rot[pi].astype(np.float64))
    sign = np.where(dots < 0, np.float32(-1), np.float32(1)).astype(np.float32)
    sky = np.arange(n - skybox, n)
    out = dict(means3D=np.concatenate([lerp(means3D), means3D[sky]]),
               shs=np.concatenate([lerp(shs), shs[sky]]),
               opacities=np.concatenate([lerp(opacities), opacities[sky]]),
               scales=np.concatenate([lerp(scales), scales[sky]]),
               rotations=np.concatenate([lerp(rot, sign), rot[sky]]))
    w = np.concatenate([t, np.ones(skybox, np.float32)])
    return out, w, sign


def render_post_scatter(grads_slot, n_rows, render_indices, parent_indices, ts, sign, skybox):
    """Autograd of the gather + lerp: slot gradients -> store-row gradients (summed in float64)."""
    ri = np.asarray(render_indices, dtype=np.int64)
    pi = np.asarray(parent_indices, dtype=np.int64)
    K = ri.shape[0]
    t = np.asarray(ts[:K], dtype=np.float64)
    out = {}
    for name, g in grads_slot.items():
        g = np.asarray(g, dtype=np.float64)
        flat = g.reshape(g.shape[0], -1)
        acc = np.zeros((n_rows, flat.shape[1]), np.float64)
        s = sign.astype(np.float64).reshape(K, 1) if name == "rotations" else 1.0
        np.add.at(acc, ri, flat[:K] * t.reshape(K, 1))
        np.add.at(acc, pi, flat[:K] * (1.0 - t).reshape(K, 1) * s)   # pi == -1 -> last row, weight 0
        if skybox:
            acc[n_rows - skybox:] += flat[K:K + skybox]
        out[name] = acc.reshape((n_rows,) + g.shape[1:]).astype(np.float32)
    return out


# ---------------------------------------------------------------------------------------------------
# hierarchy cut
# ---------------------------------------------------------------------------------------------------
def expand_to_size(nodes, boxes, size, viewpoint):
    nodes, boxes, vp = _i32(nodes), _f32(boxes), _f32(viewpoint)
    N = nodes.shape[0]
    cap = int(nodes[:, 3].sum(dtype=np.int64) + nodes[:, 4].sum(dtype=np.int64)) + 1
    ri, pi, ni = (np.zeros(cap, np.int32) for _ in range(3))
    K = lib().orc_expand_to_size(C.c_int(N), C.c_float(size), _p(nodes), _p(boxes), _p(vp), _p(ri), _p(pi), _p(ni))
    return K, ri[:K], pi[:K], ni[:K]


def get_interpolation_weights(node_indices, size, nodes, boxes, viewpoint):
    idx, nodes, boxes, vp = _i32(node_indices), _i32(nodes), _f32(boxes), _f32(viewpoint)
    K = idx.shape[0]
    ts, kids = np.zeros(K, np.float32), np.zeros(K, np.int32)
    lib().orc_interpolation_weights(C.c_int(K), _p(idx), C.c_float(size), _p(nodes), _p(boxes), _p(vp), _p(ts),
                                    _p(kids))
    return ts, kids


def expand_to_target(nodes, target):
    nodes = _i32(nodes)
    N = nodes.shape[0]
    n = lib().orc_expand_to_target(C.c_int(N), _p(nodes), C.c_int(target), None, C.c_int64(0))
    out = np.zeros(max(n, 1), np.int32)
    lib().orc_expand_to_target(C.c_int(N), _p(nodes), C.c_int(target), _p(out), C.c_int64(n))
    return out[:n]


# ---------------------------------------------------------------------------------------------------
# sparse Adam
# ---------------------------------------------------------------------------------------------------
def sparse_adam(param, grad, exp_avg, exp_avg_sq, relevant, lr, step, betas=(0.9, 0.999), eps=1e-15):
    """In place on float32 C-contiguous arrays [rows, ...]; `step` is the incremented step count."""
    rows = param.shape[0]
    width = int(np.prod(param.shape[1:])) if param.ndim > 1 else 1
    rel = None if relevant is None or len(relevant) == 0 else np.ascontiguousarray(relevant, dtype=np.int64)
    for a in (param, exp_avg, exp_avg_sq):
        assert a.dtype == np.float32 and a.flags.c_contiguous
    lib().orc_sparse_adam(_p(param), _p(_f32(grad)), _p(exp_avg), _p(exp_avg_sq), C.c_int64(rows), C.c_int(width),
                          _p(rel), C.c_int64(0 if rel is None else rel.shape[0]), C.c_double(lr),
                          C.c_double(betas[0]), C.c_double(betas[1]), C.c_double(eps), C.c_double(step))


# ---------------------------------------------------------------------------------------------------
# one-call checker used by __graft_entry__.smoke() and the GPU tests
# ---------------------------------------------------------------------------------------------------
def render_post_reference(cam, attrs, skybox, limit, dL_dcolor=None, dL_dinvdepth=None, bg=None, do_depth=True):
    """Whole hot path on the host for torch inputs (any device): cut at `limit` for camera `cam`, gather +
    lerp, rasterize forward (+ backward when dL_dcolor is given).  Returns numpy arrays:
    dict(count, render_indices, parent_indices, ts, kids, render (clamped), depth, radii, grads{...})."""
    import math

    def n(t):
        return t.detach().cpu().numpy() if hasattr(t, "detach") else np.asarray(t)

    a = {k: n(v) for k, v in attrs.items()}
    vp = n(cam.camera_center)
    W, H = int(cam.image_width), int(cam.image_height)
    K, ri, pi, ni = expand_to_size(a["nodes"], a["boxes"], float(limit), vp)
    ts, kids = get_interpolation_weights(ni, float(limit), a["nodes"], a["boxes"], vp)
    at, w, sign = render_post_gather(a["means3D"], a["shs"], a["opacities"], a["scales"], a["rotations"], ri, pi, ts,
                                     skybox)
    kids_full = np.concatenate([kids, np.ones(skybox, np.int32)])
    bgv = np.zeros(3, np.float32) if bg is None else n(bg)
    o = rasterize_forward(at["means3D"], at["opacities"], at["scales"], at["rotations"],
                          n(cam.world_view_transform).ravel(), n(cam.full_proj_transform).ravel(), vp, bgv, W, H,
                          math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), shs=at["shs"], ts=w, kids=kids_full,
                          do_depth=do_depth)
    out = dict(count=K, render_indices=ri, parent_indices=pi, ts=ts, kids=kids_full,
               render=np.clip(o["color"], 0, 1), depth=o["invdepth"], radii=o["radii"],
               num_rendered=o["num_rendered"])
    if dL_dcolor is not None:
        dcol = n(dL_dcolor) * ((o["color"] >= 0) & (o["color"] <= 1))     # clamp(0,1) backward
        gr = rasterize_backward(o, dcol, None if dL_dinvdepth is None else n(dL_dinvdepth))
        store = render_post_scatter(dict(means3D=gr["dL_dmeans3D"], shs=gr["dL_dsh"], opacities=gr["dL_dopacity"],
                                         scales=gr["dL_dscales"], rotations=gr["dL_drotations"]),
                                    a["means3D"].shape[0], ri, pi, ts, sign, skybox)
        store["means2D"] = gr["dL_dmeans2D"]
        out["grads"] = store
    return out
