"""Access to the reference builds under oracle/_ref (TEST INFRASTRUCTURE, never on the product path).

* ``ref_hr()`` / ``ref_gh()`` import the pybind modules compiled from the unmodified reference sources
  by oracle/build_ref.py.
* ``RefRasterizer`` restates the thin Python wrapper of HR/diff_gaussian_rasterization/__init__.py:39-144
  around the reference ``_C`` functions; ``ref_render_post`` restates the gather + lerp glue of
  gaussian_renderer/__init__.py:138-304 (torch ops, autograd) on top of it.  Reference Python files cannot
  travel to the GPU box, the compiled extension can.
* ``parse_ref_buffers`` decodes the reference's opaque byte buffers with the layouts of
  HR/cuda_rasterizer/rasterizer_impl.cu:159-199.
"""
import importlib.machinery
import importlib.util
import math
import os

from typing import NamedTuple

import torch

from . import ROOT


class RefSettings(NamedTuple):
    """Field-for-field GaussianRasterizationSettings (HR/diff_gaussian_rasterization/__init__.py:146-163)."""
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    render_indices: torch.Tensor
    parent_indices: torch.Tensor
    interpolation_weights: torch.Tensor
    num_node_kids: torch.Tensor
    do_depth: bool

REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def _load(name):
    path = os.path.join(REF_DIR, name + ".so")
    if not os.path.exists(path):
        return None
    loader = importlib.machinery.ExtensionFileLoader(name, path)
    spec = importlib.util.spec_from_loader(name, loader)
    mod = importlib.util.module_from_spec(spec)
    loader.exec_module(mod)
    return mod


_cache = {}


def ref_hr():
    if "hr" not in _cache:
        _cache["hr"] = _load("ref_hr_C")
    return _cache["hr"]


def ref_gh():
    if "gh" not in _cache:
        _cache["gh"] = _load("ref_gh_C")
    return _cache["gh"]


class _RefRasterize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, rs):
        C = ref_hr()
        args = (rs.bg, rs.render_indices, rs.parent_indices, rs.interpolation_weights, rs.num_node_kids,
                means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
                rs.sh_degree, rs.campos, rs.prefiltered, rs.debug, rs.do_depth)
        num_rendered, color, radii, geom, binning, img, invdepths = C.rasterize_gaussians(*args)
        ctx.rs = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities,
                              geom, binning, img)
        ctx.mark_non_differentiable(radii)
        _RefRasterize.last = dict(num_rendered=num_rendered, geom=geom, binning=binning, img=img)
        return color, radii, invdepths

    @staticmethod
    def backward(ctx, grad_out_color, _, grad_out_depth):
        C = ref_hr()
        rs = ctx.rs
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities, geom, binning, img = \
            ctx.saved_tensors
        if grad_out_depth is None:
            grad_out_depth = torch.empty(0, device=grad_out_color.device)
        args = (rs.bg, rs.render_indices, rs.parent_indices, rs.interpolation_weights, rs.num_node_kids,
                means3D, radii, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color.contiguous(),
                grad_out_depth.contiguous(), sh, rs.sh_degree, rs.campos, geom, ctx.num_rendered, binning, img,
                rs.debug)
        g2d, gcol, gop, g3d, gcov, gsh, gsc, grot = C.rasterize_gaussians_backward(*args)
        return g3d, g2d, gsh, gcol, gop, gsc, grot, gcov, None


def ref_rasterize(rs, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                  cov3D_precomp=None):
    e = torch.Tensor([])
    return _RefRasterize.apply(means3D, means2D, e if shs is None else shs,
                               e if colors_precomp is None else colors_precomp, opacities,
                               e if scales is None else scales, e if rotations is None else rotations,
                               e if cov3D_precomp is None else cov3D_precomp, rs)


def make_settings(Settings, cam, bg, sh_degree=3, debug=False, do_depth=True, scale_modifier=1.0,
                  render_indices=None, parent_indices=None, interpolation_weights=None, num_node_kids=None):
    dev = bg.device
    ei = torch.empty(0, dtype=torch.int32, device=dev)
    ef = torch.empty(0, dtype=torch.float32, device=dev)
    return Settings(
        image_height=int(cam.image_height), image_width=int(cam.image_width),
        tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), bg=bg,
        scale_modifier=scale_modifier, viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
        sh_degree=sh_degree, campos=cam.camera_center, prefiltered=False, debug=debug,
        render_indices=ei if render_indices is None else render_indices,
        parent_indices=ei if parent_indices is None else parent_indices,
        interpolation_weights=ef if interpolation_weights is None else interpolation_weights,
        num_node_kids=ei if num_node_kids is None else num_node_kids, do_depth=do_depth)


def ref_render_post(Settings, cam, pc, bg, render_indices, parent_indices, interpolation_weights, num_node_kids,
                    do_depth=True, debug=False):
    """gaussian_renderer/__init__.py:138-304 with interp_python=True, restated on the reference rasterizer."""
    screenspace_points = torch.zeros_like(pc.get_xyz, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    means3D, means2D, opacity = pc.get_xyz, screenspace_points, pc.get_opacity
    scales, rotations, shs = pc.get_scaling, pc.get_rotation, pc.get_features
    render_inds = render_indices.long()
    num_entries = render_indices.size(0)
    interps = interpolation_weights[:num_entries].unsqueeze(1)
    interps_inv = (1 - interpolation_weights[:num_entries]).unsqueeze(1)
    parent_inds = parent_indices[:num_entries].long()
    means3D_base = (interps * means3D[render_inds] + interps_inv * means3D[parent_inds]).contiguous()
    scales_base = (interps * scales[render_inds] + interps_inv * scales[parent_inds]).contiguous()
    shs_base = (interps.unsqueeze(2) * shs[render_inds] + interps_inv.unsqueeze(2) * shs[parent_inds]).contiguous()
    parents = rotations[parent_inds]
    rots = rotations[render_inds]
    dots = torch.bmm(rots.unsqueeze(1), parents.unsqueeze(2)).flatten()
    parents[dots < 0] *= -1
    rotations_base = ((interps * rots) + interps_inv * parents).contiguous()
    opacity_base = (interps * opacity[render_inds] + interps_inv * opacity[parent_inds]).contiguous()
    n = pc._xyz.size(0)
    skybox_inds = torch.arange(n - pc.skybox_points, n, device=means3D.device).long()
    means3D = torch.cat((means3D_base, means3D[skybox_inds])).contiguous()
    shs = torch.cat((shs_base, shs[skybox_inds])).contiguous()
    opacity = torch.cat((opacity_base, opacity[skybox_inds])).contiguous()
    rotations = torch.cat((rotations_base, rotations[skybox_inds])).contiguous()
    means2D = means2D[:(num_entries + pc.skybox_points)].contiguous()
    scales = torch.cat((scales_base, scales[skybox_inds])).contiguous()
    interpolation_weights = interpolation_weights.clone().detach()
    interpolation_weights[num_entries:num_entries + pc.skybox_points] = 1.0
    num_node_kids[num_entries:num_entries + pc.skybox_points] = 1
    rs = make_settings(Settings, cam, bg, sh_degree=pc.active_sh_degree, debug=debug, do_depth=do_depth,
                       render_indices=torch.Tensor([]).int(), parent_indices=torch.Tensor([]).int(),
                       interpolation_weights=interpolation_weights, num_node_kids=num_node_kids)
    image, radii, depth = ref_rasterize(rs, means3D, means2D, opacity, shs=shs, scales=scales, rotations=rotations)
    image = image.clamp(0, 1)
    vis = radii > 0
    return {"render": image, "viewspace_points": screenspace_points, "visibility_filter": vis,
            "radii": radii[vis], "depth": depth}


def ref_render(Settings, cam, pc, bg, use_trained_exp=False, debug=False):
    """gaussian_renderer/__init__.py:20-135 (plain render, indices=None, SH evaluated by the rasterizer) restated
    on the reference rasterizer: full-store activations through the model's get_* properties, do_depth=True,
    learnable exposure, clamp; visibility_filter as global ids."""
    screenspace_points = torch.zeros_like(pc.get_xyz, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    dev = bg.device
    rs = make_settings(Settings, cam, bg, sh_degree=pc.active_sh_degree, debug=debug, do_depth=True,
                       render_indices=torch.empty(0).int().to(dev), parent_indices=torch.empty(0).int().to(dev),
                       interpolation_weights=torch.empty(0).float().to(dev),
                       num_node_kids=torch.empty(0).int().to(dev))
    image, radii, depth = ref_rasterize(rs, pc.get_xyz, screenspace_points, pc.get_opacity, shs=pc.get_features,
                                        scales=pc.get_scaling, rotations=pc.get_rotation)
    if use_trained_exp:
        exposure = pc.get_exposure_from_name(cam.image_name)
        image = torch.matmul(image.permute(1, 2, 0), exposure[:3, :3]).permute(2, 0, 1) + exposure[:3, 3, None, None]
    image = image.clamp(0, 1)
    vis = radii > 0
    return {"render": image, "depth": depth, "viewspace_points": screenspace_points,
            "visibility_filter": vis.nonzero().flatten().long(), "radii": radii[vis]}


def _align(x, a=128):
    return (x + a - 1) // a * a


def parse_ref_buffers(geom, binning, img, P, R, W, H):
    """Typed views into the reference's buffers (rasterizer_impl.cu:159-199).  Offsets are relative to
    128-byte-aligned bases (torch allocations are)."""
    out = {}
    o = 0

    def take(buf, nbytes, dtype):
        nonlocal o
        o = _align(o)
        v = buf[o:o + nbytes].view(dtype)
        o += nbytes
        return v

    if geom is not None and geom.numel():
        assert geom.data_ptr() % 128 == 0
        o = 0
        out["depths"] = take(geom, 4 * P, torch.float32)
        out["clamped"] = take(geom, 3 * P, torch.uint8)
        take(geom, 3 * P, torch.uint8)
        take(geom, 4 * P, torch.int32)
        out["means2D"] = take(geom, 8 * P, torch.float32).view(P, 2)
        out["cov3D"] = take(geom, 24 * P, torch.float32).view(P, 6)
        out["conic_opacity"] = take(geom, 16 * P, torch.float32).view(P, 4)
        out["rgb"] = take(geom, 12 * P, torch.float32).view(P, 3)
        out["tiles_touched"] = take(geom, 4 * P, torch.int32)
    if binning is not None and binning.numel() and R > 0:
        assert binning.data_ptr() % 128 == 0
        o = 0
        out["point_list"] = take(binning, 4 * R, torch.int32)
        take(binning, 4 * R, torch.int32)
        out["point_list_keys"] = take(binning, 8 * R, torch.int64)
    if img is not None and img.numel():
        assert img.data_ptr() % 128 == 0
        o = 0
        N = W * H
        out["final_T"] = take(img, 4 * N, torch.float32)
        out["n_contrib"] = take(img, 4 * N, torch.int32)
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        out["ranges"] = take(img, 8 * N, torch.int32)[: 2 * tiles]
    return out
