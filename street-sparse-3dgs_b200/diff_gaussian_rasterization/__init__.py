"""Drop-in for the reference package ``diff_gaussian_rasterization``
(HR/diff_gaussian_rasterization/__init__.py:17-215): ``GaussianRasterizationSettings``,
``GaussianRasterizer``, ``rasterize_gaussians`` and the ``_C`` submodule, backed by hand-written sm_100a
kernels through the C ABI of include/h3dgs_b200.h.  There is no CPU or PyTorch fallback.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C
from ._C import SparseGradients  # noqa: F401


def _with_count(out):
    """(color, radii, invdepth) as the reference returns them; the number of visible slots, which the forward knows
    anyway, rides on ``radii.num_visible`` so that callers can compact per-slot results without a device->host
    round trip (``radii[radii > 0]`` in the reference's render glue blocks on the GPU)."""
    color, radii, invdepth, num_visible = out
    radii.num_visible = num_visible
    return color, radii, invdepth


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _with_count(_RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                                 cov3Ds_precomp, raster_settings, 0, None))


def rasterize_gaussians_fused(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                              cov3Ds_precomp, raster_settings, skybox_points, sparse=None):
    """Hierarchy-cut render with the child/parent interpolation of ``render_post`` done inside the
    kernels.  Attribute tensors are the whole model store; ``raster_settings.render_indices`` /
    ``parent_indices`` / ``interpolation_weights`` select and blend rows; the last ``skybox_points``
    store rows are appended un-interpolated.  ``means2D`` has one row per rendered slot.
    ``sparse`` (a ``SparseGradients``): the backward leaves the store-row gradients there in compact form instead
    of returning N-row tensors (the parameters' ``.grad`` stay untouched)."""
    return _with_count(_RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                                 cov3Ds_precomp, raster_settings, int(skybox_points), sparse))


def rasterize_gaussians_raw(xyz, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw,
                            raster_settings, skybox_points, activations, sparse=None):
    """``rasterize_gaussians_fused`` on the RAW parameter store of a reference ``GaussianModel``
    (scene/gaussian_model.py:121-145): the kernels apply ``activations`` = (scale, rotation, opacity) codes
    (``h3dgs_native.ACT_*``) and the dc/rest concatenation to the rows they gather instead of torch
    activating all N rows per render; gradients come back for the raw parameters."""
    return _with_count(_RasterizeGaussiansRaw.apply(xyz, means2D, features_dc, features_rest, opacity_raw,
                                                    scaling_raw, rotation_raw, raster_settings, int(skybox_points),
                                                    tuple(activations), sparse))


class _RasterizeGaussiansRaw(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, means2D, f_dc, f_rest, opacity, scaling, rotation, raster_settings, skyboxnum, acts,
                sparse=None):
        rs = raster_settings
        e = torch.Tensor([])
        args = (rs.bg, rs.render_indices, rs.parent_indices, rs.interpolation_weights, rs.num_node_kids,
                xyz, e, opacity, scaling, rotation, rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix,
                rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, f_dc, rs.sh_degree, rs.campos,
                rs.prefiltered, rs.debug, rs.do_depth)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepths, ctx.aux = \
            _C.rasterize_gaussians_ex(*args, skyboxnum, f_rest, acts,
                                      means2D.size(0) if (any(ctx.needs_input_grad) and sparse is None) else -1)
        ctx.sparse = sparse
        if sparse is not None:
            sparse.params = {"means3D": xyz, "sh": f_dc, "sh_rest": f_rest, "opacity": opacity, "scales": scaling,
                             "rot": rotation}
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.skyboxnum = skyboxnum
        ctx.acts = acts
        ctx.means2D_rows = means2D.size(0)
        ctx.save_for_backward(xyz, f_dc, f_rest, opacity, scaling, rotation, radii, geomBuffer, binningBuffer,
                              imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii, invdepths, ctx.aux.V

    @staticmethod
    def backward(ctx, grad_out_color, _, grad_out_depth, _v=None):
        rs = ctx.raster_settings
        xyz, f_dc, f_rest, opacity, scaling, rotation, radii, geomBuffer, binningBuffer, imgBuffer = ctx.saved_tensors
        if grad_out_depth is None:
            grad_out_depth = torch.empty(0, device=grad_out_color.device)
        e = torch.Tensor([])
        args = (rs.bg, rs.render_indices, rs.parent_indices, rs.interpolation_weights, rs.num_node_kids,
                xyz, radii, e, opacity, scaling, rotation, rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix,
                rs.tanfovx, rs.tanfovy, grad_out_color, grad_out_depth, f_dc, rs.sh_degree, rs.campos,
                geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, rs.debug)
        prezero, ctx.aux.prezero = ctx.aux.prezero, None   # a second backward through this node allocates afresh
        (grad_means2D, _gc, grad_opacity, grad_xyz, _gcov, grad_sh, grad_scaling, grad_rotation) = \
            _C.rasterize_gaussians_backward_ex(*args, ctx.skyboxnum, f_rest, ctx.acts, prezero, ctx.sparse, ctx.aux.V)
        grad_dc, grad_rest = grad_sh
        if grad_means2D.size(0) != ctx.means2D_rows:
            full = torch.zeros((ctx.means2D_rows, 3), dtype=grad_means2D.dtype, device=grad_means2D.device)
            n = min(ctx.means2D_rows, grad_means2D.size(0))
            full[:n] = grad_means2D[:n]
            grad_means2D = full
        return (grad_xyz, grad_means2D, grad_dc, grad_rest, grad_opacity, grad_scaling, grad_rotation,
                None, None, None, None)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, skyboxnum, sparse=None):
        rs = raster_settings
        # argument order of the C++ entry point (HR/rasterize_points.h:18-43)
        args = (rs.bg, rs.render_indices, rs.parent_indices, rs.interpolation_weights, rs.num_node_kids,
                means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
                rs.sh_degree, rs.campos, rs.prefiltered, rs.debug, rs.do_depth)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepths, ctx.aux = \
            _C.rasterize_gaussians_ex(*args, skyboxnum, None, None,
                                      means2D.size(0) if (any(ctx.needs_input_grad) and sparse is None) else -1)
        ctx.sparse = sparse
        if sparse is not None:
            sparse.params = {"means3D": means3D, "sh": sh, "colors": colors_precomp, "opacity": opacities,
                             "scales": scales, "rot": rotations, "cov3D": cov3Ds_precomp}
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.skyboxnum = skyboxnum
        ctx.means2D_rows = means2D.size(0)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              opacities, geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii, invdepths, ctx.aux.V

    @staticmethod
    def backward(ctx, grad_out_color, _, grad_out_depth, _v=None):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities, geomBuffer,
         binningBuffer, imgBuffer) = ctx.saved_tensors
        if grad_out_depth is None:
            grad_out_depth = torch.empty(0, device=grad_out_color.device)
        args = (rs.bg, rs.render_indices, rs.parent_indices, rs.interpolation_weights, rs.num_node_kids,
                means3D, radii, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color,
                grad_out_depth, sh, rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered, binningBuffer,
                imgBuffer, rs.debug)
        prezero, ctx.aux.prezero = ctx.aux.prezero, None   # a second backward through this node allocates afresh
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh,
         grad_scales, grad_rotations) = _C.rasterize_gaussians_backward_ex(*args, ctx.skyboxnum, None, None,
                                                                           prezero, ctx.sparse, ctx.aux.V)

        if grad_means2D.size(0) != ctx.means2D_rows:
            # index mode called with a store-sized means2D: slots map to its leading rows
            full = torch.zeros((ctx.means2D_rows, 3), dtype=grad_means2D.dtype, device=grad_means2D.device)
            n = min(ctx.means2D_rows, grad_means2D.size(0))
            full[:n] = grad_means2D[:n]
            grad_means2D = full

        def opt(g, x):
            return g if (g is not None and x.numel() != 0) else None

        return (grad_means3D, grad_means2D, opt(grad_sh, sh), opt(grad_colors_precomp, colors_precomp),
                grad_opacities, opt(grad_scales, scales), opt(grad_rotations, rotations),
                opt(grad_cov3Ds_precomp, cov3Ds_precomp), None, None, None)


class GaussianRasterizationSettings(NamedTuple):
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


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        # boolean mask of points in front of the camera's near plane
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        rs = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, rs)
