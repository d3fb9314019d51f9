"""Drop-in for the reference's ``gaussian_renderer`` (gaussian_renderer/__init__.py:20-417):
``render``, ``render_post`` and ``render_coarse`` with the same signatures and returned dict keys.

What changes underneath: where the reference gathers rows and interpolates child/parent attributes
with a dozen torch kernels (and autograd scatters N-sized zero tensors back, :202-235), the cut
indices and interpolation weights go straight into the rasterizer, whose preprocess kernels do the
gather + lerp (forward) and the t / (1-t) scatter (backward).  Results follow the Python arithmetic
(three separately rounded ops per lerp), so images match the reference path to the last bit of the
projection.
"""
import math

import torch

import os

import h3dgs_native as nat
from diff_gaussian_rasterization import (GaussianRasterizationSettings, GaussianRasterizer,
                                         rasterize_gaussians_fused, rasterize_gaussians_raw)



class _ExposureClamp(torch.autograd.Function):
    """clamp(matmul(image.permute(1, 2, 0), E[:3, :3]).permute(2, 0, 1) + E[:3, 3, None, None], 0, 1) -- the tail of
    the reference's render() / render_post() (gaussian_renderer/__init__.py:115-118, 277-281) -- as one kernel forward
    and one backward (h3dgs_exposure_clamp_*), instead of permute, matmul, permute, add, clamp and their autograd."""

    @staticmethod
    def forward(ctx, image, exposure):
        image = image.contiguous()
        E = None if exposure is None else exposure[:3, :4].contiguous()
        out = torch.empty_like(image)
        nat.check(nat.lib().h3dgs_exposure_clamp_forward(image.size(1), image.size(2), image.data_ptr(), nat.ptr(E),
                                                         out.data_ptr(), nat.stream_handle()), "exposure_clamp")
        ctx.save_for_backward(image, E if E is not None else torch.empty(0, device=image.device))
        ctx.has_e = E is not None
        return out

    @staticmethod
    def backward(ctx, grad_out):
        image, E = ctx.saved_tensors
        E = E if ctx.has_e else None
        grad_out = grad_out.contiguous()
        grad_image = torch.empty_like(image)
        want_e = ctx.has_e and ctx.needs_input_grad[1]
        grad_e = torch.zeros((3, 4), dtype=torch.float32, device=image.device) if want_e else None
        nat.check(nat.lib().h3dgs_exposure_clamp_backward(image.size(1), image.size(2), image.data_ptr(), nat.ptr(E),
                                                          grad_out.data_ptr(), grad_image.data_ptr(), nat.ptr(grad_e),
                                                          nat.stream_handle()), "exposure_clamp backward")
        return grad_image, grad_e


def exposure_clamp(image, exposure=None):
    """Fused exposure affine (``exposure``: [3,4] or None) + clamp to [0, 1] of a rendered [3,H,W] image."""
    if image.dim() != 3 or image.size(0) != 3 or not image.is_cuda or image.dtype != torch.float32:
        raise RuntimeError("exposure_clamp expects a float32 CUDA image [3, H, W]")
    if exposure is not None and (exposure.dim() != 2 or exposure.size(0) < 3 or exposure.size(1) < 4):
        raise RuntimeError("exposure must be a [3, 4] matrix")
    return _ExposureClamp.apply(image, exposure)


def _raw_store(pc, pipe, override_color):
    """The raw parameter tensors + activation codes of a reference-style ``GaussianModel``
    (scene/gaussian_model.py:40-48,121-145,411-412) when the kernels can apply its activations themselves;
    None for any other model object (then its activated ``get_*`` properties are used, as in the reference)."""
    if override_color is not None or pipe.convert_SHs_python or pipe.compute_cov3D_python:
        return None
    if os.environ.get("H3DGS_RAW_STORE", "1") == "0":
        return None
    try:
        t = (pc._xyz, pc._features_dc, pc._features_rest, pc._opacity, pc._scaling, pc._rotation)
        sa, ra, oa = pc.scaling_activation, pc.rotation_activation, pc.opacity_activation
    except AttributeError:
        return None
    if sa is not torch.exp or ra is not torch.nn.functional.normalize:
        return None
    if oa is torch.sigmoid:
        oc = nat.ACT_SIGMOID
    elif oa is torch.abs:
        oc = nat.ACT_ABS
    else:
        return None
    if any((not x.is_cuda) or x.dtype != torch.float32 or not x.is_contiguous() for x in t):
        return None
    if t[1].dim() != 3 or t[1].size(1) != 1 or t[2].dim() != 3 or t[2].size(1) < 1:
        return None
    return t, (nat.ACT_EXP, nat.ACT_NORMALIZE, oc)


def _eval_sh(deg, sh, dirs):
    """utils/sh_utils.py:57-112 (used only for pipe.convert_SHs_python)."""
    C0 = 0.28209479177387814
    C1 = 0.4886025119029199
    C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
    C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
          -0.4570457994644658, 1.445305721320277, -0.5900435899266435]
    result = C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[..., 0:1], dirs[..., 1:2], dirs[..., 2:3]
        result = result - C1 * y * sh[..., 1] + C1 * z * sh[..., 2] - C1 * x * sh[..., 3]
        if deg > 1:
            xx, yy, zz = x * x, y * y, z * z
            xy, yz, xz = x * y, y * z, x * z
            result = (result + C2[0] * xy * sh[..., 4] + C2[1] * yz * sh[..., 5] +
                      C2[2] * (2.0 * zz - xx - yy) * sh[..., 6] + C2[3] * xz * sh[..., 7] +
                      C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                result = (result + C3[0] * y * (3 * xx - yy) * sh[..., 9] + C3[1] * xy * z * sh[..., 10] +
                          C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] +
                          C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12] +
                          C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + C3[5] * z * (xx - yy) * sh[..., 14] +
                          C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return result


def _screenspace_points(pc):
    # zero tensor whose gradient is the 2D mean gradient (densification statistics).  The reference builds
    # it as ``zeros(requires_grad=True) + 0`` with retain_grad(); a leaf gives callers the same ``.grad``
    # without the N-row add (and without autograd cloning the gradient into the retained copy).
    xyz = pc.get_xyz
    return torch.zeros(xyz.shape, dtype=xyz.dtype, requires_grad=True, device="cuda")


def _appearance(viewpoint_camera, pc, pipe, override_color):
    shs, colors_precomp = None, None
    if override_color is None:
        if pipe.convert_SHs_python:
            shs_view = pc.get_features.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
            dir_pp = (pc.get_xyz - viewpoint_camera.camera_center.repeat(pc.get_features.shape[0], 1))
            dir_pp_normalized = dir_pp / dir_pp.norm(dim=1, keepdim=True)
            sh2rgb = _eval_sh(pc.active_sh_degree, shs_view, dir_pp_normalized)
            colors_precomp = torch.clamp_min(sh2rgb + 0.5, 0.0)
        else:
            shs = pc.get_features
    else:
        colors_precomp = override_color
    return shs, colors_precomp


def _geometry(pc, pipe, scaling_modifier):
    if pipe.compute_cov3D_python:
        return None, None, pc.get_covariance(scaling_modifier)
    return pc.get_scaling, pc.get_rotation, None


def _settings(viewpoint_camera, pc, bg_color, scaling_modifier, debug, do_depth, render_indices,
              parent_indices, interpolation_weights, num_node_kids):
    return GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height),
        image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5),
        tanfovy=math.tan(viewpoint_camera.FoVy * 0.5),
        bg=bg_color,
        scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center,
        prefiltered=False,
        debug=debug,
        render_indices=render_indices,
        parent_indices=parent_indices,
        interpolation_weights=interpolation_weights,
        num_node_kids=num_node_kids,
        do_depth=do_depth)


def _visible(radii):
    """(mask, indices) of the slots with radii > 0.  The forward already knows how many there are
    (``radii.num_visible``, set by the rasterizer wrappers), so the indices come from ``nonzero_static`` -- no
    device->host sync, the host keeps queueing the loss and the backward while the blend kernel runs (the
    reference's ``radii[radii > 0]`` blocks on the GPU)."""
    mask = radii > 0
    n = getattr(radii, "num_visible", None)
    if n is not None:
        return mask, torch.nonzero_static(mask, size=int(n)).flatten()
    return mask, mask.nonzero().flatten()


def _empty(dtype):
    return torch.empty(0, dtype=dtype, device="cuda")


def _rasterize(settings, means3D, means2D, opacity, shs, colors_precomp, scales, rotations, cov3D_precomp,
               skybox_points=None, sparse=None):
    e = torch.Tensor([])
    if skybox_points is None:
        return GaussianRasterizer(raster_settings=settings)(
            means3D=means3D, means2D=means2D, shs=shs, colors_precomp=colors_precomp, opacities=opacity,
            scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp)
    return rasterize_gaussians_fused(
        means3D, means2D, e if shs is None else shs, e if colors_precomp is None else colors_precomp,
        opacity, e if scales is None else scales, e if rotations is None else rotations,
        e if cov3D_precomp is None else cov3D_precomp, settings, skybox_points, sparse)


def _as_index(t):
    return t.to(device="cuda", dtype=torch.int32).contiguous()


def render(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, override_color=None,
           indices=None, use_trained_exp=False):
    """Render the scene (gaussian_renderer/__init__.py:20-135).  Background tensor must be on GPU.

    On a reference-style ``GaussianModel`` (train_single.py:111) the RAW parameters go to the kernels, which apply
    exp / normalize / sigmoid and the dc/rest split to the rows they read: the reference activates all N rows
    with four torch kernels and a 192 B/row ``cat`` per call, and autograd runs their backward over N rows."""
    screenspace_points = _screenspace_points(pc)
    none_i, none_f = _empty(torch.int32), _empty(torch.float32)
    raw = _raw_store(pc, pipe, override_color) if indices is None else None
    if raw is not None:
        settings = _settings(viewpoint_camera, pc, bg_color, scaling_modifier, pipe.debug, True, none_i, none_i,
                             none_f, none_i)
        (xyz, f_dc, f_rest, op_raw, sc_raw, rot_raw), acts = raw
        rendered_image, radii, depth_image = rasterize_gaussians_raw(xyz, screenspace_points, f_dc, f_rest, op_raw,
                                                                     sc_raw, rot_raw, settings, 0, acts)
    else:
        means3D, means2D, opacity = pc.get_xyz, screenspace_points, pc.get_opacity
        scales, rotations, cov3D_precomp = _geometry(pc, pipe, scaling_modifier)
        shs, colors_precomp = _appearance(viewpoint_camera, pc, pipe, override_color)
        if indices is not None:
            settings = _settings(viewpoint_camera, pc, bg_color, scaling_modifier, pipe.debug, True,
                                 _as_index(indices), none_i, none_f, none_i)
            rendered_image, radii, depth_image = _rasterize(settings, means3D, means2D[indices].contiguous(), opacity,
                                                            shs, colors_precomp, scales, rotations, cov3D_precomp, 0)
        else:
            settings = _settings(viewpoint_camera, pc, bg_color, scaling_modifier, pipe.debug, True, none_i, none_i,
                                 none_f, none_i)
            rendered_image, radii, depth_image = _rasterize(settings, means3D, means2D, opacity, shs, colors_precomp,
                                                            scales, rotations, cov3D_precomp)

    # exposure affine + clamp (gaussian_renderer/__init__.py:115-118) in one kernel forward and one backward
    exposure = pc.get_exposure_from_name(viewpoint_camera.image_name) if use_trained_exp else None
    rendered_image = exposure_clamp(rendered_image, exposure)

    subfilter, visible = _visible(radii)
    if indices is not None:
        vis_filter = torch.zeros(pc._xyz.size(0), dtype=bool, device="cuda")
        w = vis_filter[indices]
        w[subfilter] = True
        vis_filter[indices] = w
        vis_ids = vis_filter.nonzero().flatten().long()
    else:
        vis_ids = visible.long()
    return {"render": rendered_image,
            "depth": depth_image,
            "viewspace_points": screenspace_points,
            "visibility_filter": vis_ids,
            "radii": radii[visible]}


def render_post(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, override_color=None,
                render_indices=torch.Tensor([]).int(), parent_indices=torch.Tensor([]).int(),
                interpolation_weights=torch.Tensor([]).float(), num_node_kids=torch.Tensor([]).int(),
                interp_python=True, use_trained_exp=False, do_depth=False, sparse_grads=None):
    """Render the scene from a hierarchy cut (gaussian_renderer/__init__.py:138-304).

    ``sparse_grads`` (extension; a ``diff_gaussian_rasterization.SparseGradients``): the backward leaves the
    gradients of the store parameters there in compact form -- one row per store row the render touched -- instead of
    materialising N-row ``.grad`` tensors (4 GB of zero fill per step on a 17 M-row store, plus the optimizer's scan
    for the touched rows).  ``scene.OurAdam.Adam.step_sparse`` consumes it.  ``viewspace_points`` then has one row
    per rendered slot instead of one per store row."""
    if sparse_grads is not None and not (render_indices.size(0) != 0 and interp_python):
        raise RuntimeError("sparse_grads needs a hierarchy cut rendered with interp_python=True")
    if sparse_grads is not None:
        n_slots = render_indices.size(0) + int(pc.skybox_points)
        screenspace_points = torch.zeros((n_slots, 3), dtype=pc.get_xyz.dtype, requires_grad=True, device="cuda")
    else:
        screenspace_points = _screenspace_points(pc)
    means2D = screenspace_points
    # Hierarchy cut with Python-style interpolation on a reference GaussianModel: hand the RAW parameters to
    # the kernels, which activate only the rows of the cut (the reference activates all N rows per render).
    raw = _raw_store(pc, pipe, override_color) if (render_indices.size(0) != 0 and interp_python) else None
    if raw is None:
        means3D, opacity = pc.get_xyz, pc.get_opacity
        scales, rotations, cov3D_precomp = _geometry(pc, pipe, scaling_modifier)
        shs, colors_precomp = _appearance(viewpoint_camera, pc, pipe, override_color)

    skybox = None
    if render_indices.size(0) != 0:
        num_entries = render_indices.size(0)
        idx = _as_index(render_indices)
        if interp_python:
            skybox = int(pc.skybox_points)
            par = _as_index(parent_indices[:num_entries])
            # the reference marks the appended skybox slots t = 1 / kids = 1, mutating the caller's kids
            if skybox:
                num_node_kids[num_entries:num_entries + skybox] = 1
        else:
            skybox = 0
            par = _empty(torch.int32)
        # interp_python: slots map to the leading rows of the store-sized tensor (the reference slices it,
        # :239); the rasterizer writes those rows of a full-size gradient itself
        if not interp_python:
            means2D = means2D[render_indices.long()].contiguous()
        render_indices, parent_indices = idx, par
        if interpolation_weights.numel():
            interpolation_weights = interpolation_weights.to("cuda")
        if num_node_kids.numel():
            num_node_kids = num_node_kids.to("cuda")

    settings = _settings(viewpoint_camera, pc, bg_color, scaling_modifier, pipe.debug, do_depth, render_indices,
                         parent_indices, interpolation_weights, num_node_kids)
    if raw is not None:
        (xyz, f_dc, f_rest, op_raw, sc_raw, rot_raw), acts = raw
        rendered_image, radii, depth_image = rasterize_gaussians_raw(xyz, means2D, f_dc, f_rest, op_raw, sc_raw,
                                                                     rot_raw, settings, skybox, acts, sparse_grads)
    else:
        rendered_image, radii, depth_image = _rasterize(settings, means3D, means2D, opacity, shs, colors_precomp,
                                                        scales, rotations, cov3D_precomp, skybox, sparse_grads)

    exposure = None
    if use_trained_exp and pc.pretrained_exposures:
        try:
            exposure = pc.pretrained_exposures[viewpoint_camera.image_name]
        except Exception:
            print(f"Exposures should be optimized in single. Missing exposure for image {viewpoint_camera.image_name}")
    rendered_image = exposure_clamp(rendered_image, exposure)   # :277-281, one kernel each way

    vis_filter, visible = _visible(radii)
    out = {"render": rendered_image,
           "viewspace_points": screenspace_points,
           "visibility_filter": vis_filter,
           "radii": radii[visible]}
    if do_depth:
        out["depth"] = depth_image
    return out


def render_coarse(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, zfar=0.0,
                  override_color=None, indices=None):
    """Render for the coarse optimisation (gaussian_renderer/__init__.py:306-417): debug on, no depth."""
    screenspace_points = _screenspace_points(pc)
    means3D, means2D, opacity = pc.get_xyz, screenspace_points, pc.get_opacity
    scales, rotations, cov3D_precomp = _geometry(pc, pipe, scaling_modifier)
    shs, colors_precomp = _appearance(viewpoint_camera, pc, pipe, override_color)

    if indices is not None:
        idx = _as_index(indices)
        settings = _settings(viewpoint_camera, pc, bg_color, scaling_modifier, True, False, idx,
                             _empty(torch.int32), _empty(torch.float32), _empty(torch.int32))
        rendered_image, radii, _ = _rasterize(settings, means3D, means2D[indices].contiguous(), opacity, shs,
                                              colors_precomp, scales, rotations, cov3D_precomp, 0)
    else:
        settings = _settings(viewpoint_camera, pc, bg_color, scaling_modifier, True, False, _empty(torch.int32),
                             _empty(torch.int32), _empty(torch.float32), _empty(torch.int32))
        rendered_image, radii, _ = _rasterize(settings, means3D, means2D, opacity, shs, colors_precomp, scales,
                                              rotations, cov3D_precomp)

    subfilter, visible = _visible(radii)
    if indices is not None:
        vis_filter = torch.zeros(pc._xyz.size(0), dtype=bool, device="cuda")
        w = vis_filter[indices]
        w[subfilter] = True
        vis_filter[indices] = w
    else:
        vis_filter = subfilter
    return {"render": rendered_image,
            "viewspace_points": screenspace_points,
            "visibility_filter": vis_filter,
            "radii": radii[visible]}
