"""Densification for single-chunk training on the fused kernels of csrc/densify.cu: drop-in
replacements of three pieces of the reference's training loop that operate on a reference-style ``GaussianModel``
(attributes ``_xyz _features_dc _features_rest _opacity _scaling _rotation``, ``optimizer`` with one parameter per
named group, ``xyz_gradient_accum``, ``denom``, ``max_radii2D``, ``percent_dense``, ``scaffold_points``):

    update_max_radii_and_stats(model, viewspace_point_tensor, visibility_filter, radii)
        = train_single.py:193-194: ``max_radii2D[vis] = max(max_radii2D[vis], radii)`` +
          ``GaussianModel.add_densification_stats`` (scene/gaussian_model.py:780-794), one kernel
    densify_and_prune(model, max_grad, min_opacity, extent)
        = ``GaussianModel.densify_and_prune`` (scene/gaussian_model.py:736-778): clone, split, prune -- one selection
          kernel, three scans and ONE rebuild pass over the six parameter tensors and their Adam moments instead of
          three cat / mask rebuilds of all eighteen tensors.  The split samples are drawn by the same
          ``torch.normal(mean=zeros, std=stds)`` call as the reference's, so a seeded run consumes the random stream
          identically.  The optional ground-truth point-cloud constraint (:757-758, FAISS) is not implemented.

There is no CPU fallback: the functions raise if the model does not live on a CUDA device."""
import ctypes

import torch
from torch import nn

import h3dgs_native as nat

_GROUPS = (("xyz", "_xyz"), ("f_dc", "_features_dc"), ("f_rest", "_features_rest"), ("opacity", "_opacity"),
           ("scaling", "_scaling"), ("rotation", "_rotation"))


def _activation_code(model):
    act = getattr(model, "opacity_activation", torch.sigmoid)
    if act is torch.sigmoid:
        return 3   # H3DGS_ACT_SIGMOID
    if act is torch.abs:
        return 4   # H3DGS_ACT_ABS
    raise NotImplementedError("opacity activation must be torch.sigmoid or torch.abs")


def update_max_radii_and_stats(model, viewspace_point_tensor, visibility_filter, radii=None, is_depth_only=False):
    """train_single.py:193-194.  ``visibility_filter``: int64 row indices (what render() returns); ``radii``: the
    radii of exactly those rows (render()'s ``radii``), or None to update the gradient statistics only."""
    grad = viewspace_point_tensor.grad
    if grad is None or not grad.is_cuda:
        raise RuntimeError("viewspace_point_tensor.grad must be a CUDA tensor (no CPU fallback)")
    vis = visibility_filter
    if vis.dtype == torch.bool:
        vis = vis.nonzero().flatten()
    vis = vis.long().contiguous()
    r = None if radii is None else radii.to(torch.int32).contiguous()
    if r is not None and r.numel() != vis.numel():
        raise RuntimeError("radii must hold one entry per visible row")
    for t in (model.xyz_gradient_accum, model.denom, model.max_radii2D):
        if not t.is_contiguous() or t.dtype != torch.float32:
            raise RuntimeError("densification statistics must be contiguous float32 tensors")
    nat.check(nat.lib().h3dgs_densify_stats(vis.numel(), nat.ptr(vis), nat.ptr(r), grad.contiguous().data_ptr(),
                                            model.xyz_gradient_accum.data_ptr(), model.denom.data_ptr(),
                                            model.max_radii2D.data_ptr() if r is not None else None,
                                            nat.stream_handle()), "h3dgs_densify_stats")


def densify_and_prune(model, max_grad, min_opacity, extent, gt_point_cloud_constraints=False, samples=None):
    """scene/gaussian_model.py:736-778.  ``samples`` (tests): the [2 S, 3] normal draws to use instead of drawing."""
    if gt_point_cloud_constraints:
        raise NotImplementedError("the ground-truth point-cloud constraint (FAISS) is not part of this path")
    if min_opacity > 0.15:
        raise NotImplementedError("min_opacity above the selection's own opacity bound (0.15) is not supported")
    L = nat.lib()
    params = [getattr(model, attr) for _, attr in _GROUPS]
    n = params[0].size(0)
    dev = params[0].device
    if not params[0].is_cuda:
        raise RuntimeError("densify_and_prune needs a CUDA model (no CPU fallback)")
    data = [p.detach().contiguous() for p in params]
    widths = [d.numel() // max(n, 1) for d in data]
    ws = torch.empty(L.h3dgs_densify_workspace_bytes(n), dtype=torch.uint8, device=dev)
    counts = (ctypes.c_int32 * 3)()
    scaffold = model.scaffold_points if getattr(model, "scaffold_points", None) is not None else 0
    accum = model.xyz_gradient_accum.contiguous()
    nat.check(L.h3dgs_densify_select(n, accum.data_ptr(), model.max_radii2D.contiguous().data_ptr(), data[3].data_ptr(),
                                     _activation_code(model), data[4].data_ptr(), float(max_grad), float(min_opacity),
                                     float(model.percent_dense * extent), int(scaffold), ws.data_ptr(), ws.numel(),
                                     counts, nat.stream_handle()), "h3dgs_densify_select")
    K, C, S = int(counts[0]), int(counts[1]), int(counts[2])
    if samples is None:
        stds = torch.empty((2 * S, 3), dtype=torch.float32, device=dev)
        nat.check(L.h3dgs_densify_split_std(n, data[4].data_ptr(), ws.data_ptr(), S, nat.ptr(stds), nat.stream_handle()),
                  "h3dgs_densify_split_std")
        samples = torch.normal(mean=torch.zeros((2 * S, 3), device=dev), std=stds)   # gaussian_model.py:687-689
    samples = samples.to(dev, torch.float32).contiguous()
    if samples.shape != (2 * S, 3):
        raise RuntimeError(f"samples must be [{2 * S}, 3]")
    M = K + C + 2 * S
    opt = model.optimizer
    group_of = {g["name"]: g for g in opt.param_groups}
    arr = (nat.DensifyTensor * len(_GROUPS))()
    new_params, new_state = [], []
    for i, (name, attr) in enumerate(_GROUPS):
        old = params[i]
        st = opt.state.get(group_of[name]["params"][0], None) if name in group_of else None
        npar = torch.empty((M,) + tuple(old.shape[1:]), dtype=torch.float32, device=dev)
        arr[i].param, arr[i].new_param, arr[i].width = data[i].data_ptr(), npar.data_ptr(), widths[i]
        if st is not None and "exp_avg" in st:
            m_new, v_new = torch.empty_like(npar), torch.empty_like(npar)
            m_old, v_old = st["exp_avg"].contiguous(), st["exp_avg_sq"].contiguous()
            arr[i].exp_avg, arr[i].exp_avg_sq = m_old.data_ptr(), v_old.data_ptr()
            arr[i].new_exp_avg, arr[i].new_exp_avg_sq = m_new.data_ptr(), v_new.data_ptr()
            new_state.append((st, m_new, v_new, m_old, v_old))
        else:
            new_state.append(None)
        new_params.append(npar)
    nat.check(L.h3dgs_densify_build(n, arr, len(_GROUPS), 0, 4, 5, ws.data_ptr(), K, C, S, nat.ptr(samples),
                                    nat.stream_handle()), "h3dgs_densify_build")
    # hand the new tensors to the optimizer exactly as cat_tensors_to_optimizer / _prune_optimizer do (:560-650)
    for i, (name, attr) in enumerate(_GROUPS):
        par = nn.Parameter(new_params[i].requires_grad_(True))
        if name in group_of:
            g = group_of[name]
            old = g["params"][0]
            if new_state[i] is not None:
                st, m_new, v_new, _, _ = new_state[i]
                st["exp_avg"], st["exp_avg_sq"] = m_new, v_new
                del opt.state[old]
                opt.state[par] = st
            elif old in opt.state:
                del opt.state[old]
            g["params"][0] = par
        setattr(model, attr, par)
    model.xyz_gradient_accum = torch.zeros((M, 1), device=dev)
    if hasattr(model, "xyz_gradient_accum_depth"):
        model.xyz_gradient_accum_depth = torch.zeros((M, 1), device=dev)
    model.denom = torch.zeros((M, 1), device=dev)
    model.max_radii2D = torch.zeros((M,), device=dev)
    # the reference's newly_split_points_mask (:707-710) only feeds the ground-truth constraint, which is not
    # implemented here; the split children are the last 2 S rows
    model.newly_split_points_mask = torch.zeros(M, dtype=torch.bool, device=dev)
    if S:
        model.newly_split_points_mask[-2 * S:] = True
    return {"kept": K, "clones": C, "splits": S, "rows": M, "samples": samples}
