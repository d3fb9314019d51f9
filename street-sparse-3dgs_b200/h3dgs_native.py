"""ctypes binding of libh3dgs_b200.so (include/h3dgs_b200.h).

This is the only place that knows both torch and the C ABI: tensors go in as raw device pointers
plus the current CUDA stream.  There is no fallback: if the library is missing or was not built for
this machine, importing the op packages raises.
"""
import ctypes
import os
from ctypes import (POINTER, Structure, byref, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t,
                    c_void_p)

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libh3dgs_b200.so")


class H3dgsError(RuntimeError):
    pass


class View(Structure):
    _fields_ = [("image_width", c_int32), ("image_height", c_int32),
                ("tan_fovx", c_float), ("tan_fovy", c_float),
                ("viewmatrix", c_void_p), ("projmatrix", c_void_p),
                ("campos", c_void_p), ("background", c_void_p)]


class Gaussians(Structure):
    _fields_ = [("P", c_int32), ("store_rows", c_int32), ("sh_degree", c_int32), ("sh_coeffs", c_int32),
                ("skyboxnum", c_int32), ("scale_modifier", c_float),
                ("means3D", c_void_p), ("shs", c_void_p), ("colors_precomp", c_void_p),
                ("opacities", c_void_p), ("scales", c_void_p), ("rotations", c_void_p),
                ("cov3D_precomp", c_void_p),
                ("render_indices", c_void_p), ("parent_indices", c_void_p),
                ("ts", c_void_p), ("kids", c_void_p),
                ("shs_rest", c_void_p), ("scale_activation", c_int32), ("rotation_activation", c_int32),
                ("opacity_activation", c_int32)]


ACT_NONE, ACT_EXP, ACT_NORMALIZE, ACT_SIGMOID, ACT_ABS = 0, 1, 2, 3, 4


class Grads(Structure):
    _fields_ = [("dL_dmeans2D", c_void_p), ("dL_dmeans3D", c_void_p), ("dL_dsh", c_void_p),
                ("dL_dcolors", c_void_p), ("dL_dopacity", c_void_p), ("dL_dscales", c_void_p),
                ("dL_drotations", c_void_p), ("dL_dcov3D", c_void_p), ("dL_dsh_rest", c_void_p),
                ("dirty_rows", c_void_p), ("rows_prezeroed", c_int32),
                ("compact_rows", c_void_p), ("compact_count", c_void_p), ("row_map", c_void_p),
                ("compact_capacity", c_int32)]


class RowTensor(Structure):
    _fields_ = [("data", c_void_p), ("width", c_int32)]


class Span(Structure):
    _fields_ = [("data", c_void_p), ("count", c_int64)]


class DebugView(Structure):
    _fields_ = [("point_list_keys", c_void_p), ("point_list", c_void_p), ("ranges", c_void_p),
                ("final_T", c_void_p), ("n_contrib", c_void_p), ("depths", c_void_p),
                ("tiles_touched", c_void_p), ("point_offsets", c_void_p), ("records", c_void_p),
                ("cov3D", c_void_p), ("tile_rects", c_void_p), ("order_fwd", c_void_p), ("order_bwd", c_void_p),
                ("tile_last", c_void_p)]


class AdamTensor(Structure):
    _fields_ = [("param", c_void_p), ("grad", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p),
                ("width", c_int32), ("step_size", c_float), ("bias_correction2_sqrt", c_float),
                ("beta1", c_float), ("beta2", c_float), ("eps", c_float),
                ("one_minus_beta1", c_float), ("one_minus_beta2", c_float)]


class DensifyTensor(Structure):
    _fields_ = [("param", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p), ("new_param", c_void_p),
                ("new_exp_avg", c_void_p), ("new_exp_avg_sq", c_void_p), ("width", c_int32)]


ADAM_MAX_TENSORS = 8

# name -> (restype, argtypes); every symbol include/h3dgs_b200.h declares
SIGNATURES = {
    "h3dgs_abi_version": (c_int, []),
    "h3dgs_last_error": (c_char_p, []),
    "h3dgs_geom_bytes": (c_size_t, [c_int32]),
    "h3dgs_image_bytes": (c_size_t, [c_int32, c_int32]),
    "h3dgs_binning_bytes": (c_size_t, [c_int64, c_int32, c_int32]),
    "h3dgs_backward_workspace_bytes": (c_size_t, [c_int32]),
    "h3dgs_forward_begin": (c_int, [POINTER(View), POINTER(Gaussians), c_void_p, c_void_p, c_size_t,
                                     c_void_p, c_void_p, c_int, c_void_p]),
    "h3dgs_forward_finish": (c_int, [POINTER(View), POINTER(Gaussians), c_void_p, c_void_p, c_size_t,
                                      c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "h3dgs_backward": (c_int, [POINTER(View), POINTER(Gaussians), c_void_p, c_void_p, c_void_p, c_void_p,
                                c_int64, c_void_p, c_void_p, POINTER(Grads), c_void_p, c_size_t, c_int,
                                c_void_p]),
    "h3dgs_zero_rows": (c_int, [c_void_p, c_int64, POINTER(RowTensor), c_int32, c_void_p]),
    "h3dgs_zero_fill": (c_int, [POINTER(Span), c_int32, c_int32, c_void_p]),
    "h3dgs_mark_visible": (c_int, [c_int32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "h3dgs_debug_view": (c_int, [c_int32, c_int32, c_int32, c_int64, c_void_p, c_void_p, c_void_p,
                                  POINTER(DebugView)]),
    "h3dgs_expand_workspace_bytes": (c_size_t, [c_int32]),
    "h3dgs_expand_to_size": (c_int, [c_int32, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_size_t, POINTER(c_int32), c_void_p]),
    "h3dgs_expand_to_size_static_max_nodes": (c_int32, []),
    "h3dgs_expand_to_size_static": (c_int, [c_int32, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                             c_void_p, POINTER(c_int32), c_void_p]),
    "h3dgs_interpolation_weights": (c_int, [c_int32, c_void_p, c_float, c_void_p, c_void_p, c_float,
                                             c_float, c_float, c_void_p, c_void_p, c_void_p]),
    "h3dgs_cut_step_workspace_bytes": (c_size_t, [c_int32]),
    "h3dgs_cut_step": (c_int, [c_int32, c_void_p, c_void_p, c_int32, c_void_p, c_void_p, c_void_p, c_float, c_void_p,
                                c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, POINTER(c_int32), c_void_p]),
    "h3dgs_expand_to_target": (c_int64, [c_int32, c_void_p, c_int32, c_void_p, c_int64]),
    "h3dgs_hier_read_header": (c_int, [c_char_p, POINTER(c_int64), POINTER(c_int64), POINTER(c_int)]),
    "h3dgs_hier_read": (c_int, [c_char_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                 c_void_p]),
    "h3dgs_hier_write": (c_int, [c_char_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_int]),
    "h3dgs_dp_pack": (c_int, [POINTER(RowTensor), c_int32, c_void_p, c_void_p, c_int32, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_int32, c_void_p, c_void_p]),
    "h3dgs_dp_accumulate": (c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int32, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_int32, c_void_p, c_void_p]),
    "h3dgs_dp_adam": (c_int, [POINTER(AdamTensor), c_int32, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_int32,
                               c_void_p, c_void_p, c_void_p, c_void_p]),
    "h3dgs_dp_scatter_params": (c_int, [POINTER(AdamTensor), c_int32, c_void_p, c_void_p, c_void_p, c_int32, c_void_p]),
    "h3dgs_densify_stats": (c_int, [c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "h3dgs_densify_workspace_bytes": (c_size_t, [c_int64]),
    "h3dgs_densify_select": (c_int, [c_int32, c_void_p, c_void_p, c_void_p, c_int32, c_void_p, c_float, c_float, c_float,
                                      c_int32, c_void_p, c_size_t, POINTER(c_int32), c_void_p]),
    "h3dgs_densify_split_std": (c_int, [c_int32, c_void_p, c_void_p, c_int32, c_void_p, c_void_p]),
    "h3dgs_densify_build": (c_int, [c_int32, POINTER(DensifyTensor), c_int32, c_int32, c_int32, c_int32, c_void_p,
                                     c_int32, c_int32, c_int32, c_void_p, c_void_p]),
    "h3dgs_hier_build_workspace_bytes": (c_size_t, [c_int64]),
    "h3dgs_hier_build": (c_int, [c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                  c_void_p]),
    "h3dgs_sparse_adam": (c_int, [POINTER(AdamTensor), c_int32, c_void_p, c_int64, c_int64, c_void_p]),
    "h3dgs_sparse_adam_compact": (c_int, [POINTER(AdamTensor), c_int32, c_void_p, c_void_p, c_int32, c_void_p, c_void_p]),
    "h3dgs_pack_rows": (c_int, [POINTER(RowTensor), c_int32, c_void_p, c_int64, c_void_p, c_void_p]),
    "h3dgs_unpack_rows": (c_int, [POINTER(RowTensor), c_int32, c_void_p, c_int64, c_void_p, c_float, c_void_p]),
    "h3dgs_ssim_forward": (c_int, [c_int32, c_int32, c_int32, c_void_p, c_void_p, c_void_p, POINTER(c_float), c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "h3dgs_ssim_backward": (c_int, [c_int32, c_int32, c_int32, c_void_p, c_void_p, c_void_p, POINTER(c_float), c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float, c_void_p, c_void_p]),
    "h3dgs_exposure_clamp_forward": (c_int, [c_int32, c_int32, c_void_p, c_void_p, c_void_p, c_void_p]),
    "h3dgs_exposure_clamp_backward": (c_int, [c_int32, c_int32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "h3dgs_masked_l1_forward": (c_int, [c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "h3dgs_masked_l1_backward": (c_int, [c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_void_p]),
    "h3dgs_blend_stats": (c_int, [c_void_p]),
    "h3dgs_measure_peaks": (c_int, [POINTER(ctypes.c_double), c_void_p, c_int32, c_void_p]),
    "h3dgs_debug_set": (c_int, [c_int, c_int]),
    "h3dgs_launch_count": (c_int64, []),
    "h3dgs_profile_enable": (c_int, [c_int]),
    "h3dgs_profile_read": (c_int, [POINTER(ctypes.c_double), POINTER(c_int64), c_int]),
    "h3dgs_profile_name": (c_char_p, [c_int]),
}
PROF_KERNELS = 20
ABI_VERSION = 8


def measure_peaks(reps=5):
    """{"fp32_tflops", "fp32x2_tflops", "mufu_gops", "issue_ginst"} measured on the current device
    (h3dgs_measure_peaks)."""
    out = (ctypes.c_double * 4)()
    scratch = torch.zeros(4, device="cuda")
    check(lib().h3dgs_measure_peaks(out, scratch.data_ptr(), reps, stream_handle()), "h3dgs_measure_peaks")
    return {"fp32_tflops": out[0], "fp32x2_tflops": out[1], "mufu_gops": out[2], "issue_ginst": out[3]}


def profile_read(reset=True):
    """{kernel name: (total ms, launches)} accumulated while profiling was enabled."""
    L = lib()
    ms = (ctypes.c_double * PROF_KERNELS)()
    n = (c_int64 * PROF_KERNELS)()
    check(L.h3dgs_profile_read(ms, n, 1 if reset else 0), "h3dgs_profile_read")
    return {L.h3dgs_profile_name(i).decode(): (ms[i], int(n[i])) for i in range(PROF_KERNELS)}

_lib = None


def lib():
    """Load the shared library once; fail loudly when it is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise H3dgsError(
                f"{LIB_PATH} not found: build it with `python street-sparse-3dgs_b200/build.py` "
                "(there is no CPU or PyTorch fallback)")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        if handle.h3dgs_abi_version() != ABI_VERSION:
            raise H3dgsError("libh3dgs_b200.so ABI version mismatch")
        _lib = handle
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().h3dgs_last_error().decode("utf-8", "replace")
        raise H3dgsError(f"{what} failed (code {rc}): {msg}")


def ptr(t):
    """Device (or host) pointer of a tensor; None (NULL) for an empty / absent tensor."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def stream_handle():
    return torch.cuda.current_stream().cuda_stream


def as_f32(t, name):
    if t is None or t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32, got {t.dtype}")
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor")
    return t.contiguous()


def as_i32(t, name):
    if t is None or t.numel() == 0:
        return t
    if t.dtype != torch.int32:
        raise TypeError(f"{name} must be int32, got {t.dtype}")
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor")
    return t.contiguous()
