"""Drop-in for the reference's pybind module ``diff_gaussian_rasterization._C``
(HR/ext.cpp:15-18, HR/rasterize_points.h:18-73): same function names, positional arguments, return
tuples and error behaviour, implemented on top of the C ABI in include/h3dgs_b200.h.

Differences that a caller can observe, all deliberate:
  * kernels run on torch's current stream (the reference used the legacy default stream);
  * ``mark_visible`` exists (the reference's Python calls it but its binding is missing);
  * with ``indices`` given, the semantics are those of ``gaussian_renderer.render_post``'s Python
    interpolation (complete gradients, ``dL_dmeans2D`` per slot) -- the reference's native index
    path has an incomplete backward and is never reached from its own Python.
"""
import ctypes
import os

import torch

import h3dgs_native as nat

NUM_CHANNELS = 3

# Host path of the two hot calls.  "ext" (default when lib/_h3dgs_torch.so exists): csrc/torch_binding.cpp, one C++
# call per forward / backward.  "ctypes": the Python binding below, kept as the reference binding of the C ABI.
# Both drive the same libh3dgs_b200.so entry points in the same order; tests/test_gpu_host_paths.py compares them.
HOST = os.environ.get("H3DGS_HOST", "auto")
_ext_mod = None


def ext():
    """The torch extension module, or None when the ctypes host path is selected / the extension is not built."""
    global _ext_mod
    if HOST == "ctypes":
        return None
    if _ext_mod is None:
        path = os.path.join(os.path.dirname(nat.LIB_PATH), "_h3dgs_torch.so")
        if not os.path.exists(path):
            if HOST == "ext":
                raise nat.H3dgsError(f"{path} not found: build it with `python street-sparse-3dgs_b200/build.py`")
            return None
        import importlib.machinery
        import importlib.util
        nat.lib()   # libh3dgs_b200.so first (and its ABI check)
        loader = importlib.machinery.ExtensionFileLoader("_h3dgs_torch", path)
        mod = importlib.util.module_from_spec(importlib.util.spec_from_loader("_h3dgs_torch", loader))
        loader.exec_module(mod)
        if mod.abi_version() != nat.ABI_VERSION:
            raise nat.H3dgsError("_h3dgs_torch.so was built against another ABI version")
        _ext_mod = mod
    return _ext_mod


_none = {}


def _t(x, like):
    """None -> an empty tensor (the extension's 'absent')."""
    if x is not None:
        return x
    key = like.device
    e = _none.get(key)
    if e is None:
        e = _none[key] = torch.empty(0, device=like.device)
    return e


def _null(t):
    return t is None or t.numel() == 0


def _scene(bg, indices, parent_indices, ts, kids, means3D, colors, opacity, scales, rotations,
           scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, sh, degree,
           campos, skyboxnum, sh_rest=None, acts=None):
    """Build the two C structs; returns (view, gaussians, keepalive list).  With ``sh_rest`` / ``acts`` the
    attribute tensors are the RAW parameter store (``sh`` = _features_dc, ``sh_rest`` = _features_rest,
    ``acts`` = (scale, rotation, opacity) activation codes applied inside the kernels)."""
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    keep = []

    def f32(t, name):
        t = nat.as_f32(t, name)
        keep.append(t)
        return nat.ptr(t)

    def i32(t, name):
        t = nat.as_i32(t, name)
        keep.append(t)
        return nat.ptr(t)

    rows = means3D.size(0)
    fused = not _null(indices)
    P = (indices.size(0) + skyboxnum) if fused else rows
    view = nat.View(int(W), int(H), float(tan_fovx), float(tan_fovy), f32(viewmatrix, "viewmatrix"),
                    f32(projmatrix, "projmatrix"), f32(campos, "campos"), f32(bg, "bg"))
    M = 0 if _null(sh) else sh.size(1)
    if not _null(sh_rest):
        if _null(sh) or sh.size(1) != 1 or sh_rest.size(0) != sh.size(0):
            raise RuntimeError("sh_rest needs sh of shape (rows, 1, 3) with the same number of rows")
        M = 1 + sh_rest.size(1)
    g = nat.Gaussians()
    g.P = P
    g.store_rows = rows
    g.sh_degree = int(degree)
    g.sh_coeffs = M
    g.skyboxnum = int(skyboxnum)
    g.scale_modifier = float(scale_modifier)
    g.means3D = f32(means3D, "means3D")
    g.shs = f32(sh, "sh")
    g.colors_precomp = f32(colors, "colors_precomp")
    g.opacities = f32(opacity, "opacities")
    g.scales = f32(scales, "scales")
    g.rotations = f32(rotations, "rotations")
    g.cov3D_precomp = f32(cov3D_precomp, "cov3D_precomp")
    g.render_indices = i32(indices, "render_indices") if fused else None
    g.parent_indices = i32(parent_indices, "parent_indices") if fused else None
    g.ts = f32(ts, "interpolation_weights")
    g.kids = i32(kids, "num_node_kids")
    g.shs_rest = f32(sh_rest, "sh_rest") if not _null(sh_rest) else None
    g.scale_activation, g.rotation_activation, g.opacity_activation = acts if acts is not None else (0, 0, 0)
    if fused and not _null(parent_indices) and parent_indices.size(0) < indices.size(0):
        raise RuntimeError("parent_indices is shorter than render_indices")
    for name, t in (("interpolation_weights", ts), ("num_node_kids", kids)):
        if not _null(t) and t.size(0) < P - skyboxnum:
            raise RuntimeError(f"{name} must have at least one entry per rendered Gaussian")
    return view, g, keep


def _grad_shapes(rows, M, sh, sh_rest, colors, scales, rotations, cov3D_precomp):
    """Per-store-row gradient tensors of one backward call; None = that input is absent."""
    split = not _null(sh_rest)
    return {"means3D": (rows, 3), "opacity": (rows, 1),
            "sh": (rows, 1 if split else M, 3) if not _null(sh) else None,
            "sh_rest": (rows, M - 1, 3) if split else None,
            "colors": (rows, NUM_CHANNELS) if not _null(colors) else None,
            "scales": (rows, 3) if not _null(scales) else None,
            "rot": (rows, 4) if not _null(rotations) else None,
            "cov3D": (rows, 6) if not _null(cov3D_precomp) else None}


class _Prezero:
    """Gradient tensors of a fused-cut render, allocated by the FORWARD and zero-filled on a side stream
    forked right before the blend kernel.  The 4 GB fill of a 17 M-row store (0.6 ms of pure HBM writes) then
    runs under the issue-bound forward blend and in the host-side gap between forward and backward instead of
    in front of the backward kernels.  ``H3DGS_PREZERO=0`` (or ``_C.PREZERO = False``) disables."""

    _streams = {}

    def __init__(self, tensors, means2D):
        self.tensors, self.means2D, self.done = tensors, means2D, None

    @classmethod
    def side_stream(cls, dev):
        s = cls._streams.get(dev.index)
        if s is None:
            s = cls._streams[dev.index] = torch.cuda.Stream(device=dev)
        return s

    @classmethod
    def allocate(cls, dev, shapes, means2D_rows, P):
        opts = dict(dtype=torch.float32, device=dev)
        tensors = {k: torch.empty(v, **opts) for k, v in shapes.items() if v is not None}
        return cls(tensors, torch.empty((max(means2D_rows, P), 3), **opts))

    def fill(self, dev, fork_event, P):
        """Enqueue the zero fill on the side stream, ordered after ``fork_event``."""
        L = nat.lib()
        side = self.side_stream(dev)
        side.wait_event(fork_event)
        spans = [(t.data_ptr(), t.numel()) for t in self.tensors.values()]
        m2 = self.means2D
        if m2.size(0) > P:   # rows of a store-sized means2D beyond the rendered slots get no gradient
            spans.append((m2.data_ptr() + P * 3 * 4, (m2.size(0) - P) * 3))
        for i in range(0, len(spans), 8):
            chunk = spans[i:i + 8]
            arr = (nat.Span * len(chunk))(*[nat.Span(a, n) for a, n in chunk])
            nat.check(L.h3dgs_zero_fill(arr, len(chunk), 0, side.cuda_stream), "h3dgs_zero_fill")
        for t in list(self.tensors.values()) + [m2]:
            t.record_stream(side)
        self.done = torch.cuda.Event()
        self.done.record(side)


PREZERO = os.environ.get("H3DGS_PREZERO", "1") != "0"


class _Aux:
    """What a forward hands to its own backward besides the opaque buffers: the pre-zeroed gradient tensors (or
    None) and the number of visible slots."""

    def __init__(self, prezero, V):
        self.prezero, self.V = prezero, V


class SparseGradients:
    """Compact gradients of one fused-cut render (H3dgsGrads.compact_*): ``rows`` int64 [capacity] holds the store
    row of each compact row, ``count`` (device int32 [1]) how many are valid, ``grads`` maps a name to its
    [capacity, ...] tensor and ``params`` the same name to the parameter tensor it belongs to.  Filled by the
    backward of a render made with ``sparse=<this object>``; consumed by ``scene.OurAdam.Adam.step_sparse``."""

    def __init__(self):
        self.rows = self.count = None
        self.grads, self.params = {}, {}
        self.capacity = 0
        self.ready = False

    def dense(self, name):
        """The equivalent dense [N, ...] gradient (tests / debugging; costs a sync)."""
        g, p = self.grads[name], self.params[name]
        U = int(self.count.item())
        out = torch.zeros_like(p)
        out.index_add_(0, self.rows[:U], g[:U].reshape((U,) + tuple(p.shape[1:])))
        return out


_row_maps = {}


def _row_map(dev, rows):
    """int32 [rows] scratch of the compact-gradient mode, all -1 between calls (the library restores it)."""
    E = ext()
    if E is not None:
        return E.row_map(torch.empty(0, device=dev), rows)
    key = (dev.index, rows)
    m = _row_maps.get(key)
    if m is None:
        for k in [k for k in _row_maps if k[0] == dev.index]:
            del _row_maps[k]
        m = _row_maps[key] = torch.full((rows,), -1, dtype=torch.int32, device=dev)
    return m


class _Counts:
    """Pinned host words + CUDA event through which a forward's (num_rendered, num_visible) reach the host
    (h3dgs_forward_begin).  A small free list: renders in flight on different streams each hold their own."""

    _free = []

    def __init__(self):
        self.host = torch.zeros(2, dtype=torch.int64).pin_memory()
        self.event = torch.cuda.Event()

    @classmethod
    def get(cls):
        c = cls._free.pop() if cls._free else cls()
        c.event.record()   # materialises the CUDA event; the library re-records it behind the copy
        return c

    def read(self):
        """Wait for the copy alone (the GPU keeps working on the depth sort) and return (R, V)."""
        self.event.synchronize()
        R, V = int(self.host[0]), int(self.host[1])
        _Counts._free.append(self)
        return R, V



def rasterize_gaussians_ex(bg, indices, parent_indices, ts, kids, means3D, colors, opacity, scales,
                           rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx,
                           tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug,
                           do_depth, skyboxnum=0, sh_rest=None, acts=None, prezero_means2D_rows=None):
    """``prezero_means2D_rows`` (fused-cut mode, set by the autograd wrapper when a backward will follow):
    also allocate the backward's gradient tensors and start their zero fill; an 8th return value carries them."""
    E = ext()
    if E is not None and means3D.is_cuda:
        a = acts if acts is not None else (0, 0, 0)
        want = prezero_means2D_rows is not None and prezero_means2D_rows >= 0 and PREZERO and not _arena.enabled()
        m = means3D
        R, color, radii, geom, binning, img, invdepth, V, pre = E.rasterize_forward(
            _t(bg, m), _t(indices, m), _t(parent_indices, m), _t(ts, m), _t(kids, m), m, _t(colors, m),
            _t(opacity, m), _t(scales, m), _t(rotations, m), float(scale_modifier), _t(cov3D_precomp, m), viewmatrix,
            projmatrix, float(tan_fovx), float(tan_fovy), int(image_height), int(image_width), _t(sh, m), int(degree),
            campos, bool(prefiltered), bool(debug), bool(do_depth), int(skyboxnum), _t(sh_rest, m), int(a[0]),
            int(a[1]), int(a[2]), int(prezero_means2D_rows) if want else -1)
        if prezero_means2D_rows is not None:
            return R, color, radii, geom, binning, img, invdepth, _Aux(pre, V)
        return R, color, radii, geom, binning, img, invdepth
    L = nat.lib()
    dev = means3D.device
    H, W = int(image_height), int(image_width)
    view, g, keep = _scene(bg, indices, parent_indices, ts, kids, means3D, colors, opacity, scales,
                           rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx,
                           tan_fovy, H, W, sh, degree, campos, skyboxnum, sh_rest, acts)
    P = g.P
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    geom = torch.empty((L.h3dgs_geom_bytes(P),), dtype=torch.uint8, device=dev)
    stream = nat.stream_handle()
    counts = _Counts.get()
    nat.check(L.h3dgs_forward_begin(ctypes.byref(view), ctypes.byref(g), nat.ptr(radii), nat.ptr(geom),
                                    geom.numel(), counts.host.data_ptr(), counts.event.cuda_event,
                                    int(bool(debug)), stream), "h3dgs_forward_begin")
    # everything that does not need R is allocated while the preprocess kernels run; then wait for the counts
    # alone -- the depth sort of the visible slots is already queued behind them
    img = torch.empty((L.h3dgs_image_bytes(W, H),), dtype=torch.uint8, device=dev)
    out_color = torch.empty((NUM_CHANNELS, H, W), dtype=torch.float32, device=dev)
    out_invdepth = torch.empty((1 if do_depth else 0, H, W), dtype=torch.float32, device=dev)
    R, V = counts.read()
    if R >= (1 << 30):
        raise nat.H3dgsError("more than 2^30 tile instances")
    binning = torch.empty((L.h3dgs_binning_bytes(R, W, H) if R > 0 else 0,), dtype=torch.uint8, device=dev)
    pre, fork = None, None
    if prezero_means2D_rows is not None and prezero_means2D_rows >= 0 and PREZERO and not _null(indices) \
            and R > 0 and P > 0 and not _arena.enabled():
        # allocated BEFORE the fork point is recorded: whatever used these blocks last was enqueued earlier
        pre = _Prezero.allocate(dev, _grad_shapes(g.store_rows, g.sh_coeffs, sh, sh_rest, colors, scales,
                                                  rotations, cov3D_precomp), prezero_means2D_rows, P)
        fork = torch.cuda.Event()
        fork.record()   # creates the CUDA event; the library re-records it in front of the blend kernel
    nat.check(L.h3dgs_forward_finish(ctypes.byref(view), ctypes.byref(g), nat.ptr(geom), nat.ptr(binning),
                                     binning.numel(), nat.ptr(img), R, V, nat.ptr(out_color),
                                     nat.ptr(out_invdepth), int(bool(debug)), stream,
                                     fork.cuda_event if fork is not None else None),
              "h3dgs_forward_finish")
    if pre is not None:
        pre.fill(dev, fork, P)
    if prezero_means2D_rows is not None:
        return R, out_color, radii, geom, binning, img, out_invdepth, _Aux(pre, V)
    return R, out_color, radii, geom, binning, img, out_invdepth


def rasterize_gaussians(bg, indices, parent_indices, ts, kids, means3D, colors, opacity, scales, rotations,
                        scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                        image_height, image_width, sh, degree, campos, prefiltered, debug, do_depth):
    """-> (num_rendered, color[3,H,W], radii[P], geomBuffer, binningBuffer, imgBuffer, invdepth)
    (HR/rasterize_points.cu:35-136)."""
    return rasterize_gaussians_ex(bg, indices, parent_indices, ts, kids, means3D, colors, opacity, scales,
                                  rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                                  tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                                  prefiltered, debug, do_depth, 0)


class _Lease:
    """Per-row gradient tensors handed to one backward call; ``release`` turns them back into bare storages
    so that autograd can take the tensors over without a copy."""

    def __init__(self, entry, tensors, dirty):
        self.entry, self.tensors, self.dirty = entry, tensors, dirty

    def release(self):
        e = self.entry
        e["storages"] = {k: v.untyped_storage() for k, v in self.tensors.items()}
        e["dirty"] = self.dirty
        self.tensors = None


class _GradArena:
    """Recycles the N-row gradient tensors of fused-cut renders (render_post).

    A render adds into the rows of its cut only (a few % of a 17 M-row store), but autograd needs dense
    [N, ...] tensors: the reference zero-fills 4 GB per backward for them (HR/rasterize_points.cu:168-198 plus
    the index_put scatters of its Python glue).  Here a gradient set is allocated and zero-filled once; the
    backward kernels record which rows they added into (``H3dgsGrads.dirty_rows``); when every tensor of the
    set has been dropped by its users (``optimizer.zero_grad(set_to_none=True)``, ``p.grad = None``) the set is
    handed out again after ``h3dgs_zero_rows`` cleared just those rows.

    Contract: while they hold such a gradient, callers may scale it or zero parts of it (train_post.py:171-182
    zeroes skybox and anchor rows), but must not write non-zero values into rows the render did not touch.
    Off by default: clearing scattered rows is partial-sector read-modify-write traffic, and at the density of
    the benchmark cut (15 % of the store rows touched per render) it costs as much as the streaming 4 GB
    fill, which the library now overlaps with the blend kernel instead.  It pays when a render touches a few
    % of the rows or less.  ``H3DGS_GRAD_ARENA=1`` enables recycling; ``H3DGS_GRAD_ARENA=check`` additionally
    verifies on every reuse that the recycled set is all zero."""

    MAX_SETS = 4

    def __init__(self):
        self.entries = []

    @staticmethod
    def enabled():
        import os
        return os.environ.get("H3DGS_GRAD_ARENA", "0") in ("1", "check") and hasattr(torch._C, "_storage_Use_Count")

    def release_all(self):
        self.entries.clear()

    def lease(self, dev, shapes, P):
        import os
        L = nat.lib()
        key = (dev.index, tuple(sorted((k, v) for k, v in shapes.items() if v is not None)))
        opts = dict(dtype=torch.float32, device=dev)
        dirty = torch.empty((1 + 2 * P,), dtype=torch.int32, device=dev)
        for e in self.entries:
            if e["key"] != key or e["storages"] is None:
                continue
            if any(torch._C._storage_Use_Count(st._cdata) != 1 for st in e["storages"].values()):
                continue   # somebody still holds (a view of) one of these gradients
            tensors = {k: torch.empty(0, **opts).set_(st, 0, shapes[k]) for k, st in e["storages"].items()}
            e["storages"] = None
            names = list(tensors)
            arr = (nat.RowTensor * len(names))()
            for i, k in enumerate(names):
                arr[i].data = tensors[k].data_ptr()
                arr[i].width = tensors[k][0].numel()
            nat.check(L.h3dgs_zero_rows(nat.ptr(e["dirty"]), e["dirty"].numel() - 1, arr, len(names),
                                        nat.stream_handle()), "h3dgs_zero_rows")
            if os.environ.get("H3DGS_GRAD_ARENA") == "check":
                for k, v in tensors.items():
                    if bool(v.any()):
                        raise nat.H3dgsError(f"gradient arena: recycled '{k}' gradient is not all zero -- a caller "
                                             "wrote into rows outside the render's cut")
            return _Lease(e, tensors, dirty)
        tensors = {k: torch.zeros(v, **opts) for k, v in shapes.items() if v is not None}
        e = {"key": key, "storages": None, "dirty": None}
        self.entries.append(e)
        if len(self.entries) > self.MAX_SETS:
            self.entries.pop(0)
        return _Lease(e, tensors, dirty)


_arena = _GradArena()


def release_gradient_arena():
    """Drop the recycled gradient sets (frees their device memory once the caller's references are gone)."""
    _arena.release_all()


def rasterize_gaussians_backward_ex(bg, indices, parent_indices, ts, kids, means3D, radii, colors,
                                    opacities, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                                    projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_invdepth, sh,
                                    degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug,
                                    skyboxnum=0, sh_rest=None, acts=None, prezero=None, sparse=None, num_visible=0):
    """With ``sh_rest`` the returned dL_dsh is the pair (dL_dfeatures_dc, dL_dfeatures_rest).  ``prezero``: the
    gradient tensors the forward allocated and zero-filled (``_Prezero``); dL_dmeans2D then has its rows.
    ``sparse`` (a ``SparseGradients``, fused-cut mode): the per-store-row gradients are produced in compact form
    into it -- no N-row tensors exist -- and the returned per-row gradients are None."""
    E = ext()
    if E is not None and means3D.is_cuda and not isinstance(prezero, _Prezero) and not _arena.enabled():
        return _backward_ext(E, bg, indices, parent_indices, ts, kids, means3D, radii, colors, opacities, scales,
                             rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                             dL_dout_color, dL_dout_invdepth, sh, degree, campos, geomBuffer, R, binningBuffer,
                             imageBuffer, debug, skyboxnum, sh_rest, acts, prezero, sparse, num_visible)
    if prezero is not None and not isinstance(prezero, _Prezero):
        prezero = None   # made by the extension's forward; this path fills its own
    L = nat.lib()
    dev = means3D.device
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    view, g, keep = _scene(bg, indices, parent_indices, ts, kids, means3D, colors, opacities, scales,
                           rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx,
                           tan_fovy, H, W, sh, degree, campos, skyboxnum, sh_rest, acts)
    P, rows, M = g.P, g.store_rows, g.sh_coeffs
    split = not _null(sh_rest)
    opts = dict(dtype=torch.float32, device=dev)
    # per-store-row gradients; only those of inputs that exist are materialised, the others come back empty
    shapes = _grad_shapes(rows, M, sh, sh_rest, colors, scales, rotations, cov3D_precomp)
    fused = not _null(indices)
    if sparse is not None:
        if not fused:
            raise RuntimeError("sparse gradients need a hierarchy-cut render (render_indices)")
        return _backward_sparse(L, dev, view, g, shapes, sparse, num_visible, radii, geomBuffer, binningBuffer,
                                imageBuffer, R, dL_dout_color, dL_dout_invdepth, debug, split)
    if prezero is not None and not (fused and P > 0 and R > 0 and prezero.done is not None):
        prezero = None
    dL_dmeans2D = prezero.means2D if prezero is not None else torch.empty((P, 3), **opts)
    lease = _arena.lease(dev, shapes, P) if (prezero is None and fused and P > 0 and R > 0 and _arena.enabled()) else None
    if prezero is not None:
        torch.cuda.current_stream().wait_event(prezero.done)
        t = prezero.tensors
    elif lease is not None:
        t = lease.tensors
    else:
        t = {k: torch.empty(v, **opts) for k, v in shapes.items() if v is not None}
    empty = {"sh": (0, 0, 3), "colors": (0, 3), "scales": (0, 3), "rot": (0, 4), "cov3D": (0, 6)}
    dL_dmeans3D, dL_dopacity = t["means3D"], t["opacity"]
    dL_dsh = t["sh"] if "sh" in t else torch.empty(empty["sh"], **opts)
    dL_dsh_rest = t.get("sh_rest")
    dL_dcolors = t["colors"] if "colors" in t else torch.empty(empty["colors"], **opts)
    dL_dscales = t["scales"] if "scales" in t else torch.empty(empty["scales"], **opts)
    dL_drot = t["rot"] if "rot" in t else torch.empty(empty["rot"], **opts)
    dL_dcov3D = t["cov3D"] if "cov3D" in t else torch.empty(empty["cov3D"], **opts)
    grads = nat.Grads(nat.ptr(dL_dmeans2D), nat.ptr(dL_dmeans3D), nat.ptr(dL_dsh), nat.ptr(dL_dcolors),
                      nat.ptr(dL_dopacity), nat.ptr(dL_dscales), nat.ptr(dL_drot), nat.ptr(dL_dcov3D),
                      nat.ptr(dL_dsh_rest), nat.ptr(lease.dirty) if lease is not None else None,
                      1 if (lease is not None or prezero is not None) else 0)
    ws = torch.empty((L.h3dgs_backward_workspace_bytes(P),), dtype=torch.uint8, device=dev)
    dcol = nat.as_f32(dL_dout_color, "dL_dout_color")
    ddep = nat.as_f32(dL_dout_invdepth, "dL_dout_invdepth")
    nat.check(L.h3dgs_backward(ctypes.byref(view), ctypes.byref(g), nat.ptr(radii), nat.ptr(geomBuffer),
                               nat.ptr(binningBuffer), nat.ptr(imageBuffer), int(R), nat.ptr(dcol),
                               nat.ptr(ddep), ctypes.byref(grads), nat.ptr(ws), ws.numel(),
                               int(bool(debug)), nat.stream_handle()),
              "h3dgs_backward")
    if lease is not None:
        lease.release()
        del lease, t
    if split:
        dL_dsh = (dL_dsh, dL_dsh_rest)
    return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drot


_SPARSE_NAMES = ("means3D", "sh", "sh_rest", "colors", "opacity", "scales", "rot", "cov3D")


def _backward_ext(E, bg, indices, parent_indices, ts, kids, means3D, radii, colors, opacities, scales, rotations,
                  scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color,
                  dL_dout_invdepth, sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug, skyboxnum,
                  sh_rest, acts, prezero, sparse, num_visible):
    a = acts if acts is not None else (0, 0, 0)
    m = means3D
    split = not _null(sh_rest)
    if sparse is not None and _null(indices):
        raise RuntimeError("sparse gradients need a hierarchy-cut render (render_indices)")
    (m2, dcolors, dopac, dmeans3D, dcov, dsh, dsh_rest, dscales, drot, crow, ccount) = E.rasterize_backward(
        _t(bg, m), _t(indices, m), _t(parent_indices, m), _t(ts, m), _t(kids, m), m, _t(radii, m), _t(colors, m),
        _t(opacities, m), _t(scales, m), _t(rotations, m), float(scale_modifier), _t(cov3D_precomp, m), viewmatrix,
        projmatrix, float(tan_fovx), float(tan_fovy), dL_dout_color, _t(dL_dout_invdepth, m), _t(sh, m), int(degree),
        campos, geomBuffer, int(R), binningBuffer, imageBuffer, bool(debug), int(skyboxnum), _t(sh_rest, m),
        int(a[0]), int(a[1]), int(a[2]), prezero, 1 if sparse is not None else 0, int(num_visible))
    if sparse is not None:
        present = {"means3D": dmeans3D, "opacity": dopac, "sh": dsh if not _null(sh) else None, "sh_rest": dsh_rest,
                   "colors": dcolors if not _null(colors) else None, "scales": dscales if not _null(scales) else None,
                   "rot": drot if not _null(rotations) else None, "cov3D": dcov if not _null(cov3D_precomp) else None}
        sparse.grads = {k: v for k, v in present.items() if v is not None}
        sparse.rows, sparse.count, sparse.capacity = crow, ccount, crow.numel()
        sparse.ready = True
        return m2, None, None, None, None, (None, None) if split else None, None, None
    if split:
        dsh = (dsh, dsh_rest)
    return m2, dcolors, dopac, dmeans3D, dcov, dsh, dscales, drot


def _backward_sparse(L, dev, view, g, shapes, sparse, V, radii, geomBuffer, binningBuffer, imageBuffer, R,
                     dL_dout_color, dL_dout_invdepth, debug, split):
    P, rows = g.P, g.store_rows
    opts = dict(dtype=torch.float32, device=dev)
    cap = max(2 * int(V), 1)   # a visible slot adds into at most two store rows (child, parent)
    t = {k: torch.empty((cap,) + tuple(v[1:]), **opts) for k, v in shapes.items() if v is not None}
    sparse.rows = torch.empty((cap,), dtype=torch.int64, device=dev)
    sparse.count = torch.empty((1,), dtype=torch.int32, device=dev)
    sparse.capacity = cap
    sparse.grads = t
    dL_dmeans2D = torch.empty((P, 3), **opts)
    grads = nat.Grads(nat.ptr(dL_dmeans2D), nat.ptr(t["means3D"]), nat.ptr(t.get("sh")), nat.ptr(t.get("colors")),
                      nat.ptr(t["opacity"]), nat.ptr(t.get("scales")), nat.ptr(t.get("rot")), nat.ptr(t.get("cov3D")),
                      nat.ptr(t.get("sh_rest")), None, 0,
                      sparse.rows.data_ptr(), sparse.count.data_ptr(), _row_map(dev, rows).data_ptr(), cap)
    ws = torch.empty((L.h3dgs_backward_workspace_bytes(P),), dtype=torch.uint8, device=dev)
    dcol = nat.as_f32(dL_dout_color, "dL_dout_color")
    ddep = nat.as_f32(dL_dout_invdepth, "dL_dout_invdepth")
    nat.check(L.h3dgs_backward(ctypes.byref(view), ctypes.byref(g), nat.ptr(radii), nat.ptr(geomBuffer),
                               nat.ptr(binningBuffer), nat.ptr(imageBuffer), int(R), nat.ptr(dcol), nat.ptr(ddep),
                               ctypes.byref(grads), nat.ptr(ws), ws.numel(), int(bool(debug)), nat.stream_handle()),
              "h3dgs_backward")
    sparse.ready = True
    return dL_dmeans2D, None, None, None, None, (None, None) if split else None, None, None


def rasterize_gaussians_backward(bg, indices, parent_indices, ts, kids, means3D, radii, colors, opacities,
                                 scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                                 tan_fovx, tan_fovy, dL_dout_color, dL_dout_invdepth, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, debug):
    """-> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)
    (HR/rasterize_points.cu:138-242)."""
    return rasterize_gaussians_backward_ex(bg, indices, parent_indices, ts, kids, means3D, radii, colors,
                                           opacities, scales, rotations, scale_modifier, cov3D_precomp,
                                           viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color,
                                           dL_dout_invdepth, sh, degree, campos, geomBuffer, R,
                                           binningBuffer, imageBuffer, debug, 0)


def mark_visible(positions, viewmatrix, projmatrix):
    """bool[P]: in front of the near plane (HR/cuda_rasterizer/rasterizer_impl.cu:54-66,145-157)."""
    L = nat.lib()
    positions = nat.as_f32(positions, "positions")
    P = positions.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=positions.device)
    if P:
        v = nat.as_f32(viewmatrix, "viewmatrix")
        p = nat.as_f32(projmatrix, "projmatrix")
        nat.check(L.h3dgs_mark_visible(P, nat.ptr(positions), nat.ptr(v), nat.ptr(p), nat.ptr(present),
                                       nat.stream_handle()), "h3dgs_mark_visible")
    return present


def debug_view(P, W, H, R, geom, binning, img):
    """Typed views into the opaque buffers (parity tests): dict of torch tensors aliasing them."""
    L = nat.lib()
    dv = nat.DebugView()
    nat.check(L.h3dgs_debug_view(P, W, H, R, nat.ptr(geom), nat.ptr(binning), nat.ptr(img),
                                 ctypes.byref(dv)), "h3dgs_debug_view")

    def view_of(buf, address, count, dtype):
        if not address or buf is None or buf.numel() == 0:
            return None
        off = address - buf.data_ptr()
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        return buf[off:off + nbytes].view(dtype)

    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    return {
        "point_list_keys": view_of(binning, dv.point_list_keys, R, torch.int64),
        "point_list": view_of(binning, dv.point_list, R, torch.int32),
        "ranges": view_of(img, dv.ranges, tiles * 2, torch.int32),
        "final_T": view_of(img, dv.final_T, W * H, torch.float32),
        "n_contrib": view_of(img, dv.n_contrib, W * H, torch.int32),
        "order_fwd": view_of(img, dv.order_fwd, tiles, torch.int32),
        "order_bwd": view_of(img, dv.order_bwd, tiles, torch.int32),
        "tile_last": view_of(img, dv.tile_last, tiles, torch.int32),
        "depths": view_of(geom, dv.depths, P, torch.float32),
        "tiles_touched": view_of(geom, dv.tiles_touched, P, torch.int32),
        "point_offsets": view_of(geom, dv.point_offsets, P, torch.int32),
        "records": view_of(geom, dv.records, P * 16, torch.float32),
        "cov3D": view_of(geom, dv.cov3D, P * 6, torch.float32),
        "tile_rects": view_of(geom, dv.tile_rects, P * 4, torch.int16),
    }
