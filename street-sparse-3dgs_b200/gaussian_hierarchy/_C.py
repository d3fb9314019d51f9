"""Drop-in for ``gaussian_hierarchy._C`` (GH/ext.cpp:15-21, GH/torch/torch_interface.cpp:8-120):
``load_hierarchy``, ``write_hierarchy``, ``expand_to_target``, ``expand_to_size``,
``get_interpolation_weights`` -- same names, argument order, in-place output tensors and return values.
The cut runs as sm_100a kernels through include/h3dgs_b200.h on torch's current stream.
"""
import ctypes

import torch

import h3dgs_native as nat


def load_hierarchy(filename):
    """-> (pos[P,3], shs[P,16,3], alpha[P,1], scale[P,3], rot[P,4], nodes[N,7] int32, boxes[N,2,4]), all CPU
    (GH/torch/torch_interface.cpp:8-38).  Raises RuntimeError("File not found!") like the reference."""
    L = nat.lib()
    P = ctypes.c_int64(0)
    N = ctypes.c_int64(0)
    comp = ctypes.c_int(0)
    path = str(filename).encode()
    nat.check(L.h3dgs_hier_read_header(path, ctypes.byref(P), ctypes.byref(N), ctypes.byref(comp)),
              "load_hierarchy")
    P, N = int(P.value), int(N.value)
    pos = torch.empty((P, 3), dtype=torch.float32)
    shs = torch.empty((P, 16, 3), dtype=torch.float32)
    alpha = torch.empty((P, 1), dtype=torch.float32)
    scale = torch.empty((P, 3), dtype=torch.float32)
    rot = torch.empty((P, 4), dtype=torch.float32)
    nodes = torch.empty((N, 7), dtype=torch.int32)
    boxes = torch.empty((N, 2, 4), dtype=torch.float32)
    nat.check(L.h3dgs_hier_read(path, pos.data_ptr(), shs.data_ptr(), alpha.data_ptr(), scale.data_ptr(),
                                rot.data_ptr(), nodes.data_ptr(), boxes.data_ptr()), "load_hierarchy")
    return pos, shs, alpha, scale, rot, nodes, boxes


def write_hierarchy(filename, pos, shs, opacities, log_scales, rotations, nodes, boxes, compressed=True):
    """Writes the fp16-compressed form by default, as the reference binding does
    (GH/torch/torch_interface.cpp:40-67, GH/hierarchy_writer.h:32)."""
    L = nat.lib()

    def host(t, dtype):
        return t.detach().to(device="cpu", dtype=dtype).contiguous()

    pos_h, shs_h = host(pos, torch.float32), host(shs, torch.float32)
    op_h, sc_h, rot_h = host(opacities, torch.float32), host(log_scales, torch.float32), host(rotations, torch.float32)
    nodes_h, boxes_h = host(nodes, torch.int32), host(boxes, torch.float32)
    nat.check(L.h3dgs_hier_write(str(filename).encode(), pos_h.size(0), nodes_h.size(0), pos_h.data_ptr(),
                                 shs_h.data_ptr(), op_h.data_ptr(), sc_h.data_ptr(), rot_h.data_ptr(),
                                 nodes_h.data_ptr(), boxes_h.data_ptr(), 1 if compressed else 0),
              "write_hierarchy")


def expand_to_target(nodes, target):
    """DFS pre-order cut at depth `target`; returns a new CPU int32 tensor
    (GH/torch/torch_interface.cpp:69-75, GH/traversal.cpp:15-38)."""
    L = nat.lib()
    nodes_h = nodes.detach().to(device="cpu", dtype=torch.int32).contiguous()
    N = nodes_h.size(0)
    count = L.h3dgs_expand_to_target(N, nodes_h.data_ptr() if N else None, int(target), None, 0)
    if count < 0:
        nat.check(int(count), "expand_to_target")
    out = torch.empty((int(count),), dtype=torch.int32)
    if count:
        got = L.h3dgs_expand_to_target(N, nodes_h.data_ptr(), int(target), out.data_ptr(), int(count))
        if got < 0:
            nat.check(int(got), "expand_to_target")
    return out


_ws_cache = {}


def _workspace(nbytes, device):
    key = (device.index if device.index is not None else torch.cuda.current_device())
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty((nbytes,), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


def expand_to_size(nodes, boxes, size, viewpoint, viewdir, render_indices, parent_indices,
                   nodes_for_render_indices):
    """Granularity cut; writes the three index tensors in place and returns the entry count
    (GH/torch/torch_interface.cpp:77-98).  `viewpoint` is a CUDA tensor, `viewdir` is unused by the math."""
    L = nat.lib()
    nodes = nat.as_i32(nodes, "nodes")
    boxes = nat.as_f32(boxes, "boxes")
    viewpoint = nat.as_f32(viewpoint, "viewpoint")
    for name, t in (("render_indices", render_indices), ("parent_indices", parent_indices),
                    ("nodes_for_render_indices", nodes_for_render_indices)):
        if t.dtype != torch.int32 or not t.is_cuda or not t.is_contiguous():
            raise TypeError(f"{name} must be a contiguous CUDA int32 tensor")
    N = nodes.size(0)
    ws = _workspace(L.h3dgs_expand_workspace_bytes(N), nodes.device)
    count = ctypes.c_int32(0)
    nat.check(L.h3dgs_expand_to_size(N, float(size), nat.ptr(nodes), nat.ptr(boxes), nat.ptr(viewpoint),
                                     nat.ptr(render_indices), nat.ptr(parent_indices),
                                     nat.ptr(nodes_for_render_indices), nat.ptr(ws), ws.numel(),
                                     ctypes.byref(count), nat.stream_handle()), "expand_to_size")
    return int(count.value)


def get_interpolation_weights(indices, size, nodes, boxes, viewpoint, viewdir, ts, num_kids):
    """Per-entry interpolation weight and sibling count, written in place into the first len(indices)
    slots (GH/torch/torch_interface.cpp:100-120).  `viewpoint` is read on the host, as in the reference;
    a CUDA tensor is accepted too (it costs a sync)."""
    L = nat.lib()
    indices = nat.as_i32(indices, "indices")
    nodes = nat.as_i32(nodes, "nodes")
    boxes = nat.as_f32(boxes, "boxes")
    vp = viewpoint.detach().to("cpu", torch.float32).flatten()
    if ts.dtype != torch.float32 or not ts.is_cuda or not ts.is_contiguous():
        raise TypeError("ts must be a contiguous CUDA float32 tensor")
    if num_kids.dtype != torch.int32 or not num_kids.is_cuda or not num_kids.is_contiguous():
        raise TypeError("num_kids must be a contiguous CUDA int32 tensor")
    K = indices.size(0)
    if ts.numel() < K or num_kids.numel() < K:
        raise RuntimeError("ts / num_kids are shorter than indices")
    nat.check(L.h3dgs_interpolation_weights(K, nat.ptr(indices), float(size), nat.ptr(nodes), nat.ptr(boxes),
                                            float(vp[0]), float(vp[1]), float(vp[2]), nat.ptr(ts),
                                            nat.ptr(num_kids), nat.stream_handle()),
              "get_interpolation_weights")
