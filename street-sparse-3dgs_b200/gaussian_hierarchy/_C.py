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


def build_hierarchy(positions, shs, opacities, scales, rotations, align=True, return_debug=False):
    """GPU replacement of the tree stages of the reference's offline ``GaussianHierarchyCreator`` executable
    (GH/mainHierarchyCreator.cpp:130-141: KD-tree, ClusterMerger, RotationAligner, Writer::makeHierarchy); the
    reference has no Python binding for them.  Takes N ACTIVATED Gaussians as CUDA tensors (scales not log,
    rotations (w, x, y, z), shs [N,16,3] or [N,48]) and returns the seven tensors ``load_hierarchy`` returns and
    ``write_hierarchy`` takes -- (pos[2N-1,3], shs[2N-1,16,3], alpha[2N-1,1], log_scale[2N-1,3], rot[2N-1,4],
    nodes[2N-1,7] int32, boxes[2N-1,2,4]) -- on the device.  With ``return_debug`` also the moment-matched covariances
    [2N-1,6] and the input index of every leaf row (-1 for merged rows)."""
    L = nat.lib()
    positions = nat.as_f32(positions, "positions")
    N = positions.size(0)
    shs = nat.as_f32(shs, "shs").reshape(N, 48)
    opacities = nat.as_f32(opacities, "opacities").reshape(N)
    scales = nat.as_f32(scales, "scales")
    rotations = nat.as_f32(rotations, "rotations")
    if N == 0:
        raise RuntimeError("build_hierarchy: no Gaussians")
    if scales.shape != (N, 3) or rotations.shape != (N, 4) or positions.shape != (N, 3):
        raise RuntimeError("build_hierarchy: expected positions [N,3], scales [N,3], rotations [N,4]")
    dev = positions.device
    nn = 2 * N - 1
    f = dict(dtype=torch.float32, device=dev)
    o_pos, o_shs, o_op = torch.empty((nn, 3), **f), torch.empty((nn, 16, 3), **f), torch.empty((nn, 1), **f)
    o_ls, o_rot = torch.empty((nn, 3), **f), torch.empty((nn, 4), **f)
    o_nodes = torch.empty((nn, 7), dtype=torch.int32, device=dev)
    o_boxes = torch.empty((nn, 2, 4), **f)
    o_cov = torch.empty((nn, 6), **f) if return_debug else None
    o_leaf = torch.empty((nn,), dtype=torch.int32, device=dev) if return_debug else None
    ws = torch.empty((L.h3dgs_hier_build_workspace_bytes(N),), dtype=torch.uint8, device=dev)
    nat.check(L.h3dgs_hier_build(N, nat.ptr(positions), nat.ptr(shs), nat.ptr(opacities), nat.ptr(scales),
                                 nat.ptr(rotations), 1 if align else 0, nat.ptr(o_pos), nat.ptr(o_shs), nat.ptr(o_op),
                                 nat.ptr(o_ls), nat.ptr(o_rot), nat.ptr(o_nodes), nat.ptr(o_boxes), nat.ptr(o_cov),
                                 nat.ptr(o_leaf), nat.ptr(ws), ws.numel(), nat.stream_handle()), "build_hierarchy")
    out = (o_pos, o_shs, o_op, o_ls, o_rot, o_nodes, o_boxes)
    return out + (o_cov, o_leaf) if return_debug else out


class IncrementalCut:
    """The reference viewer's frame-to-frame cut maintenance (Switching::changeToSizeStep, GH/runtime_switching.cu:283-400;
    no Python binding in the reference): keeps a node list and its split flags on the device and moves the cut one
    level per ``step`` towards the cut of ``expand_to_size`` for the current viewpoint."""

    def __init__(self, nodes, boxes, max_nodes=None):
        self.nodes = nat.as_i32(nodes, "nodes")
        self.boxes = nat.as_f32(boxes, "boxes")
        dev = self.nodes.device
        n = self.nodes.size(0)
        self.max_nodes = int(max_nodes if max_nodes is not None else n)
        i32 = dict(dtype=torch.int32, device=dev)
        self.node_indices = [torch.zeros(self.max_nodes, **i32), torch.zeros(self.max_nodes, **i32)]
        self.split = torch.zeros(n, **i32)
        self.render_indices = torch.zeros(n, **i32)
        self.parent_indices = torch.zeros(n, **i32)
        self.nodes_of_render_indices = torch.zeros(n, **i32)
        self.nodes_to_expand = torch.zeros(self.max_nodes, **i32)
        self.ws = torch.empty(nat.lib().h3dgs_cut_step_workspace_bytes(self.max_nodes), dtype=torch.uint8, device=dev)
        self.cur, self.count, self.rendered, self.need_expansion = 0, 1, 0, 0   # the list starts as {root}

    def step(self, target_size, viewpoint):
        """One refinement step; ``viewpoint``: CUDA float tensor [3].  -> (success, node count, render entries)."""
        vp = nat.as_f32(viewpoint, "viewpoint")
        out = (ctypes.c_int32 * 4)()
        src, dst = self.node_indices[self.cur], self.node_indices[self.cur ^ 1]
        nat.check(nat.lib().h3dgs_cut_step(self.count, nat.ptr(src), nat.ptr(dst), self.max_nodes, nat.ptr(self.nodes),
                                           nat.ptr(self.boxes), nat.ptr(vp), float(target_size), nat.ptr(self.split),
                                           nat.ptr(self.render_indices), nat.ptr(self.parent_indices),
                                           nat.ptr(self.nodes_of_render_indices), nat.ptr(self.nodes_to_expand),
                                           nat.ptr(self.ws), self.ws.numel(), out, nat.stream_handle()), "cut_step")
        self.need_expansion = int(out[3])
        if out[0]:
            self.cur ^= 1
            self.count, self.rendered = int(out[1]), int(out[2])
        return bool(out[0]), self.count, self.rendered

    @property
    def current_nodes(self):
        return self.node_indices[self.cur][:self.count]
