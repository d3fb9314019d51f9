"""Seeded synthetic "street corridor" scenes, street-view cameras and flat hierarchies.

Distributions follow SURVEY.md section 8(d): positions uniform in x in [-100,100] m (street axis),
y in [-12,12], z in [0,15] with 60 % of the points snapped to within 0.3 m of the planes y = +-10
(facades) and z = 0 (road); log-scales U(ln 0.01, ln 0.4); rotations normalised N(0,1)^4; raw
opacity N(0,1.5); SH dc N(0,0.5), rest N(0,0.05); cameras at 1.6-2.5 m height looking along the
cube-face yaws, FoVx 90 degrees, znear 0.01 / zfar 100, matrices built like scene/cameras.py:96-99.
"""
import math
from types import SimpleNamespace

import numpy as np
import torch


def street_gaussians(n, seed=0, device="cpu", extent=(100.0, 12.0, 15.0), sh_coeffs=16, free_corridor=False):
    """``free_corridor``: keep the drivable space empty, as a captured street is.  The 40 % of the points that
    are not snapped to a surface would otherwise fill the air around the cameras with up-to-0.4 m blobs, and a
    few thousand of them right in front of the lens occlude the whole scene (a render then back-propagates
    into < 1 % of its visible Gaussians).  With the flag those that fall between the facades are moved into
    the building slabs behind them (9.7 m <= |y| <= ``extent[1]``); x and z are kept."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    ex, ey, ez = extent
    pos = torch.rand((n, 3), generator=g) * torch.tensor([2 * ex, 2 * ey, ez]) - torch.tensor([ex, ey, 0.0])
    u = torch.rand((n,), generator=g)
    jitter = (torch.rand((n,), generator=g) - 0.5) * 0.6
    left = u < 0.2
    right = (u >= 0.2) & (u < 0.4)
    road = (u >= 0.4) & (u < 0.6)
    pos[left, 1] = -10.0 + jitter[left]
    pos[right, 1] = 10.0 + jitter[right]
    pos[road, 2] = jitter[road].abs()
    if free_corridor:
        air = (u >= 0.6) & (pos[:, 1].abs() < 9.7)
        side = torch.where(pos[:, 1] < 0, -1.0, 1.0)
        depth = (pos[:, 1].abs() / 9.7) * (ey - 9.7)          # uniform again inside the slab
        pos[air, 1] = (side * (9.7 + depth))[air]
    log_scales = torch.rand((n, 3), generator=g) * (math.log(0.4) - math.log(0.01)) + math.log(0.01)
    rot = torch.randn((n, 4), generator=g)
    rot = rot / rot.norm(dim=1, keepdim=True)
    raw_opacity = torch.randn((n, 1), generator=g) * 1.5
    sh = torch.randn((n, sh_coeffs, 3), generator=g) * 0.05
    sh[:, 0, :] = torch.randn((n, 3), generator=g) * 0.5
    out = dict(means3D=pos.float(), log_scales=log_scales.float(), scales=log_scales.exp().float(),
               rotations=rot.float(), raw_opacity=raw_opacity.float(),
               opacities=torch.sigmoid(raw_opacity).float(), shs=sh.float())
    return {k: v.to(device).contiguous() for k, v in out.items()}


def skybox_gaussians(n, seed=1, device="cpu", radius=1000.0, sh_coeffs=16):
    """Upper-hemisphere shell like scene/gaussian_model.py:186-201 (skybox points)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    theta = 2.0 * math.pi * torch.rand((n,), generator=g)
    phi = torch.arccos(1.0 - 1.4 * torch.rand((n,), generator=g))
    pos = torch.stack([radius * torch.cos(theta) * torch.sin(phi), radius * torch.sin(theta) * torch.sin(phi),
                       radius * torch.cos(phi)], dim=1)
    scale = torch.full((n, 3), radius * 2.2 / math.sqrt(max(n, 1)))
    rot = torch.zeros((n, 4)); rot[:, 0] = 1.0
    sh = torch.zeros((n, sh_coeffs, 3)); sh[:, 0, :] = torch.rand((n, 3), generator=g)
    out = dict(means3D=pos.float(), scales=scale.float(), log_scales=scale.log().float(), rotations=rot.float(),
               opacities=torch.full((n, 1), 0.7), raw_opacity=torch.full((n, 1), 0.8473), shs=sh.float())
    return {k: v.to(device).contiguous() for k, v in out.items()}


def _projection(znear, zfar, fovx, fovy, primx=0.5, primy=0.5):
    # utils/graphics_utils.py:51-77
    tan_y, tan_x = math.tan(fovy / 2), math.tan(fovx / 2)
    top = tan_y * znear
    bottom = (1 - primy) * 2 * -top
    top = primy * 2 * top
    right = tan_x * znear
    left = (1 - primx) * 2 * -right
    right = primx * 2 * right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def street_camera(width, height, seed=0, device="cpu", x=None, yaw_deg=None, height_m=None, pitch_deg=0.0,
                  fovx_deg=90.0):
    """A pinhole street-view camera; world up is +z, the street runs along x."""
    rng = np.random.RandomState(seed)
    if x is None:
        x = float(rng.uniform(-60, 60))
    if yaw_deg is None:
        yaw_deg = float(rng.choice([0.0, 90.0, 180.0, 270.0]) + rng.uniform(-10, 10))
    if height_m is None:
        height_m = float(rng.uniform(1.6, 2.5))
    y = float(rng.uniform(-3, 3))
    yaw, pitch = math.radians(yaw_deg), math.radians(pitch_deg)
    fwd = np.array([math.cos(yaw) * math.cos(pitch), math.sin(yaw) * math.cos(pitch), math.sin(pitch)])
    up = np.array([0.0, 0.0, 1.0])
    right = np.cross(fwd, up); right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    R_wc = np.stack([right, down, fwd], axis=0)            # world -> camera (x right, y down, z forward)
    C = np.array([x, y, height_m])
    W2C = np.eye(4); W2C[:3, :3] = R_wc; W2C[:3, 3] = -R_wc @ C
    fovx = math.radians(fovx_deg)
    fovy = 2.0 * math.atan(math.tan(fovx / 2) * height / width)
    view = torch.tensor(np.float32(W2C)).transpose(0, 1)
    proj = _projection(0.01, 100.0, fovx, fovy).transpose(0, 1)
    full = view.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0)
    center = view.inverse()[3, :3]
    return SimpleNamespace(image_width=width, image_height=height, FoVx=fovx, FoVy=fovy,
                           world_view_transform=view.to(device).contiguous(),
                           full_proj_transform=full.to(device).contiguous(),
                           camera_center=center.to(device).contiguous(), image_name=f"cam{seed}",
                           znear=0.01, zfar=100.0)


# ------------------------------------------------------------------------------------------------
# flat hierarchy (GH/types.h:51-60, GH/writer.cpp:19-84): complete binary KD-tree, heap-numbered nodes,
# Gaussians stored in DFS pre-order (one merged Gaussian per interior node, one leaf Gaussian per leaf)
# ------------------------------------------------------------------------------------------------
def _preorder_index(level, j, levels):
    """Pre-order rank of node j of `level` in a complete binary tree with `levels` levels below the root."""
    idx = torch.zeros_like(j)
    for l in range(level):
        bit = (j >> (level - 1 - l)) & 1          # l-th step from the root: 0 = left, 1 = right
        left_size = (1 << (levels - l)) - 1        # nodes in the left subtree hanging off depth l
        idx = idx + 1 + bit * left_size
    return idx


def build_hierarchy(levels, seed=0, device="cpu", extent=(100.0, 12.0, 15.0), leaf=None, free_corridor=False):
    """Balanced KD hierarchy over 2**levels leaves.

    Returns dict(means3D, log_scales, scales, rotations, opacities, shs) for all 2**(levels+1)-1
    Gaussians in DFS pre-order plus nodes int32 [N,7] and boxes fp32 [N,2,4].
    """
    n = 1 << levels
    if leaf is None:
        leaf = street_gaussians(n, seed=seed, device=device, extent=extent, free_corridor=free_corridor)
    pos = leaf["means3D"]
    # median splits along the longest axis, level by level (one global sort per level)
    order = torch.arange(n, device=device)
    for l in range(levels):
        seg = 1 << l
        size = n >> l
        p = pos[order].view(seg, size, 3)
        ext = p.max(dim=1).values - p.min(dim=1).values
        axis = ext.argmax(dim=1)                                        # [seg]
        coord = torch.gather(p, 2, axis.view(seg, 1, 1).expand(seg, size, 1)).squeeze(2)
        sub = coord.argsort(dim=1)
        order = torch.gather(order.view(seg, size), 1, sub).reshape(-1)
    leaf = {k: v[order] for k, v in leaf.items()}

    N = 2 * n - 1
    M = leaf["shs"].shape[1]
    gpos = torch.empty((N, 3), device=device); glog = torch.empty((N, 3), device=device)
    grot = torch.empty((N, 4), device=device); gop = torch.empty((N, 1), device=device)
    gsh = torch.empty((N, M, 3), device=device)
    nodes = torch.empty((N, 7), dtype=torch.int32, device=device)
    boxes = torch.zeros((N, 2, 4), device=device)

    # leaves (level == levels)
    j = torch.arange(n, device=device)
    nid = (n - 1) + j
    gid = _preorder_index(levels, j, levels)
    cur = dict(pos=leaf["means3D"], scale2=leaf["scales"] ** 2, rot=leaf["rotations"], op=leaf["opacities"],
               sh=leaf["shs"])
    ext3 = 3.0 * leaf["scales"].max(dim=1, keepdim=True).values
    cur_min, cur_max = leaf["means3D"] - ext3, leaf["means3D"] + ext3
    gpos[gid], glog[gid], grot[gid], gop[gid], gsh[gid] = cur["pos"], leaf["log_scales"], cur["rot"], cur["op"], cur["sh"]
    nodes[nid] = torch.stack([torch.zeros_like(j), (nid - 1) // 2, gid, torch.ones_like(j), torch.zeros_like(j),
                              torch.zeros_like(j), torch.zeros_like(j)], dim=1).int()
    boxes[nid, 0, :3], boxes[nid, 1, :3] = cur_min, cur_max            # leaf size (w) stays 0
    # interior levels, bottom-up: moment-matched merge of the two children
    for l in range(levels - 1, -1, -1):
        m = 1 << l
        j = torch.arange(m, device=device)
        nid = (m - 1) + j
        gid = _preorder_index(l, j, levels)
        w = cur["op"].view(m, 2, 1) + 1e-6
        wn = w / w.sum(dim=1, keepdim=True)
        p2 = cur["pos"].view(m, 2, 3)
        mean = (wn * p2).sum(dim=1)
        var = (wn * (cur["scale2"].view(m, 2, 3) + (p2 - mean.unsqueeze(1)) ** 2)).sum(dim=1)
        new = dict(pos=mean, scale2=var, rot=cur["rot"].view(m, 2, 4)[:, 0],
                   op=cur["op"].view(m, 2, 1).mean(dim=1).clamp(max=0.95),
                   sh=(wn.unsqueeze(3) * cur["sh"].view(m, 2, M, 3)).sum(dim=1))
        new_min = torch.minimum(cur_min.view(m, 2, 3)[:, 0], cur_min.view(m, 2, 3)[:, 1])
        new_max = torch.maximum(cur_max.view(m, 2, 3)[:, 0], cur_max.view(m, 2, 3)[:, 1])
        gpos[gid], glog[gid], grot[gid], gop[gid], gsh[gid] = new["pos"], 0.5 * var.log(), new["rot"], new["op"], new["sh"]
        depth = torch.full_like(j, levels - l)
        parent = torch.where(nid > 0, (nid - 1) // 2, torch.full_like(nid, -1))
        nodes[nid] = torch.stack([depth, parent, gid, torch.zeros_like(j), torch.ones_like(j), 2 * nid + 1,
                                  torch.full_like(j, 2)], dim=1).int()
        size = (new_max - new_min).max(dim=1).values
        boxes[nid, 0, :3], boxes[nid, 1, :3] = new_min, new_max
        boxes[nid, 0, 3], boxes[nid, 1, 3] = size, size
        cur, cur_min, cur_max = new, new_min, new_max
    return dict(means3D=gpos.contiguous(), log_scales=glog.contiguous(), scales=glog.exp().contiguous(),
                rotations=grot.contiguous(), opacities=gop.contiguous(), shs=gsh.contiguous(),
                nodes=nodes.contiguous(), boxes=boxes.contiguous())


class Pipe:
    convert_SHs_python = False
    compute_cov3D_python = False
    debug = False


class StoreModel:
    """Minimal stand-in for scene.GaussianModel: activated attribute tensors with requires_grad, the
    properties gaussian_renderer reads, hierarchy tensors and a skybox count."""

    def __init__(self, attrs, skybox_points=0, active_sh_degree=3, requires_grad=True):
        self._xyz = attrs["means3D"].clone().requires_grad_(requires_grad)
        self._scaling = attrs["scales"].clone().requires_grad_(requires_grad)
        self._rotation = attrs["rotations"].clone().requires_grad_(requires_grad)
        self._opacity = attrs["opacities"].clone().requires_grad_(requires_grad)
        self._features = attrs["shs"].clone().requires_grad_(requires_grad)
        self.skybox_points = skybox_points
        self.active_sh_degree = active_sh_degree
        self.max_sh_degree = 3
        self.pretrained_exposures = None
        self.nodes = attrs.get("nodes")
        self.boxes = attrs.get("boxes")

    get_xyz = property(lambda self: self._xyz)
    get_scaling = property(lambda self: self._scaling)
    get_rotation = property(lambda self: self._rotation)
    get_opacity = property(lambda self: self._opacity)
    get_features = property(lambda self: self._features)

    def parameters(self):
        return [self._xyz, self._features, self._opacity, self._scaling, self._rotation]

    def zero_grad(self):
        for p in self.parameters():
            p.grad = None
