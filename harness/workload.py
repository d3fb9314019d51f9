"""The benchmark workload (BASELINE.json config 4): hierarchy post-optimisation on a cut of ~5 M
Gaussians with interpolation weights and a skybox, 1920x1080 street-view cube-face cameras.

Both arms run the same step through their own public API:
    cut (expand_to_size + get_interpolation_weights) -> render_post -> L1 loss -> backward -> sparse Adam
"ours" = the drop-in packages of this repository; "reference" = the unmodified reference CUDA extensions
(oracle/_ref) under a restatement of the reference's Python glue (harness/refglue.py, harness/ref_adam.py).
"""
import math

import torch

from . import refglue, scenes


class HierModel:
    """Parameter store with the activations of scene/gaussian_model.py in hierarchy mode (:125-145,
    :411-412): exp scales, normalised rotations, abs opacities, dc/rest features concatenated."""

    def __init__(self, attrs, skybox_points):
        f = attrs["shs"]
        self._xyz = attrs["means3D"].clone().requires_grad_(True)
        self._features_dc = f[:, :1, :].clone().contiguous().requires_grad_(True)
        self._features_rest = f[:, 1:, :].clone().contiguous().requires_grad_(True)
        self._opacity = attrs["opacities"].clone().requires_grad_(True)
        self._scaling = attrs["log_scales"].clone().requires_grad_(True)
        self._rotation = attrs["rotations"].clone().requires_grad_(True)
        self.nodes = attrs["nodes"]
        self.boxes = attrs["boxes"]
        self.skybox_points = skybox_points
        self.active_sh_degree = 3
        self.max_sh_degree = 3
        self.pretrained_exposures = None
        # scene/gaussian_model.py:40-48 (setup_functions) and :411-412 (hierarchy mode: abs opacity)
        self.scaling_activation = torch.exp
        self.rotation_activation = torch.nn.functional.normalize
        self.opacity_activation = torch.abs

    get_xyz = property(lambda self: self._xyz)
    get_scaling = property(lambda self: self.scaling_activation(self._scaling))
    get_rotation = property(lambda self: self.rotation_activation(self._rotation))
    get_opacity = property(lambda self: self.opacity_activation(self._opacity))
    get_features = property(lambda self: torch.cat((self._features_dc, self._features_rest), dim=1))

    def param_groups(self, lr=1e-4):
        # scene/gaussian_model.py:286-296
        return [{"params": [self._xyz], "lr": lr, "name": "xyz"},
                {"params": [self._features_dc], "lr": 0.0025, "name": "f_dc"},
                {"params": [self._features_rest], "lr": 0.0025 / 20.0, "name": "f_rest"},
                {"params": [self._opacity], "lr": 0.05, "name": "opacity"},
                {"params": [self._scaling], "lr": 0.005, "name": "scaling"},
                {"params": [self._rotation], "lr": 0.001, "name": "rotation"}]

    def parameters(self):
        return [self._xyz, self._features_dc, self._features_rest, self._opacity, self._scaling, self._rotation]


class SingleModel:
    """Parameter store of single-chunk training (train_single.py; scene/gaussian_model.py:40-48,121-153,186-201):
    raw parameters with exp / normalize / SIGMOID activations, dc/rest features, `skybox_points` skybox rows at the
    FRONT of the store, one learnable 3x4 exposure per camera."""

    def __init__(self, attrs, skybox_points, camera_names=("cam0",)):
        f = attrs["shs"]
        self._xyz = attrs["means3D"].clone().requires_grad_(True)
        self._features_dc = f[:, :1, :].clone().contiguous().requires_grad_(True)
        self._features_rest = f[:, 1:, :].clone().contiguous().requires_grad_(True)
        self._opacity = torch.logit(attrs["opacities"].clamp(1e-4, 1 - 1e-4)).requires_grad_(True)
        self._scaling = attrs["log_scales"].clone().requires_grad_(True)
        self._rotation = attrs["rotations"].clone().requires_grad_(True)
        self.skybox_points = skybox_points
        self.active_sh_degree = 3
        self.max_sh_degree = 3
        self.pretrained_exposures = None
        self.scaling_activation = torch.exp
        self.rotation_activation = torch.nn.functional.normalize
        self.opacity_activation = torch.sigmoid
        dev = self._xyz.device
        self.exposure_mapping = {n: i for i, n in enumerate(camera_names)}
        self._exposure = torch.eye(3, 4, device=dev)[None].repeat(len(camera_names), 1, 1).requires_grad_(True)

    get_xyz = property(lambda self: self._xyz)
    get_scaling = property(lambda self: self.scaling_activation(self._scaling))
    get_rotation = property(lambda self: self.rotation_activation(self._rotation))
    get_opacity = property(lambda self: self.opacity_activation(self._opacity))
    get_features = property(lambda self: torch.cat((self._features_dc, self._features_rest), dim=1))

    def get_exposure_from_name(self, image_name):
        return self._exposure[self.exposure_mapping[image_name]]

    param_groups = HierModel.param_groups

    def parameters(self):
        return [self._xyz, self._features_dc, self._features_rest, self._opacity, self._scaling, self._rotation]


def single_chunk_scene(n, skybox, seed, device):
    """Config 2 scene: `skybox` skybox Gaussians FIRST (scene/gaussian_model.py:186-201), then n street Gaussians."""
    sc = scenes.street_gaussians(n, seed=seed, device=device, free_corridor=True)
    if skybox:
        sky = scenes.skybox_gaussians(skybox, seed=seed + 1, device=device)
        sc = {k: torch.cat((sky[k], sc[k])).contiguous() for k in sc}
    return sc


def ours_render(cam, model, bg, use_trained_exp=True):
    from gaussian_renderer import render
    return render(cam, model, scenes.Pipe, bg, use_trained_exp=use_trained_exp)


def ref_render(cam, model, bg, use_trained_exp=True):
    return refglue.ref_render(refglue.RefSettings, cam, model, bg, use_trained_exp=use_trained_exp)


class ActivatedModel:
    """Same store, attributes already activated (leaf tensors): isolates the rasterizer path."""

    def __init__(self, attrs, skybox_points):
        self._xyz = attrs["means3D"].clone().requires_grad_(True)
        self._f = attrs["shs"].clone().requires_grad_(True)
        self._o = attrs["opacities"].clone().requires_grad_(True)
        self._s = attrs["scales"].clone().requires_grad_(True)
        self._r = attrs["rotations"].clone().requires_grad_(True)
        self.nodes, self.boxes = attrs["nodes"], attrs["boxes"]
        self.skybox_points = skybox_points
        self.active_sh_degree = 3
        self.max_sh_degree = 3
        self.pretrained_exposures = None

    get_xyz = property(lambda self: self._xyz)
    get_scaling = property(lambda self: self._s)
    get_rotation = property(lambda self: self._r)
    get_opacity = property(lambda self: self._o)
    get_features = property(lambda self: self._f)

    def parameters(self):
        return [self._xyz, self._f, self._o, self._s, self._r]


def build_scene(levels, skybox, seed, device, free_corridor=False):
    """Hierarchy of 2**levels leaves with `skybox` skybox Gaussians appended at the END
    (scene/gaussian_model.py:389-401).  ``free_corridor`` (the benchmark scene): no Gaussians floating in the
    drivable space, see scenes.street_gaussians."""
    h = scenes.build_hierarchy(levels, seed=seed, device=device, free_corridor=free_corridor)
    if skybox:
        sky = scenes.skybox_gaussians(skybox, seed=seed + 1, device=device)
        for k in ("means3D", "log_scales", "scales", "rotations", "opacities", "shs"):
            h[k] = torch.cat((h[k], sky[k])).contiguous()
    return h


def cameras(width, height, count, device, seed=0):
    """Street-view cube faces: yaw cycles 0/90/180/270 (+ jitter) at positions along the street."""
    cams = []
    for i in range(count):
        yaw = [0.0, 90.0, 180.0, 270.0][i % 4] + ((i * 37) % 11 - 5)
        x = -40.0 + 80.0 * ((i * 0.6180339887) % 1.0)
        cams.append(scenes.street_camera(width, height, seed=seed + i, device=device, x=x, yaw_deg=yaw,
                                         height_m=1.6 + 0.9 * ((i * 0.37) % 1.0)))
    return cams


def threshold_for_tau(cam, tau):
    # render_hierarchy.py:60-61
    return (2.0 * (tau + 0.5)) * math.tan(cam.FoVx * 0.5) / (0.5 * cam.image_width)


class Cut:
    """Pre-allocated cut tensors (train_post.py:59-63) and the two cut calls (train_post.py:91-113)."""

    def __init__(self, n_store, device):
        self.render_indices = torch.zeros(n_store, dtype=torch.int32, device=device)
        self.parent_indices = torch.zeros(n_store, dtype=torch.int32, device=device)
        self.nodes_for_render_indices = torch.zeros(n_store, dtype=torch.int32, device=device)
        self.interpolation_weights = torch.zeros(n_store, dtype=torch.float32, device=device)
        self.num_siblings = torch.zeros(n_store, dtype=torch.int32, device=device)
        self.count = 0

    def compute(self, gh, model, cam, limit, campos_host=None):
        zeros3 = torch.zeros(3)
        self.count = gh.expand_to_size(model.nodes, model.boxes, limit, cam.camera_center, zeros3,
                                       self.render_indices, self.parent_indices, self.nodes_for_render_indices)
        host = cam.camera_center.cpu() if campos_host is None else campos_host
        gh.get_interpolation_weights(self.nodes_for_render_indices[:self.count].contiguous(), limit, model.nodes,
                                     model.boxes, host, zeros3, self.interpolation_weights, self.num_siblings)
        return self.count


def find_limit(gh, model, cam, target, device, iters=30):
    """Bisection on the granularity threshold so that the cut has ~target entries for `cam`."""
    cut = Cut(model.nodes.size(0) + 1, device)
    lo, hi = 1e-6, 10.0
    for _ in range(iters):
        mid = math.sqrt(lo * hi)
        n = cut.compute(gh, model, cam, mid)
        if n > target:
            lo = mid
        else:
            hi = mid
    return math.sqrt(lo * hi)


def ours_render_post(cam, model, bg, cut, do_depth=True, sparse_grads=None):
    from gaussian_renderer import render_post
    K = cut.count
    return render_post(cam, model, scenes.Pipe, bg, render_indices=cut.render_indices[:K],
                       parent_indices=cut.parent_indices[:K], interpolation_weights=cut.interpolation_weights,
                       num_node_kids=cut.num_siblings, use_trained_exp=False, do_depth=do_depth,
                       sparse_grads=sparse_grads)


def ref_render_post(cam, model, bg, cut, do_depth=True):
    K = cut.count
    return refglue.ref_render_post(refglue.RefSettings, cam, model, bg, cut.render_indices[:K],
                                   cut.parent_indices[:K], cut.interpolation_weights, cut.num_siblings,
                                   do_depth=do_depth)


def interpolated_inputs(model, cut, skybox):
    """The P-row tensors the reference's render_post hands to its rasterizer (gaussian_renderer/__init__.py:202-243):
    child/parent lerp of the cut rows + the skybox rows, detached -- the inputs of the rasterizer-only timing."""
    with torch.no_grad():
        K = cut.count
        ri, pi = cut.render_indices[:K].long(), cut.parent_indices[:K].long()
        t = cut.interpolation_weights[:K].unsqueeze(1)
        ti = 1 - t
        n = model.get_xyz.size(0)
        sky = torch.arange(n - skybox, n, device=ri.device)

        def lerp(x, tt=t, tti=ti):
            return torch.cat((tt * x[ri] + tti * x[pi], x[sky])).contiguous()

        rot = model.get_rotation
        par, ch = rot[pi], rot[ri]
        par[(ch * par).sum(dim=1) < 0] *= -1
        out = dict(means3D=lerp(model.get_xyz), scales=lerp(model.get_scaling), opacities=lerp(model.get_opacity),
                   rotations=torch.cat((t * ch + ti * par, rot[sky])).contiguous(),
                   shs=lerp(model.get_features, t.unsqueeze(2), ti.unsqueeze(2)))
        ts = cut.interpolation_weights[:K + skybox].clone()
        ts[K:] = 1.0
        kids = cut.num_siblings[:K + skybox].clone()
        kids[K:] = 1
        out["ts"], out["kids"] = ts, kids
    return out


def rasterizer_only(impl, cams, cuts, model, bg, skybox, dcol, ddep, reps=3):
    """BASELINE.md section 2 / north_star "reference CUDA rasterizer": plain rasterize_gaussians forward and backward
    on pre-interpolated P-row tensors (no gather / lerp glue, no N-row scatter), timed separately with CUDA events.
    impl: "ours" (diff_gaussian_rasterization of this repository) or "reference" (oracle/_ref)."""
    if impl == "ours":
        import diff_gaussian_rasterization as dgr
        Settings = dgr.GaussianRasterizationSettings

        def raster(rs, p, m2d):
            return dgr.GaussianRasterizer(rs)(means3D=p["means3D"], means2D=m2d, opacities=p["opacities"], shs=p["shs"],
                                              scales=p["scales"], rotations=p["rotations"])
    else:
        Settings = refglue.RefSettings

        def raster(rs, p, m2d):
            return refglue.ref_rasterize(rs, p["means3D"], m2d, p["opacities"], shs=p["shs"], scales=p["scales"],
                                         rotations=p["rotations"])
    fwd_ms, bwd_ms = [], []
    for cam, cut in zip(cams, cuts):
        inp = interpolated_inputs(model, cut, skybox)
        rs = refglue.make_settings(Settings, cam, bg, do_depth=True, interpolation_weights=inp["ts"],
                                   num_node_kids=inp["kids"])
        p = {k: inp[k].requires_grad_(True) for k in ("means3D", "scales", "opacities", "rotations", "shs")}
        m2d = torch.zeros_like(p["means3D"], requires_grad=True)
        for r in range(reps + 1):
            for v in list(p.values()) + [m2d]:
                v.grad = None
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record()
            color, radii, depth = raster(rs, p, m2d)
            e[1].record()
            torch.autograd.backward([color, depth], [dcol, ddep])
            e[2].record()
            torch.cuda.synchronize()
            if r > 0:   # first repetition warms the allocator
                fwd_ms.append(e[0].elapsed_time(e[1]))
                bwd_ms.append(e[1].elapsed_time(e[2]))
        del inp, p, m2d, color, radii, depth
    f, b = sum(fwd_ms) / len(fwd_ms), sum(bwd_ms) / len(bwd_ms)
    return {"fwd_ms": f, "bwd_ms": b, "fwd_bwd_per_s": 1000.0 / (f + b), "cameras": len(cams), "reps": reps,
            "what": "plain rasterize_gaussians forward / backward on the pre-interpolated P-row tensors of each camera's "
                    "cut (interpolation weights + kids given, depth on), CUDA events around each call"}
