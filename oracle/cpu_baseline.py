"""TEST / BENCH INFRASTRUCTURE ONLY -- the timed CPU baseline of bench.py (`cpu_baseline` key).

The reference has no CPU rasterizer; the host baseline is the oracle port
(oracle/h3dgs_oracle.c, OpenMP over Gaussians / pixels) running the same step as the bench's `value`:
hierarchy cut -> render_post gather + lerp -> rasterizer forward + backward -> scatter to store rows,
on a BOUNDED sample of the workload: same image size and scene generator, a smaller hierarchy and cut.
"""
import math
import os
import time

import numpy as np


def run_bounded(levels=18, skybox=5000, target=250_000, width=1920, height=1080, min_seconds=8.0, max_reps=5):
    import torch
    from harness import scenes, workload
    from . import cpu_reference as cr

    cores = os.cpu_count() or 1
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    attrs = workload.build_scene(levels, skybox, 0, "cpu", free_corridor=True)
    cam = workload.cameras(width, height, 1, "cpu", seed=0)[0]
    a = {k: v.numpy() for k, v in attrs.items()}
    vp = cam.camera_center.numpy()
    view = cam.world_view_transform.numpy().ravel()
    proj = cam.full_proj_transform.numpy().ravel()
    bg = np.zeros(3, np.float32)
    tanx, tany = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
    # granularity threshold giving ~target entries (same bisection as harness.workload.find_limit)
    lo, hi = 1e-6, 10.0
    for _ in range(24):
        mid = math.sqrt(lo * hi)
        K = cr.expand_to_size(a["nodes"], a["boxes"], mid, vp)[0]
        lo, hi = (mid, hi) if K > target else (lo, mid)
    limit = math.sqrt(lo * hi)
    g = torch.Generator().manual_seed(1234)
    dcol = (torch.randn(3, height, width, generator=g) / (width * height)).numpy()
    ddep = (torch.randn(1, height, width, generator=g) / (width * height)).numpy()
    n_rows = a["means3D"].shape[0]

    def step():
        K, ri, pi, ni = cr.expand_to_size(a["nodes"], a["boxes"], limit, vp)
        ts, kids = cr.get_interpolation_weights(ni, limit, a["nodes"], a["boxes"], vp)
        at, w, sign = cr.render_post_gather(a["means3D"], a["shs"], a["opacities"], a["scales"], a["rotations"],
                                            ri, pi, ts, skybox)
        kids_full = np.concatenate([kids, np.ones(skybox, np.int32)])
        o = cr.rasterize_forward(at["means3D"], at["opacities"], at["scales"], at["rotations"], view, proj, vp, bg,
                                 width, height, tanx, tany, shs=at["shs"], ts=w, kids=kids_full, do_depth=True)
        gr = cr.rasterize_backward(o, dcol, ddep)
        cr.render_post_scatter(dict(means3D=gr["dL_dmeans3D"], shs=gr["dL_dsh"], opacities=gr["dL_dopacity"],
                                    scales=gr["dL_dscales"], rotations=gr["dL_drotations"]), n_rows, ri, pi, ts,
                               sign, skybox)
        return K, int((o["radii"] > 0).sum()), o["num_rendered"]

    K, V, R = step()  # warm-up (page faults, OpenMP pool)
    reps, t0 = 0, time.perf_counter()
    while reps < max_reps and (reps == 0 or time.perf_counter() - t0 < min_seconds):
        step()
        reps += 1
    dt = (time.perf_counter() - t0) / reps
    return {"value": 1.0 / dt, "unit": "renders/s", "cores": cores, "kind": "port",
            "sample": (f"oracle port (C + OpenMP, {cores} threads): cut + render_post fwd+bwd at {width}x{height} on a "
                       f"{K + skybox}-slot cut of a {n_rows}-row hierarchy (visible {V}, instances {R}); "
                       f"{reps} reps, {dt:.2f} s each.  The bench workload has ~5.1 M slots: about "
                       f"{5.1e6 / max(K + skybox, 1):.0f}x the slots of this sample"),
            "sample_slots": int(K + skybox), "sample_visible": int(V), "sample_instances": int(R)}
