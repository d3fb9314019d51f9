"""Diagnostic: run-to-run spread of the gradients with the side-stream zero fill on / off (same inputs)."""
import sys

import torch

from . import add_paths

add_paths()
from . import scenes, workload  # noqa: E402

DEV = torch.device("cuda", 0)


def run(mode, attrs, sky, cam, bg, dcol):
    import diff_gaussian_rasterization as dgr
    import gaussian_hierarchy
    dgr._C.PREZERO = mode
    n = attrs["means3D"].size(0)
    model = workload.ActivatedModel(attrs, sky)
    limit = workload.find_limit(gaussian_hierarchy._C, model, cam, 6000, DEV)
    cut = workload.Cut(n, DEV)
    cut.compute(gaussian_hierarchy._C, model, cam, limit)
    out = workload.ours_render_post(cam, model, bg, cut, do_depth=False)
    (out["render"] * dcol).sum().backward()
    torch.cuda.synchronize()
    return [p.grad.clone() for p in model.parameters()] + [out["viewspace_points"].grad.clone()]


def main():
    sky = 300
    attrs = workload.build_scene(14, sky, 4, DEV)
    cam = scenes.street_camera(800, 600, seed=2, device=DEV, yaw_deg=3.0, x=-5.0)
    bg = torch.zeros(3, device=DEV)
    g = torch.Generator().manual_seed(5)
    dcol = (torch.randn(3, 600, 800, generator=g) / 480000).to(DEV)
    base = run(False, attrs, sky, cam, bg, dcol)
    names = [f"p{i}" for i in range(len(base) - 1)] + ["vp"]
    for mode in (False, True):
        worst = [0.0] * len(base)
        where = [None] * len(base)
        for it in range(40):
            cur = run(mode, attrs, sky, cam, bg, dcol)
            for k, (a, b) in enumerate(zip(cur, base)):
                d = (a - b).abs()
                r = float(d.max()) / (float(b.abs().max()) + 1e-30)
                if r > worst[k]:
                    worst[k] = r
                    i = int(d.flatten().argmax())
                    where[k] = (it, i // max(1, d[0].numel()), float(a.flatten()[i]), float(b.flatten()[i]),
                                float(b.abs().max()), int((d > 1e-4 * b.abs().max()).sum()))
        print("mode", mode)
        for k in range(len(base)):
            print(f"  {names[k]:4s} shape {tuple(base[k].shape)} worst rel {worst[k]:.3e} at {where[k]}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
