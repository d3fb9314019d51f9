"""Diagnostic: torch.profiler table of the end-to-end step of bench.py (cut + render_post + L1 + backward +
sparse Adam + loss.item()).   python -m harness.e2e_profile [--impl ours|reference] [--levels 23]"""
import argparse
import sys
from types import SimpleNamespace

import torch

from . import add_paths

add_paths()
from . import refglue, workload  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--levels", type=int, default=23)
    ap.add_argument("--target", type=int, default=5_000_000)
    ap.add_argument("--skybox", type=int, default=100_000)
    ap.add_argument("--rows", type=int, default=45)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    W, H = 1920, 1080
    if args.impl == "ours":
        import gaussian_hierarchy
        gh = gaussian_hierarchy._C
        render_post = workload.ours_render_post
        from scene.OurAdam import Adam
    else:
        gh = refglue.ref_gh()
        render_post = workload.ref_render_post
    attrs = workload.build_scene(args.levels, args.skybox, 0, dev, free_corridor=True)
    n_store = attrs["means3D"].size(0)
    cams = workload.cameras(W, H, 4, dev, seed=0)
    bg = torch.zeros(3, device=dev)
    hm = workload.HierModel(attrs, args.skybox)
    limit = workload.find_limit(gh, hm, cams[0], args.target, dev)
    if args.impl == "ours":
        opt = Adam(hm.param_groups(), lr=0.0, eps=1e-15)
    else:
        from harness.ref_adam import RefSparseAdam
        opt = RefSparseAdam(hm.param_groups(), eps=1e-15)
    cut = workload.Cut(n_store, dev)
    g = torch.Generator().manual_seed(1)
    gt_host = torch.rand(3, H, W, generator=g).pin_memory()
    gt_dev = torch.empty(3, H, W, device=dev)
    cam_host = [torch.cat([c.world_view_transform.flatten().cpu(), c.full_proj_transform.flatten().cpu(),
                           c.camera_center.flatten().cpu()]).pin_memory() for c in cams]
    cam_dev = torch.empty(35, device=dev)

    def step(i):
        c0 = cams[i % len(cams)]
        with torch.profiler.record_function("h2d"):
            cam_dev.copy_(cam_host[i % len(cams)], non_blocking=True)
            gt_dev.copy_(gt_host, non_blocking=True)
        c = SimpleNamespace(image_width=W, image_height=H, FoVx=c0.FoVx, FoVy=c0.FoVy,
                            world_view_transform=cam_dev[:16].view(4, 4),
                            full_proj_transform=cam_dev[16:32].view(4, 4), camera_center=cam_dev[32:35],
                            image_name=c0.image_name)
        with torch.profiler.record_function("cut"):
            cut.compute(gh, hm, c, limit, campos_host=cam_host[i % len(cams)][32:35])
        for p in hm.parameters():
            p.grad = None
        with torch.profiler.record_function("render_post"):
            out = render_post(c, hm, bg, cut, do_depth=False)
        with torch.profiler.record_function("loss"):
            loss = torch.abs(out["render"] - gt_dev).mean()
        with torch.profiler.record_function("backward"):
            loss.backward()
        with torch.profiler.record_function("adam"):
            relevant = (hm._opacity.grad.flatten() != 0).nonzero().flatten().long()
            opt.step(relevant)
        with torch.profiler.record_function("item"):
            return loss.item()

    for i in range(3):
        step(i)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    n = 4
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for i in range(n):
            step(3 + i)
        torch.cuda.synchronize()
    print(f"# {args.impl}: {n} e2e steps, store rows {n_store}, cut {cut.count}")
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=args.rows, max_name_column_width=70))


if __name__ == "__main__":
    sys.exit(main())
