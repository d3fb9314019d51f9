"""Diagnostic: GPU timeline (kernel / memcpy / memset start, duration, idle gap before) of ONE step of
bench.py's workload, from torch.profiler (Kineto/CUPTI; nsys is not in the image).
    python -m harness.timeline [--e2e] [--levels 23] [--step-index 2]
Prints every device activity of the chosen step in start order with the idle time that preceded it, and the
totals (busy, idle, span): the place to look for host-side bubbles between kernels."""
import argparse
import json
import os
import sys
import tempfile
from types import SimpleNamespace

import torch

from . import add_paths

add_paths()
from . import workload  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--levels", type=int, default=23)
    ap.add_argument("--target", type=int, default=5_000_000)
    ap.add_argument("--skybox", type=int, default=100_000)
    ap.add_argument("--e2e", action="store_true")
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--step-index", type=int, default=2)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    W, H = 1920, 1080
    import gaussian_hierarchy
    gh = gaussian_hierarchy._C
    attrs = workload.build_scene(args.levels, args.skybox, 0, dev, free_corridor=True)
    n_store = attrs["means3D"].size(0)
    cams = workload.cameras(W, H, 4, dev, seed=0)
    bg = torch.zeros(3, device=dev)
    g = torch.Generator().manual_seed(1234)
    if args.e2e:
        from scene.OurAdam import Adam
        model = workload.HierModel(attrs, args.skybox)
        opt = Adam(model.param_groups(), lr=0.0, eps=1e-15)
    else:
        model = workload.ActivatedModel(attrs, args.skybox)
    limit = workload.find_limit(gh, model, cams[0], args.target, dev)
    cut = workload.Cut(n_store, dev)
    cuts = []
    if not args.e2e:
        for c in cams:
            k = workload.Cut(n_store, dev)
            k.compute(gh, model, c, limit)
            cuts.append(k)
    dcol = (torch.randn(3, H, W, generator=g) / (W * H)).to(dev)
    ddep = (torch.randn(1, H, W, generator=g) / (W * H)).to(dev)
    gt_host = torch.rand(3, H, W, generator=g).pin_memory()
    gt_dev = torch.empty(3, H, W, device=dev)
    cam_host = [torch.cat([c.world_view_transform.flatten().cpu(), c.full_proj_transform.flatten().cpu(),
                           c.camera_center.flatten().cpu()]).pin_memory() for c in cams]
    cam_dev = torch.empty(35, device=dev)

    # the e2e step mirrors bench.py's: fused L1 + SSIM loss, ground truth prefetched on a copy stream, the cut of
    # step i+1 launched on a side stream right after step i's backward
    if args.e2e:
        from utils.loss_utils import fused_l1_ssim_loss
        gt_dev2 = [torch.empty(3, H, W, device=dev) for _ in range(2)]
        cam_dev2 = [torch.empty(35, device=dev) for _ in range(2)]
        cuts2 = [workload.Cut(n_store, dev) for _ in range(2)]
        copy_stream, cut_stream = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        ready = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]
        cut_done = [torch.cuda.Event() for _ in range(2)]
        cut_free = [torch.cuda.Event() for _ in range(2)]
        for ev in consumed + cut_free:
            ev.record()
        state = {"gt": -1, "cut": -1}

        def view(k, buf):
            c0 = cams[k]
            return SimpleNamespace(image_width=W, image_height=H, FoVx=c0.FoVx, FoVy=c0.FoVy,
                                   world_view_transform=buf[:16].view(4, 4), full_proj_transform=buf[16:32].view(4, 4),
                                   camera_center=buf[32:35], image_name=c0.image_name)

        def prefetch(j):
            if state["gt"] >= j:
                return
            copy_stream.wait_event(consumed[j % 2])
            with torch.cuda.stream(copy_stream):
                gt_dev2[j % 2].copy_(gt_host, non_blocking=True)
                ready[j % 2].record(copy_stream)
            state["gt"] = j

        def launch_cut(j):
            if state["cut"] >= j:
                return
            k, b = j % len(cams), j % 2
            cut_stream.wait_event(cut_free[b])
            with torch.cuda.stream(cut_stream):
                cam_dev2[b].copy_(cam_host[k], non_blocking=True)
                cuts2[b].compute(gh, model, view(k, cam_dev2[b]), limit, campos_host=cam_host[k][32:35])
                cut_done[b].record(cut_stream)
            state["cut"] = j

    def step(i):
        with torch.profiler.record_function(f"STEP{i}"):
            c0 = cams[i % len(cams)]
            for p in model.parameters():
                p.grad = None
            if not args.e2e:
                out = workload.ours_render_post(c0, model, bg, cuts[i % len(cams)], do_depth=True)
                loss = (out["render"] * dcol).sum() + (out["depth"] * ddep).sum()
                loss.backward()
                return
            b = i % 2
            prefetch(i)
            prefetch(i + 1)
            launch_cut(i)
            torch.cuda.current_stream().wait_event(cut_done[b])
            out = workload.ours_render_post(view(i % len(cams), cam_dev2[b]), model, bg, cuts2[b], do_depth=False)
            torch.cuda.current_stream().wait_event(ready[b])
            loss = fused_l1_ssim_loss(out["render"], gt_dev2[b], 0.2)[0]
            loss.backward()
            consumed[b].record()
            cut_free[b].record()
            launch_cut(i + 1)
            relevant = (model._opacity.grad.flatten() != 0).nonzero().flatten().long()
            opt.step(relevant)
            loss.item()

    for i in range(4):
        step(i)
    torch.cuda.synchronize()
    base_i = 4
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for i in range(args.steps):
            step(base_i + i)
        torch.cuda.synchronize()
    path = os.path.join(tempfile.mkdtemp(), "trace.json")
    prof.export_chrome_trace(path)
    ev = json.load(open(path))["traceEvents"]
    steps = sorted([e for e in ev if e.get("name", "").startswith("STEP") and e.get("cat") == "user_annotation"],
                   key=lambda e: e["ts"])
    dev_ev = sorted([e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")], key=lambda e: e["ts"])
    # device work is attributed to a step by the CPU-side launch window via correlation ids
    launches = {e["args"].get("correlation"): e["ts"] for e in ev
                if e.get("cat") in ("cuda_runtime", "cuda_driver") and "args" in e and "correlation" in e["args"]}
    st = steps[args.step_index]
    t0, t1 = st["ts"], st["ts"] + st["dur"]
    mine = [e for e in dev_ev if t0 <= launches.get(e["args"].get("correlation"), -1) <= t1]
    print(f"# step {args.step_index}: host span {st['dur']:.0f} us; {len(mine)} device activities")
    print(f"# {'start':>8s} {'gap':>7s} {'dur':>8s} {'stream':>6s}  name")
    base = mine[0]["ts"]
    end = base
    busy = 0.0
    for e in mine:
        gap = e["ts"] - end
        print(f"  {e['ts'] - base:8.1f} {gap:7.1f} {e['dur']:8.1f} {str(e['args'].get('stream', '')):>6s}  {e['name'][:90]}")
        end = max(end, e["ts"] + e["dur"])
        busy += e["dur"]
    span = end - base
    print(f"# device span {span:.0f} us, sum of activity durations {busy:.0f} us (overlap counts twice)")
    # next step's first device activity - this step's first = the steady-state period
    if args.step_index + 1 < len(steps):
        nx = steps[args.step_index + 1]
        nmine = [e for e in dev_ev if nx["ts"] <= launches.get(e["args"].get("correlation"), -1) <= nx["ts"] + nx["dur"]]
        if nmine:
            print(f"# period to the next step's first activity: {nmine[0]['ts'] - base:.0f} us")
    return 0


if __name__ == "__main__":
    sys.exit(main())
