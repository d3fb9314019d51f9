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
    ap.add_argument("--dp", action="store_true", help="under torchrun: the data-parallel step, rank 0 prints its timeline")
    ap.add_argument("--dp-capacity", type=int, default=1_500_000)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--step-index", type=int, default=2)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if args.dp:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    W, H = 1920, 1080
    import gaussian_hierarchy
    gh = gaussian_hierarchy._C
    attrs = workload.build_scene(args.levels, args.skybox, 0, dev, free_corridor=True)
    n_store = attrs["means3D"].size(0)
    cams = workload.cameras(W, H, 4, dev, seed=0)
    bg = torch.zeros(3, device=dev)
    g = torch.Generator().manual_seed(1234)
    if args.e2e or args.dp:
        from scene.OurAdam import Adam
        model = workload.HierModel(attrs, args.skybox)
        opt = Adam(model.param_groups(), lr=0.0, eps=1e-15)
    else:
        model = workload.ActivatedModel(attrs, args.skybox)
    limit = workload.find_limit(gh, model, cams[0], args.target, dev)
    cut = workload.Cut(n_store, dev)
    cuts = []
    if not (args.e2e or args.dp):
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

    # the e2e step mirrors bench.py's: host pipeline from the package (h3dgs_pipeline: image prefetch, cut one step
    # ahead, late loss read-back), compact gradients, fused L1 + SSIM loss, Adam.step_sparse; with --dp (under torchrun)
    # the optimizer step is the peer-memory exchange of h3dgs_parallel.CompactRowExchange
    if args.e2e or args.dp:
        import diff_gaussian_rasterization as dgr
        import h3dgs_pipeline as pl
        from utils.loss_utils import fused_l1_ssim_loss
        pf = pl.HostPrefetcher({"gt": ((3, H, W), torch.float32)}, dev)
        ca = pl.CutAhead(gh, model.nodes, model.boxes, n_store, dev)
        rb = pl.LossReadback()
        ex = None
        if args.dp:
            import torch.distributed as dist
            import h3dgs_parallel as par
            ex = par.CompactRowExchange(opt, n_store, args.dp_capacity, group=dist.group.WORLD)

        def cam_of(i):
            return (i * world + rank) % len(cams) if args.dp else i % len(cams)

    def step(i):
        with torch.profiler.record_function(f"STEP{i}"):
            for p in model.parameters():
                p.grad = None
            if not (args.e2e or args.dp):
                out = workload.ours_render_post(cams[i % len(cams)], model, bg, cuts[i % len(cams)], do_depth=True)
                loss = (out["render"] * dcol).sum() + (out["depth"] * ddep).sum()
                loss.backward()
                return
            pf.fetch(i, {"gt": gt_host})
            pf.fetch(i + 1, {"gt": gt_host})
            ca.launch(i, cams[cam_of(i)], cam_host[cam_of(i)], limit)
            c, this_cut = ca.get(i)
            sg = dgr.SparseGradients()
            out = workload.ours_render_post(c, model, bg, this_cut, do_depth=False, sparse_grads=sg)
            loss = fused_l1_ssim_loss(out["render"], pf.get(i)["gt"], 0.2)[0]
            loss.backward()
            pf.done(i)
            ca.done(i)
            ca.launch(i + 1, cams[cam_of(i + 1)], cam_host[cam_of(i + 1)], limit)
            if ex is not None:
                ex.step(sg)
            else:
                opt.step_sparse(sg)
            rb.push(loss)
            rb.pop_ready(keep_newest=1)

    for i in range(4):
        step(i)
    torch.cuda.synchronize()
    base_i = 4
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for i in range(args.steps):
            step(base_i + i)
        torch.cuda.synchronize()
    if rank != 0:
        return 0
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
