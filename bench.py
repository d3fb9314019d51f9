#!/usr/bin/env python
"""Headline benchmark: fwd+bwd renders/sec at 1920x1080 on a ~5 M-Gaussian hierarchy cut (BASELINE.json
metric, config 4: interpolation weights + skybox), one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

`value`  : render_post forward + backward per second with everything resident in HBM (cut tensors
           precomputed per camera, fixed upstream gradient), CUDA-event timed, max over ranks.
`e2e`    : the train_post iteration body through the public Python API with HOST inputs each step: camera
           matrices + ground-truth image copied from pinned memory, cut (expand_to_size +
           get_interpolation_weights), render_post, L1 loss, backward, sparse Adam, loss read back.
`--impl reference` runs the same two measurements on the UNMODIFIED reference CUDA extensions built for
sm_100a under oracle/_ref (the reference has no CPU rasterizer), plus the CPU oracle as
`cpu_baseline`.  N > 1 is weak scaling: every rank owns an independent street chunk (no collective on the
data path); on more than one GPU the line also carries `dp`: one chunk replicated, one camera per rank, gradients
exchanged through NVLink peer memory (harness/bench_dp.py); `--mode dp` makes that the headline value.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from harness import add_paths  # noqa: E402

add_paths()


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=4, choices=[1, 2, 3, 4],
                    help="BASELINE.json config: 4 = train_post step (the headline), 2 = train_single step, "
                         "3 = render_hierarchy forward-only over a batch of views")
    ap.add_argument("--mode", default="chunks", choices=["chunks", "dp"])
    ap.add_argument("--levels", type=int, default=23, help="hierarchy has 2**levels leaves")
    ap.add_argument("--target", type=int, default=5_000_000, help="cut size")
    ap.add_argument("--skybox", type=int, default=100_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--cameras", type=int, default=8)
    ap.add_argument("--loss", default="l1ssim", choices=["l1ssim", "l1"],
                    help="e2e loss: train_post.py's 0.8 L1 + 0.2 (1 - SSIM), or L1 alone")
    ap.add_argument("--build-gaussians", type=int, default=100_000, help="config 1: Gaussians in the chunk")
    ap.add_argument("--single-gaussians", type=int, default=1_000_000, help="config 2: Gaussians in the chunk")
    ap.add_argument("--views", type=int, default=64, help="config 3: views per batch")
    ap.add_argument("--tau", type=float, default=6.0, help="config 3: granularity in pixels")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--serial-cut", action="store_true",
                    help="e2e: compute each step's cut at the top of the step instead of one step ahead")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-raster-only", action="store_true")
    ap.add_argument("--e2e-dropin", action="store_true",
                    help="e2e (ours): time the unchanged train_post.py loop body (dense gradients) as the headline e2e")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--dp-shard", type=int, default=-1, help="dp: 1 / 0 = sharded / replicated optimizer; -1 = by world size")
    ap.add_argument("--no-dp", action="store_true", help="on more than one GPU: skip the camera-batch data-parallel measurement")
    return ap.parse_args()


from harness.clocks import ClockSampler  # noqa: E402


def main():
    args = parse()
    if args.config != 4:
        from harness import bench_configs
        return bench_configs.run(args)
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device: the product path has no CPU fallback"}))
        return 1
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    from harness import refglue, workload
    if args.impl == "ours":
        import gaussian_hierarchy
        gh = gaussian_hierarchy._C
        render_post = workload.ours_render_post
    else:
        gh = refglue.ref_gh()
        if gh is None or refglue.ref_hr() is None:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/*.so not built (run oracle/build_ref.py where /root/reference exists)"}))
            return 0
        render_post = workload.ref_render_post

    W, H = args.width, args.height
    seed = rank
    attrs = workload.build_scene(args.levels, args.skybox, seed, dev, free_corridor=True)
    n_store = attrs["means3D"].size(0)
    cams = workload.cameras(W, H, args.cameras, dev, seed=seed * 100)
    bg = torch.zeros(3, device=dev)

    # ---------------- device-resident measurement ----------------
    model = workload.ActivatedModel(attrs, args.skybox)
    limit = workload.find_limit(gh, model, cams[0], args.target, dev)
    cuts = []
    for c in cams:
        cut = workload.Cut(n_store, dev)
        cut.compute(gh, model, c, limit)
        # keep only what a render needs, trimmed to the cut (saves memory with many cameras)
        K = cut.count
        cut.render_indices = cut.render_indices[:K].clone()
        cut.parent_indices = cut.parent_indices[:K].clone()
        cut.interpolation_weights = cut.interpolation_weights[:K + args.skybox].clone()
        cut.num_siblings = cut.num_siblings[:K + args.skybox].clone()
        cut.nodes_for_render_indices = None
        cuts.append(cut)
    g = torch.Generator(device="cpu").manual_seed(1234)
    dcol = (torch.randn(3, H, W, generator=g) / (W * H)).to(dev)
    ddep = (torch.randn(1, H, W, generator=g) / (W * H)).to(dev)

    stats = {}

    # camera-batch data parallelism is measured by harness/bench_dp.py after the chunk-parallel numbers (every run on
    # more than one GPU carries both); `--mode dp` makes it the line's headline value

    # ours: compact gradients (render_post(sparse_grads=...)): the per-row gradients of a render are produced for the
    # rows of its cut only; `value_dropin` below is the same step with the dense N-row .grad tensors of the drop-in API
    res_mode = {"sparse": args.impl == "ours"}
    if args.impl == "ours":
        import diff_gaussian_rasterization as dgr_mod

    def resident_step(i, collect=False):
        # camera-batch data parallel: every rank takes a different camera of the same chunk
        k = i % len(cams)
        c = cams[k]
        cut = cuts[k]
        for p in model.parameters():
            p.grad = None
        sg = dgr_mod.SparseGradients() if res_mode["sparse"] else None
        if sg is not None:
            out = render_post(c, model, bg, cut, do_depth=True, sparse_grads=sg)
        else:
            out = render_post(c, model, bg, cut, do_depth=True)
        loss = (out["render"] * dcol).sum() + (out["depth"] * ddep).sum()
        loss.backward()
        if collect and sg is not None:
            stats.setdefault("P", []).append(cut.count + args.skybox)
            stats.setdefault("V", []).append(int(out["visibility_filter"].sum()))
            U = int(sg.count.item())
            stats.setdefault("compact_rows", []).append(U)
            stats.setdefault("grad_rows", []).append(int((sg.grads["opacity"][:U] != 0).sum()))
            return
        if collect:
            stats.setdefault("P", []).append(cut.count + args.skybox)
            stats.setdefault("V", []).append(int(out["visibility_filter"].sum()))
            stats.setdefault("grad_rows", []).append(int((model.parameters()[2].grad.flatten() != 0).sum()))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def launch_count():
        if args.impl != "ours":
            return 0
        import h3dgs_native
        return h3dgs_native.lib().h3dgs_launch_count()

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        barrier()
        timed.launches = launch_count()
        sampler = ClockSampler(local_rank)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(warmup + i)
        e1.record()
        timed.launches = launch_count() - timed.launches
        barrier()
        clocks = sampler.stop()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, clocks

    measured = None
    if args.impl == "ours":
        import h3dgs_native as nat
        # FP32 / MUFU / issue ceilings of THIS device, from dependent-chain micro-kernels
        measured = nat.measure_peaks(reps=5)
    ms_total, clocks = timed(resident_step, args.steps, max(args.warmup, 3))
    gpu_launches = int(timed.launches) if args.impl == "ours" else None
    ms_per_step = ms_total / args.steps
    value = world * 1000.0 / ms_per_step if args.mode == "chunks" else 1000.0 / ms_per_step * world
    value_dropin = None
    if res_mode["sparse"]:
        res_mode["sparse"] = False
        ms_d, _ = timed(resident_step, args.steps, max(args.warmup, 3))
        value_dropin = {"value": world * 1000.0 / (ms_d / args.steps), "unit": "renders/s", "ms_per_step": ms_d / args.steps,
                        "what": "the same step with the dense N-row .grad tensors of the drop-in API (zero fill of "
                                "236 B x store rows per backward, overlapped with the forward blend)"}
        res_mode["sparse"] = True
        for p in model.parameters():
            p.grad = None
        torch.cuda.empty_cache()

    # workload statistics (untimed): P, V, R per camera
    for i in range(len(cams)):
        resident_step(i, collect=True)
    torch.cuda.synchronize()

    # ---------------- per-kernel breakdown + roofline of the dominant kernel ----------------
    roofline, kernels = None, None
    if args.impl == "ours" and not args.no_profile:
        nat.lib().h3dgs_profile_enable(1)
        nat.profile_read(reset=True)
        R_list = []
        import diff_gaussian_rasterization as dgr
        n_prof = len(cams)
        prezero_was, dgr._C.PREZERO = dgr._C.PREZERO, False   # no side-stream overlap while timing kernels one by one
        for i in range(n_prof):
            resident_step(i)
        torch.cuda.synchronize()
        dgr._C.PREZERO = prezero_was
        prof = nat.profile_read(reset=True)
        nat.lib().h3dgs_profile_enable(0)
        kernels = {k: {"ms_per_step": v[0] / n_prof, "launches_per_step": v[1] / n_prof}
                   for k, v in prof.items() if v[1] > 0}
        stats["kernel_ms_per_step_sum"] = sum(v["ms_per_step"] for v in kernels.values())
        top = max(kernels, key=lambda k: kernels[k]["ms_per_step"])
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
        stats["R"], stats["n_contrib"] = workload_R(cams, cuts, model, bg, args, render_post)
        Rm = sum(stats["R"]) / len(stats["R"])
        Pm = sum(stats["P"]) / len(stats["P"])
        Vm = sum(stats["V"]) / len(stats["V"])
        npix = W * H
        # algorithmic bytes per step of each kernel id (DESIGN.md "Kernels"): fp32, every array touched once
        # the blend kernels by SURVEY.md 8(d)'s contract: 52 B gathered per instance (this implementation stages a
        # 64-byte record + a 4-byte id, "own_layout_bytes" below), 28 / 24 B per pixel, 40 B of reduced gradient
        own_layout = {"blend_bwd": 64 * Rm + 4 * Rm + 24 * npix + 40 * Rm, "blend_fwd": 64 * Rm + 4 * Rm + 28 * npix}
        alg = {
            "blend_bwd": 52 * Rm + 24 * npix + 40 * Rm,
            "blend_fwd": 52 * Rm + 28 * npix,
            # two-level sort: 4 depth passes over V (key, slot) pairs + 2 tile passes over R (u16 tile, slot) pairs,
            # the last one writing the 64-bit key next to the slot id
            "sort_onesweep_pass": 64 * Vm + 32 * Rm,
            "emit_instances": 12 * Pm + 20 * Vm + 6 * Rm,
            "tile_ranges": 8 * Rm + 16 * ((W + 15) // 16) * ((H + 15) // 16),
            "preprocess_fwd": 64 * Pm + 2 * 236 * Vm + 107 * Vm,
            "preprocess_bwd": 48 * Vm + 2 * 236 * Vm + 3 * 236 * Vm + 12 * Pm,
            "scan": 8 * Pm + 8 * Vm,
            "memset": 236 * n_store,
        }
        if res_mode["sparse"]:   # compact gradients: no N-row fill; what is left are the per-slot accumulators and sort counters
            alg["memset"] = 48 * Pm
            alg["preprocess_bwd"] += 2 * 236 * avg_or(stats.get("compact_rows"), 0)
        # instrumented pass (untimed): how far every evaluated (pixel, entry) pair gets -> FLOP of the blend kernels
        # by the contract of SURVEY.md 8(d) (interp + depth variants): forward 11 / 30 / 32 / 49 FLOP for a pair
        # rejected at power > 0 / at alpha < 1/255 / at the T test / blended; backward 11 / 30 / 100
        cnt = torch.zeros(16 + 4 * ((W + 15) // 16) * ((H + 15) // 16), dtype=torch.int64, device=dev)
        nat.lib().h3dgs_blend_stats(cnt.data_ptr())
        for i in range(n_prof):
            resident_step(i)
        torch.cuda.synchronize()
        nat.lib().h3dgs_blend_stats(None)
        cs = [x / n_prof for x in cnt[:9].tolist()]
        flop = {"blend_fwd": 11 * cs[0] + 30 * cs[1] + 32 * cs[2] + 49 * cs[3],
                "blend_bwd": 11 * cs[5] + 30 * cs[6] + 100 * cs[7]}
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        fp32_peak = float(measured["fp32_tflops"])   # TFLOP/s, three-register FFMA chains filling every SM
        stats["pairs"] = {"fwd_rejected_power": cs[0], "fwd_rejected_alpha": cs[1], "fwd_saturating": cs[2],
                          "fwd_blended": cs[3], "fwd_lane_slots": cs[4], "bwd_rejected_power": cs[5],
                          "bwd_rejected_alpha": cs[6], "bwd_contributing": cs[7], "bwd_lane_slots": cs[8]}
        for k, v in kernels.items():
            if k in alg:
                v["hbm_gbs"] = alg[k] / (v["ms_per_step"] * 1e-3) / 1e9
                v["hbm_frac"] = v["hbm_gbs"] / peak_gbs
            if k in flop:
                v["fp32_tflops"] = flop[k] / (v["ms_per_step"] * 1e-3) / 1e12
                v["fp32_frac"] = v["fp32_tflops"] / fp32_peak
                v["bound"] = "issue slots / FP32 pipe (ncu: profiles/)"
                v["own_layout_bytes"] = own_layout[k]
        dur_ms = kernels[top]["ms_per_step"] / max(kernels[top]["launches_per_step"], 1)
        bytes_per_launch = alg.get(top)
        if bytes_per_launch is not None:
            ach = bytes_per_launch / (dur_ms * 1e-3) / 1e9
            roofline = {"kernel": top, "bound": "hbm", "achieved": ach, "peak": peak_gbs, "unit": "GB/s",
                        "frac": ach / peak_gbs, "traffic": None,
                        "algorithmic_bytes_model": "SURVEY.md 8(d): blend_bwd 52 R + 24 Npix + 40 R, blend_fwd 52 R + 28 Npix "
                                                   "(R = tile instances of the render); other kernels: DESIGN.md section 3",
                        "algorithmic_bytes": bytes_per_launch,
                        "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650",
                        "note": "the blend kernels are bound by issue slots / the FP32 pipe, not by HBM (ncu: issue "
                                "slots ~85 % busy, DRAM ~1 %): the HBM fraction is reported as the contract asks, the "
                                "FP32 figure next to it is the meaningful one"}
            try:   # DRAM traffic of the same kernel from the committed ncu --set full capture (per launch)
                tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
                roofline["traffic"] = tr[top]["dram_bytes_per_launch"]
                roofline["traffic_source"] = tr["_source"]
            except Exception:
                pass
            if top in flop:
                roofline["fp32"] = {"achieved": kernels[top]["fp32_tflops"], "peak": fp32_peak, "unit": "TFLOP/s",
                                    "frac": kernels[top]["fp32_frac"],
                                    "peak_source": "measured", "peak_how": "h3dgs_measure_peaks at bench start: 3-register "
                                    "FFMA dependent chains, 8 per thread, 8 x 256 threads per SM, best of 5",
                                    "datasheet_peak": sms * 128 * 2 * float(peaks.get("sm_max_mhz", 1965.0)) * 1e6 / 1e12,
                                    "ffma2_peak": measured["fp32x2_tflops"], "mufu_ex2_gops": measured["mufu_gops"],
                                    "issue_peak_ginst": measured["issue_ginst"]}

    # ---------------- rasterizer alone (north_star's ">= 3x the reference CUDA rasterizer") ----------------
    raster_only = None
    if not args.no_raster_only:
        n_ro = min(len(cams), 4)
        raster_only = workload.rasterizer_only(args.impl, cams[:n_ro], cuts[:n_ro], model, bg, args.skybox, dcol, ddep)
        torch.cuda.empty_cache()

    # ---------------- end-to-end measurement (host inputs every step) ----------------
    e2e, e2e_dropin = None, None
    if not args.no_e2e:
        del model
        torch.cuda.empty_cache()
        hm = workload.HierModel(attrs, args.skybox)
        if args.impl == "ours":
            from scene.OurAdam import Adam
            opt = Adam(hm.param_groups(), lr=0.0, eps=1e-15)
        else:
            from harness.ref_adam import RefSparseAdam
            opt = RefSparseAdam(hm.param_groups(), eps=1e-15)
        cut = workload.Cut(n_store, dev)
        gt_host = [torch.rand(3, H, W, generator=g).pin_memory() for _ in range(2)]
        cam_host = []
        for c in cams:
            buf = torch.cat([c.world_view_transform.flatten().cpu(), c.full_proj_transform.flatten().cpu(),
                             c.camera_center.flatten().cpu()]).pin_memory()
            cam_host.append(buf)
        cam_dev = torch.empty(35, device=dev)
        # host pipeline: the ground-truth image of step i+1 crosses PCIe on a copy stream while
        # step i computes; every step still issues exactly one 24.9 MB host->device copy inside the timed region
        import h3dgs_pipeline as pl   # street-sparse-3dgs_b200/h3dgs_pipeline.py: the host pipeline lives in the package
        pf = pl.HostPrefetcher({"gt": ((3, H, W), torch.float32)}, dev)
        pipe_state = {"n": 0, "primed": False}

        def prefetch(j):
            pf.fetch(j, {"gt": gt_host[j % 2]})
        from types import SimpleNamespace
        h2d = gt_host[0].numel() * 4 + 35 * 4
        losses = []

        # host pipeline, second part (ours only; the reference arm keeps its serial loop): the cut of step i+1
        # depends on the camera and the hierarchy's boxes, not on the parameters, so it is launched on a side
        # stream right after step i's backward and its bandwidth-bound kernels run under the issue-bound blend.
        # Every step still runs exactly one cut and copies its own camera from pinned host memory.
        cut_ahead = args.impl == "ours" and not args.serial_cut
        if cut_ahead:
            ca = pl.CutAhead(gh, hm.nodes, hm.boxes, n_store, dev)

        def cam_index(i):
            return i % len(cams)

        def cam_view(k, buf):
            return pl.camera_view(cams[k], buf)

        def launch_cut(i):
            k = cam_index(i)
            ca.launch(i, cams[k], cam_host[k], limit)

        # the loss of train_post.py:132-137: (1 - lambda_dssim) L1 + lambda_dssim (1 - SSIM), lambda_dssim = 0.2
        if args.loss == "l1":
            def loss_fn(image, gt):
                return torch.abs(image - gt).mean()
        elif args.impl == "ours":
            from utils.loss_utils import fused_l1_ssim_loss

            def loss_fn(image, gt):
                return fused_l1_ssim_loss(image, gt, 0.2)[0]
        else:
            from oracle import loss_reference   # the reference's torch formulas (utils/loss_utils.py:17-63), on the GPU

            def loss_fn(image, gt):
                return loss_reference.l1_ssim_loss(image, gt, 0.2)[0]

        # ours: compact gradients (render_post(sparse_grads=...) + Adam.step_sparse: no N-row gradient tensors, no
        # zero fill, no row scan) and the loss value read back through pinned memory one step late, so the host never
        # waits for the step it has just queued; --e2e-dropin times the unchanged train_post.py loop body instead
        # (dense .grad tensors, relevant = opacity.grad != 0, loss.item()).  The reference arm is always the latter.
        mode = {"sparse": args.impl == "ours"}
        readback = pl.LossReadback()
        if args.impl == "ours":
            import diff_gaussian_rasterization as dgr

        host_prof = {} if os.environ.get("H3DGS_HOST_PROFILE") else None
        import time as _time

        def _mark(name, t0):
            if host_prof is not None:
                host_prof[name] = host_prof.get(name, 0.0) + (_time.perf_counter() - t0)
            return _time.perf_counter()

        def e2e_step(i):
            t0 = _time.perf_counter()
            step_i = i
            k = cam_index(i)
            c0 = cams[k]
            i = k
            n = pipe_state["n"]
            pipe_state["n"] = n + 1
            if not pipe_state["primed"]:
                prefetch(n)
                pipe_state["primed"] = True
            prefetch(n + 1)
            if cut_ahead:
                launch_cut(step_i)            # a no-op when the previous step already launched it
                c, this_cut = ca.get(step_i)
            else:
                cam_dev.copy_(cam_host[i % len(cams)], non_blocking=True)
                c, this_cut = cam_view(k, cam_dev), cut
                cut.compute(gh, hm, c, limit, campos_host=cam_host[i % len(cams)][32:35])
            t0 = _mark("prefetch+cut_get", t0)
            for p in hm.parameters():
                p.grad = None
            sg = dgr.SparseGradients() if mode["sparse"] else None
            if sg is not None:
                out = render_post(c, hm, bg, this_cut, do_depth=False, sparse_grads=sg)
            else:
                out = render_post(c, hm, bg, this_cut, do_depth=False)
            t0 = _mark("render_post", t0)
            loss = loss_fn(out["render"], pf.get(n)["gt"])
            t0 = _mark("loss", t0)
            loss.backward()
            t0 = _mark("backward", t0)
            pf.done(n)
            if cut_ahead:
                ca.done(step_i)
                launch_cut(step_i + 1)
            t0 = _mark("cut_launch", t0)
            if sg is not None:
                opt.step_sparse(sg)
                t0 = _mark("adam", t0)
                # device -> host read of this step's loss: asynchronous copy into pinned memory, picked up one step
                # later (the last one after the timed region's final synchronize)
                readback.push(loss)
                losses.extend(readback.pop_ready(keep_newest=1))
                _mark("readback", t0)
                return
            relevant = (hm._opacity.grad.flatten() != 0).nonzero().flatten().long()
            opt.step(relevant)
            losses.append(loss.item())

        e2e_dropin = None
        if mode["sparse"]:
            mode["sparse"] = False
            ms_d, _ = timed(e2e_step, args.steps, max(args.warmup, 3))
            e2e_dropin = {"value": world * 1000.0 / (ms_d / args.steps), "unit": "renders/s",
                          "ms_per_step": ms_d / args.steps,
                          "step": "the unchanged train_post.py loop body on the drop-in packages: dense N-row .grad "
                                  "tensors, relevant = (opacity.grad != 0).nonzero(), Adam.step(relevant), loss.item()"}
            mode["sparse"] = not args.e2e_dropin
            for p in hm.parameters():
                p.grad = None
            torch.cuda.empty_cache()
        if host_prof is not None:
            host_prof.clear()
        ms_e2e, clocks_e2e = timed(e2e_step, args.steps, max(args.warmup, 3))
        if host_prof is not None:
            calls = args.steps + max(args.warmup, 3)
            print("host ms per e2e step: " + ", ".join(f"{k} {1000.0 * v / calls:.3f}" for k, v in host_prof.items()),
                  file=sys.stderr)
        e2e_value = world * 1000.0 / (ms_e2e / args.steps)
        e2e = {"value": e2e_value, "unit": "renders/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
               "ms_per_step": ms_e2e / args.steps, "clocks": clocks_e2e,
               "step": "cut + render_post + " + ("0.8 L1 + 0.2 (1 - SSIM)" if args.loss == "l1ssim" else "L1") +
                       " + backward + sparse Adam + loss read-back" +
                       (" (compact gradients: render_post(sparse_grads=...) + Adam.step_sparse; the loss value crosses "
                        "through pinned memory and is read one step late)" if mode["sparse"] else " (loss.item())") +
                       "; ground-truth image prefetched one step ahead on a copy stream (one 24.9 MB H2D per step)" +
                       ("; the cut of the next step runs on a side stream under this step's backward"
                        if cut_ahead else "")}

    # ---------------- camera-batch data parallelism (SURVEY.md 8e mode B) ----------------
    dp_line = None
    if world > 1 and args.impl == "ours" and not args.no_dp:
        from harness import bench_dp
        if not args.no_e2e:
            del hm, opt, cut
        model = None
        torch.cuda.empty_cache()
        dp_line = bench_dp.run(args, dev, rank, world, attrs, timed, dcol, ddep)

    # ---------------- CPU baseline (oracle port, bounded sample) ----------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            from oracle import cpu_baseline as cb
            cpu_baseline = cb.run_bounded()
        except Exception as ex:  # the oracle is test infrastructure; its absence must not break the bench
            cpu_baseline = {"error": f"{type(ex).__name__}: {ex}"}

    if rank == 0:
        def avg(x):
            return sum(x) / len(x) if x else None
        line = {
            "metric": "fwd+bwd renders/sec at 1920x1080, 5M-Gaussian cut",
            "value": value, "unit": "renders/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "impl": args.impl,
            "config": {"workload": "config4: render_post fwd+bwd on a hierarchy cut with interpolation weights + skybox",
                       "image": [W, H], "hierarchy_leaves": 1 << args.levels, "store_rows": n_store,
                       "cut_target": args.target, "skybox": args.skybox, "cameras": len(cams),
                       "scene": "street corridor, drivable space free of Gaussians (harness/scenes.py free_corridor)",
                       "l2": "inputs larger than L2 (parameter store %.1f GB, cut cycles over %d cameras)" %
                             (n_store * 236 / 1e9, len(cams)),
                       "multi_gpu": "independent chunk per rank, no collective" if args.mode == "chunks" else
                                    "one chunk replicated, one camera per rank, NCCL all-reduce of the union of touched gradient rows"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": gpu_launches, "roofline": roofline,
            "kernels": kernels, "cpu_baseline": cpu_baseline, "rasterizer_only": raster_only,
            "e2e_dropin": e2e_dropin, "value_dropin": value_dropin, "dp": dp_line,
            # run statistics of the workload (measured, not configuration): slots, visible slots, tile instances, ...
            "workload_stats": {
                "P_mean": avg(stats.get("P")), "V_mean": avg(stats.get("V")), "R_mean": avg(stats.get("R")),
                "mean_tile_list_length": (avg(stats["R"]) / (((W + 15) // 16) * ((H + 15) // 16))) if stats.get("R") else None,
                "mean_last_contributor_per_pixel": avg(stats.get("n_contrib")),
                "rows_with_gradient_mean": avg(stats.get("grad_rows")),
                "compact_rows_mean": avg(stats.get("compact_rows")),
                "pairs_per_render": stats.get("pairs")},
            "measured_peaks": measured,
        }
        if args.mode == "dp" and dp_line is not None:
            line["chunks"] = {"value": line["value"], "ms_per_step": line["ms_per_step"], "e2e": line["e2e"]}
            line["value"], line["ms_per_step"], line["e2e"] = dp_line["value"], dp_line["ms_per_step"], None
            line["scaling"] = "weak"
            line["config"]["multi_gpu"] = ("one chunk replicated, one camera per rank and step, gradients of the union of "
                                           "touched rows exchanged through NVLink peer memory (see `dp`)")
        if args.impl == "reference":
            line["reference_kind"] = "unmodified reference CUDA extensions (oracle/_ref, sm_100a) under a restatement of its Python glue"
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def avg_or(x, default):
    return sum(x) / len(x) if x else default


def workload_R(cams, cuts, model, bg, args, render_post):
    """num_rendered and the mean per-pixel depth of blending (position of the last contributing list entry)
    per camera, read from the forward pass (untimed)."""
    import torch
    import diff_gaussian_rasterization as dgr
    Rs, nc = [], []
    orig = dgr._C.rasterize_gaussians_ex

    def spy(*a, **k):
        out = orig(*a, **k)
        Rs.append(int(out[0]))
        dv = dgr._C.debug_view(out[2].numel(), args.width, args.height, int(out[0]), out[3], out[4], out[5])
        nc.append(float(dv["n_contrib"].float().mean()))
        return out

    dgr._C.rasterize_gaussians_ex = spy
    try:
        with torch.no_grad():
            for c, cut in zip(cams, cuts):
                render_post(c, model, bg, cut, do_depth=True)
    finally:
        dgr._C.rasterize_gaussians_ex = orig
    return Rs, nc


if __name__ == "__main__":
    sys.exit(main())
