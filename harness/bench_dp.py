"""Camera-batch data parallelism of BASELINE.json config 4 (SURVEY.md 8e mode B), measured inside bench.py whenever
it runs on more than one GPU: one chunk replicated on every rank, one camera per rank and step,
render_post(sparse_grads) forward + backward, gradients of the rows any rank touched summed over NVLink peer memory
(h3dgs_parallel.CompactRowExchange / csrc/dp_exchange.cu) and ONE sparse Adam step on every replica.

Also measured, so that the efficiency can be read off one line: the same step on one GPU alone (Adam.step_sparse on
the rank's own compact gradients, no exchange), and a check that the replicas are still bit-identical at the end."""
import hashlib

import torch
import torch.distributed as dist

from harness import workload


def _checksum(params):
    h = hashlib.sha256()
    for p in params:
        h.update(p.detach().view(torch.int32).sum(dtype=torch.int64).cpu().numpy().tobytes())
        h.update(p.detach().double().sum().cpu().numpy().tobytes())
    return h.hexdigest()


def run(args, dev, rank, world, attrs, timed, dcol, ddep):
    import diff_gaussian_rasterization as dgr
    import gaussian_hierarchy
    import h3dgs_parallel as par
    from scene.OurAdam import Adam
    gh = gaussian_hierarchy._C
    W, H = args.width, args.height
    # replicas start from rank 0's parameters, bit for bit
    for k in sorted(attrs):
        if torch.is_tensor(attrs[k]) and attrs[k].is_cuda:
            dist.broadcast(attrs[k], src=0)
    n_store = attrs["means3D"].size(0)
    n_cams = max(args.cameras, 2 * world)
    n_cams += (-n_cams) % world
    cams = workload.cameras(W, H, n_cams, dev, seed=0)
    bg = torch.zeros(3, device=dev)
    hm = workload.HierModel(attrs, args.skybox)
    opt = Adam(hm.param_groups(), lr=0.0, eps=1e-15)
    limit = workload.find_limit(gh, hm, cams[0], args.target, dev)
    cuts = []
    for c in cams:
        cut = workload.Cut(n_store, dev)
        cut.compute(gh, hm, c, limit)
        K = cut.count
        cut.render_indices = cut.render_indices[:K].clone()
        cut.parent_indices = cut.parent_indices[:K].clone()
        cut.interpolation_weights = cut.interpolation_weights[:K + args.skybox].clone()
        cut.num_siblings = cut.num_siblings[:K + args.skybox].clone()
        cut.nodes_for_render_indices = None
        cuts.append(cut)

    def render(k):
        for p in hm.parameters():
            p.grad = None
        sg = dgr.SparseGradients()
        out = workload.ours_render_post(cams[k], hm, bg, cuts[k], do_depth=True, sparse_grads=sg)
        ((out["render"] * dcol).sum() + (out["depth"] * ddep).sum()).backward()
        return sg

    # cost of every camera (ms of its forward + backward, warm) and the size of its gated gradient block
    costs, gated = [], []
    for k in range(n_cams):
        render(k)
        best = float("inf")
        for _ in range(2):   # the smaller of two: the first timed render of a process still grows the allocator's pools
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            sg = render(k)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        costs.append(best)
        U = int(sg.count)
        gated.append(int((sg.grads["opacity"][:U] != 0).sum()))
    t = torch.tensor(costs + [float(max(gated))], device=dev, dtype=torch.float64)
    dist.broadcast(t, src=0)           # every rank must build the same sampler
    costs = t[:-1].tolist()
    cap_t = torch.tensor([max(gated)], device=dev, dtype=torch.int64)
    dist.all_reduce(cap_t, op=dist.ReduceOp.MAX)
    capacity = int(cap_t.item() * 1.25) + 1024
    sampler = par.CostBucketSampler(costs, world, seed=0)
    ex = par.CompactRowExchange(opt, n_store, capacity, group=dist.group.WORLD,
                                shard_optimizer=None if args.dp_shard < 0 else bool(args.dp_shard))

    def single_step(i):   # what one GPU does alone with the same kernels: no exchange
        sg = render(i % n_cams)
        opt.step_sparse(sg)

    def dp_step(i):
        sg = render(sampler.camera(i, rank))
        ex.step(sg)

    steps, warm = args.steps, max(args.warmup, 3)
    ms_single, _ = timed(single_step, steps, warm)
    # the single-GPU steps moved each replica differently: start the DP run from identical replicas again
    with torch.no_grad():
        for p in hm.parameters():
            dist.broadcast(p, src=0)
        for st in opt.state.values():
            st["exp_avg"].zero_()
            st["exp_avg_sq"].zero_()
    ms_dp, clocks = timed(dp_step, steps, warm)
    torch.cuda.synchronize()
    ex.check_overflow()
    sums = [None] * world
    dist.all_gather_object(sums, _checksum(hm.parameters()))

    # where the exchange time goes: render / pack / barrier (includes waiting for the slowest rank) / accumulate per
    # peer / adam, CUDA events on one step of the heaviest and one of the lightest camera group
    def stage_times(step_index):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(7)]
        dist.barrier()
        torch.cuda.synchronize()
        ev[0].record()
        sg = render(sampler.camera(step_index, rank))
        ev[1].record()
        ex.pack(sg)
        ev[2].record()
        ex.arena.barrier()
        ev[3].record()
        ex.accumulate(ex.arena.base_ptrs)
        ev[4].record()
        ex.apply()
        ev[5].record()
        if ex.sharded:
            ex.pblock.barrier()
            ex.pull_params()
        ev[6].record()
        ex.parity ^= 1
        torch.cuda.synchronize()
        out = {n: ev[i].elapsed_time(ev[i + 1])
               for i, n in enumerate(("render", "pack", "barrier", "accumulate", "adam", "barrier2+pull_params"))}
        out["union_rows"] = int(ex.flags[0])
        out["cameras"] = sampler.batch(step_index)
        return out
    group_cost = [sum(costs[c] for c in sampler.groups[g]) for g in sampler.visit]
    heavy_step = max(range(len(group_cost)), key=lambda i: group_cost[i])
    light_step = min(range(len(group_cost)), key=lambda i: group_cost[i])
    stages = {"heaviest_group": stage_times(heavy_step), "lightest_group": stage_times(light_step)}
    union = stages["heaviest_group"]["union_rows"]

    single = 1000.0 * steps / ms_single
    value = world * 1000.0 * steps / ms_dp
    return {
        "value": value, "unit": "renders/s", "ms_per_step": ms_dp / steps, "n_gpus": world,
        "single_gpu_same_step": {"value": single, "ms_per_step": ms_single / steps,
                                 "what": "render_post(sparse_grads) fwd+bwd + Adam.step_sparse on one GPU alone, max over ranks"},
        "efficiency_vs_single": value / (world * single),
        "cameras": n_cams, "camera_cost_ms": costs,
        "sampler": "CostBucketSampler: cameras of similar cost share a step, rotated over the ranks",
        "exchange": "all-gather of gated compact gradient rows through symmetric (NVLink peer) memory, summed in rank "
                    "order while read, one sparse Adam over the union (csrc/dp_exchange.cu); NCCL only for setup",
        "optimizer": "sharded over rows ((row >> 6) % world == rank): each rank sums and updates the rows it owns and copies its "
                     "peers' updated rows; moments live on the owner" if ex.sharded else
                     "replicated: every rank sums and updates the whole union",
        "exchange_stages_ms": stages,
        "gated_rows_per_render_max": max(gated), "arena_capacity_rows": capacity, "union_rows_last_step": union,
        "nvlink_bytes_per_rank_per_step": (world - 1) * (sum(gated) / len(gated)) * (4 * ex.width + 4),
        "replicas_identical": len(set(sums)) == 1,
        "clocks": clocks,
    }
