"""Camera-batch data parallelism over REAL peer memory, checked against the dense formulation it replaces.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        -m harness.dp_check [--steps 4]

Every rank holds the same model twice.  Copy A takes the product path: render_post(sparse_grads) on the rank's camera,
CompactRowExchange.step (pack -> symmetric-memory barrier -> accumulate every peer's block over NVLink in rank order ->
averaged sparse Adam).  Copy B takes the formulation of SURVEY.md 8e: the SAME compact gradients scattered into dense
N-row tensors, NCCL all-reduce(sum), / world, relevant = rows whose opacity gradient is non-zero on some rank,
Adam.step(relevant).  After every step the two copies must agree bit for bit (world a power of two), and all replicas
must be bit-identical with each other."""
import argparse
import hashlib
import json
import os

import torch
import torch.distributed as dist

from harness import add_paths, scenes, workload

add_paths()


def digest(params):
    h = hashlib.sha256()
    for p in params:
        h.update(p.detach().cpu().numpy().tobytes())
    return h.hexdigest()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--levels", type=int, default=17)
    ap.add_argument("--target", type=int, default=60000)
    ap.add_argument("--shard", type=int, default=-1, help="1 / 0: sharded / replicated optimizer; -1: the default for this world size")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    import diff_gaussian_rasterization as dgr
    import gaussian_hierarchy
    import h3dgs_parallel as par
    from scene.OurAdam import Adam
    gh = gaussian_hierarchy._C
    sky, W, H = 2000, 1024, 768
    attrs = workload.build_scene(args.levels, sky, 0, dev, free_corridor=True)
    for k in sorted(attrs):
        if torch.is_tensor(attrs[k]) and attrs[k].is_cuda:
            dist.broadcast(attrs[k], src=0)
    n = attrs["means3D"].size(0)
    A, B = workload.HierModel(attrs, sky), workload.HierModel(attrs, sky)
    optA = Adam(A.param_groups(), lr=0.0, eps=1e-15)
    optB = Adam(B.param_groups(), lr=0.0, eps=1e-15)
    cams = [scenes.street_camera(W, H, seed=40 + i, device=dev, yaw_deg=15.0 + 83.0 * i, x=-30.0 + 7.0 * i)
            for i in range(2 * world)]
    limit = workload.find_limit(gh, A, cams[0], args.target, dev)
    bg = torch.zeros(3, device=dev)
    g = torch.Generator().manual_seed(3)
    dcol = (torch.randn(3, H, W, generator=g) / (W * H)).to(dev)
    ex = par.CompactRowExchange(optA, n, capacity=400_000, group=dist.group.WORLD,
                                shard_optimizer=None if args.shard < 0 else bool(args.shard))
    names = ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
    report = {"world": world, "rows": n, "sharded_optimizer": ex.sharded, "steps": []}
    for step in range(args.steps):
        cam = cams[(step * world + rank) % len(cams)]
        cut = workload.Cut(n, dev)
        cut.compute(gh, A, cam, limit)
        sg = dgr.SparseGradients()
        out = workload.ours_render_post(cam, A, bg, cut, do_depth=False, sparse_grads=sg)
        (out["render"] * dcol).sum().backward()
        # copy B: dense tensors from the same compact gradients, NCCL all-reduce, dense-list Adam
        U = int(sg.count)
        gate = sg.grads["opacity"][:U].flatten() != 0
        rows = sg.rows[:U][gate]
        by_ptr = {p.data_ptr(): i for i, p in enumerate(A.parameters())}
        relevant = torch.zeros(n, dtype=torch.float32, device=dev)
        relevant[rows] = 1.0
        dense = [torch.zeros_like(p) for p in B.parameters()]
        for k in sg.grads:
            i = by_ptr[sg.params[k].data_ptr()]
            dense[i].index_add_(0, rows, sg.grads[k][:U][gate].reshape((rows.numel(),) + tuple(dense[i].shape[1:])))
        # sum in rank order, as the exchange does (NCCL's ring/tree order is its own: gather, then add 0, 1, 2, ...)
        for i, d in enumerate(dense):
            parts = [torch.empty_like(d) for _ in range(world)]
            dist.all_gather(parts, d)
            acc = parts[0].clone()
            for r in range(1, world):
                acc += parts[r]
            acc *= 1.0 / world
            B.parameters()[i].grad = acc
        dist.all_reduce(relevant, op=dist.ReduceOp.MAX)
        optB.step(relevant.nonzero().flatten())
        # copy A: the product path
        ex.step(sg)
        torch.cuda.synchronize()
        ex.check_overflow()
        same = all(torch.equal(a.detach().view(torch.int32), b.detach().view(torch.int32))
                   for a, b in zip(A.parameters(), B.parameters()))
        # sharded: a row's moments live on its owner
        own = (par.owner_of(torch.arange(n, device=dev), world) == rank) if ex.sharded else slice(None)
        mom = all(torch.equal(optA.state[a][key][own].view(torch.int32), optB.state[b][key][own].view(torch.int32))
                  for a, b in zip(A.parameters(), B.parameters()) for key in ("exp_avg", "exp_avg_sq"))
        sums = [None] * world
        dist.all_gather_object(sums, digest(A.parameters()))
        report["steps"].append({"step": step, "gated_rows": int(gate.sum()), "union_rows": int(ex.flags[0]),
                                "matches_dense_allreduce_adam": bool(same), "moments_match": bool(mom),
                                "replicas_identical": len(set(sums)) == 1})
        for p in A.parameters():
            p.grad = None
    ok = all(s["matches_dense_allreduce_adam"] and s["moments_match"] and s["replicas_identical"] for s in report["steps"])
    oks = [None] * world
    dist.all_gather_object(oks, ok)
    if rank == 0:
        report["ok"] = all(oks)
        print(json.dumps(report))
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    raise SystemExit(main())
