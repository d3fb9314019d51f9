"""World-size-2 gloo tests (CPU) of the multi-GPU host logic: chunk assignment and the sparse-row
gradient all-reduce used by camera-batch data parallelism."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "street-sparse-3dgs_b200"))
    import h3dgs_parallel as hp
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(100 + rank)
        n = 1000
        # every rank touches its own (overlapping) set of rows, zero elsewhere -- like a cut's gradient
        mask = torch.zeros(n, dtype=torch.bool)
        mask[torch.randperm(n)[:300]] = True
        tensors = [torch.randn(n, w) * mask.unsqueeze(1) for w in (3, 48, 1, 4)]
        dense = [t.clone() for t in tensors]
        rows = hp.allreduce_sparse_rows(tensors, mask, average=True)
        hp.allreduce_dense(dense, average=True)
        gathered = [torch.zeros(n, dtype=torch.int32) for _ in range(world)]
        dist.all_gather(gathered, mask.to(torch.int32))
        union = torch.stack(gathered).max(dim=0).values.nonzero().flatten()
        ok = torch.equal(rows, union)
        for a, b in zip(tensors, dense):
            ok = ok and torch.allclose(a, b, rtol=0, atol=1e-7)
            ok = ok and float(a[~torch.stack(gathered).max(dim=0).values.bool()].abs().sum()) == 0.0
        names = [f"chunk{i}" for i in range(5)]
        mine = hp.chunks_for_rank(names, rank, world)
        os.makedirs(os.path.join(out_dir, "chunk1"), exist_ok=True)
        open(os.path.join(out_dir, "chunk1", "hierarchy.hier_opt"), "w").close()
        dist.barrier()
        mine_skip = hp.chunks_for_rank(names, rank, world, output_dir=out_dir, skip_if_exists=True)
        torch.save({"ok": ok, "mine": mine, "mine_skip": mine_skip, "rows": rows.numel()},
                   os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_sparse_allreduce_and_chunk_assignment_world2(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(str(tmp_path), f"r{r}.pt")) for r in range(world)]
    assert all(r["ok"] for r in res)
    assert res[0]["rows"] == res[1]["rows"] and 300 <= res[0]["rows"] <= 600
    assert res[0]["mine"] == ["chunk0", "chunk2", "chunk4"] and res[1]["mine"] == ["chunk1", "chunk3"]
    # chunk1 is finished: it is dropped before the round-robin
    assert sorted(res[0]["mine_skip"] + res[1]["mine_skip"]) == ["chunk0", "chunk2", "chunk3", "chunk4"]


def test_single_process_is_a_noop():
    import h3dgs_parallel as hp
    t = [torch.ones(4, 3)]
    hp.allreduce_dense(t)
    rows = hp.allreduce_sparse_rows(t, torch.tensor([True, False, True, False]))
    assert rows.tolist() == [0, 2] and torch.equal(t[0], torch.ones(4, 3))


def test_cost_bucket_sampler_groups_neighbours():
    import h3dgs_parallel as par
    costs = [5.0, 1.0, 5.2, 1.1, 4.9, 0.9, 5.1, 1.2]
    s = par.CostBucketSampler(costs, 2, seed=1)
    seen = []
    for step in range(8):
        b = s.batch(step)
        assert len(b) == 2 and abs(costs[b[0]] - costs[b[1]]) < 0.5, "a step mixes a heavy and a light camera"
        seen += b
    assert sorted(seen) == sorted(list(range(8)) * 2)
    assert {s.camera(0, 0), s.camera(0, 1)} == {s.camera(4, 0), s.camera(4, 1)} and s.camera(0, 0) == s.camera(4, 1)
