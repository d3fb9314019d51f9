"""Multi-GPU plumbing, one process per GPU over torch.distributed (NCCL on NVLink; gloo in CPU tests).

The reference has no in-process parallelism: chunks are separate OS processes launched by
scripts/full_train.py:171-232.  Two modes are provided:

* chunk-parallel  -- independent street chunks, one per rank, no collective on the data path
  (`chunks_for_rank`, with the reference's skip-if-exists behaviour);
* camera-batch data parallel inside one chunk -- every rank renders a different camera of the same
  model; gradients of the parameter rows touched by ANY rank's cut are summed with one all-reduce
  per tensor over the compacted union of rows (`allreduce_sparse_rows`), instead of all-reducing the
  whole [N, 59]-float store.
"""
import os
from typing import Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist


def world_info():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def chunks_for_rank(chunk_names: Sequence[str], rank: int, world: int, output_dir: Optional[str] = None,
                    done_marker: str = "hierarchy.hier_opt", skip_if_exists: bool = False) -> List[str]:
    """Static round-robin assignment of chunks to ranks; with `skip_if_exists` a chunk whose output
    already holds `done_marker` is dropped before the assignment (scripts/full_train.py:58-59,176,198-201)."""
    names = list(chunk_names)
    if skip_if_exists and output_dir is not None:
        names = [c for c in names if not os.path.exists(os.path.join(output_dir, c, done_marker))]
    return [c for i, c in enumerate(names) if i % world == rank]


def allreduce_dense(tensors: Iterable[torch.Tensor], average: bool = True, group=None):
    """Sum (or average) every tensor over the ranks: the plain data-parallel gradient exchange."""
    rank, world = world_info()
    if world == 1:
        return
    works = [dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group, async_op=True) for t in tensors]
    for w in works:
        w.wait()
    if average:
        for t in tensors:
            t.div_(world)


def union_rows(row_mask: torch.Tensor, group=None) -> torch.Tensor:
    """Row ids (int64, ascending) that are set in ANY rank's boolean mask (one byte per row on the wire)."""
    rank, world = world_info()
    m = row_mask.to(torch.uint8)
    if world > 1:
        dist.all_reduce(m, op=dist.ReduceOp.MAX, group=group)
    return m.nonzero().flatten()


def allreduce_sparse_rows(tensors: Sequence[torch.Tensor], row_mask: torch.Tensor, average: bool = True,
                          group=None) -> torch.Tensor:
    """All-reduce only the rows touched by some rank.

    `tensors` are per-row gradient tensors [N, ...] that are zero outside this rank's `row_mask`.
    Returns the union row ids (the `relevant` argument of the sparse Adam step).  Communication volume
    is |union| * row_bytes instead of N * row_bytes.
    """
    rank, world = world_info()
    rows = union_rows(row_mask, group=group)
    if world == 1 or rows.numel() == 0:
        return rows
    if all(t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() for t in tensors) and len(tensors) <= 8:
        # device path: ONE gather kernel into a packed [U, sum of widths] payload, ONE all-reduce, ONE scatter
        # kernel that also applies the 1/world average (h3dgs_pack_rows / h3dgs_unpack_rows)
        import h3dgs_native as nat
        L = nat.lib()
        arr = (nat.RowTensor * len(tensors))()
        width = 0
        for i, t in enumerate(tensors):
            arr[i].data = t.data_ptr()
            arr[i].width = t[0].numel()
            width += arr[i].width
        U = rows.numel()
        packed = torch.empty((U, width), dtype=torch.float32, device=rows.device)
        stream = nat.stream_handle()
        nat.check(L.h3dgs_pack_rows(arr, len(tensors), rows.data_ptr(), U, packed.data_ptr(), stream), "h3dgs_pack_rows")
        dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group)
        nat.check(L.h3dgs_unpack_rows(arr, len(tensors), rows.data_ptr(), U, packed.data_ptr(),
                                      1.0 / world if average else 1.0, stream), "h3dgs_unpack_rows")
        return rows
    # host tensors (the gloo tests of this logic)
    packed = [t.index_select(0, rows) for t in tensors]
    works = [dist.all_reduce(p, op=dist.ReduceOp.SUM, group=group, async_op=True) for p in packed]
    for w in works:
        w.wait()
    for t, p in zip(tensors, packed):
        if average:
            p.div_(world)
        t.index_copy_(0, rows, p)
    return rows


# ---------------------------------------------------------------------------------------------------------------
# Camera-batch data parallelism on compact gradients, exchanged through NVLink peer memory (csrc/dp_exchange.cu)
# ---------------------------------------------------------------------------------------------------------------

OWNER_SHIFT = 6   # H3DGS_DP_OWNER_SHIFT: runs of 64 consecutive rows share an owner


def owner_of(rows, world):
    """Rank that owns each store row under the sharded optimizer."""
    return (rows >> OWNER_SHIFT) % world


class RowArena:
    """One rank's exchange block: int32 header (count, overflow flag), int32 store rows and the packed
    [capacity, width] gradients, double buffered by step parity.  Allocated in torch symmetric memory when a process
    group is given (every peer then maps it and the accumulate kernels load from it directly over NVLink), as a plain
    CUDA tensor otherwise (world 1, and the single-process tests that stand several "ranks" side by side)."""

    HEADER = 64   # int32 words in front of each buffer (count at 0, overflow at 1; the rest keeps 256-byte alignment)

    def __init__(self, capacity, width, device, group=None):
        self.capacity, self.width = int(capacity), int(width)
        self.words = self.HEADER + self.capacity + self.capacity * self.width
        self.words += (-self.words) % 64
        self.handle = None
        if group is not None and dist.get_world_size(group) > 1:
            import torch.distributed._symmetric_memory as symm_mem
            self.buf = symm_mem.empty(2 * self.words, dtype=torch.int32, device=device)
            self.handle = symm_mem.rendezvous(self.buf, group)
            self.base_ptrs = [int(p) for p in self.handle.buffer_ptrs]
            self.rank = dist.get_rank(group)
        else:
            self.buf = torch.empty(2 * self.words, dtype=torch.int32, device=device)
            self.base_ptrs = [self.buf.data_ptr()]
            self.rank = 0
        self.buf.zero_()

    def ptrs(self, base, parity):
        """(count, overflow, rows, packed) device addresses of one buffer of the block that starts at `base`."""
        b = base + 4 * self.words * parity
        return b, b + 4, b + 4 * self.HEADER, b + 4 * (self.HEADER + self.capacity)

    def local(self, parity):
        return self.ptrs(self.base_ptrs[self.rank], parity)

    def barrier(self):
        if self.handle is not None:
            self.handle.barrier(channel=0)


class CompactRowExchange:
    """``step(sparse_grads, optimizer)``: this rank's compact gradients (``render_post(sparse_grads=...)``) are summed
    with every peer's over the union of touched rows, averaged, and applied by ONE sparse Adam step -- the result of
    ``all_reduce(grad) / world`` + ``optimizer.step(relevant)`` with ``relevant`` = rows whose opacity gradient is
    non-zero on some rank, bit-identical on every replica.  See csrc/dp_exchange.cu for the data path.

    ``shard_optimizer`` (default on): ZeRO-1 over rows.  The replicated form makes every rank sum and update
    the WHOLE union (3.6 M scattered 236-byte rows for 8 street cameras: 6.5 ms); sharded, rank r sums and updates
    only the rows with ``(row >> 6) % world == r`` and publishes the updated parameter rows in a second symmetric-memory block
    that the peers copy into their replicas.  Parameters stay bit-identical on all ranks; the Adam moments of a row
    live on its owner only (``gather_moments()`` rebuilds full moment tensors, e.g. for a checkpoint)."""

    def __init__(self, optimizer, store_rows, capacity, group=None, union_capacity=None, shard_optimizer=None):
        import h3dgs_native as nat
        self.nat = nat
        self.opt = optimizer
        params = [p for g in optimizer.param_groups for p in g["params"]]
        if not params or any(p.size(0) != store_rows for p in params):
            raise ValueError("every parameter tensor must have `store_rows` rows")
        self.params = params
        self.widths = [p.numel() // store_rows for p in params]
        self.width = sum(self.widths)
        self.dev = params[0].device
        self.group = group
        self.world = dist.get_world_size(group) if (group is not None or dist.is_initialized()) else 1
        if self.world > 1 and group is None:
            group = dist.group.WORLD
        self.arena = RowArena(capacity, self.width, self.dev, group if self.world > 1 else None)
        self.sharded = bool(shard_optimizer if shard_optimizer is not None else True) and self.world > 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        if union_capacity is not None:
            self.ucap = int(union_capacity)
        elif self.sharded:
            self.ucap = int(min(store_rows // self.world + 1, capacity * self.world))
        else:
            self.ucap = int(min(store_rows, capacity * self.world))
        # sharded: the block of updated parameter rows this rank publishes (one buffer: two barriers fence it)
        self.pblock = RowArena(self.ucap, self.width, self.dev, group) if self.sharded else None
        self.row_map = torch.full((store_rows,), -1, dtype=torch.int32, device=self.dev)
        self.union_rows = torch.empty(self.ucap, dtype=torch.int32, device=self.dev)
        self.union_packed = torch.empty((self.ucap, self.width), dtype=torch.float32, device=self.dev)
        self.flags = torch.zeros(4, dtype=torch.int32, device=self.dev)   # [0] union count, [1] overflow
        self.parity = 0
        self.steps = 0

    def pack(self, sg, parity=None, gate="opacity"):
        """Stage 1: gated compact rows of `sg` -> this rank's arena."""
        nat = self.nat
        parity = self.parity if parity is None else parity
        by_ptr = {sg.params[k].data_ptr(): k for k in sg.grads if k in sg.params}
        arr = (nat.RowTensor * len(self.params))()
        for i, p in enumerate(self.params):
            name = by_ptr.get(p.data_ptr())
            if name is None:
                raise RuntimeError("a parameter of the optimizer has no compact gradient in this SparseGradients")
            arr[i].data = sg.grads[name].data_ptr()
            arr[i].width = self.widths[i]
        gate_t = sg.grads.get(gate) if gate is not None else None
        cnt, ovf, rows, packed = self.arena.local(parity)
        nat.check(nat.lib().h3dgs_dp_pack(arr, len(self.params), sg.rows.data_ptr(), sg.count.data_ptr(), sg.capacity,
                                          nat.ptr(gate_t), rows, packed, cnt, self.arena.capacity, ovf,
                                          nat.stream_handle()), "h3dgs_dp_pack")

    def accumulate(self, bases, parity=None):
        """Stage 3: add the blocks that start at `bases` (one per rank, rank order) into the local union."""
        nat = self.nat
        parity = self.parity if parity is None else parity
        self.flags[0:1].zero_()   # the union count; the overflow flag is sticky
        for base in bases:
            cnt, _, rows, packed = self.arena.ptrs(base, parity)
            nat.check(nat.lib().h3dgs_dp_accumulate(rows, packed, cnt, self.arena.capacity, self.width,
                                                    self.world if self.sharded else 1, self.rank,
                                                    self.row_map.data_ptr(), self.union_rows.data_ptr(),
                                                    self.union_packed.data_ptr(), self.flags.data_ptr(), self.ucap,
                                                    self.flags.data_ptr() + 4, nat.stream_handle()),
                      "h3dgs_dp_accumulate")

    def apply(self, world=None):
        """Stage 4: averaged sparse Adam over the union."""
        out = None
        if self.sharded:
            cnt, _, rows, packed = self.pblock.local(0)
            out = (rows, packed, cnt)
        self.opt.step_union(self.params, self.union_rows, self.flags, self.union_packed,
                            1.0 / (world if world is not None else self.world), self.row_map, self.ucap, out=out)

    def pull_params(self):
        """Sharded, last stage: copy every peer's block of updated parameter rows into the local replica."""
        nat = self.nat
        arr = (nat.AdamTensor * len(self.params))()
        for i, p in enumerate(self.params):
            arr[i].param = p.data_ptr()
            arr[i].width = self.widths[i]
        # staggered: rank r starts with peer r + 1, so that at any moment every GPU's NVLink egress serves ONE reader
        # (all ranks reading peer 0, then peer 1, ... shares one GPU's 900 GB/s among world - 1 readers: measured
        # 2.7 ms instead of 0.6 ms for 7 x 105 MB on 8 GPUs); the blocks hold disjoint rows, the order is free
        world = len(self.pblock.base_ptrs)
        for k in range(1, world):
            r = (self.rank + k) % world
            cnt, _, rows, packed = self.pblock.ptrs(self.pblock.base_ptrs[r], 0)
            nat.check(nat.lib().h3dgs_dp_scatter_params(arr, len(self.params), rows, packed, cnt, self.pblock.capacity,
                                                        nat.stream_handle()), "h3dgs_dp_scatter_params")

    def gather_moments(self):
        """Sharded: full exp_avg / exp_avg_sq on every rank (sum over the owners; rows a rank does not own are zero
        there unless the optimizer was also stepped outside the exchange)."""
        if not self.sharded:
            return
        for p in self.params:
            st = self.opt.state.get(p)
            if st:
                dist.all_reduce(st["exp_avg"], group=self.group)
                dist.all_reduce(st["exp_avg_sq"], group=self.group)

    def step(self, sg, gate="opacity"):
        if sg is None or not sg.ready:
            raise RuntimeError("CompactRowExchange.step: the SparseGradients object has not been filled by a backward pass")
        self.pack(sg, gate=gate)
        self.arena.barrier()          # every peer's block of this parity is complete and visible
        self.accumulate(self.arena.base_ptrs)
        self.apply()
        if self.sharded:
            self.pblock.barrier()     # every owner's block of updated rows is complete
            self.pull_params()
        self.parity ^= 1              # a peer may still be reading this buffer: the next step packs into the other one
        self.steps += 1

    def check_overflow(self):
        """Host-side check (a sync): raises when an arena or the union was too small at some step."""
        ovf_local = int(self.arena.buf[1]) | int(self.arena.buf[self.arena.words + 1])
        if ovf_local or int(self.flags[1]):
            raise RuntimeError("CompactRowExchange: capacity exceeded (arena %d rows, union %d rows)"
                               % (self.arena.capacity, self.ucap))


class CostBucketSampler:
    """Camera order for synchronous data parallelism: a step lasts as long as its slowest rank, and street views differ
    2x in cost (looking along the street vs across it), so cameras are sorted by a cost estimate (tile instances of
    the previous epoch's render, any monotone proxy) and dealt out in groups of `world` NEIGHBOURS in that order; the
    groups are then visited in a seeded random order and rotated over the ranks.  `batch(step)` -> the camera index of
    every rank; identical on all ranks (no communication)."""

    def __init__(self, costs, world, seed=0):
        import random
        order = sorted(range(len(costs)), key=lambda i: (-float(costs[i]), i))
        self.world = world
        self.groups = [order[i:i + world] for i in range(0, len(order) - len(order) % world, world)]
        if not self.groups:   # fewer cameras than ranks: wrap around
            self.groups = [[order[i % len(order)] for i in range(world)]]
        self.visit = list(range(len(self.groups)))
        random.Random(seed).shuffle(self.visit)

    def batch(self, step):
        g = self.groups[self.visit[step % len(self.visit)]]
        r = (step // len(self.visit)) % self.world
        return g[r:] + g[:r]

    def camera(self, step, rank):
        return self.batch(step)[rank]
