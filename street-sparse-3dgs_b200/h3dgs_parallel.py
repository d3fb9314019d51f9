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
