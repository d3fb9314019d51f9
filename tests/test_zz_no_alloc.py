"""h3dgs_expand_to_size_static (include/h3dgs_b200.h): the hierarchy cut with its scratch in a static __device__
array of the library -- no workspace argument, no allocation inside the call.  No counterpart in the reference; the
check is against the existing entry point h3dgs_expand_to_size on the same inputs (integer outputs: identical), on
two calls in a row that share the array (a larger cut first, so that the second call meets stale contents).
Named test_zz_* so that pytest collects it last."""
import ctypes

import pytest
import torch

from harness import add_paths, workload

add_paths()

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


def _cut(L, nat, static, nodes, boxes, n, viewpoint, size, out):
    """One call through the C ABI; every buffer is allocated by the caller beforehand."""
    count = ctypes.c_int32(-1)
    stream = nat.stream_handle()
    if static:
        rc = L.h3dgs_expand_to_size_static(n, float(size), nat.ptr(nodes), nat.ptr(boxes), nat.ptr(viewpoint),
                                           nat.ptr(out[0]), nat.ptr(out[1]), nat.ptr(out[2]),
                                           ctypes.byref(count), stream)
    else:
        ws = out[3]
        rc = L.h3dgs_expand_to_size(n, float(size), nat.ptr(nodes), nat.ptr(boxes), nat.ptr(viewpoint),
                                    nat.ptr(out[0]), nat.ptr(out[1]), nat.ptr(out[2]),
                                    nat.ptr(ws), ws.numel(), ctypes.byref(count), stream)
    nat.check(rc, "expand_to_size")
    return count.value


def test_static_scratch_cut_matches_the_workspace_cut_twice_in_a_row():
    import h3dgs_native as nat
    L = nat.lib()
    assert L.h3dgs_expand_to_size_static_max_nodes() == 1 << 20
    attrs = workload.build_scene(15, 500, 2, DEV)
    nodes, boxes = attrs["nodes"].contiguous(), attrs["boxes"].contiguous()
    n = nodes.size(0)
    assert 0 < n <= 1 << 20

    def buffers(with_ws):
        b = [torch.full((n,), -7, dtype=torch.int32, device=DEV) for _ in range(3)]
        if with_ws:
            b.append(torch.empty(L.h3dgs_expand_workspace_bytes(n), dtype=torch.uint8, device=DEV))
        return b

    # a fine cut first (many entries), then a coarse one from another viewpoint: the second call finds the first
    # call's counts, offsets and look-back words in the shared array
    calls = [(torch.tensor([0.0, 1.5, 0.0], device=DEV), 0.002),
             (torch.tensor([25.0, 1.5, -3.0], device=DEV), 0.05)]
    counts = []
    for viewpoint, size in calls:
        want, got = buffers(True), buffers(False)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(DEV)
        k_got = _cut(L, nat, True, nodes, boxes, n, viewpoint, size, got)
        assert torch.cuda.memory_allocated(DEV) == before, "the static entry point went through torch's allocator"
        k_want = _cut(L, nat, False, nodes, boxes, n, viewpoint, size, want)
        torch.cuda.synchronize()
        assert k_got == k_want and k_got > 0
        for a, b in zip(got[:3], want[:3]):
            assert torch.equal(a, b)          # integer outputs: identical, including the untouched tail
        counts.append(k_got)
    assert counts[0] > counts[1], counts      # the two calls really were different cuts


def test_static_scratch_cut_rejects_sizes_beyond_its_maximum():
    import h3dgs_native as nat
    L = nat.lib()
    n = (1 << 20) + 1
    dummy = torch.zeros(64, dtype=torch.int32, device=DEV)
    count = ctypes.c_int32(-1)
    rc = L.h3dgs_expand_to_size_static(n, 0.01, nat.ptr(dummy), nat.ptr(dummy), nat.ptr(dummy), nat.ptr(dummy),
                                       None, None, ctypes.byref(count), nat.stream_handle())
    assert rc != 0 and count.value == 0
    assert b"H3DGS_STATIC_CUT_MAX_NODES" in L.h3dgs_last_error()
    with pytest.raises(Exception):
        nat.check(rc, "expand_to_size_static")
