"""Shared body of the .hier-file parity tests: `gaussian_hierarchy._C.write_hierarchy / load_hierarchy` against
tests/golden/hier_l6.npz, which holds files written and read back by the REFERENCE's own HierarchyWriter /
HierarchyLoader (tests/golden/make_golden_hier.py -> oracle/_ref/ref_gh_tool)."""
import os

import numpy as np
import pytest
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hier_l6.npz")
ORDER = ("means3D", "shs", "opacities", "log_scales", "rotations", "nodes", "boxes")


def check(tmp_path, device):
    import gaussian_hierarchy
    import h3dgs_native as nat
    G = np.load(GOLD)
    t = {k: torch.from_numpy(G[f"in_{k}"]) for k in ORDER}
    t["shs"] = t["shs"].view(-1, 16, 3)
    t["boxes"] = t["boxes"].view(-1, 2, 4)
    args = [t[k].to(device) for k in ORDER]
    # the Python binding writes the compressed form (GH/torch/torch_interface.cpp:55-66 omits the flag)
    for name, kw in (("compressed", {}), ("plain", {"compressed": False})):
        path = str(tmp_path / f"{name}.hier")
        gaussian_hierarchy._C.write_hierarchy(path, *args, **kw)
        got = np.frombuffer(open(path, "rb").read(), np.uint8)
        assert got.size == G[f"{name}_bytes"].size, f"{name}: file size {got.size} != reference {G[f'{name}_bytes'].size}"
        assert np.array_equal(got, G[f"{name}_bytes"]), f"{name}: bytes differ from the reference writer's file"
        # and reading the REFERENCE's file gives what the reference loader gives
        ref_path = str(tmp_path / f"ref_{name}.hier")
        with open(ref_path, "wb") as fh:
            fh.write(G[f"{name}_bytes"].tobytes())
        pos, shs, alpha, scale, rot, nodes, boxes = gaussian_hierarchy._C.load_hierarchy(ref_path)
        for k, v in zip(ORDER, (pos, shs.reshape(-1, 48), alpha, scale, rot, nodes, boxes.reshape(-1, 8))):
            want = G[f"{name}_load_{k}"]
            assert not v.is_cuda and v.numpy().dtype == want.dtype
            assert np.array_equal(v.numpy().view(np.int32), want.view(np.int32)), f"{name}: loaded {k} differs"
    # counters above 32000 do not fit the fp16 container: the reference throws "Would lose information!"
    bad = [a.clone() for a in args]
    bad[5][9, 6] = 32001
    with pytest.raises(nat.H3dgsError, match="lose information"):
        gaussian_hierarchy._C.write_hierarchy(str(tmp_path / "bad.hier"), *bad)
