"""The torch restatement of the reference's densification path (oracle/densify_reference.py) against golden vectors
produced by the reference's own GaussianModel methods (tests/golden/make_golden_densify.py)."""
import os

import numpy as np
import pytest
import torch

from harness import ROOT
from oracle import densify_reference as dr


def load_case(tag, device="cpu"):
    z = np.load(os.path.join(ROOT, "tests", "golden", "densify_ref.npz"))
    t = lambda k: torch.from_numpy(z[k]).to(device)
    has_state = f"{tag}_in_m_xyz" in z.files
    state = {"params": {k: t(f"{tag}_in_{k}") for k in dr.NAMES},
             "exp_avg": {k: t(f"{tag}_in_m_{k}") if has_state else None for k in dr.NAMES},
             "exp_avg_sq": {k: t(f"{tag}_in_v_{k}") if has_state else None for k in dr.NAMES},
             "xyz_gradient_accum": t(f"{tag}_in_accum"), "denom": t(f"{tag}_in_denom"), "max_radii2D": t(f"{tag}_in_radii")}
    scaffold = int(z[f"{tag}_in_scaffold"])
    extra = {"vis": t(f"{tag}_in_vis"), "vs_grad": t(f"{tag}_in_vs_grad"), "vis_radii": t(f"{tag}_in_vis_radii"),
             "scaffold": None if scaffold < 0 else scaffold, "samples": t(f"{tag}_samples"), "has_state": has_state}
    want = {k: t(k2) for k, k2 in ((k[len(tag) + 1:], k) for k in z.files if k.startswith(f"{tag}_out_") or k.startswith(f"{tag}_stats_"))}
    return state, extra, want


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_restatement_reproduces_the_reference_methods(tag):
    state, x, want = load_case(tag)
    dr.add_stats(state, x["vs_grad"], x["vis"], x["vis_radii"])
    assert torch.equal(state["xyz_gradient_accum"].nan_to_num(-7.0), want["stats_accum"].nan_to_num(-7.0))
    assert torch.equal(state["denom"], want["stats_denom"]) and torch.equal(state["max_radii2D"], want["stats_radii"])
    dr.densify_and_prune(state, 0.0002, 0.005, 4.0, 0.01, x["scaffold"], normal=lambda mean, std: x["samples"])
    for k in dr.NAMES:
        assert torch.equal(state["params"][k], want[f"out_{k}"]), k
        if x["has_state"]:
            assert torch.equal(state["exp_avg"][k], want[f"out_m_{k}"]) and torch.equal(state["exp_avg_sq"][k], want[f"out_v_{k}"])
    assert torch.equal(state["max_radii2D"], want["out_radii"]) and torch.equal(state["xyz_gradient_accum"], want["out_accum"])
    assert torch.equal(state["newly_split_points_mask"], want["out_newly_split"])
