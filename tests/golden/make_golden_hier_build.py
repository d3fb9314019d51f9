#!/usr/bin/env python
"""Golden vectors for the hierarchy build, made by the REFERENCE ITSELF: oracle/_ref/ref_gh_tool
`build` = the reference's PointbasedKdTreeGenerator + ClusterMerger + RotationAligner + Writer::makeHierarchy compiled
from /root/reference/submodules/gaussianhierarchy by oracle/build_ref.py.  Run where /root/reference exists:

    python tests/golden/make_golden_hier_build.py      -> tests/golden/hier_build_601.npz
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from harness import ref_creator  # noqa: E402
from oracle import build_ref  # noqa: E402
from tests.hier_build_util import synthetic_gaussians  # noqa: E402


def main():
    build_ref.build_tool(verbose=False)
    n = 601
    g = synthetic_gaussians(n, seed=20241)
    aligned, _ = ref_creator.run_build(**g, align=True)
    merged, _ = ref_creator.run_build(**g, align=False)
    for k in ("positions", "shs", "opacities", "nodes", "boxes"):
        assert np.array_equal(aligned[k], merged[k]), k     # the aligner touches rotations and scales only
    out = {"in_" + k: v for k, v in g.items()}
    out.update({"out_" + k: v for k, v in aligned.items()})
    out["merged_log_scales"], out["merged_rotations"] = merged["log_scales"], merged["rotations"]
    path = os.path.join(ROOT, "tests", "golden", "hier_build_601.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
