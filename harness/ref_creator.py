"""TEST INFRASTRUCTURE ONLY -- drives oracle/_ref/ref_gh_tool `build`: the reference Creator's own tree stages
(PointbasedKdTreeGenerator, ClusterMerger, RotationAligner, Writer::makeHierarchy) compiled from the reference
sources by oracle/build_ref.py.  Used to make the golden fixtures, by the parity tests when the binary travelled to
the GPU box, and as the timed CPU baseline of the hierarchy build (bench.py --config 1)."""
import os
import subprocess
import tempfile

import numpy as np

from harness import ROOT

TOOL = os.path.join(ROOT, "oracle", "_ref", "ref_gh_tool")


def available():
    return os.path.exists(TOOL)


def read_blob(path):
    raw = open(path, "rb").read()
    P, N = np.frombuffer(raw, np.int32, 2)
    off = 8
    out = {}
    for name, shape, dt in (("positions", (P, 3), np.float32), ("shs", (P, 48), np.float32), ("opacities", (P,), np.float32),
                            ("log_scales", (P, 3), np.float32), ("rotations", (P, 4), np.float32),
                            ("nodes", (N, 7), np.int32), ("boxes", (N, 8), np.float32)):
        cnt = int(np.prod(shape))
        out[name] = np.frombuffer(raw, dt, cnt, off).reshape(shape).copy()
        off += cnt * 4
    assert off == len(raw)
    return out


def run_build(pos, shs, opacity, scale, rot, align=True):
    """-> (arrays as Writer::makeHierarchy lays them out, {"kd": s, "merge": s, "align": s})."""
    n = len(pos)
    with tempfile.TemporaryDirectory() as d:
        gb, ob = os.path.join(d, "in.gblob"), os.path.join(d, "out.blob")
        with open(gb, "wb") as f:
            f.write(np.int32(n).tobytes())
            for a, w in ((pos, 3), (shs, 48), (opacity, 1), (scale, 3), (rot, 4)):
                a = np.ascontiguousarray(np.asarray(a, np.float32).reshape(n, w))
                f.write(a.tobytes())
        res = subprocess.run([TOOL, "build", gb, ob, "1" if align else "0"], check=True, capture_output=True, text=True)
        secs = [float(x) for x in res.stdout.strip().split()[1:4]]
        return read_blob(ob), dict(zip(("kd", "merge", "align"), secs))
