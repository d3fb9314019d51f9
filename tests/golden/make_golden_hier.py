"""Generates tests/golden/hier_l6.npz with the REFERENCE's HierarchyWriter / HierarchyLoader (GH/hierarchy_writer.cpp:27-118,
GH/hierarchy_loader.cpp:26-128), compiled from the reference sources into oracle/_ref/ref_gh_tool by oracle/build_ref.py
(driver: oracle/reftool/ref_gh_tool.cpp).  The fixture holds the inputs, the bytes of the compressed and uncompressed
.hier files the reference writes for them, and what the reference loader reads back from each.
Run once:  python tests/golden/make_golden_hier.py"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from harness import scenes  # noqa: E402

TOOL = os.path.join(ROOT, "oracle", "_ref", "ref_gh_tool")
ORDER = ("means3D", "shs", "opacities", "log_scales", "rotations", "nodes", "boxes")


def write_blob(path, a):
    with open(path, "wb") as f:
        np.asarray([a["means3D"].shape[0], a["nodes"].shape[0]], np.int32).tofile(f)
        for k in ORDER:
            np.ascontiguousarray(a[k]).tofile(f)


def read_blob(path):
    raw = open(path, "rb").read()
    P, N = np.frombuffer(raw, np.int32, 2)
    off, out = 8, {}
    for k, (dt, w) in zip(ORDER, ((np.float32, 3), (np.float32, 48), (np.float32, 1), (np.float32, 3), (np.float32, 4),
                                  (np.int32, 7), (np.float32, 8))):
        n = (N if k in ("nodes", "boxes") else P) * w
        out[k] = np.frombuffer(raw, dt, n, off).reshape(-1, w).copy()
        off += n * 4
    assert off == len(raw)
    return out


def main():
    h = scenes.build_hierarchy(6, seed=4, device="cpu")
    a = {k: h[k].numpy() for k in ORDER}
    a["shs"] = a["shs"].reshape(-1, 48)
    a["boxes"] = a["boxes"].reshape(-1, 8)
    # values that exercise the fp16 container: large and tiny magnitudes, negative zero, a node counter near the limit
    a["log_scales"][3] = [-11.5, 4.25, -0.0]
    a["shs"][5, :4] = [65000.0, -70000.0, 1e-8, 3.14159]
    a["nodes"][7, 0] = 31999
    out = {f"in_{k}": v for k, v in a.items()}
    with tempfile.TemporaryDirectory() as d:
        blob = os.path.join(d, "in.blob")
        write_blob(blob, a)
        for name, flag in (("compressed", "1"), ("plain", "0")):
            hier, back = os.path.join(d, name + ".hier"), os.path.join(d, name + ".blob")
            subprocess.run([TOOL, "hier-write", blob, hier, flag], check=True, stdout=subprocess.DEVNULL)
            subprocess.run([TOOL, "hier-load", hier, back], check=True, stdout=subprocess.DEVNULL)
            out[f"{name}_bytes"] = np.frombuffer(open(hier, "rb").read(), np.uint8)
            for k, v in read_blob(back).items():
                out[f"{name}_load_{k}"] = v
        # "Would lose information!": a counter above 32000 must make the compressed writer throw
        bad = dict(a)
        bad["nodes"] = a["nodes"].copy()
        bad["nodes"][9, 6] = 32001
        write_blob(blob, bad)
        r = subprocess.run([TOOL, "hier-write", blob, os.path.join(d, "bad.hier"), "1"], capture_output=True, text=True)
        assert r.returncode == 1 and "lose information" in r.stderr, r
    np.savez_compressed(os.path.join(HERE, "hier_l6.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
