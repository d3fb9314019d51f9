import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from harness import add_paths  # noqa: E402

add_paths()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def pytest_terminal_summary(terminalreporter, exitstatus, config):
    """Gradient-parity table of the session: measured error, the reference's own run-to-run spread and the bound
    each comparison used (tests/util.py assert_grad_close)."""
    try:
        from tests import util
    except Exception:
        return
    if not util.GRAD_LOG:
        return
    import json
    tr = terminalreporter
    tr.section("gradient parity (ours vs mean of reference runs)")
    tr.write_line(f"{'tensor':<18}{'relerr':>11}{'ref spread':>12}{'bound':>11}  branch       test")
    for r in util.GRAD_LOG:
        tr.write_line(f"{r['tensor']:<18}{r['relerr']:>11.3e}{r['spread']:>12.3e}{r['bound']:>11.3e}  "
                      f"{r['branch']:<12} {r['test'].split('::')[-1]}")
    out = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, "grad_parity.jsonl"), "w") as fh:
            for r in util.GRAD_LOG:
                fh.write(json.dumps(r) + "\n")
    except OSError:
        pass
