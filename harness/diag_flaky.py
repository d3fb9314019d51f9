"""Diagnostic: repeat the raw-store render_post parity check and print the gradient errors of every run."""
import sys

import torch

from . import add_paths

add_paths()
from tests import test_gpu_hierarchy as T, util  # noqa: E402


def main():
    errs = []
    orig = util.assert_grad_close

    def spy(ours, ref, ref_again, name, more_refs=()):
        refs = [ref.double(), ref_again.double()] + [r.double() for r in more_refs]
        mean = sum(refs) / len(refs)
        den = float(mean.abs().max()) + 1e-30
        noise = max(float((r - mean).abs().max()) for r in refs) / den
        err = float((ours.double() - mean).abs().max()) / den
        errs.append((name, err, noise))
        return err

    util.assert_grad_close = spy
    T.util.assert_grad_close = spy
    for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
        errs.clear()
        T._fused_vs_glue(15, 1000, 800, 600, 15000, seed=7, depth=True, raw=torch.abs)
        print(i, " ".join(f"{n}:{e:.2e}/{z:.1e}" for n, e, z in errs), flush=True)


if __name__ == "__main__":
    main()
