#!/usr/bin/env python
"""Overlay the drop-in modules onto a checkout of iacopoermacora/Street-sparse-3DGS.

Python puts the directory of the script being run (`train_post.py`, ...) ahead of PYTHONPATH, so the reference's own
`gaussian_renderer/`, `scene/` and `utils/` packages win over same-named packages found later on the path.  The two
extension packages the reference installs with pip (`diff_gaussian_rasterization`, `gaussian_hierarchy`) are replaced
through PYTHONPATH; the three pure-Python files are replaced IN the checkout:

    gaussian_renderer/__init__.py     render / render_post / render_coarse on the fused kernels
    scene/OurAdam.py                  fused sparse Adam
    utils/loss_utils.py               fused L1 + SSIM, masked losses (ssim_masked, create_window kept)

    python street-sparse-3dgs_b200/install_into_reference.py /path/to/Street-sparse-3DGS          # install
    python street-sparse-3dgs_b200/install_into_reference.py /path/to/Street-sparse-3DGS --undo   # restore
    PYTHONPATH=/path/to/street-sparse-3dgs_b200 python train_post.py ...                             # unchanged script

Originals are kept next to the overlaid files as `<name>.reference_orig`.  Nothing else in the checkout is touched
(`scene/__init__.py`, `scene/gaussian_model.py`, `utils/general_utils.py`, ... stay the reference's)."""
import argparse
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = ("gaussian_renderer/__init__.py", "scene/OurAdam.py", "utils/loss_utils.py")
MARK = "# overlaid by street-sparse-3dgs_b200/install_into_reference.py\n"


def install(root):
    done = []
    for rel in FILES:
        dst = os.path.join(root, rel)
        src = os.path.join(HERE, rel)
        if not os.path.exists(dst):
            raise SystemExit(f"{dst} not found: is {root} a Street-sparse-3DGS checkout?")
        backup = dst + ".reference_orig"
        if not os.path.exists(backup):
            shutil.copy2(dst, backup)
        with open(src) as f:
            body = f.read()
        with open(dst, "w") as f:
            f.write(MARK + body)
        done.append(rel)
    return done


def undo(root):
    done = []
    for rel in FILES:
        dst = os.path.join(root, rel)
        backup = dst + ".reference_orig"
        if os.path.exists(backup):
            shutil.move(backup, dst)
            done.append(rel)
    return done


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("reference_root")
    ap.add_argument("--undo", action="store_true")
    a = ap.parse_args(argv)
    files = undo(a.reference_root) if a.undo else install(a.reference_root)
    print(("restored: " if a.undo else "overlaid: ") + ", ".join(files))
    if not a.undo:
        print(f"now run the reference's scripts with PYTHONPATH={HERE}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
