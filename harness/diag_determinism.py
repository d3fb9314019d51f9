"""Diagnostic: is the forward bit-deterministic from run to run?  Renders the same scene N times (low-level call:
keys, lists, ranges, n_contrib, final_T, image hashed separately; then through the autograd API with a backward,
image only) and reports which component ever differs from the first run.
    python -m harness.diag_determinism [N]"""
import hashlib
import sys

import torch

from . import add_paths

add_paths()
from . import refglue  # noqa: E402
from tests import util  # noqa: E402


def h(t):
    return hashlib.sha256(t.detach().contiguous().cpu().numpy().tobytes()).hexdigest()[:12]


def main():
    n_iter = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    import diff_gaussian_rasterization as dgr
    n, w, hgt = 60000, 800, 600
    sc, cam, ts, kids = util.make_scene(n, w, hgt, seed=2, dev="cuda", interp=True)
    rs = refglue.make_settings(dgr.GaussianRasterizationSettings, cam, torch.zeros(3, device="cuda"), sh_degree=3,
                               do_depth=True, interpolation_weights=ts, num_node_kids=kids)
    args = util.forward_args(rs, sc)
    first, bad = None, {}
    for it in range(n_iter):
        R, color, radii, geom, binning, img, invd = dgr._C.rasterize_gaussians(*args)
        dv = dgr._C.debug_view(n, w, hgt, R, geom, binning, img)
        cur = {"R": str(R), "radii": h(radii), "keys": h(dv["point_list_keys"]), "list": h(dv["point_list"]),
               "ranges": h(dv["ranges"]), "n_contrib": h(dv["n_contrib"]), "final_T": h(dv["final_T"]),
               "image": h(color), "invdepth": h(invd), "records": h(dv["records"])}
        if first is None:
            first = cur
        for k in cur:
            if cur[k] != first[k]:
                bad.setdefault(k, []).append(it)
    print("low-level forward:", {k: v[:8] for k, v in bad.items()} or "all identical", "over", n_iter, "runs")
    first, bad = None, []
    for it in range(n_iter):
        p = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        m2d = torch.zeros_like(p["means3D"], requires_grad=True)
        c, r, d = dgr.GaussianRasterizer(rs)(means3D=p["means3D"], means2D=m2d, opacities=p["opacities"],
                                             shs=p["shs"], scales=p["scales"], rotations=p["rotations"])
        (c.sum() + d.sum()).backward()
        cur = h(c) + h(d)
        if first is None:
            first = cur
        if cur != first:
            bad.append(it)
    print("autograd forward+backward: image differs in runs", bad[:16] or "none", "of", n_iter)
    return 0


if __name__ == "__main__":
    sys.exit(main())
