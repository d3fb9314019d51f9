"""Diagnostic: does the raw-store path (kernels apply exp / normalize / abs / dc+rest to the gathered rows)
reproduce torch's own activation kernels bit for bit?   python -m harness.diag_raw
Run once per H3DGS_NORM_VARIANT (0..3) to see which summation order torch's F.normalize reduction uses."""
import os
import sys

import torch

from . import add_paths

add_paths()
from . import scenes, workload  # noqa: E402


def main():
    import diff_gaussian_rasterization as dgr
    import gaussian_hierarchy
    dev = torch.device("cuda", 0)
    attrs = workload.build_scene(17, 3000, 3, dev)
    m = workload.HierModel(attrs, 3000)
    g = torch.Generator().manual_seed(5)
    n = m._xyz.size(0)
    with torch.no_grad():
        m._rotation.mul_((0.25 + 3.0 * torch.rand(n, 1, generator=g)).to(dev))
        m._opacity.mul_((torch.rand(n, 1, generator=g) < 0.5).to(dev) * 2.0 - 1.0)
    cam = scenes.street_camera(1024, 768, seed=1, device=dev, yaw_deg=4.0, x=-10.0)
    bg = torch.zeros(3, device=dev)
    limit = workload.find_limit(gaussian_hierarchy._C, m, cam, 60000, dev)
    cut = workload.Cut(n, dev)
    cut.compute(gaussian_hierarchy._C, m, cam, limit)
    views = {}
    orig = dgr._C.rasterize_gaussians_ex

    def spy(*a, **k):
        out = orig(*a, **k)
        P = out[2].numel()
        dv = dgr._C.debug_view(P, 1024, 768, out[0], out[3], out[4], out[5])
        views[spy.tag] = {k2: v.clone() for k2, v in dv.items() if v is not None}
        return out

    dgr._C.rasterize_gaussians_ex = spy
    res = {}
    for tag, env in (("raw", "1"), ("activated", "0")):
        os.environ["H3DGS_RAW_STORE"] = env
        spy.tag = tag
        with torch.no_grad():
            res[tag] = workload.ours_render_post(cam, m, bg, cut, do_depth=True)
    torch.cuda.synchronize()
    a, b = views["raw"], views["activated"]
    print("norm variant", os.environ.get("H3DGS_NORM_VARIANT", "0"), "slots", cut.count)
    vis = a["tiles_touched"] > 0
    P = vis.numel()
    for d in (a, b):
        d["records"] = d["records"].view(P, 16)[vis]
        d["cov3D"] = d["cov3D"].view(P, 6)[vis]
    for k in ("records", "cov3D", "tiles_touched", "point_list_keys", "point_list"):
        same = a[k].shape == b[k].shape and bool(torch.equal(a[k].contiguous().view(torch.uint8), b[k].contiguous().view(torch.uint8)))
        extra = ""
        if not same and a[k].shape == b[k].shape and a[k].dtype == torch.float32:
            extra = " mismatching elements: %d of %d" % (int((a[k].view(torch.int32) != b[k].view(torch.int32)).sum()),
                                                          a[k].numel())
        print(f"  {k}: bit-equal={same}{extra}")
    print("  image bit-equal:", bool(torch.equal(res["raw"]["render"], res["activated"]["render"])))


if __name__ == "__main__":
    sys.exit(main())
