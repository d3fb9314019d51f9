"""Diagnostic: backward blend variant 1 (one pixel per lane) against variant 2 (two pixels per lane) on the same
forward state: where do the mean-2D gradients differ?"""
import torch

from harness import add_paths, refglue, scenes
from tests import util

add_paths()
DEV = "cuda"


def main(n=100000, w=1024, h=768):
    import diff_gaussian_rasterization as dgr
    import h3dgs_native as nat
    sc, cam, ts, kids = util.make_scene(n, w, h, 3, DEV, interp=True)
    bg = torch.zeros(3, device=DEV)
    rs = refglue.make_settings(dgr.GaussianRasterizationSettings, cam, bg, do_depth=True, interpolation_weights=ts,
                               num_node_kids=kids)
    g = torch.Generator().manual_seed(1)
    dcol = (torch.randn(3, h, w, generator=g) / (w * h)).to(DEV)
    ddep = (torch.randn(1, h, w, generator=g) / (w * h)).to(DEV)
    res = {}
    for variant in (1, 2, 1, 2):
        nat.lib().h3dgs_debug_set(1, variant)
        p = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "shs", "scales", "rotations")}
        m2d = torch.zeros_like(p["means3D"], requires_grad=True)
        color, radii, depth = dgr.GaussianRasterizer(rs)(means3D=p["means3D"], means2D=m2d, opacities=p["opacities"],
                                                         shs=p["shs"], scales=p["scales"], rotations=p["rotations"])
        torch.autograd.backward([color, depth], [dcol, ddep])
        torch.cuda.synchronize()
        res.setdefault(variant, []).append({"m2d": m2d.grad.double(), "op": p["opacities"].grad.double(),
                                            "sh": p["shs"].grad.double(), "radii": radii})
    for key in ("m2d", "op", "sh"):
        a1, b1 = res[1][0][key], res[1][1][key]
        a2, b2 = res[2][0][key], res[2][1][key]
        den = float(a1.abs().max())
        print(f"{key}: v1 run-to-run {float((a1 - b1).abs().max()) / den:.3e}  v2 run-to-run {float((a2 - b2).abs().max()) / den:.3e}"
              f"  v1 vs v2 {float((a1 - a2).abs().max()) / den:.3e}")
    a1, a2 = res[1][0]["m2d"], res[2][0]["m2d"]
    den = float(a1.abs().max())
    err = (a1 - a2).abs().max(dim=1).values
    worst = err.topk(10).indices.tolist()
    view, proj = cam.world_view_transform, cam.full_proj_transform
    for i in worst:
        ph = torch.cat((sc["means3D"][i], torch.ones(1, device=DEV))) @ proj
        x = ((ph[0] / ph[3] + 1) * w - 1) * 0.5
        y = ((ph[1] / ph[3] + 1) * h - 1) * 0.5
        print(f"slot {i}: err/max {float(err[i]) / den:.3e} |g| {float(a1[i].abs().max()) / den:.3e} of max  v1 {a1[i].tolist()}  v2 {a2[i].tolist()}"
              f"  pixel ({float(x):.2f}, {float(y):.2f}) radius {int(res[1][0]['radii'][i])} opacity {float(sc['opacities'][i]):.3f} t {float(ts[i]):.3f}")
    # list lengths of the tiles around the worst slot
    out = dgr._C.rasterize_gaussians_ex(*util.forward_args(rs, sc))
    dv = dgr._C.debug_view(out[2].numel(), w, h, int(out[0]), out[3], out[4], out[5])
    rng = dv["ranges"].view(-1, 2).long()
    lens = (rng[:, 1] - rng[:, 0]).view((h + 15) // 16, (w + 15) // 16)
    print("max tile list length", int(lens.max()), "tiles with > 1024 entries", int((lens > 1024).sum()))
    i = worst[0]
    ph = torch.cat((sc["means3D"][i], torch.ones(1, device=DEV))) @ proj
    tx, ty = int(((ph[0] / ph[3] + 1) * w - 1) * 0.5) // 16, int(((ph[1] / ph[3] + 1) * h - 1) * 0.5) // 16
    print("tile of the worst slot", tx, ty, "list lengths around it:")
    print(lens[max(ty - 2, 0):ty + 3, max(tx - 3, 0):tx + 4])
    rel = err / den
    for thr in (1e-4, 1e-5, 1e-6):
        print(f"slots with err/max > {thr:g}: {int((rel > thr).sum())} of {int((a1.abs().sum(1) > 0).sum())} with a gradient")


if __name__ == "__main__":
    main()
    print("=" * 80)
    main(n=20000, w=640, h=480)
