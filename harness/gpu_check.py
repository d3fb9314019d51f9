"""Diagnostic comparison of the B200 kernels against the reference CUDA build (oracle/_ref) on one
seeded scene; prints where (if anywhere) results diverge.  Usage on the GPU box:
    python -m harness.gpu_check [--n 100000] [--w 1024] [--h 768] [--interp] [--seed 0]
"""
import argparse
import sys

import torch

from . import add_paths

add_paths()
from . import refglue, scenes  # noqa: E402


def relerr(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / (b.abs().max() + 1e-30))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100000)
    ap.add_argument("--w", type=int, default=1024)
    ap.add_argument("--h", type=int, default=768)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--interp", action="store_true")
    ap.add_argument("--nodepth", action="store_true")
    ap.add_argument("--yaw", type=float, default=None)
    args = ap.parse_args()
    import diff_gaussian_rasterization as dgr
    dev = "cuda"
    sc = scenes.street_gaussians(args.n, seed=args.seed, device=dev)
    cam = scenes.street_camera(args.w, args.h, seed=args.seed, device=dev, yaw_deg=args.yaw)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    ts = kids = None
    if args.interp:
        g = torch.Generator().manual_seed(5)
        ts = torch.rand(args.n, generator=g).to(dev)
        ts[::3] = 1.0
        kids = torch.randint(1, 6, (args.n,), generator=g, dtype=torch.int32).to(dev)
    rs = refglue.make_settings(dgr.GaussianRasterizationSettings, cam, bg, do_depth=not args.nodepth,
                               interpolation_weights=ts, num_node_kids=kids)
    g = torch.Generator().manual_seed(7)
    dcol = (torch.randn(3, args.h, args.w, generator=g) / (args.w * args.h)).to(dev)
    ddep = (torch.randn(1, args.h, args.w, generator=g) / (args.w * args.h)).to(dev)

    def run(fn):
        p = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        m2d = torch.zeros_like(p["means3D"], requires_grad=True)
        color, radii, depth = fn(p, m2d)
        loss = (color * dcol).sum()
        if depth.numel():
            loss = loss + (depth * ddep).sum()
        loss.backward()
        return color.detach(), radii, depth.detach(), {k: v.grad for k, v in p.items()}, m2d.grad

    def ours(p, m2d):
        out = dgr.GaussianRasterizer(rs)(means3D=p["means3D"], means2D=m2d, opacities=p["opacities"], shs=p["shs"],
                                         scales=p["scales"], rotations=p["rotations"])
        return out

    def ref(p, m2d):
        return refglue.ref_rasterize(rs, p["means3D"], m2d, p["opacities"], shs=p["shs"], scales=p["scales"],
                                     rotations=p["rotations"])

    # forward buffers
    C = dgr._C
    e = torch.Tensor([])
    fa = (rs.bg, rs.render_indices, rs.parent_indices, rs.interpolation_weights, rs.num_node_kids, sc["means3D"], e,
          sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
          rs.tanfovy, rs.image_height, rs.image_width, sc["shs"], 3, rs.campos, False, False, rs.do_depth)
    R, color, radii, geom, binning, img, invd = C.rasterize_gaussians(*fa)
    dv = C.debug_view(args.n, args.w, args.h, R, geom, binning, img)
    Rr, color_r, radii_r, geom_r, binning_r, img_r, invd_r = refglue.ref_hr().rasterize_gaussians(*fa)
    rv = refglue.parse_ref_buffers(geom_r, binning_r, img_r, args.n, Rr, args.w, args.h)
    torch.cuda.synchronize()
    print(f"P={args.n} visible ours={(radii > 0).sum().item()} ref={(radii_r > 0).sum().item()}  R ours={R} ref={Rr}")
    print("radii equal:", bool((radii == radii_r).all()), " mismatches:", int((radii != radii_r).sum()))
    vis = radii_r > 0
    print("depth bits equal (visible):", bool((dv["depths"][vis].view(torch.int32) == rv["depths"][vis].view(torch.int32)).all()))
    rec = dv["records"].view(args.n, 16)
    print("xy bits equal:", bool((rec[vis][:, 0:2].contiguous().view(torch.int32) == rv["means2D"][vis].contiguous().view(torch.int32)).all()))
    print("conic_opacity bits equal:", bool((rec[vis][:, 4:8].contiguous().view(torch.int32) == rv["conic_opacity"][vis].contiguous().view(torch.int32)).all()),
          " relerr", relerr(rec[vis][:, 4:8], rv["conic_opacity"][vis]))
    print("rgb bits equal:", bool((rec[vis][:, 8:11].contiguous().view(torch.int32) == rv["rgb"][vis].contiguous().view(torch.int32)).all()),
          " relerr", relerr(rec[vis][:, 8:11], rv["rgb"][vis]))
    print("cov3D bits equal:", bool((dv["cov3D"].view(args.n, 6)[vis].contiguous().view(torch.int32) == rv["cov3D"][vis].contiguous().view(torch.int32)).all()))
    print("tiles_touched equal:", bool((dv["tiles_touched"] == rv["tiles_touched"]).all()))
    if R == Rr and R > 0:
        print("keys equal:", bool((dv["point_list_keys"] == rv["point_list_keys"]).all()),
              " point_list equal:", bool((dv["point_list"] == rv["point_list"]).all()),
              " ranges equal:", bool((dv["ranges"] == rv["ranges"]).all()))
        print("n_contrib equal:", bool((dv["n_contrib"] == rv["n_contrib"]).all()), " mismatches", int((dv["n_contrib"] != rv["n_contrib"]).sum()),
              " final_T bits equal:", bool((dv["final_T"].view(torch.int32) == rv["final_T"].view(torch.int32)).all()))
    print("color bits equal:", bool((color.view(torch.int32) == color_r.view(torch.int32)).all()), " relerr", relerr(color, color_r),
          " maxabs", float((color - color_r).abs().max()))
    if invd.numel():
        print("invdepth bits equal:", bool((invd.view(torch.int32) == invd_r.view(torch.int32)).all()), " relerr", relerr(invd, invd_r))

    co, ro, do_, go, g2o = run(ours)
    cr, rr, dr, gr, g2r = run(ref)
    torch.cuda.synchronize()
    cr2, rr2, dr2, gr2, g2r2 = run(ref)
    torch.cuda.synchronize()
    for k in go:
        print(f"grad {k:10s} relerr {relerr(go[k], gr[k]):.3e}  ref-vs-ref {relerr(gr2[k], gr[k]):.3e}  max|ref| {float(gr[k].abs().max()):.3e}")
    print(f"grad means2D   relerr {relerr(g2o, g2r):.3e}  ref-vs-ref {relerr(g2r2, g2r):.3e}")
    d = (go["scales"] - gr["scales"]).abs().max(dim=1).values
    top = d.topk(6).indices
    for i in top.tolist():
        print("worst scales idx", i, "ours", go["scales"][i].tolist(), "ref", gr["scales"][i].tolist(), "ref2", gr2["scales"][i].tolist(),
              "radius", int(ro[i]), "t", float(ts[i]) if ts is not None else None, "kids", int(kids[i]) if kids is not None else None,
              "opac", float(sc["opacities"][i]), "scale", sc["scales"][i].tolist())
    return 0


if __name__ == "__main__":
    sys.exit(main())
