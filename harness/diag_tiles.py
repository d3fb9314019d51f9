"""Diagnostic: per-tile work distribution of the bench workload (list length, how far the forward walks,
last contributor per pixel) -- the load-balance picture behind the blend kernels' time."""
import sys
import torch
from . import add_paths
add_paths()
from . import workload  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    import diff_gaussian_rasterization as dgr
    import gaussian_hierarchy
    gh = gaussian_hierarchy._C
    W, H = 1920, 1080
    attrs = workload.build_scene(23, 100_000, 0, dev, free_corridor=True)
    cams = workload.cameras(W, H, 8, dev, seed=0)
    model = workload.ActivatedModel(attrs, 100_000)
    limit = workload.find_limit(gh, model, cams[0], 5_000_000, dev)
    bg = torch.zeros(3, device=dev)
    cap = {}
    orig = dgr._C.rasterize_gaussians_ex

    def spy(*a, **k):
        out = orig(*a, **k)
        cap["out"] = out
        return out

    dgr._C.rasterize_gaussians_ex = spy
    for i, c in enumerate(cams):
        cut = workload.Cut(attrs["means3D"].size(0), dev)
        cut.compute(gh, model, c, limit)
        with torch.no_grad():
            workload.ours_render_post(c, model, bg, cut, do_depth=True)
        R, color, radii, geom, binning, img = cap["out"][:6]
        dv = dgr._C.debug_view(radii.numel(), W, H, R, geom, binning, img)
        rg = dv["ranges"].view(-1, 2).long()
        n = (rg[:, 1] - rg[:, 0]).float()
        nc = dv["n_contrib"].view(H, W).float()
        fT = dv["final_T"].view(H, W)
        pad_h = (16 - H % 16) % 16
        ncp = torch.nn.functional.pad(nc, (0, 0, 0, pad_h))
        tile_last = ncp.view((H + pad_h) // 16, 16, W // 16, 16).amax(dim=(1, 3)).flatten()
        sat = (fT < 1e-3).float().mean()
        q = torch.tensor([0.5, 0.9, 0.99, 1.0], device=dev)
        print(f"cam {i}: P {radii.numel()} V {int((radii > 0).sum())} R {R}  list len mean {n.mean():.0f} "
              f"q50/90/99/max {[int(x) for x in torch.quantile(n, q)]}  last-contrib/pixel mean {nc.mean():.0f} "
              f"tile-max-last mean {tile_last.mean():.0f} q50/90/99/max {[int(x) for x in torch.quantile(tile_last, q)]} "
              f"pixels with T<1e-3: {sat:.2f}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
