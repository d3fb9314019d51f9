"""Diagnostic: per-tile start / end times of the blend kernels (instrumented COUNT variants) for one camera of
the bench workload: where the tail of each kernel comes from.   python -m harness.diag_balance [--cam 0]"""
import argparse
import sys
import torch
from . import add_paths
add_paths()
from . import workload  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cam", type=int, default=0)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    import diff_gaussian_rasterization as dgr
    import gaussian_hierarchy
    import h3dgs_native as nat
    gh = gaussian_hierarchy._C
    W, H = 1920, 1080
    T = (W // 16) * ((H + 15) // 16)
    attrs = workload.build_scene(23, 100_000, 0, dev, free_corridor=True)
    cams = workload.cameras(W, H, 8, dev, seed=0)
    model = workload.ActivatedModel(attrs, 100_000)
    limit = workload.find_limit(gh, model, cams[0], 5_000_000, dev)
    bg = torch.zeros(3, device=dev)
    c = cams[args.cam]
    cut = workload.Cut(attrs["means3D"].size(0), dev)
    cut.compute(gh, model, c, limit)
    g = torch.Generator().manual_seed(1234)
    dcol = (torch.randn(3, H, W, generator=g) / (W * H)).to(dev)
    ddep = (torch.randn(1, H, W, generator=g) / (W * H)).to(dev)
    cap = {}
    orig = dgr._C.rasterize_gaussians_ex

    def spy(*a, **k):
        out = orig(*a, **k)
        cap["out"] = out
        return out

    dgr._C.rasterize_gaussians_ex = spy
    cnt = torch.zeros(16 + 4 * T, dtype=torch.int64, device=dev)
    for it in range(2):
        cnt.zero_()
        nat.lib().h3dgs_blend_stats(cnt.data_ptr())
        out = workload.ours_render_post(c, model, bg, cut, do_depth=True)
        ((out["render"] * dcol).sum() + (out["depth"] * ddep).sum()).backward()
        torch.cuda.synchronize()
        nat.lib().h3dgs_blend_stats(None)
    R, color, radii, geom, binning, img = cap["out"][:6]
    dv = dgr._C.debug_view(radii.numel(), W, H, R, geom, binning, img)
    rg = dv["ranges"].view(-1, 2).long()
    n = (rg[:, 1] - rg[:, 0]).cpu()
    t = cnt[16:].view(T, 4).cpu()
    print("# backward: since the list cuts (csrc/common.cuh BLEND_SEG) the per-tile times cover the FIRST work item of "
          "each tile (list positions < 1024); the deeper items lead the grid and are not timed individually")
    for name, a, b in (("forward", 0, 1), ("backward", 2, 3)):
        st, en = t[:, a], t[:, b]
        ok = en > 0
        t0 = int(st[ok].min())
        span = int(en[ok].max()) - t0
        dur = (en - st)[ok].double()
        print(f"{name}: kernel span {span / 1e3:.0f} us, {int(ok.sum())} tiles, CTA time sum {dur.sum() / 1e3:.0f} us "
              f"= {dur.sum() / span:.0f} CTAs running on average (capacity 148 x 5 = 740)")
        # concurrency over time in 20 slices
        rs, re_ = (st[ok] - t0).double(), (en[ok] - t0).double()
        conc = []
        for i in range(20):
            lo, hi = span * i / 20.0, span * (i + 1) / 20.0
            ov = (re_.clamp(max=hi) - rs.clamp(min=lo)).clamp(min=0)
            conc.append(int(ov.sum() / (hi - lo)))
        print("  CTAs running per 5 % slice of the span:", conc)
        top = torch.argsort(en - st, descending=True)[:8]
        for i in top.tolist():
            print(f"  tile {i:5d} (x {i % 120:3d}, y {i // 120:2d}) list {int(n[i]):6d}: start +{(int(st[i]) - t0) / 1e3:7.0f} us, "
                  f"runs {(int(en[i]) - int(st[i])) / 1e3:7.0f} us")
    return 0


if __name__ == "__main__":
    sys.exit(main())
