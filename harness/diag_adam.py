"""Diagnostic: rows selected by `opacity.grad != 0` after one e2e render and the time of the sparse Adam step."""
import sys
import torch
from . import add_paths
add_paths()
from . import workload  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    import gaussian_hierarchy
    from scene.OurAdam import Adam
    gh = gaussian_hierarchy._C
    attrs = workload.build_scene(23, 100_000, 0, dev, free_corridor=True)
    cams = workload.cameras(1920, 1080, 2, dev, seed=0)
    hm = workload.HierModel(attrs, 100_000)
    limit = workload.find_limit(gh, hm, cams[0], 5_000_000, dev)
    cut = workload.Cut(attrs["means3D"].size(0), dev)
    opt = Adam(hm.param_groups(), lr=0.0, eps=1e-15)
    bg = torch.zeros(3, device=dev)
    gt = torch.rand(3, 1080, 1920, device=dev)
    for i in range(3):
        cut.compute(gh, hm, cams[i % 2], limit)
        for p in hm.parameters():
            p.grad = None
        out = workload.ours_render_post(cams[i % 2], hm, bg, cut, do_depth=False)
        loss = torch.abs(out["render"] - gt).mean()
        loss.backward()
        relevant = (hm._opacity.grad.flatten() != 0).nonzero().flatten().long()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        opt.step(relevant)
        e1.record()
        torch.cuda.synchronize()
        print(f"cut {cut.count} visible {int(out['visibility_filter'].sum())} relevant {relevant.numel()} "
              f"xyz.grad rows != 0: {int((hm._xyz.grad.abs().sum(1) != 0).sum())}  adam {e0.elapsed_time(e1):.3f} ms")
    return 0


if __name__ == "__main__":
    sys.exit(main())
