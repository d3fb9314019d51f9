"""Diagnostic: fused render_post vs the reference glue -- where do the gradients differ most?
    python -m harness.diag_fused LEVELS SKY W H TARGET"""
import sys

import torch

from . import add_paths

add_paths()
from . import refglue, scenes, workload  # noqa: E402


def main():
    levels, sky, w, h, target = (int(x) for x in sys.argv[1:6])
    import gaussian_hierarchy
    DEV = "cuda"
    seed = levels
    attrs = workload.build_scene(levels, sky, seed, DEV)
    cam = scenes.street_camera(w, h, seed=seed, device=DEV, yaw_deg=4.0, x=-10.0)
    bg = torch.tensor([0.05, 0.1, 0.15], device=DEV)
    models = [workload.ActivatedModel(attrs, sky) for _ in range(4)]
    limit = workload.find_limit(gaussian_hierarchy._C, models[0], cam, target, DEV)
    n = attrs["means3D"].size(0)
    g = torch.Generator().manual_seed(11)
    dcol = (torch.randn(3, h, w, generator=g) / (w * h)).to(DEV)
    ddep = (torch.randn(1, h, w, generator=g) / (w * h)).to(DEV)

    def run(model, fn, gh):
        cut = workload.Cut(n, DEV)
        cut.compute(gh, model, cam, limit)
        out = fn(cam, model, bg, cut, do_depth=True)
        ((out["render"] * dcol).sum() + (out["depth"] * ddep).sum()).backward()
        return out, cut

    o1, cut = run(models[0], workload.ours_render_post, gaussian_hierarchy._C)
    o2, _ = run(models[1], workload.ours_render_post, gaussian_hierarchy._C)
    r1, _ = run(models[2], workload.ref_render_post, refglue.ref_gh())
    r2, _ = run(models[3], workload.ref_render_post, refglue.ref_gh())
    torch.cuda.synchronize()
    K = cut.count
    print("K", K, "visible", int(o1["visibility_filter"].sum()))
    names = ["xyz", "features", "opacity", "scaling", "rotation"]
    for i, name in enumerate(names):
        a, a2, b, b2 = (m.parameters()[i].grad.double() for m in models)
        den = b.abs().max()
        print(f"{name:9s} ours-ref {float((a - b).abs().max() / den):.3e}  ours-ours {float((a - a2).abs().max() / den):.3e}"
              f"  ref-ref {float((b - b2).abs().max() / den):.3e}  max|ref| {float(den):.3e}")
        d = (a - b).abs().reshape(a.size(0), -1).max(1).values
        row = int(d.argmax())
        ri = cut.render_indices[:K]
        pi = cut.parent_indices[:K]
        as_child = (ri == row).nonzero().flatten()
        as_parent = (pi == row).nonzero().flatten()
        print("   worst row", row, "ours", a[row].flatten()[:4].tolist(), "ref", b[row].flatten()[:4].tolist(),
              "ref2", b2[row].flatten()[:4].tolist())
        print("   as child slots", as_child[:4].tolist(), "t", cut.interpolation_weights[as_child[:4]].tolist(),
              "| as parent of", as_parent.numel(), "slots", "t", cut.interpolation_weights[as_parent[:6]].tolist())
        if name == "rotation":
            print("   scales", attrs["scales"][row].tolist(), "rot", attrs["rotations"][row].tolist())
            for s in as_parent[:6].tolist():
                print("     child slot", s, "radii", int(o1_radii(o1, s)))


def o1_radii(o, s):
    return o["visibility_filter"][s]


if __name__ == "__main__":
    main()
