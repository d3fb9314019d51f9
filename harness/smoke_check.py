"""smoke(): hierarchy cut -> render_post fwd+bwd -> sparse Adam on a tiny scene, checked against the
CPU oracle (oracle/) and, when it was built (oracle/_ref), against the reference CUDA extensions."""
import torch

from . import refglue, scenes, workload


def run(dev):
    import gaussian_hierarchy
    from scene.OurAdam import Adam
    import math
    from oracle import cpu_reference
    gh = gaussian_hierarchy._C
    # The scene, the camera and the granularity threshold are made on the HOST (numpy / torch-CPU, the cut
    # bisection with the CPU oracle) and uploaded, so that the first kernels this process launches on the GPU
    # are the product's own: cut, preprocess, sort, blend, backward (the driver's launch list is capped).
    attrs_cpu = workload.build_scene(12, 256, 0, "cpu")
    cam_cpu = scenes.street_camera(640, 360, seed=2, device="cpu", yaw_deg=3.0, x=-20.0)
    nodes_np, boxes_np = attrs_cpu["nodes"].numpy(), attrs_cpu["boxes"].numpy()
    vp = cam_cpu.camera_center.numpy()
    lo, hi = 1e-6, 10.0
    for _ in range(30):   # harness.workload.find_limit, on the host
        mid = math.sqrt(lo * hi)
        lo, hi = (mid, hi) if cpu_reference.expand_to_size(nodes_np, boxes_np, mid, vp)[0] > 2000 else (lo, mid)
    limit = math.sqrt(lo * hi)
    attrs = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in attrs_cpu.items()}
    cam = scenes.street_camera(640, 360, seed=2, device=dev, yaw_deg=3.0, x=-20.0)
    bg = torch.zeros(3, device=dev)
    model = workload.ActivatedModel(attrs, 256)
    cut = workload.Cut(attrs["means3D"].size(0), dev)
    cut.compute(gh, model, cam, limit)
    out = workload.ours_render_post(cam, model, bg, cut, do_depth=True)
    loss = out["render"].mean() + out["depth"].mean()
    loss.backward()
    torch.cuda.synchronize()
    assert torch.isfinite(out["render"]).all() and out["render"].abs().sum() > 0
    grads = [p.grad.clone() for p in model.parameters()]
    out_radii = torch.zeros(cut.count + 256, dtype=torch.int32, device=dev)
    out_radii[out["visibility_filter"]] = out["radii"]
    assert all(torch.isfinite(g).all() for g in grads)

    # ---- checker 1: the CPU oracle (oracle/, pinned against the reference's golden vectors) ----
    n_pix = 640 * 360
    ones = torch.full((3, 360, 640), 1.0 / (3 * n_pix))
    oned = torch.full((1, 360, 640), 1.0 / n_pix)
    ref = cpu_reference.render_post_reference(cam, attrs, 256, limit, dL_dcolor=ones, dL_dinvdepth=oned, bg=bg)
    assert ref["count"] == cut.count
    assert (torch.from_numpy(ref["render_indices"]).to(dev) == cut.render_indices[:cut.count]).all()
    assert (torch.from_numpy(ref["radii"]).to(dev) == out_radii).all(), "radii differ from the CPU oracle"
    img_err = float((torch.from_numpy(ref["render"]).to(dev) - out["render"]).abs().max())
    assert img_err < 2e-6, f"image differs from the CPU oracle: {img_err:.3e}"
    for g, key in zip(grads, ("means3D", "shs", "opacities", "scales", "rotations")):
        r = torch.from_numpy(ref["grads"][key]).to(dev)
        err = float((g - r).abs().max() / (r.abs().max() + 1e-30))
        assert err < 2e-4, f"gradient {key} mismatch vs CPU oracle: {err:.3e}"

    # ---- checker 2 (when oracle/_ref was built): the unmodified reference CUDA extensions ----
    if refglue.ref_hr() is not None:
        ref_model = workload.ActivatedModel(attrs, 256)
        rcut = workload.Cut(attrs["means3D"].size(0), dev)
        rcut.compute(refglue.ref_gh(), ref_model, cam, limit)
        assert rcut.count == cut.count
        assert torch.equal(rcut.render_indices[:cut.count], cut.render_indices[:cut.count])
        rout = workload.ref_render_post(cam, ref_model, bg, rcut, do_depth=True)
        (rout["render"].mean() + rout["depth"].mean()).backward()
        torch.cuda.synchronize()
        assert torch.equal(rout["render"], out["render"]), "image differs from the reference build"
        for g, p in zip(grads, ref_model.parameters()):
            err = (g - p.grad).abs().max() / (p.grad.abs().max() + 1e-30)
            assert err < 2e-4, f"gradient mismatch vs reference: {float(err):.3e}"

    opt = Adam([{"params": [model._xyz], "lr": 1e-3}], lr=0.0, eps=1e-15)
    before = model._xyz.detach().clone()
    relevant = (model._o.grad.flatten() != 0).nonzero().flatten().long()
    opt.step(relevant)
    torch.cuda.synchronize()
    assert not torch.equal(before, model._xyz.detach())
    print(f"smoke ok: cut={cut.count} visible={int(out['visibility_filter'].sum())}")
