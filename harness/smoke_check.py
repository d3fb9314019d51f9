"""smoke(): hierarchy cut -> render_post fwd+bwd -> sparse Adam on a tiny scene, compared with the
reference build when it is present (oracle/_ref) and with the CPU oracle otherwise."""
import torch

from . import refglue, scenes, workload


def run(dev):
    import gaussian_hierarchy
    from scene.OurAdam import Adam
    gh = gaussian_hierarchy._C
    attrs = workload.build_scene(12, 256, 0, dev)
    cam = scenes.street_camera(640, 360, seed=2, device=dev, yaw_deg=3.0, x=-20.0)
    bg = torch.zeros(3, device=dev)
    model = workload.ActivatedModel(attrs, 256)
    limit = workload.find_limit(gh, model, cam, 2000, dev)
    cut = workload.Cut(attrs["means3D"].size(0), dev)
    cut.compute(gh, model, cam, limit)
    out = workload.ours_render_post(cam, model, bg, cut, do_depth=True)
    loss = out["render"].mean() + out["depth"].mean()
    loss.backward()
    torch.cuda.synchronize()
    assert torch.isfinite(out["render"]).all() and out["render"].abs().sum() > 0
    grads = [p.grad.clone() for p in model.parameters()]
    assert all(torch.isfinite(g).all() for g in grads)

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
    else:
        from oracle import cpu_reference
        cpu_reference.check_render_post(cam, attrs, cut, out, grads)

    opt = Adam([{"params": [model._xyz], "lr": 1e-3}], lr=0.0, eps=1e-15)
    before = model._xyz.detach().clone()
    relevant = (model._o.grad.flatten() != 0).nonzero().flatten().long()
    opt.step(relevant)
    torch.cuda.synchronize()
    assert not torch.equal(before, model._xyz.detach())
    print(f"smoke ok: cut={cut.count} visible={int(out['visibility_filter'].sum())}")
