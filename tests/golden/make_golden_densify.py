#!/usr/bin/env python
"""Golden vectors of the densification path written by the REFERENCE ITSELF: the methods of its GaussianModel
(/root/reference/scene/gaussian_model.py: add_densification_stats, densify_and_prune and everything they call) are
executed from the reference's own source with the device literal "cuda" replaced by "cpu" (the methods hard-code the
device; nothing else is touched) on seeded inputs, inside a stub environment for the imports the module makes at the
top (simple_knn, plyfile, faiss: not used by these methods).  Runs in the build container only (needs /root/reference).

    python tests/golden/make_golden_densify.py      ->  tests/golden/densify_ref.npz
"""
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("H3DGS_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference_class():
    for name in ("simple_knn", "simple_knn._C", "plyfile", "faiss", "gaussian_hierarchy", "gaussian_hierarchy._C"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["simple_knn._C"].distCUDA2 = None
    sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = None
    sys.modules["gaussian_hierarchy._C"].load_hierarchy = sys.modules["gaussian_hierarchy._C"].write_hierarchy = None
    pkgs = {}
    for pkg in ("utils", "scene"):            # packages without running their __init__ (scene/__init__ imports the world)
        m = types.ModuleType(pkg)
        m.__path__ = [os.path.join(REF, pkg)]
        sys.modules[pkg] = m
        pkgs[pkg] = m

    def exec_module(modname, relpath):
        src = open(os.path.join(REF, relpath)).read().replace('"cuda"', '"cpu"').replace("'cuda'", "'cpu'")
        m = types.ModuleType(modname)
        m.__file__ = os.path.join(REF, relpath)
        sys.modules[modname] = m
        exec(compile(src, m.__file__, "exec"), m.__dict__)
        return m

    exec_module("utils.general_utils", "utils/general_utils.py")
    for dep in ("utils.system_utils", "utils.sh_utils", "utils.graphics_utils"):
        exec_module(dep, dep.replace(".", "/") + ".py")
    exec_module("scene.OurAdam", "scene/OurAdam.py")
    gm = exec_module("scene.gaussian_model", "scene/gaussian_model.py")
    return gm.GaussianModel, sys.modules["scene.OurAdam"].Adam


def make_case(n, seed, scaffold, with_state):
    g = torch.Generator().manual_seed(seed)
    p = {"xyz": torch.randn(n, 3, generator=g) * 5,
         "f_dc": torch.randn(n, 1, 3, generator=g),
         "f_rest": torch.randn(n, 15, 3, generator=g) * 0.1,
         "opacity": torch.randn(n, 1, generator=g) * 2.5,
         "scaling": torch.log(torch.rand(n, 3, generator=g) * 0.05 + 0.002),
         "rotation": torch.randn(n, 4, generator=g)}
    accum = torch.rand(n, 1, generator=g) * 0.001
    accum[::97] = float("nan")
    radii = torch.randint(0, 60, (n,), generator=g).float()
    denom = torch.randint(0, 5, (n, 1), generator=g).float()
    moments = {k: (torch.randn_like(v) * 0.01, torch.rand_like(v) * 0.001) for k, v in p.items()} if with_state else None
    # one statistics step in front
    vis = torch.randperm(n, generator=g)[: n // 3].sort().values
    vs_grad = torch.randn(n, 3, generator=g) * 0.0005
    vis_radii = torch.randint(1, 80, (vis.numel(),), generator=g).int()
    return p, accum, radii, denom, moments, vis, vs_grad, vis_radii, scaffold


def main():
    GaussianModel, Adam = load_reference_class()
    out = {}
    cases = [("a", 400, 1, None, True), ("b", 300, 2, 40, True), ("c", 160, 3, None, False)]
    for tag, n, seed, scaffold, with_state in cases:
        p, accum, radii, denom, moments, vis, vs_grad, vis_radii, scaffold = make_case(n, seed, scaffold, with_state)
        m = GaussianModel(3)
        m._xyz = torch.nn.Parameter(p["xyz"].clone())
        m._features_dc = torch.nn.Parameter(p["f_dc"].clone())
        m._features_rest = torch.nn.Parameter(p["f_rest"].clone())
        m._opacity = torch.nn.Parameter(p["opacity"].clone())
        m._scaling = torch.nn.Parameter(p["scaling"].clone())
        m._rotation = torch.nn.Parameter(p["rotation"].clone())
        m.percent_dense = 0.01
        m.scaffold_points = scaffold
        m.xyz_gradient_accum = accum.clone()
        m.xyz_gradient_accum_depth = torch.zeros_like(accum)
        m.denom = denom.clone()
        m.max_radii2D = radii.clone()
        groups = [{"params": [getattr(m, a)], "lr": 1e-3, "name": nme} for a, nme in
                  (("_xyz", "xyz"), ("_features_dc", "f_dc"), ("_features_rest", "f_rest"), ("_opacity", "opacity"),
                   ("_scaling", "scaling"), ("_rotation", "rotation"))]
        m.optimizer = Adam(groups, lr=0.0, eps=1e-15)
        if with_state:
            for grp in groups:
                q = grp["params"][0]
                m.optimizer.state[q] = {"step": torch.tensor(3.0), "exp_avg": moments[grp["name"]][0].clone(),
                                        "exp_avg_sq": moments[grp["name"]][1].clone()}
        # the statistics step exactly as train_single.py:193-194 writes it
        vp = types.SimpleNamespace(grad=vs_grad)
        m.max_radii2D[vis] = torch.max(m.max_radii2D[vis], vis_radii)
        m.add_densification_stats(vp, vis, False)
        out[f"{tag}_stats_accum"] = m.xyz_gradient_accum.numpy().copy()
        out[f"{tag}_stats_denom"] = m.denom.numpy().copy()
        out[f"{tag}_stats_radii"] = m.max_radii2D.numpy().copy()
        # densify_and_prune with the sampler's output recorded
        drawn = {}
        orig_normal = torch.normal

        def rec_normal(*a, **k):
            r = orig_normal(*a, **k)
            drawn["samples"] = r.detach().clone()
            return r

        torch.manual_seed(100 + seed)
        torch.normal = rec_normal
        try:
            m.densify_and_prune(0.0002, 0.005, 4.0, False)
        finally:
            torch.normal = orig_normal
        for k, v in p.items():
            out[f"{tag}_in_{k}"] = v.numpy()
        out[f"{tag}_in_accum"], out[f"{tag}_in_radii"], out[f"{tag}_in_denom"] = accum.numpy(), radii.numpy(), denom.numpy()
        out[f"{tag}_in_vis"], out[f"{tag}_in_vs_grad"], out[f"{tag}_in_vis_radii"] = vis.numpy(), vs_grad.numpy(), vis_radii.numpy()
        out[f"{tag}_in_scaffold"] = np.int64(-1 if scaffold is None else scaffold)
        if with_state:
            for k in p:
                out[f"{tag}_in_m_{k}"], out[f"{tag}_in_v_{k}"] = moments[k][0].numpy(), moments[k][1].numpy()
        out[f"{tag}_samples"] = drawn["samples"].numpy()
        for a, nme in (("_xyz", "xyz"), ("_features_dc", "f_dc"), ("_features_rest", "f_rest"), ("_opacity", "opacity"),
                       ("_scaling", "scaling"), ("_rotation", "rotation")):
            q = getattr(m, a)
            out[f"{tag}_out_{nme}"] = q.detach().numpy().copy()
            st = m.optimizer.state.get(q)
            if st:
                out[f"{tag}_out_m_{nme}"], out[f"{tag}_out_v_{nme}"] = st["exp_avg"].numpy().copy(), st["exp_avg_sq"].numpy().copy()
        out[f"{tag}_out_radii"] = m.max_radii2D.numpy().copy()
        out[f"{tag}_out_accum"] = m.xyz_gradient_accum.numpy().copy()
        out[f"{tag}_out_newly_split"] = m.newly_split_points_mask.numpy().copy()
        op = torch.sigmoid(p["opacity"]).flatten()
        print(tag, "rows", n, "->", m._xyz.shape[0], "split children", int(m.newly_split_points_mask.sum()),
              "low-opacity rows", int((op < 0.005).sum()), "clones", m._xyz.shape[0] - n + int((op < 0.005).sum())
              - int(m.newly_split_points_mask.sum()) // 2)
    np.savez_compressed(os.path.join(HERE, "densify_ref.npz"), **out)


if __name__ == "__main__":
    main()
