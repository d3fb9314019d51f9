"""Densification kernels (csrc/densify.cu via h3dgs_densify.py) against the vectors the reference's own
GaussianModel methods produced (tests/golden/densify_ref.npz) and against the torch restatement on the GPU at a size
where the old / clone / split bookkeeping is exercised on a hundred thousand rows."""
import types

import pytest
import torch

from harness import add_paths
from oracle import densify_reference as dr
from tests.test_cpu_densify_oracle import load_case

add_paths()

pytestmark = pytest.mark.gpu
DEV = "cuda"
ATTR = dict(xyz="_xyz", f_dc="_features_dc", f_rest="_features_rest", opacity="_opacity", scaling="_scaling", rotation="_rotation")


def _model(state, scaffold, percent_dense=0.01):
    from scene.OurAdam import Adam
    m = types.SimpleNamespace()
    for k, a in ATTR.items():
        setattr(m, a, torch.nn.Parameter(state["params"][k].clone()))
    m.percent_dense, m.scaffold_points, m.opacity_activation = percent_dense, scaffold, torch.sigmoid
    m.xyz_gradient_accum = state["xyz_gradient_accum"].clone()
    m.xyz_gradient_accum_depth = torch.zeros_like(m.xyz_gradient_accum)
    m.denom, m.max_radii2D = state["denom"].clone(), state["max_radii2D"].clone()
    groups = [{"params": [getattr(m, a)], "lr": 1e-3, "name": k} for k, a in ATTR.items()]
    m.optimizer = Adam(groups, lr=0.0, eps=1e-15)
    for k, a in ATTR.items():
        if state["exp_avg"][k] is not None:
            m.optimizer.state[getattr(m, a)] = {"step": torch.tensor(3.0), "exp_avg": state["exp_avg"][k].clone(),
                                                "exp_avg_sq": state["exp_avg_sq"][k].clone()}
    return m


def _close(a, b, tol):
    return float((a - b).abs().max()) <= tol * max(float(b.abs().max()), 1e-30)


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_matches_the_reference_methods_fixture(tag):
    import h3dgs_densify as hd
    state, x, want = load_case(tag, DEV)
    m = _model(state, x["scaffold"])
    vp = types.SimpleNamespace(grad=x["vs_grad"])
    hd.update_max_radii_and_stats(m, vp, x["vis"], x["vis_radii"])
    a, b = m.xyz_gradient_accum, want["stats_accum"]
    assert torch.equal(a.isnan(), b.isnan()) and _close(a.nan_to_num(0), b.nan_to_num(0), 2e-7)
    assert torch.equal(m.denom, want["stats_denom"]) and torch.equal(m.max_radii2D, want["stats_radii"])
    m.xyz_gradient_accum = want["stats_accum"].clone()   # the decisions below hang on these bits: same inputs for both
    info = hd.densify_and_prune(m, 0.0002, 0.005, 4.0, samples=x["samples"])
    for k, attr in ATTR.items():
        got, ref = getattr(m, attr).detach(), want[f"out_{k}"]
        assert got.shape == ref.shape, (k, got.shape, ref.shape)
        if k in ("xyz", "scaling"):      # split children: R sample + xyz, log(exp(s) / 1.6) -- float arithmetic
            assert _close(got, ref, 2e-6), k
            keep = info["kept"] + info["clones"]
            assert torch.equal(got[:keep], ref[:keep]), k
        else:
            assert torch.equal(got, ref), k
        if x["has_state"]:
            st = m.optimizer.state[getattr(m, attr)]
            assert torch.equal(st["exp_avg"], want[f"out_m_{k}"]) and torch.equal(st["exp_avg_sq"], want[f"out_v_{k}"]), k
            assert float(st["step"]) == 3.0
        assert m.optimizer.param_groups[list(ATTR).index(k)]["params"][0] is getattr(m, attr)
    assert torch.equal(m.max_radii2D, want["out_radii"]) and torch.equal(m.xyz_gradient_accum, want["out_accum"])
    assert int(m.newly_split_points_mask.sum()) == 2 * info["splits"]


def test_matches_the_restatement_at_scale_with_its_own_samples():
    import h3dgs_densify as hd
    n = 120_000
    g = torch.Generator().manual_seed(5)
    p = {"xyz": torch.randn(n, 3, generator=g) * 30, "f_dc": torch.randn(n, 1, 3, generator=g),
         "f_rest": torch.randn(n, 15, 3, generator=g) * 0.1, "opacity": torch.randn(n, 1, generator=g) * 2.5,
         "scaling": torch.log(torch.rand(n, 3, generator=g) * 0.05 + 0.002), "rotation": torch.randn(n, 4, generator=g)}
    accum = torch.rand(n, 1, generator=g) * 0.001
    state = {"params": {k: v.to(DEV) for k, v in p.items()},
             "exp_avg": {k: (torch.randn_like(v) * 0.01).to(DEV) for k, v in p.items()},
             "exp_avg_sq": {k: (torch.rand_like(v) * 0.001).to(DEV) for k, v in p.items()},
             "xyz_gradient_accum": accum.to(DEV), "denom": torch.ones(n, 1, device=DEV),
             "max_radii2D": torch.randint(0, 60, (n,), generator=g).float().to(DEV)}
    m = _model(state, 5000)
    torch.manual_seed(77)
    info = hd.densify_and_prune(m, 0.0002, 0.005, 4.0)
    torch.manual_seed(77)
    drawn = dr.densify_and_prune(state, 0.0002, 0.005, 4.0, 0.01, 5000)
    assert torch.equal(drawn, info["samples"]), "the random stream was not consumed like the reference's"
    assert info["clones"] > 1000 and info["splits"] > 1000 and info["kept"] < n
    for k, attr in ATTR.items():
        got, ref = getattr(m, attr).detach(), state["params"][k]
        assert got.shape == ref.shape
        assert _close(got, ref, 2e-6) if k in ("xyz", "scaling") else torch.equal(got, ref), k
        st = m.optimizer.state[getattr(m, attr)]
        assert torch.equal(st["exp_avg"], state["exp_avg"][k]) and torch.equal(st["exp_avg_sq"], state["exp_avg_sq"][k]), k


def test_statistics_step_matches_torch_ops():
    import h3dgs_densify as hd
    n = 300_000
    g = torch.Generator().manual_seed(9)
    state = {"xyz_gradient_accum": (torch.rand(n, 1, generator=g) * 1e-3).to(DEV), "denom": torch.zeros(n, 1, device=DEV),
             "max_radii2D": torch.randint(0, 50, (n,), generator=g).float().to(DEV)}
    m = types.SimpleNamespace(**{k: v.clone() for k, v in state.items()})
    vis = torch.randperm(n, generator=g)[:90_000].sort().values.to(DEV)
    grad = (torch.randn(n, 3, generator=g) * 1e-3).to(DEV)
    radii = torch.randint(1, 99, (90_000,), generator=g).int().to(DEV)
    hd.update_max_radii_and_stats(m, types.SimpleNamespace(grad=grad), vis, radii)
    dr.add_stats(state, grad, vis, radii)
    assert _close(m.xyz_gradient_accum, state["xyz_gradient_accum"], 2e-7)
    assert torch.equal(m.denom, state["denom"]) and torch.equal(m.max_radii2D, state["max_radii2D"])


def test_nothing_selected_and_everything_split():
    """Edge cases of the selection: a threshold nothing reaches (only the opacity prune acts; no samples are drawn) and
    a threshold everything reaches with a zero dense extent (every unpruned row above the opacity bound splits)."""
    import h3dgs_densify as hd
    state, x, want = load_case("c", DEV)
    n = state["params"]["xyz"].shape[0]
    op = torch.sigmoid(state["params"]["opacity"]).flatten()
    m = _model(state, None)
    info = hd.densify_and_prune(m, 1e9, 0.005, 4.0)
    assert info["clones"] == 0 and info["splits"] == 0 and info["samples"].shape == (0, 3)
    keep = op >= 0.005
    assert info["kept"] == int(keep.sum()) == m._xyz.shape[0]
    assert torch.equal(m._xyz.detach(), state["params"]["xyz"][keep])
    assert m._features_rest not in m.optimizer.state   # case c carries no optimizer state: nothing to hand over
    m = _model(state, None, percent_dense=0.0)
    state["xyz_gradient_accum"][:] = 1.0
    m.xyz_gradient_accum = state["xyz_gradient_accum"].clone()
    m.max_radii2D = torch.full((n,), 10.0, device=DEV)
    torch.manual_seed(1)
    info = hd.densify_and_prune(m, 1e-9, 0.005, 4.0)
    sel = op > 0.15
    assert info["splits"] == int(sel.sum()) and info["clones"] == 0
    assert info["kept"] == int(((op >= 0.005) & ~sel).sum())
    assert m._xyz.shape[0] == info["kept"] + 2 * info["splits"]
    assert torch.isfinite(m._xyz).all() and torch.isfinite(m._scaling).all()
