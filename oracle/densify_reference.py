"""TEST INFRASTRUCTURE ONLY -- torch restatement (device-agnostic) of the reference's densification path, the checker
of csrc/densify.cu.  Never imported by the product path.

Restates, operation by operation (SGM = /root/reference/scene/gaussian_model.py):
  add_stats           train_single.py:193 + SGM:780-794
  densify_and_prune   SGM:736-778 -> densify_and_clone (:712-734), densify_and_split (:672-710),
                      cat_tensors_to_optimizer (:610-650), prune_points / _prune_optimizer (:560-608)
on a plain dict state: params {name: tensor}, exp_avg / exp_avg_sq {name: tensor or None}, xyz_gradient_accum [N,1],
denom [N,1], max_radii2D [N].  `normal` is the sampler of the split children (torch.normal in the reference); the
fixtures record what it returned so that every implementation can be fed the same draws.

Pinned: tests/golden/make_golden_densify.py executes the reference's OWN GaussianModel methods (source read from
/root/reference, "cuda" -> "cpu") on seeded inputs; tests/test_cpu_densify_oracle.py compares this file with them."""
import torch

NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def build_rotation(r):                                   # utils/general_utils.py:82-103
    norm = torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = torch.zeros((q.size(0), 3, 3), device=r.device)
    r_, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - r_ * z)
    R[:, 0, 2] = 2 * (x * z + r_ * y)
    R[:, 1, 0] = 2 * (x * y + r_ * z)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - r_ * x)
    R[:, 2, 0] = 2 * (x * z - r_ * y)
    R[:, 2, 1] = 2 * (y * z + r_ * x)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def add_stats(state, viewspace_grad, update_filter, radii=None):
    """train_single.py:193 (when `radii` is given) and SGM:780-794."""
    if radii is not None:
        state["max_radii2D"][update_filter] = torch.max(state["max_radii2D"][update_filter], radii)
    norm_value = torch.norm(viewspace_grad[update_filter, :2], dim=-1, keepdim=True)
    max_value = torch.max(norm_value, state["xyz_gradient_accum"][update_filter])
    state["xyz_gradient_accum"][update_filter] = max_value
    state["denom"][update_filter] += 1


def _cat(state, new):                                    # SGM:610-650, 652-671
    for k in NAMES:
        ext = new[k]
        if state["exp_avg"][k] is not None:
            state["exp_avg"][k] = torch.cat((state["exp_avg"][k], torch.zeros_like(ext)), dim=0)
            state["exp_avg_sq"][k] = torch.cat((state["exp_avg_sq"][k], torch.zeros_like(ext)), dim=0)
        state["params"][k] = torch.cat((state["params"][k], ext), dim=0)
    n = state["params"]["xyz"].shape[0]
    dev = state["params"]["xyz"].device
    state["xyz_gradient_accum"] = torch.zeros((n, 1), device=dev)
    state["denom"] = torch.zeros((n, 1), device=dev)
    state["max_radii2D"] = torch.cat((state["max_radii2D"], torch.zeros((new["xyz"].shape[0]), device=dev)))


def _prune(state, mask):                                 # SGM:560-608
    valid = ~mask
    for k in NAMES:
        if state["exp_avg"][k] is not None:
            state["exp_avg"][k] = state["exp_avg"][k][valid]
            state["exp_avg_sq"][k] = state["exp_avg_sq"][k][valid]
        state["params"][k] = state["params"][k][valid]
    state["xyz_gradient_accum"] = state["xyz_gradient_accum"][valid]
    state["denom"] = state["denom"][valid]
    state["max_radii2D"] = state["max_radii2D"][valid]


def densify_and_prune(state, max_grad, min_opacity, extent, percent_dense, scaffold_points=None,
                      opacity_activation=torch.sigmoid, normal=torch.normal):
    """SGM:736-778 without the optional ground-truth point-cloud constraint (SGM:757-758, needs FAISS)."""
    p = state["params"]
    dev = p["xyz"].device
    grads = state["xyz_gradient_accum"]
    grads[grads.isnan()] = 0.0
    get_opacity = lambda: opacity_activation(state["params"]["opacity"])
    get_scaling = lambda: torch.exp(state["params"]["scaling"])
    # densify_and_clone (:712-734)
    sel = torch.where(torch.norm(grads, dim=-1) * state["max_radii2D"] * torch.pow(get_opacity().flatten(), 1 / 5.0) >= max_grad,
                      True, False)
    sel = torch.logical_and(sel, get_opacity().flatten() > 0.15)
    sel = torch.logical_and(sel, torch.max(get_scaling(), dim=1).values <= percent_dense * extent)
    if scaffold_points is not None:
        sel[:scaffold_points] = False
    _cat(state, {k: state["params"][k][sel] for k in NAMES})
    # densify_and_split (:672-710)
    N = 2
    n_init = state["params"]["xyz"].shape[0]
    padded = torch.zeros((n_init), device=dev)
    padded[:grads.shape[0]] = grads.squeeze()
    sel = torch.where(padded * state["max_radii2D"] * torch.pow(get_opacity().flatten(), 1 / 5.0) >= max_grad, True, False)
    sel = torch.logical_and(sel, get_opacity().flatten() > 0.15)
    sel = torch.logical_and(sel, torch.max(get_scaling(), dim=1).values > percent_dense * extent)
    if scaffold_points is not None:
        sel[:scaffold_points] = False
    stds = get_scaling()[sel].repeat(N, 1)
    means = torch.zeros((stds.size(0), 3), device=dev)
    samples = normal(mean=means, std=stds)
    rots = build_rotation(state["params"]["rotation"][sel]).repeat(N, 1, 1)
    new = {"xyz": torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + state["params"]["xyz"][sel].repeat(N, 1),
           "scaling": torch.log(get_scaling()[sel].repeat(N, 1) / (0.8 * N)),
           "rotation": state["params"]["rotation"][sel].repeat(N, 1),
           "f_dc": state["params"]["f_dc"][sel].repeat(N, 1, 1),
           "f_rest": state["params"]["f_rest"][sel].repeat(N, 1, 1),
           "opacity": state["params"]["opacity"][sel].repeat(N, 1)}
    num_new = N * int(sel.sum())
    _cat(state, new)
    _prune(state, torch.cat((sel, torch.zeros(N * int(sel.sum()), device=dev, dtype=bool))))
    newly_split = torch.zeros(state["params"]["xyz"].shape[0], dtype=bool, device=dev)
    if num_new > 0:
        newly_split[-num_new:] = True
    # opacity prune (:752-755, :772)
    prune_mask = (get_opacity() < min_opacity).squeeze()
    if scaffold_points is not None:
        prune_mask[:scaffold_points] = False
    _prune(state, prune_mask)
    state["max_radii2D"] = torch.zeros((state["params"]["xyz"].shape[0]), device=dev)
    state["newly_split_points_mask"] = newly_split   # as the reference leaves it: sized BEFORE the opacity prune
    return samples
