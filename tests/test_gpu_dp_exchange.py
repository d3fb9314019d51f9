"""Data-parallel gradient exchange (csrc/dp_exchange.cu, h3dgs_parallel.CompactRowExchange) on ONE GPU: several
"ranks" stand side by side in one process (the accumulate kernel only loads from the blocks it is given, so a peer's
block in the same memory exercises the same code as one behind NVLink).  Checked against the dense formulation it
replaces: sum of the ranks' dense gradients, / world, relevant = rows with a non-zero opacity gradient on some rank,
Adam.step(relevant) (train_post.py:171-190) -- parameters and moments bit for bit (world a power of two: the
average is then an exact scaling on both sides).  The 2-GPU run of the same comparison over real peer memory is
harness/dp_check.py (torchrun; log under profiles/)."""
import pytest
import torch

from harness import add_paths, scenes, workload

add_paths()

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _renders(world, levels=13, sky=300, w=640, h=480, target=4000):
    import diff_gaussian_rasterization as dgr
    import gaussian_hierarchy
    from scene.OurAdam import Adam
    attrs = workload.build_scene(levels, sky, 3, DEV)
    model = workload.HierModel(attrs, sky)
    opt = Adam(model.param_groups(), lr=0.0, eps=1e-15)
    cams = [scenes.street_camera(w, h, seed=20 + r, device=DEV, yaw_deg=10.0 + 70.0 * r, x=-20.0 + 9.0 * r)
            for r in range(world)]
    limit = workload.find_limit(gaussian_hierarchy._C, model, cams[0], target, DEV)
    n = attrs["means3D"].size(0)
    bg = torch.zeros(3, device=DEV)
    g = torch.Generator().manual_seed(5)
    dcol = (torch.randn(3, h, w, generator=g) / (w * h)).to(DEV)
    sgs = []
    for c in cams:
        cut = workload.Cut(n, DEV)
        cut.compute(gaussian_hierarchy._C, model, c, limit)
        sg = dgr.SparseGradients()
        out = workload.ours_render_post(c, model, bg, cut, do_depth=False, sparse_grads=sg)
        (out["render"] * dcol).sum().backward()
        assert sg.ready and int(sg.count) > 100
        sgs.append(sg)
    return model, opt, sgs, n


def _dense_step(model, sgs, world):
    """The formulation being replaced, on a deep copy of the model, from the SAME compact gradients."""
    import copy
    from scene.OurAdam import Adam
    ref = copy.copy(model)
    names = ["_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"]
    for nme in names:
        setattr(ref, nme, getattr(model, nme).detach().clone().requires_grad_(True))
    opt = Adam(ref.param_groups(), lr=0.0, eps=1e-15)
    by_ptr = {getattr(model, nme).data_ptr(): nme for nme in names}
    for nme in names:
        getattr(ref, nme).grad = torch.zeros_like(getattr(ref, nme))
    relevant = torch.zeros(ref._xyz.size(0), dtype=torch.bool, device=DEV)
    for sg in sgs:   # rank order
        for k in sg.grads:
            nme = by_ptr[sg.params[k].data_ptr()]
            U = int(sg.count)
            gate = sg.grads["opacity"][:U].flatten() != 0
            rows = sg.rows[:U][gate]
            gk = sg.grads[k][:U][gate].reshape((rows.numel(),) + tuple(getattr(ref, nme).shape[1:]))
            getattr(ref, nme).grad.index_add_(0, rows, gk)   # rows are unique within a rank: a plain add per element
            relevant[rows] = True
    for nme in names:
        getattr(ref, nme).grad.mul_(1.0 / world)
    opt.step(relevant.nonzero().flatten())
    return ref, opt, int(relevant.sum())


@pytest.mark.parametrize("world", [1, 2, 4])
def test_exchange_matches_dense_allreduce_and_adam(world):
    import h3dgs_parallel as par
    model, opt, sgs, n = _renders(world)
    ref, ref_opt, n_rel = _dense_step(model, sgs, world)
    cap = max(sg.capacity for sg in sgs)
    ex = par.CompactRowExchange(opt, n, cap, group=None, union_capacity=min(n, cap * world))
    arenas = [ex.arena] + [par.RowArena(cap, ex.width, torch.device(DEV)) for _ in range(world - 1)]
    own = ex.arena
    for r, sg in enumerate(sgs):
        ex.arena = arenas[r]
        ex.pack(sg, parity=0)
    ex.arena = own
    ex.accumulate([a.base_ptrs[0] for a in arenas], parity=0)
    assert int(ex.flags[0]) == n_rel, "union size differs from the dense `relevant`"
    ex.apply(world=world)
    torch.cuda.synchronize()
    ex.check_overflow()
    assert bool((ex.row_map == -1).all()), "row map not restored"
    names = ["_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"]
    for nme in names:
        a, b = getattr(model, nme).detach(), getattr(ref, nme).detach()
        assert torch.equal(a.view(torch.int32), b.view(torch.int32)), f"{nme}: parameters differ from the dense step"
        sa, sb = opt.state[getattr(model, nme)], ref_opt.state[getattr(ref, nme)]
        assert torch.equal(sa["exp_avg"].view(torch.int32), sb["exp_avg"].view(torch.int32)), nme
        assert torch.equal(sa["exp_avg_sq"].view(torch.int32), sb["exp_avg_sq"].view(torch.int32)), nme


def test_sharded_optimizer_exchange_matches_dense(world=2):
    """ZeRO-1 over rows: rank r sums and updates the rows it owns ((row >> 6) % world == r), publishes the updated parameter rows,
    and copies its peers' blocks.  Two replicas side by side in one process: both must end with the parameters of the
    dense formulation, and every row's moments must be right on its owner."""
    import copy
    import h3dgs_parallel as par
    from scene.OurAdam import Adam
    model, opt, sgs, n = _renders(world)
    ref, ref_opt, n_rel = _dense_step(model, sgs, world)
    names = ["_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"]
    cap = max(sg.capacity for sg in sgs)
    replicas, exs = [], []
    for r in range(world):
        m = copy.copy(model)
        for nme in names:
            setattr(m, nme, getattr(model, nme).detach().clone().requires_grad_(True))
        o = Adam(m.param_groups(), lr=0.0, eps=1e-15)
        ex = par.CompactRowExchange(o, n, cap, group=None, union_capacity=n // world + 1)
        ex.sharded, ex.world, ex.rank = True, world, r
        ex.pblock = par.RowArena(ex.ucap, ex.width, torch.device(DEV))
        replicas.append((m, o))
        exs.append(ex)
    # stage 1 on every "rank": its own gradients into its own arena (the compact gradients refer to `model`'s
    # parameters: the pack only needs their widths and order, taken from the replica's optimizer)
    for r, ex in enumerate(exs):
        sg = sgs[r]
        remap = copy.copy(sg)
        remap.params = {k: getattr(replicas[r][0], [nme for nme in names if getattr(model, nme).data_ptr() == v.data_ptr()][0])
                        for k, v in sg.params.items()}
        ex.pack(remap, parity=0)
    bases = [ex.arena.base_ptrs[0] for ex in exs]
    for ex in exs:
        ex.accumulate(bases, parity=0)
        ex.apply(world=world)
    assert sum(int(ex.flags[0]) for ex in exs) == n_rel
    pbases = [ex.pblock.base_ptrs[0] for ex in exs]
    for ex in exs:
        ex.pblock.base_ptrs = pbases
        ex.pull_params()
    torch.cuda.synchronize()
    for r, (m, o) in enumerate(replicas):
        for nme in names:
            a, b = getattr(m, nme).detach(), getattr(ref, nme).detach()
            assert torch.equal(a.view(torch.int32), b.view(torch.int32)), f"replica {r} {nme}: parameters differ"
            sa, sb = o.state[getattr(m, nme)], ref_opt.state[getattr(ref, nme)]
            own = par.owner_of(torch.arange(n, device=DEV), world) == r
            for key in ("exp_avg", "exp_avg_sq"):
                assert torch.equal(sa[key][own].view(torch.int32), sb[key][own].view(torch.int32)), (r, nme, key)
        assert bool((exs[r].row_map == -1).all())


def test_exchange_reports_a_too_small_arena():
    import h3dgs_parallel as par
    model, opt, sgs, n = _renders(1)
    ex = par.CompactRowExchange(opt, n, 64, group=None, union_capacity=64)
    ex.step(sgs[0])
    torch.cuda.synchronize()
    with pytest.raises(RuntimeError, match="capacity exceeded"):
        ex.check_overflow()
