"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy, float32) of the tree stages of the reference's hierarchy
Creator, the checker of the GPU hierarchy build.  Never imported by the product path.

Restates, in the reference's order (GH = /root/reference/submodules/gaussianhierarchy):
  kd_tree      GH/PointbasedKdTreeGenerator.cpp:16-73   recursive median split along the longest side of the
                                                       node's bounds (positions +- 3 max scale)
  merge        GH/ClusterMerger.cpp:23-165              bottom-up moment-matched merge of the two children
  align        GH/rotation_aligner.cpp:14-114           top-down exhaustive rotation matching child -> parent
  flatten      GH/writer.cpp:18-115                     pre-order node table / Gaussian rows / boxes

Pinned against the reference itself: tests/golden/make_golden_hier_build.py runs oracle/_ref/ref_gh_tool `build`
(the reference's own classes compiled from its sources) on seeded inputs and tests/test_cpu_hier_build.py compares.

Two places where the reference's result is not a function of its input alone, and what is done about them:
  * std::nth_element leaves ties at the median in an implementation-defined order; here (and in the CUDA build)
    equal coordinates are ordered by Gaussian index.  The fixtures have no ties at any median.
  * Eigen's SelfAdjointEigenSolver fixes the signs of the eigenvectors by the path its QL iteration takes;
    numpy's LAPACK solver takes another.  A merged Gaussian is the same ellipsoid either way (compare covariances);
    after `align` every non-root rotation is canonical relative to its parent, so whole trees differ at most by one
    signed axis permutation inherited from the root (tests/hier_build_util.py resolves it).
"""
import numpy as np

F = np.float32


def surface(s):                                        # ClusterMerger.cpp:16-21 ellipseSurface
    return F(F(F(s[0] * s[1]) + F(s[0] * s[2])) + F(s[1] * s[2]))


def matrix_from_quat(q):                               # common.h:86-102 (w, x, y, z), not normalised
    s, x, y, z = (F(v) for v in q)
    one, two = F(1), F(2)
    return np.array([[one - two * (y * y + z * z), two * (x * y - s * z), two * (x * z + s * y)],
                     [two * (x * y + s * z), one - two * (x * x + z * z), two * (y * z - s * x)],
                     [two * (x * z - s * y), two * (y * z + s * x), one - two * (x * x + y * y)]], dtype=F)


def covariance(scale, rot):                            # common.h:104-133 computeCovariance
    T = matrix_from_quat(rot) * np.asarray(scale, dtype=F)[None, :]
    C = (T @ T.T).astype(F)
    return np.array([C[0, 0], C[0, 1], C[0, 2], C[1, 1], C[1, 2], C[2, 2]], dtype=F)


def eigen_quat(m):
    """Eigen::Quaternionf(Matrix3f) -> (w, x, y, z): the published conversion of Eigen/src/Geometry/Quaternion.h
    (the dependency is vendored by the reference under GH/dependencies/eigen)."""
    m = m.astype(F)
    t = F(m[0, 0] + m[1, 1] + m[2, 2])
    q = np.zeros(4, dtype=F)   # w, x, y, z
    if t > 0:
        t = F(np.sqrt(F(t + F(1))))
        q[0] = F(0.5) * t
        t = F(0.5) / t
        q[1] = (m[2, 1] - m[1, 2]) * t
        q[2] = (m[0, 2] - m[2, 0]) * t
        q[3] = (m[1, 0] - m[0, 1]) * t
    else:
        i = 0
        if m[1, 1] > m[0, 0]:
            i = 1
        if m[2, 2] > m[i, i]:
            i = 2
        j = (i + 1) % 3
        k = (j + 1) % 3
        t = F(np.sqrt(F(m[i, i] - m[j, j] - m[k, k] + F(1))))
        q[1 + i] = F(0.5) * t
        t = F(0.5) / t
        q[0] = (m[k, j] - m[j, k]) * t
        q[1 + j] = (m[j, i] + m[i, j]) * t
        q[1 + k] = (m[k, i] + m[i, k]) * t
    return q


class Tree:
    """Explicit tree in arrays, nodes in the reference's flattened order (writer.cpp:18-92 populateRec): node 0 is
    the root, the children of the k-th internal node met in pre-order are nodes 2k+1 and 2k+2, and the Gaussian of a
    node (leaf Gaussian or merged Gaussian) is row `start` = the node's pre-order rank."""

    def __init__(self, n_nodes):
        z = lambda *s, dt=F: np.zeros(s, dtype=dt)
        self.depth = z(n_nodes, dt=np.int32)
        self.parent = np.full(n_nodes, -1, dtype=np.int32)
        self.start = z(n_nodes, dt=np.int32)
        self.count_leafs = z(n_nodes, dt=np.int32)
        self.count_merged = z(n_nodes, dt=np.int32)
        self.start_children = z(n_nodes, dt=np.int32)
        self.count_children = z(n_nodes, dt=np.int32)
        self.boxes = z(n_nodes, 8)
        self.leaf_index = np.full(n_nodes, -1, dtype=np.int32)   # input Gaussian of a leaf node
        self.preorder = []                                       # node ids in visiting order

    def node_table(self):
        return np.stack([self.depth, self.parent, self.start, self.count_leafs, self.count_merged,
                         self.start_children, self.count_children], axis=1).astype(np.int32)


def kd_tree(pos, scale):
    """PointbasedKdTreeGenerator.cpp:16-73 + the node numbering of writer.cpp populateRec."""
    pos = np.asarray(pos, dtype=F)
    n = pos.shape[0]
    r = (F(3.0) * np.asarray(scale, dtype=F).max(axis=1)).astype(F)
    lo_all, hi_all = (pos - r[:, None]).astype(F), (pos + r[:, None]).astype(F)
    t = Tree(2 * n - 1)
    n_alloc = 1
    rows = 0
    stack = [(0, np.arange(n, dtype=np.int64))]
    while stack:
        nid, idx = stack.pop()
        t.preorder.append(nid)
        lo, hi = lo_all[idx].min(axis=0), hi_all[idx].max(axis=0)
        t.boxes[nid, 0:3], t.boxes[nid, 4:7] = lo, hi
        t.start[nid] = rows
        rows += 1
        t.start_children[nid] = n_alloc
        if idx.size == 1:
            t.depth[nid] = 0
            t.count_leafs[nid] = 1
            t.leaf_index[nid] = idx[0]
            continue
        t.count_merged[nid] = 1
        t.count_children[nid] = 2
        ext = (hi - lo).astype(F)
        axis, greatest = 0, F(0)
        for a in range(3):                              # :43-53, strict >
            if ext[a] > greatest:
                greatest, axis = ext[a], a
        order = idx[np.lexsort((idx, pos[idx, axis]))]  # ties at the median: by Gaussian index (see header)
        half = idx.size // 2                            # pivot + 1, pivot = num / 2 - 1
        c0, c1 = n_alloc, n_alloc + 1
        n_alloc += 2
        t.parent[c0] = t.parent[c1] = nid
        stack.append((c1, order[half:]))
        stack.append((c0, order[:half]))                # child 0 is visited first
    # depth = max(children) + 1, leaves 0 (:66) -- children have larger ids than their parent
    for nid in range(2 * n - 2, -1, -1):
        if t.count_children[nid]:
            c = t.start_children[nid]
            t.depth[nid] = max(t.depth[c], t.depth[c + 1]) + 1
    return t


def _eigh(m):
    w, v = np.linalg.eigh(m.astype(F))                  # ascending, like SelfAdjointEigenSolver
    return w.astype(F), v.astype(F)


def merge_pair(g0, g1):
    """ClusterMerger.cpp:44-139 for the two representatives of a node's children.  g = dict(position, shs, opacity,
    scale, rotation, covariance)."""
    gs = (g0, g1)
    w = [F(g["opacity"] * surface(g["scale"])) for g in gs]
    wsum = F(F(F(0) + w[0]) + w[1])
    a = [F(x / wsum) for x in w]
    position = np.zeros(3, dtype=F)
    shs = np.zeros(48, dtype=F)
    for g, ai in zip(gs, a):
        position = (position + ai * g["position"]).astype(F)
        shs = (shs + ai * g["shs"]).astype(F)
    cov = np.zeros(6, dtype=F)
    for g, ai in zip(gs, a):
        d = (g["position"] - position).astype(F)
        outer = np.array([d[0] * d[0], d[1] * d[0], d[2] * d[0], d[1] * d[1], d[2] * d[1], d[2] * d[2]], dtype=F)
        cov = (cov + ai * (g["covariance"] + outer).astype(F)).astype(F)
    m = np.array([[cov[0], cov[1], cov[2]], [cov[1], cov[3], cov[4]], [cov[2], cov[4], cov[5]]], dtype=F)
    lam, vec = _eigh(m)
    if np.isnan(lam).any():
        raise RuntimeError("Found Nans!")
    eps = np.finfo(F).eps
    loops = 0
    while (lam == 0).any():                             # :90-104
        for i in range(3):
            m[i, i] = m[i, i] + max(F(m[i, i] * F(0.0001)), eps)
        lam, vec = _eigh(m)
        loops += 1
        if loops > 1000:
            raise RuntimeError("degenerate covariance")
    if np.dot(np.cross(vec[:, 0], vec[:, 1]), vec[:, 2]) < 0:   # :110-112
        vec[:, 2] = -vec[:, 2]
    scale = np.sqrt(np.abs(lam)).astype(F)
    rotation = eigen_quat(vec)                          # :127-128 (the quatFromMatrix result of :122-125 is overwritten)
    opacity = F(wsum / surface(scale))
    return dict(position=position, shs=shs, opacity=opacity, scale=scale, rotation=rotation, covariance=cov)


def match_exhaustive(ref_rot, rot, scale):
    """rotation_aligner.cpp:23-88: the right-handed signed column permutation of R(rot) closest to R(ref_rot).
    -> (rotation, scale) of the matched Gaussian."""
    Rref = matrix_from_quat(ref_rot)
    Rm = matrix_from_quat(rot)
    best_score, best, perm = F(0), None, None
    for i in range(3):
        for j in range(3):
            if j == i:
                continue
            for k in range(3):
                if k == j or k == i:
                    continue
                base = np.stack([Rm[:, i], Rm[:, j], Rm[:, k]], axis=1)
                for state in range(8):
                    sign = np.array([-1 if state & (1 << w) else 1 for w in range(3)], dtype=F)
                    test = (base * sign[None, :]).astype(F)
                    if np.dot(np.cross(test[:, 0], test[:, 1]), test[:, 2]) < 0:
                        continue
                    score = F(0)
                    for r_ in range(3):                 # frobenius(): row-major accumulation order
                        for c_ in range(3):
                            score = F(score + F(test[r_, c_] * Rref[r_, c_]))
                    if score > best_score:
                        best_score, best, perm = score, test, (i, j, k)
    if best is None:
        raise RuntimeError("no candidate with a positive score")
    return eigen_quat(best), np.array([scale[perm[0]], scale[perm[1]], scale[perm[2]]], dtype=F)


def build(pos, shs, opacity, scale, rot, align=True):
    """The Creator's tree stages on ACTIVATED Gaussians -> dict of the arrays Writer::makeHierarchy produces
    (writer.cpp:94-115): positions / shs / opacities / log_scales / rotations [2N-1, ...], nodes int32 [2N-1, 7],
    boxes float32 [2N-1, 8], plus `covariances` (the moment-matched covariance each merge accumulated; leaves: from
    scale and rotation) and `leaf_order` (input index of every leaf row, -1 for merged rows)."""
    pos, shs, opacity = np.asarray(pos, F), np.asarray(shs, F).reshape(len(pos), 48), np.asarray(opacity, F).reshape(-1)
    scale, rot = np.asarray(scale, F), np.asarray(rot, F)
    n = pos.shape[0]
    t = kd_tree(pos, scale)
    nn = 2 * n - 1
    rep = [None] * nn
    for nid in range(nn):
        if t.count_leafs[nid]:
            i = t.leaf_index[nid]
            rep[nid] = dict(position=pos[i].copy(), shs=shs[i].copy(), opacity=F(opacity[i]), scale=scale[i].copy(),
                            rotation=rot[i].copy(), covariance=covariance(scale[i], rot[i]))
    for nid in range(nn - 1, -1, -1):                   # children have larger ids: bottom-up
        if t.count_children[nid]:
            c = t.start_children[nid]
            rep[nid] = merge_pair(rep[c], rep[c + 1])
            d = (t.boxes[nid, 4:7] - t.boxes[nid, 0:3]).astype(F)   # :161-163
            t.boxes[nid, 3] = t.boxes[nid, 7] = max(d[0], d[1], d[2])
    if align:
        for nid in t.preorder:                          # rotation_aligner.cpp:90-109, top-down
            if t.count_children[nid]:
                c = t.start_children[nid]
                for ch in (c, c + 1):
                    rep[ch]["rotation"], rep[ch]["scale"] = match_exhaustive(rep[nid]["rotation"], rep[ch]["rotation"],
                                                                             rep[ch]["scale"])
    out = dict(positions=np.zeros((nn, 3), F), shs=np.zeros((nn, 48), F), opacities=np.zeros(nn, F),
               log_scales=np.zeros((nn, 3), F), rotations=np.zeros((nn, 4), F), covariances=np.zeros((nn, 6), F),
               leaf_order=np.full(nn, -1, np.int32))
    for nid in range(nn):
        r, g = t.start[nid], rep[nid]
        out["positions"][r], out["shs"][r], out["opacities"][r] = g["position"], g["shs"], g["opacity"]
        out["log_scales"][r], out["rotations"][r] = np.log(g["scale"]).astype(F), g["rotation"]
        out["covariances"][r] = g["covariance"]
        out["leaf_order"][r] = t.leaf_index[nid]
    out["nodes"], out["boxes"] = t.node_table(), t.boxes
    return out
