"""Comparison of two hierarchy builds (ours vs the reference Creator / the numpy oracle).

Integer outputs (node table) must be identical and boxes bit-equal.  Gaussian rows are compared as ellipsoids
(covariance from log-scale + rotation) and, when the trees were rotation-aligned, also axis by axis: the aligner
makes every non-root rotation canonical RELATIVE TO ITS PARENT, so two correct builds differ by one signed axis
permutation D inherited from the root's eigenvectors (R_a D = R_b for every node, scales permuted alike);
D is read off the two roots."""
import numpy as np

from oracle import hier_build as hb


def covariances(log_scales, rotations):
    q = np.asarray(rotations, np.float64)
    s = np.exp(np.asarray(log_scales, np.float64))
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)], -1),
                  np.stack([2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)], -1),
                  np.stack([2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1)], -2)
    T = R * s[:, None, :]
    return T @ np.swapaxes(T, 1, 2), R


def compare_builds(a, b, aligned, tol=1e-4):
    """a, b: dicts with positions, shs, opacities, log_scales, rotations, nodes, boxes.  Returns the error report;
    raises AssertionError on a mismatch."""
    assert np.array_equal(a["nodes"], b["nodes"]), "node tables differ"
    assert np.array_equal(a["boxes"].view(np.uint32), b["boxes"].view(np.uint32)), "boxes differ"
    rep = {}
    for k in ("positions", "shs", "opacities"):
        den = max(float(np.abs(b[k]).max()), 1e-30)
        rep[k] = float(np.abs(a[k].reshape(b[k].shape) - b[k]).max()) / den
        assert rep[k] <= tol, (k, rep[k])
    ca, Ra = covariances(a["log_scales"], a["rotations"])
    cb, Rb = covariances(b["log_scales"], b["rotations"])
    den = np.abs(cb).reshape(len(cb), -1).max(axis=1)
    rep["covariance"] = float((np.abs(ca - cb).reshape(len(cb), -1).max(axis=1) / den).max())
    assert rep["covariance"] <= tol, ("covariance", rep["covariance"])
    if aligned:
        D = np.rint(Ra[0].T @ Rb[0])
        assert np.array_equal(np.abs(D).sum(axis=0), np.ones(3)) and np.linalg.det(D) > 0, D
        perm = np.abs(D).argmax(axis=0)
        sa, sb = np.exp(a["log_scales"].astype(np.float64)), np.exp(b["log_scales"].astype(np.float64))
        # axes may only be compared where the ellipsoid has distinct axes: with two (nearly) equal scales the matched
        # permutation is decided by rounding noise
        s_sorted = np.sort(sb, axis=1)
        distinct = (np.minimum(s_sorted[:, 1] / s_sorted[:, 0], s_sorted[:, 2] / s_sorted[:, 1]) > 1.02)
        rot_err = np.abs(Ra @ D - Rb).reshape(len(Rb), -1).max(axis=1)
        sc_err = np.abs(sa[:, perm] - sb).max(axis=1) / sb.max(axis=1)
        rep["rotation_axes"] = float(rot_err[distinct].max())
        rep["scale_axes"] = float(sc_err[distinct].max())
        rep["distinct_frac"] = float(distinct.mean())
        assert rep["rotation_axes"] <= 2e-3, rep
        assert rep["scale_axes"] <= tol, rep
    return rep


def synthetic_gaussians(n, seed, extent=(100.0, 20.0, 8.0)):
    """Activated street-like Gaussians without ties in any coordinate (std::nth_element is free to order ties as it
    likes; see oracle/hier_build.py)."""
    rng = np.random.default_rng(seed)
    pos = (rng.random((n, 3)) * np.array(extent)).astype(np.float32)
    for a in range(3):
        while len(np.unique(pos[:, a])) != n:
            u, first = np.unique(pos[:, a], return_index=True)
            dup = np.setdiff1d(np.arange(n), first)
            pos[dup, a] = (rng.random(len(dup)) * extent[a]).astype(np.float32)
    scale = np.exp(rng.normal(-2.0, 0.7, (n, 3))).astype(np.float32)
    rot = rng.normal(size=(n, 4)).astype(np.float32)
    rot /= np.linalg.norm(rot, axis=1, keepdims=True)
    opacity = rng.uniform(0.05, 1.0, n).astype(np.float32)
    shs = rng.normal(size=(n, 48)).astype(np.float32)
    return dict(pos=pos, shs=shs, opacity=opacity, scale=scale, rot=rot)
