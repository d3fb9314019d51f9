"""Diagnostic: component hashes of ONE forward in a fresh process, per staging mode (H3DGS_STAGE), repeated K times.
    python -m harness.diag_stage_modes [K]"""
import os
import subprocess
import sys

CODE = r"""
import sys, hashlib, torch
from harness import add_paths; add_paths()
from harness import refglue
from tests import util
import diff_gaussian_rasterization as dgr
def h(t): return hashlib.sha256(t.detach().contiguous().cpu().numpy().tobytes()).hexdigest()[:10]
n, w, hgt = 60000, 800, 600
sc, cam, ts, kids = util.make_scene(n, w, hgt, seed=2, dev='cuda', interp=True)
rs = refglue.make_settings(dgr.GaussianRasterizationSettings, cam, torch.zeros(3, device='cuda'), sh_degree=3, do_depth=True, interpolation_weights=ts, num_node_kids=kids)
p = {k: sc[k].clone().requires_grad_(True) for k in ('means3D','shs','opacities','scales','rotations')}
m2d = torch.zeros_like(p['means3D'], requires_grad=True)
cap = {}
orig = dgr._C.rasterize_gaussians_ex
def spy(*a, **k):
    out = orig(*a, **k); cap['out'] = out; return out
dgr._C.rasterize_gaussians_ex = spy
c, r, d = dgr.GaussianRasterizer(rs)(means3D=p['means3D'], means2D=m2d, opacities=p['opacities'], shs=p['shs'], scales=p['scales'], rotations=p['rotations'])
torch.cuda.synchronize()
R, color, radii, geom, binning, img = cap['out'][:6]
dv = dgr._C.debug_view(n, w, hgt, R, geom, binning, img)
vis = radii > 0
rec = dv['records'].view(n, 16)[vis]
print('HASH', R, 'in_scales', h(sc['scales']), 'in_rot', h(sc['rotations']), 'in_opac', h(sc['opacities']), 'in_ts', h(ts), 'view', h(rs.viewmatrix), h(rs.projmatrix), 'radii', h(radii), 'rec', h(rec), 'keys', h(dv['point_list_keys']), 'list', h(dv['point_list']), 'ranges', h(dv['ranges']),
      'ncontrib', h(dv['n_contrib']), 'finalT', h(dv['final_T']), 'img', h(c), 'invd', h(d))
import os, numpy as np
os.makedirs('gpurun_out', exist_ok=True); base = 'gpurun_out/diag_rec_base.npy'
recn = dv['records'].view(n, 16).cpu().numpy().copy()
recn[~vis.cpu().numpy()] = 0
if not os.path.exists(base):
    np.save(base, recn)
else:
    b = np.load(base)
    diff = (b.view(np.uint32) != recn.view(np.uint32))
    rows = np.nonzero(diff.any(axis=1))[0]
    if rows.size:
        print('DIFF rows', rows.size, 'cols', np.nonzero(diff.any(axis=0))[0].tolist(), 'first rows', rows[:6].tolist())
        for r0 in rows[:3]:
            print('DIFF row', int(r0), 'base', b[r0].tolist(), 'now', recn[r0].tolist())
(c.sum() + d.sum()).backward()
torch.cuda.synchronize()
print('GRAD', float(p['means3D'].grad.abs().sum()))
"""


def main():
    k = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    seen = set()
    for it in range(k):
        for mode in ("ldgsts", "tma"):
            env = dict(os.environ, H3DGS_STAGE=mode, PYTHONPATH=root)
            r = subprocess.run([sys.executable, "-c", CODE], env=env, capture_output=True, text=True, timeout=300, cwd=root)
            line = [l for l in r.stdout.splitlines() if l.startswith("HASH")]
            line = line[0] if line else ("FAILED " + r.stderr[-300:])
            if line not in seen:
                print(it, mode, line, flush=True)
                seen.add(line)
            for l in r.stdout.splitlines():
                if l.startswith("DIFF"):
                    print(it, mode, l[:600], flush=True)
    print("distinct outputs:", len(seen), "over", 2 * k, "processes")


if __name__ == "__main__":
    main()
