"""How many triangle tests are re-tests: (a) cell's list identical to the previously tested non-empty cell's list,
(b) triangle already tested by this ray (perfect mailbox)."""
import sys, numpy as np
sys.path.insert(0, "/root/repo")
from oracle import oracle
from zig_raytracing_contest_b200 import rtc, scenes
which = sys.argv[1] if len(sys.argv) > 1 else "atrium"
spec = scenes.atrium() if which == "atrium" else scenes.terrain(num_tris=int(sys.argv[2]) if len(sys.argv) > 2 else 500000)
R = 128
geo = rtc.Geometry(spec.positions, spec.tri_data, (R, R, R))
desc, keep = rtc.make_scene_desc(geo, spec.materials)
cells = geo.cells.reshape(R, R, R, 2)
src = geo.tri_data["src_tri"]
cam = spec.camera(height=1080)
rng = np.random.default_rng(1)
tot = dict(rays=0, tests=0, same_as_prev=0, mailbox=0, ne_cells=0, same_cells=0)
def walk(o, d, depth):
    idx, t, u, v = oracle.trace_ray(desc, o, d)
    w = oracle.grid_walk(geo.grid, o, d)
    if w is None: return
    cs, ts = w
    n = len(cs)
    if idx >= 0:
        k = np.nonzero(ts >= np.float32(t))[0]
        n = int(k[0]) + 1 if len(k) else n
    tot["rays"] += 1
    prev = None; seen = set()
    for c in cs[:n]:
        b, e = cells[c[2], c[1], c[0]]
        if e <= b: continue
        ids = tuple(src[b:e])
        tot["ne_cells"] += 1; tot["tests"] += len(ids)
        if ids == prev: tot["same_as_prev"] += len(ids); tot["same_cells"] += 1
        tot["mailbox"] += sum(1 for i in ids if i in seen)
        seen.update(ids); prev = ids
    if idx >= 0 and depth < 3:
        p = o + d * np.float32(t)
        tp = geo.tri_pos[idx]
        nrm = np.cross(tp[1, :3], tp[2, :3]); nrm /= np.linalg.norm(nrm)
        r = rng.normal(size=3); r /= np.linalg.norm(r)
        nd = nrm + r; nd = (nd / np.linalg.norm(nd)).astype(np.float32)
        walk(p.astype(np.float32), nd, depth + 1)
for i in range(int(sys.argv[3]) if len(sys.argv) > 3 else 800):
    o, d = oracle.camera_get_ray(cam, rng.uniform(0, cam.w), rng.uniform(0, cam.h))
    walk(o, d, 0)
r = tot["rays"]
print(which, {k: round(v / r, 2) for k, v in tot.items()})
