import sys, numpy as np
sys.path.insert(0, "/root/repo")
from oracle import oracle
from zig_raytracing_contest_b200 import rtc, scenes
spec = scenes.atrium()
R = 128
geo = rtc.Geometry(spec.positions, spec.tri_data, (R, R, R))
desc, keep = rtc.make_scene_desc(geo, spec.materials)
occ = (geo.cells[:, 1] > geo.cells[:, 0]).reshape(R, R, R)
E2 = ~occ.reshape(R//4, 4, R//4, 4, R//4, 4).any(axis=(1, 3, 5))
cam = spec.camera(height=1080)
rng = np.random.default_rng(1)
tot = dict(rays=0, cells=0, eblocks=0, runs=0, ne_cells=0, occ_cells=0, ne_blocks=0)
def walk(o, d, depth):
    idx, t, u, v = oracle.trace_ray(desc, o, d)
    w = oracle.grid_walk(geo.grid, o, d)
    if w is None: return
    cs, ts = w
    n = len(cs)
    if idx >= 0:
        k = np.nonzero(ts >= np.float32(t))[0]
        n = int(k[0]) + 1 if len(k) else n
    cs = cs[:n].astype(np.int64)
    b = cs >> 2
    e = E2[b[:, 2], b[:, 1], b[:, 0]]
    key = (b[:, 2] * 64 + b[:, 1]) * 64 + b[:, 0]
    newblock = np.concatenate([[True], key[1:] != key[:-1]])
    tot["rays"] += 1; tot["cells"] += n
    tot["eblocks"] += int((newblock & e).sum())
    tot["ne_blocks"] += int((newblock & ~e).sum())
    tot["runs"] += int((e[1:] & ~e[:-1]).sum() + (1 if e[0] else 0))
    tot["ne_cells"] += int((~e).sum())
    tot["occ_cells"] += int(occ[cs[:, 2], cs[:, 1], cs[:, 0]].sum())
    if idx >= 0 and depth < 3:
        p = o + d * np.float32(t)
        tp = geo.tri_pos[idx]
        nrm = np.cross(tp[1, :3], tp[2, :3]); nrm /= np.linalg.norm(nrm)
        r = rng.normal(size=3); r /= np.linalg.norm(r)
        nd = nrm + r; nd = (nd / np.linalg.norm(nd)).astype(np.float32)
        walk(p.astype(np.float32), nd, depth + 1)
for i in range(1500):
    o, d = oracle.camera_get_ray(cam, rng.uniform(0, cam.w), rng.uniform(0, cam.h))
    walk(o, d, 0)
r = tot["rays"]
print({k: round(v / r, 2) for k, v in tot.items()})
