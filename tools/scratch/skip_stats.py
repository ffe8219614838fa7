"""How much of the DDA walk lies in empty 2^L-aligned boxes? (planning aid for block skipping)"""
import sys, numpy as np
sys.path.insert(0, "/root/repo")
from oracle import oracle
from zig_raytracing_contest_b200 import rtc, scenes
name = sys.argv[1] if len(sys.argv) > 1 else "atrium"
spec = getattr(scenes, name)() if name != "terrain" else scenes.terrain(num_tris=int(sys.argv[2]) if len(sys.argv) > 2 else 1000000)
R = 128
geo = rtc.Geometry(spec.positions, spec.tri_data, (R, R, R))
desc, keep = rtc.make_scene_desc(geo, spec.materials)
occ = (geo.cells[:, 1] > geo.cells[:, 0]).reshape(R, R, R)  # z,y,x
def emp(L):
    s = 1 << L
    o = occ.reshape(R//s, s, R//s, s, R//s, s).any(axis=(1, 3, 5))
    return ~o
E = {L: emp(L) for L in (1, 2, 3, 4)}
print("empty fraction of boxes:", {L: float(E[L].mean()) for L in E}, "cells empty", float((~occ).mean()))
cam = spec.camera(height=1080)
rng = np.random.default_rng(1)
rays = []
for i in range(1500):
    x, y = rng.uniform(0, cam.w), rng.uniform(0, cam.h)
    rays.append(oracle.camera_get_ray(cam, x, y))
tot = dict(rays=0, cells=0)
inL = {L: 0 for L in E}; runsL = {L: 0 for L in E}
best_units = 0  # hierarchical: number of skip-units if each cell goes to its largest empty box (<=L4 / <=L2)
units2 = 0
def walk(o, d, depth):
    global best_units, units2
    idx, t, u, v = oracle.trace_ray(desc, o, d)
    w = oracle.grid_walk(geo.grid, o, d)
    if w is None: return
    cs, ts = w
    n = len(cs)
    if idx >= 0:
        k = np.nonzero(ts >= np.float32(t))[0]
        n = int(k[0]) + 1 if len(k) else n
    cs = cs[:n].astype(np.int64)
    tot["rays"] += 1; tot["cells"] += n
    lvl = np.zeros(n, int)
    for L in E:
        b = cs >> L
        e = E[L][b[:, 2], b[:, 1], b[:, 0]]
        lvl[e] = L
        inL[L] += int(e.sum())
        key = (b[:, 2] * 4096 + b[:, 1]) * 4096 + b[:, 0]
        key = np.where(e, key, -1 - np.arange(n))
        runsL[L] += int(((key[1:] != key[:-1]) & e[1:]).sum() + (1 if e[0] else 0))
    for maxL, name_ in ((4, "h4"), (2, "h2")):
        l2 = np.minimum(lvl, maxL)
        b = [(tuple(c >> l), l) for c, l in zip(cs, l2)]
        u_ = 1 + sum(1 for i in range(1, n) if b[i] != b[i-1] or l2[i] == 0)
        if maxL == 4: best_units += u_
        else: units2 += u_
    if idx >= 0 and depth < 3:
        p = o + d * np.float32(t)
        tp = geo.tri_pos[idx]
        nrm = np.cross(tp[1, :3], tp[2, :3]); nrm /= np.linalg.norm(nrm)
        r = rng.normal(size=3); r /= np.linalg.norm(r)
        nd = nrm + r; nd = (nd / np.linalg.norm(nd)).astype(np.float32)
        walk(p.astype(np.float32), nd, depth + 1)
for o, d in rays: walk(o, d, 0)
print("rays", tot["rays"], "cells/ray %.1f" % (tot["cells"] / tot["rays"]))
for L in E:
    print("L%d: cells in empty boxes %.1f%%, runs/ray %.2f, cells/run %.2f" % (L, 100 * inL[L] / tot["cells"], runsL[L] / tot["rays"], inL[L] / max(1, runsL[L])))
print("hierarchical units/ray: maxL2 %.1f   maxL4 %.1f" % (units2 / tot["rays"], best_units / tot["rays"]))
# duplicate-test statistics: how many triangle tests repeat a (ray, source triangle) pair already tested
src = geo.tri_data["src_tri"] if "src_tri" in geo.tri_data.dtype.names else None
tests = dup = dup_last = 0
cell_lin = lambda c: (c[:, 2] * R + c[:, 1]) * R + c[:, 0]
def walk2(o, d, depth):
    global tests, dup, dup_last
    idx, t, u, v = oracle.trace_ray(desc, o, d)
    w = oracle.grid_walk(geo.grid, o, d)
    if w is None: return
    cs, ts = w
    n = len(cs)
    if idx >= 0:
        k = np.nonzero(ts >= np.float32(t))[0]
        n = int(k[0]) + 1 if len(k) else n
    seen = set(); last = set()
    for ci in cell_lin(cs[:n].astype(np.int64)):
        b, e = geo.cells[ci]
        if e > b:
            ids = src[b:e].tolist()
            tests += len(ids)
            dup += sum(1 for i in ids if i in seen)
            dup_last += sum(1 for i in ids if i in last)
            seen.update(ids); last = set(ids)
    if idx >= 0 and depth < 3:
        p = o + d * np.float32(t)
        tp = geo.tri_pos[idx]
        nrm = np.cross(tp[1, :3], tp[2, :3]); nrm /= np.linalg.norm(nrm)
        r = rng.normal(size=3); r /= np.linalg.norm(r)
        nd = nrm + r; nd = (nd / np.linalg.norm(nd)).astype(np.float32)
        walk2(p.astype(np.float32), nd, depth + 1)
for o, d in rays[:600]: walk2(o, d, 0)
print("tests", tests, "dup any-earlier %.1f%%  dup vs previous non-empty cell %.1f%%" % (100 * dup / tests, 100 * dup_last / tests))
