"""Procedural scenes. The reference bundles no renderable scene (its author's scenes live outside
the repository, zig_raytracing_contest.sublime-project:29-61), so BASELINE.json's configs run on
generated, seeded geometry that honours the reference loader's constraints: world-space triangle
soup with per-vertex normals + uvs and a material per triangle (src/stage1.zig:217-259), three
textures per material (src/stage1.zig:381-496), a pinhole camera node (src/stage1.zig:309-371).

Winding: the reference culls back faces (det < 1e-8 -> miss, src/linalg.zig:703-707); a triangle
is visible from the side its e1 x e2 normal points to.
"""
import numpy as np

from . import rtc

F = np.float32


class Mesh:
    def __init__(self):
        self.pos, self.nrm, self.uv, self.mat = [], [], [], []

    def add(self, p, n, uv, material):
        """p,n: [T,3,3]; uv: [T,3,2]."""
        self.pos.append(np.asarray(p, F)); self.nrm.append(np.asarray(n, F)); self.uv.append(np.asarray(uv, F))
        self.mat.append(np.full(len(p), material, np.uint32))

    def finish(self):
        pos = np.concatenate(self.pos); nrm = np.concatenate(self.nrm)
        uv = np.concatenate(self.uv); mat = np.concatenate(self.mat)
        td = np.zeros(len(pos), rtc.TRI_DATA_DTYPE)
        td["normal"], td["texcoord"], td["material"] = nrm, uv, mat
        return np.ascontiguousarray(pos, F), td


def _grid_tris(P, N, UV):
    """P,N: [nv+1,nu+1,3], UV: [nv+1,nu+1,2] vertex lattices -> two CCW triangles per quad with
    normal along dP/du x dP/dv."""
    a, b, c, d = (slice(None, -1), slice(None, -1)), (slice(None, -1), slice(1, None)), \
                 (slice(1, None), slice(1, None)), (slice(1, None), slice(None, -1))

    def tri(X, i, j, k):
        return np.stack([X[i], X[j], X[k]], axis=2).reshape(-1, 3, X.shape[-1])
    out = []
    for X in (P, N, UV):
        t1, t2 = tri(X, a, b, c), tri(X, a, c, d)
        out.append(np.stack([t1, t2], axis=1).reshape(-1, 3, X.shape[-1]))
    return out


def quad(mesh, origin, du, dv, nu, nv, material, uv_scale=(1.0, 1.0)):
    """Planar nu x nv quad patch; visible from the side du x dv points to."""
    origin, du, dv = np.asarray(origin, F), np.asarray(du, F), np.asarray(dv, F)
    s, t = np.meshgrid(np.linspace(0, 1, nu + 1, dtype=F), np.linspace(0, 1, nv + 1, dtype=F))
    P = origin + s[..., None] * du + t[..., None] * dv
    n = np.cross(du, dv); n = (n / np.linalg.norm(n)).astype(F)
    N = np.broadcast_to(n, P.shape)
    UV = np.stack([s * F(uv_scale[0]), t * F(uv_scale[1])], axis=-1)
    mesh.add(*_grid_tris(P.astype(F), N, UV.astype(F)), material)


def box(mesh, lo, hi, material, n=1, inward=False):
    lo, hi = np.asarray(lo, F), np.asarray(hi, F)
    sx, sy, sz = (hi - lo)
    X, Y, Z = np.array([sx, 0, 0], F), np.array([0, sy, 0], F), np.array([0, 0, sz], F)
    faces = [(lo, Z, Y), (lo + X, Y, Z), (lo, X, Z), (lo + Y, Z, X), (lo, Y, X), (lo + Z, X, Y)]  # -x +x -y +y -z +z outward
    for o, a, b in faces:
        if inward:
            a, b = b, a
        quad(mesh, o, a, b, n, n, material)


def sphere(mesh, center, radius, nu, nv, material):
    th, ph = np.meshgrid(np.linspace(0, 2 * np.pi, nu + 1), np.linspace(1e-3, np.pi - 1e-3, nv + 1))
    N = np.stack([np.sin(ph) * np.cos(th), np.cos(ph), np.sin(ph) * np.sin(th)], axis=-1).astype(F)
    P = (np.asarray(center, F) + F(radius) * N).astype(F)
    UV = np.stack([th / (2 * np.pi), ph / np.pi], axis=-1).astype(F)
    # dP/dth x dP/dph points inward for this parametrisation -> flip the lattice along u
    mesh.add(*_grid_tris(P[:, ::-1], N[:, ::-1], UV[:, ::-1]), material)


def cylinder(mesh, base, radius, height, nu, nv, material):
    th, h = np.meshgrid(np.linspace(0, 2 * np.pi, nu + 1), np.linspace(0, 1, nv + 1))
    N = np.stack([np.cos(th), np.zeros_like(th), np.sin(th)], axis=-1).astype(F)
    P = (np.asarray(base, F) + F(radius) * N + np.stack([0 * h, h * height, 0 * h], axis=-1)).astype(F)
    UV = np.stack([th / (2 * np.pi) * 4, h * 4], axis=-1).astype(F)
    mesh.add(*_grid_tris(P[:, ::-1], N[:, ::-1], UV[:, ::-1]), material)


def heightfield(mesh, x0, z0, x1, z1, H, material, uv_tiles=16.0):
    """H: [nz+1,nx+1] heights; visible from above (+y)."""
    nz, nx = H.shape[0] - 1, H.shape[1] - 1
    xs, zs = np.linspace(x0, x1, nx + 1, dtype=F), np.linspace(z0, z1, nz + 1, dtype=F)
    Xg, Zg = np.meshgrid(xs, zs)
    P = np.stack([Xg, H.astype(F), Zg], axis=-1).astype(F)
    dx, dz = (x1 - x0) / nx, (z1 - z0) / nz
    gx = np.gradient(H, axis=1) / dx; gz = np.gradient(H, axis=0) / dz
    N = np.stack([-gx, np.ones_like(H), -gz], axis=-1)
    N = (N / np.linalg.norm(N, axis=-1, keepdims=True)).astype(F)
    UV = np.stack([(Xg - x0) / (x1 - x0) * uv_tiles, (Zg - z0) / (z1 - z0) * uv_tiles], axis=-1).astype(F)
    # du = +x, dv = +z gives du x dv = -y; flip rows so the patch faces +y
    mesh.add(*_grid_tris(P[::-1], N[::-1], UV[::-1]), material)


# ---- textures ------------------------------------------------------------------------------------
def value_noise(n, octaves, seed):
    rng = np.random.RandomState(seed)
    out = np.zeros((n, n), np.float64); amp = 1.0; tot = 0.0
    for o in range(octaves):
        m = 2 ** (o + 2)
        g = rng.rand(m + 1, m + 1); g[-1] = g[0]; g[:, -1] = g[:, 0]
        t = np.linspace(0, m, n, endpoint=False); i = t.astype(int); f = t - i; f = f * f * (3 - 2 * f)
        r0 = g[i][:, i] * (1 - f) + g[i][:, i + 1] * f
        r1 = g[i + 1][:, i] * (1 - f) + g[i + 1][:, i + 1] * f
        out += amp * (r0 * (1 - f[:, None]) + r1 * f[:, None]); tot += amp; amp *= 0.5
    return (out / tot)


def checker(n, tiles, a, b):
    i = (np.arange(n) * tiles // n)
    m = ((i[:, None] + i[None, :]) & 1).astype(bool)
    t = np.where(m[..., None], np.asarray(a, F), np.asarray(b, F)).astype(F)
    return t


def noise_rgb(n, seed, base, var):
    v = value_noise(n, 5, seed)[..., None]
    return np.clip(np.asarray(base) + (v - 0.5) * np.asarray(var), 0.02, 0.98).astype(F)


# ---- cameras ----------------------------------------------------------------------------------------
def look_at(eye, target, up=(0, 1, 0)):
    """glTF camera node matrix (column-major [16]): looks down -Z, +Y up."""
    eye, target, up = np.asarray(eye, np.float64), np.asarray(target, np.float64), np.asarray(up, np.float64)
    z = eye - target; z /= np.linalg.norm(z)
    x = np.cross(up, z); x /= np.linalg.norm(x)
    y = np.cross(z, x)
    m = np.eye(4); m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = x, y, z, eye
    return m.T.reshape(16).astype(F)        # column-major


def load_camera(matrix, yfov, aspect_ratio, width=None, height=None):
    """stage1.loadCamera (src/stage1.zig:309-371) in fp32: size rules + pixel-unit basis."""
    from . import abi
    has_ar = aspect_ratio is not None and aspect_ratio > 0
    if not width and not height:
        raise ValueError("OutputImgSizeIsNotSpecified")
    if width and height:
        if has_ar:
            raise ValueError("CameraHasAspectRatio")
        w, h = int(width), int(height)
    else:
        if not has_ar:
            raise ValueError("CameraHasntAspectRatio")
        w = int(width) if width else int(F(height) * F(aspect_ratio))
        h = int(height) if height else int(F(width) / F(aspect_ratio))
    m = np.asarray(matrix, F).reshape(4, 4)          # rows of this array = columns of the matrix
    origin = m[3, :3]
    fwd = -m[2, :3]; fwd = (fwd * (F(1) / np.sqrt(np.sum(fwd * fwd, dtype=F), dtype=F))).astype(F)
    right = np.cross(fwd, np.array([0, 1, 0], F)).astype(F)
    right = (right * (F(1) / np.sqrt(np.sum(right * right, dtype=F), dtype=F))).astype(F)
    up = np.cross(fwd, right).astype(F)
    focal = F(F(h) / F(2)) / F(np.tan(F(yfov) / F(2), dtype=F))
    llc = (fwd * focal - right * F(F(w) / F(2)) - up * F(F(h) / F(2))).astype(F)
    cam = abi.Camera()
    cam.w, cam.h = w, h
    cam.origin[:], cam.lower_left_corner[:], cam.right[:], cam.up[:] = origin.tolist(), llc.tolist(), right.tolist(), up.tolist()
    return cam


# ---- scenes -----------------------------------------------------------------------------------------
class SceneSpec:
    """What stage1 hands to stage2/3: triangle soup, materials, and the camera node."""

    def __init__(self, name, positions, tri_data, materials, cam_matrix, yfov, aspect=16.0 / 9.0):
        self.name, self.positions, self.tri_data, self.materials = name, positions, tri_data, materials
        self.cam_matrix, self.yfov, self.aspect = cam_matrix, yfov, aspect

    def camera(self, width=None, height=None):
        return load_camera(self.cam_matrix, self.yfov, self.aspect, width, height)

    @property
    def num_tris(self):
        return int(self.positions.shape[0])


def cornell(subdiv=8, seed=3):
    """cfg1/cfg2 scene: closed room, emissive ceiling panel, two boxes, a sphere, a textured floor
    and an alpha-masked (MASK) screen. ~ (6*2 + 5*2) * subdiv^2 + ... triangles."""
    M = rtc.MaterialData
    T = rtc.TextureData
    mats = [
        M(base_color=(0.73, 0.73, 0.73)),                                           # 0 white
        M(base_color=(0.65, 0.05, 0.05)),                                           # 1 red
        M(base_color=(0.12, 0.45, 0.15)),                                           # 2 green
        M(base_color=(0.0, 0.0, 0.0), emissive=(15.0, 15.0, 15.0)),                 # 3 light
        M(base_color=T(checker(64, 8, (0.8, 0.8, 0.8), (0.2, 0.2, 0.25)))),         # 4 checker floor (REPEAT)
        M(base_color=T(noise_rgb(64, seed, (0.3, 0.4, 0.8), (0.5, 0.5, 0.3)), True, True),
          transparency=T(checker(32, 8, (1.0,), (0.0,)), True, True)),              # 5 MASK screen (CLAMP)
        M(base_color=(0.9, 0.6, 0.2), transparency=0.5),                            # 6 BLEND 50%
    ]
    m = Mesh()
    s = subdiv
    quad(m, (0, 0, 0), (0, 0, 1), (1, 0, 0), s, s, 4, (4.0, 4.0))      # floor, faces +y
    quad(m, (0, 1, 0), (1, 0, 0), (0, 0, 1), s, s, 0)                  # ceiling, faces -y
    quad(m, (0, 0, 0), (1, 0, 0), (0, 1, 0), s, s, 0)                  # back wall z=0, faces +z
    quad(m, (0, 0, 0), (0, 1, 0), (0, 0, 1), s, s, 1)                  # left wall x=0, faces +x
    quad(m, (1, 0, 0), (0, 0, 1), (0, 1, 0), s, s, 2)                  # right wall x=1, faces -x
    quad(m, (0.35, 0.999, 0.35), (0.3, 0, 0), (0, 0, 0.3), 2, 2, 3)    # light, faces -y
    box(m, (0.15, 0.0, 0.15), (0.42, 0.6, 0.42), 0, max(1, s // 4))
    box(m, (0.58, 0.0, 0.45), (0.85, 0.3, 0.72), 0, max(1, s // 4))
    sphere(m, (0.72, 0.45, 0.58), 0.15, 4 * s, 2 * s, 6)
    quad(m, (0.1, 0.05, 0.8), (0.5, 0, 0.1), (0, 0.6, 0), 2, 2, 5)     # masked screen facing +z (camera side)
    pos, td = m.finish()
    return SceneSpec("cornell", pos, td, mats, look_at((0.5, 0.5, 2.4), (0.5, 0.5, 0.0)), 0.6911)


def atrium(detail=0.87, seed=11):
    """cfg3 scene ("Sponza-class", ~2.6e5 triangles at detail=1): open-roofed colonnaded hall lit by the
    sky and emissive lamps; tessellated columns, displaced cloth banners with MASK alpha, textured
    floor heightfield, spheres. Open top so most paths end in the sky within max_bounce."""
    M, T = rtc.MaterialData, rtc.TextureData
    mats = [
        M(base_color=T(noise_rgb(256, seed, (0.62, 0.58, 0.5), (0.3, 0.3, 0.3)))),            # 0 stone
        M(base_color=T(checker(256, 16, (0.75, 0.7, 0.6), (0.35, 0.3, 0.28)))),                # 1 floor tiles
        M(base_color=T(noise_rgb(256, seed + 1, (0.7, 0.15, 0.12), (0.4, 0.2, 0.2))),
          transparency=T((value_noise(256, 4, seed + 2) > 0.38).astype(F))),                   # 2 red cloth, MASK holes
        M(base_color=T(noise_rgb(256, seed + 3, (0.15, 0.3, 0.65), (0.3, 0.3, 0.4))),
          transparency=T(checker(128, 32, (1.0,), (0.0,)))),                                   # 3 blue lattice, MASK
        M(base_color=(0.0, 0.0, 0.0), emissive=(12.0, 9.0, 5.0)),                              # 4 lamps
        M(base_color=(0.85, 0.85, 0.88)),                                                      # 5 plaster
        M(base_color=(0.8, 0.75, 0.3), transparency=0.7),                                      # 6 BLEND
    ]
    d = detail
    m = Mesh()
    L, W, Hh = 30.0, 12.0, 9.0
    n = lambda k: max(1, int(round(k * d)))
    rng = np.random.RandomState(seed)
    # floor: gently bumpy heightfield
    nf = n(160)
    Hf = 0.04 * value_noise(nf + 1, 4, seed + 5)
    heightfield(m, 0, 0, L, W, Hf[: nf + 1, : nf + 1], 1, uv_tiles=24.0)
    # walls (inside faces)
    quad(m, (0, 0, 0), (L, 0, 0), (0, Hh, 0), n(60), n(24), 5, (8, 3))            # z=0 faces +z
    quad(m, (0, 0, W), (0, Hh, 0), (L, 0, 0), n(24), n(60), 5, (3, 8))            # z=W faces -z
    quad(m, (0, 0, 0), (0, Hh, 0), (0, 0, W), n(24), n(30), 0, (3, 4))            # x=0 faces +x
    quad(m, (L, 0, 0), (0, 0, W), (0, Hh, 0), n(30), n(24), 0, (4, 3))            # x=L faces -x
    # two rows of columns + lamps + capitals
    for i in range(10):
        x = 2.0 + i * (L - 4.0) / 9
        for z in (3.0, W - 3.0):
            cylinder(m, (x, 0, z), 0.35, 6.0, n(48), n(40), 0)
            box(m, (x - 0.5, 6.0, z - 0.5), (x + 0.5, 6.4, z + 0.5), 5, n(4))
            sphere(m, (x, 6.9, z), 0.28, n(24), n(12), 4 if i % 3 == 0 else 6)
    # upper gallery slabs
    box(m, (0, 6.4, 0), (L, 6.7, 3.5), 5, n(20))
    box(m, (0, 6.4, W - 3.5), (L, 6.7, W), 5, n(20))
    # cloth banners hanging between columns (wavy, double layer so both sides are visible)
    for i in range(9):
        x0 = 2.35 + i * (L - 4.0) / 9; x1 = x0 + (L - 4.0) / 9 - 0.7
        for z, mat in ((3.0, 2), (W - 3.0, 3)):
            nu, nv = n(56), n(40)
            s, t = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1))
            wave = 0.25 * np.sin(s * 9 + i) * (1 - t) + 0.1 * np.sin(t * 7 + i * 2)
            P = np.stack([x0 + s * (x1 - x0), 2.2 + t * 3.6, z + wave], axis=-1).astype(F)
            gx = np.gradient(wave, axis=1) * nu / (x1 - x0); gy = np.gradient(wave, axis=0) * nv / 3.6
            N = np.stack([-gx, -gy, np.ones_like(wave)], axis=-1); N = (N / np.linalg.norm(N, axis=-1, keepdims=True)).astype(F)
            UV = np.stack([s * 2, t * 2], axis=-1).astype(F)
            m.add(*_grid_tris(P, N, UV), mat)                                   # faces +z
            m.add(*_grid_tris(P[:, ::-1], -N[:, ::-1], UV[:, ::-1]), mat)       # faces -z
    # clutter: spheres and boxes on the floor
    for k in range(24):
        c = (rng.uniform(3, L - 3), 0.0, rng.uniform(4.2, W - 4.2)); r = rng.uniform(0.25, 0.6)
        if k % 2:
            sphere(m, (c[0], r + 0.05, c[2]), r, n(40), n(20), (0, 5, 6)[k % 3])
        else:
            box(m, (c[0] - r, 0.02, c[2] - r), (c[0] + r, 2 * r, c[2] + r), (0, 5, 1)[k % 3], n(6))
    pos, td = m.finish()
    return SceneSpec("atrium", pos, td, mats, look_at((1.2, 2.6, W / 2 + 0.8), (L, 3.4, W / 2 - 0.5)), 0.95)


def terrain(num_tris=5_000_000, seed=7):
    """cfg4/cfg5 scene: fBm heightfield topped up with icosphere-like spheres to exactly num_tris
    source triangles; 8 materials (1 emissive, 1 MASK with a 1024^2 checker alpha, 512^2 procedural
    colour); camera inside the scene bbox."""
    M, T = rtc.MaterialData, rtc.TextureData
    mats = [M(base_color=T(noise_rgb(512, seed + i, c, (0.35, 0.35, 0.3)))) for i, c in
            enumerate([(0.35, 0.5, 0.25), (0.5, 0.45, 0.35), (0.6, 0.6, 0.62), (0.3, 0.35, 0.2), (0.7, 0.65, 0.5), (0.4, 0.4, 0.45)])]
    mats.append(M(base_color=(0.0, 0.0, 0.0), emissive=(20.0, 14.0, 8.0)))                                   # 6 emissive
    mats.append(M(base_color=T(noise_rgb(512, seed + 9, (0.8, 0.3, 0.2), (0.3, 0.3, 0.3))),
                  transparency=T(checker(1024, 64, (1.0,), (0.0,)))))                                         # 7 MASK
    m = Mesh()
    n_sph, seg = 64, (48, 24)                       # 64 spheres x 48*24*2 = 147456 triangles
    if num_tris < 1_000_000:
        n_sph, seg = 16, (16, 8)                    # small test variants: 16 x 16*8*2 = 4096 triangles
    sph_tris = n_sph * seg[0] * seg[1] * 2
    nq = int(np.floor(np.sqrt((num_tris - sph_tris) / 2)))
    S = 200.0
    H = 28.0 * (value_noise(nq + 1, 9, seed) - 0.35)
    # split into patches with one material each (the reference needs u16 indices per primitive)
    step = 255
    for z0 in range(0, nq, step):
        for x0 in range(0, nq, step):
            z1, x1 = min(nq, z0 + step), min(nq, x0 + step)
            mat = ((x0 // step) * 3 + (z0 // step) * 5) % 6
            heightfield(m, S * x0 / nq, S * z0 / nq, S * x1 / nq, S * z1 / nq, H[z0:z1 + 1, x0:x1 + 1], mat, uv_tiles=4.0)
    rng = np.random.RandomState(seed + 1)
    for k in range(n_sph):
        cx, cz = rng.uniform(10, S - 10, 2)
        iy = H[int(cz / S * nq), int(cx / S * nq)]
        r = rng.uniform(1.0, 3.5)
        sphere(m, (cx, iy + r * rng.uniform(0.6, 2.5), cz), r, seg[0], seg[1], 6 if k % 4 == 0 else (7 if k % 4 == 1 else k % 6))
    pos, td = m.finish()
    short = num_tris - len(pos)
    if short > 0:                                   # top up with small quads' halves on a ledge
        extra = Mesh()
        nx = int(np.ceil(short / 2 / 8)) + 1
        quad(extra, (5, float(H.max()) + 2, 5), (0, 0, 30), (30, 0, 0), nx, 8, 2)
        p2, t2 = extra.finish()
        pos = np.concatenate([pos, p2[:short]]); td = np.concatenate([td, t2[:short]])
    pos, td = pos[:num_tris], td[:num_tris]
    eye_h = float(H[int(0.12 * nq), int(0.1 * nq)]) + 6.0
    return SceneSpec("terrain", np.ascontiguousarray(pos), np.ascontiguousarray(td), mats,
                     look_at((0.1 * S, eye_h, 0.12 * S), (0.6 * S, eye_h - 14.0, 0.6 * S)), 0.9)


SCENES = {"cornell": cornell, "atrium": atrium, "terrain": terrain}
