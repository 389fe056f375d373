"""Procedural scenes for BASELINE.json's configs (there is no network and the reference ships no
scene: src/main.cpp:365 hard-codes an absent "mitsuba.obj").  All generators are deterministic
(seed 0x5EED), emit float32 positions + per-vertex normals + int32 index triplets + per-face material
ids + an albedo table — the exact host arrays Scene::add_triangle_mesh uploads
(src/main.cpp:288-310).  Sphere tessellation follows the lat-long scheme of Embree's own scene
generator (dependencies/embree/tutorials/common/scenegraph/geometry_creation.cpp:134-200):
numPhi rings -> 4*N*(N-1) triangles.
"""
import numpy as np

SEED = 0x5EED


class SceneData:
    def __init__(self, name, pos, idx, nrm, nidx, mat, albedo, camera):
        self.name = name
        self.pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        self.idx = np.ascontiguousarray(idx, np.int32).reshape(-1, 3)
        self.nrm = np.ascontiguousarray(nrm, np.float32).reshape(-1, 3)
        self.nidx = np.ascontiguousarray(nidx, np.int32).reshape(-1, 3)
        self.mat = np.ascontiguousarray(mat, np.int32).reshape(-1)
        self.albedo = np.ascontiguousarray(albedo, np.float32).reshape(-1, 3)
        self.camera = camera  # dict(look_from, look_at, up, fov_y)
        assert self.idx.shape == self.nidx.shape and self.mat.shape[0] == self.idx.shape[0]

    @property
    def n_tris(self):
        return self.idx.shape[0]


class _Builder:
    def __init__(self):
        self.P, self.N, self.I, self.M = [], [], [], []
        self.nv = 0

    def add(self, pos, nrm, idx, mat):
        pos = np.asarray(pos, np.float32).reshape(-1, 3)
        nrm = np.asarray(nrm, np.float32).reshape(-1, 3)
        idx = np.asarray(idx, np.int64).reshape(-1, 3)
        self.P.append(pos)
        self.N.append(nrm)
        self.I.append((idx + self.nv).astype(np.int32))
        self.M.append(np.full(idx.shape[0], mat, np.int32))
        self.nv += pos.shape[0]

    def finish(self, name, albedo, camera):
        pos = np.concatenate(self.P)
        idx = np.concatenate(self.I)
        # positions and normals share indices (one normal per vertex)
        return SceneData(name, pos, idx, np.concatenate(self.N), idx.copy(), np.concatenate(self.M), albedo, camera)


def sphere(center, radius, n_phi):
    """lat-long sphere, 4*n_phi*(n_phi-1) triangles, outward normals"""
    n_theta = 2 * n_phi
    i = np.arange(n_phi + 1)[:, None]
    j = np.arange(n_theta)[None, :]
    phi = np.pi * i / n_phi
    theta = 2.0 * np.pi * j / n_theta
    n = np.stack([np.sin(phi) * np.sin(theta), np.cos(phi) * np.ones_like(theta), np.sin(phi) * np.cos(theta)], -1)
    n = n.reshape(-1, 3)
    pos = np.asarray(center, np.float64)[None, :] + radius * n
    vid = lambda a, b: (a * n_theta + (b % n_theta))
    tris = []
    ii, jj = np.meshgrid(np.arange(1, n_phi + 1), np.arange(n_theta), indexing="ij")
    p00, p01 = vid(ii - 1, jj), vid(ii - 1, jj + 1)
    p10, p11 = vid(ii, jj), vid(ii, jj + 1)
    up = ii > 1          # not touching the north pole ring
    dn = ii < n_phi      # not touching the south pole ring
    tris.append(np.stack([p00[up], p10[up], p01[up]], -1))
    tris.append(np.stack([p01[dn], p10[dn], p11[dn]], -1))
    return pos, n, np.concatenate(tris)


def quad(p0, p1, p2, p3, normal):
    pos = np.array([p0, p1, p2, p3], np.float32)
    nrm = np.tile(np.asarray(normal, np.float32), (4, 1))
    return pos, nrm, np.array([[0, 1, 2], [0, 2, 3]])


def grid_surface(P, flip=False, wrap_u=False, wrap_v=False):
    """P: (nu, nv, 3) positions -> triangles + smooth per-vertex normals"""
    nu, nv = P.shape[:2]
    du = (np.roll(P, -1, 0) - np.roll(P, 1, 0)) if wrap_u else np.gradient(P, axis=0)
    dv = (np.roll(P, -1, 1) - np.roll(P, 1, 1)) if wrap_v else np.gradient(P, axis=1)
    n = np.cross(du, dv)
    if flip:
        n = -n
    ln = np.linalg.norm(n, axis=-1, keepdims=True)
    n = n / np.maximum(ln, 1e-20)
    mu = nu if wrap_u else nu - 1
    mv = nv if wrap_v else nv - 1
    a, b = np.meshgrid(np.arange(mu), np.arange(mv), indexing="ij")
    a1, b1 = (a + 1) % nu, (b + 1) % nv
    v00, v10, v01, v11 = a * nv + b, a1 * nv + b, a * nv + b1, a1 * nv + b1
    t0 = np.stack([v00, v10, v11], -1).reshape(-1, 3)
    t1 = np.stack([v00, v11, v01], -1).reshape(-1, 3)
    tri = np.concatenate([t0, t1])
    if flip:
        tri = tri[:, ::-1]
    return P.reshape(-1, 3), n.reshape(-1, 3), tri


_FOV = 40.0 / 180.0 * np.pi  # src/main.cpp:383


def cornell(n_phi=50, colored=False):
    """C1: open-front box (5 quads = 10 tris) + 2 spheres; n_phi=50 -> 19 610 triangles."""
    b = _Builder()
    L = 1.0
    b.add(*quad([-L, -L, -L], [L, -L, -L], [L, -L, L], [-L, -L, L], [0, 1, 0]), mat=0)    # floor
    b.add(*quad([-L, L, -L], [-L, L, L], [L, L, L], [L, L, -L], [0, -1, 0]), mat=0)       # ceiling
    b.add(*quad([-L, -L, -L], [-L, L, -L], [L, L, -L], [L, -L, -L], [0, 0, 1]), mat=0)    # back
    b.add(*quad([-L, -L, -L], [-L, -L, L], [-L, L, L], [-L, L, -L], [1, 0, 0]), mat=1)    # left
    b.add(*quad([L, -L, -L], [L, L, -L], [L, L, L], [L, -L, L], [-1, 0, 0]), mat=2)       # right
    b.add(*sphere([-0.45, -0.65, -0.30], 0.35, n_phi), mat=0)
    b.add(*sphere([0.45, -0.65, 0.35], 0.35, n_phi), mat=0)
    if colored:
        albedo = [[0.75, 0.75, 0.75], [0.75, 0.25, 0.25], [0.25, 0.25, 0.75]]
    else:
        albedo = [[1, 1, 1]] * 3   # src/main.cpp:234,244: every material is Spectrum(1.0)
    cam = dict(look_from=(0.0, 0.0, 3.9), look_at=(0.0, 0.0, 0.0), up=(0, 1, 0), fov_y=_FOV)
    return b.finish("cornell", albedo, cam)


def knob(nu=500, nv=200):
    """C2: displaced torus-knot tube (nu x nv grid, closed both ways -> 2*nu*nv tris) on a ground
    plane: ~200 K triangles."""
    b = _Builder()
    u = np.linspace(0, 2 * np.pi, nu, endpoint=False)
    v = np.linspace(0, 2 * np.pi, nv, endpoint=False)
    p, q = 2, 3
    r = 0.5 * (2 + np.cos(q * u))
    C = np.stack([r * np.cos(p * u), 0.5 * r * np.sin(q * u) * 1.2, r * np.sin(p * u)], -1)
    T = np.roll(C, -1, 0) - np.roll(C, 1, 0)
    T /= np.linalg.norm(T, axis=-1, keepdims=True)
    ref = np.array([0.0, 1.0, 0.0])
    N1 = np.cross(T, ref)
    N1 /= np.linalg.norm(N1, axis=-1, keepdims=True)
    N2 = np.cross(T, N1)
    rad = 0.28 * (1.0 + 0.12 * np.sin(9 * u)[:, None] * np.cos(7 * v)[None, :])
    P = C[:, None, :] + rad[..., None] * (np.cos(v)[None, :, None] * N1[:, None, :] + np.sin(v)[None, :, None] * N2[:, None, :])
    pos, nrm, tri = grid_surface(P, wrap_u=True, wrap_v=True)
    # orient outward
    outward = np.einsum("ij,ij->i", nrm, pos - np.repeat(C, nv, 0)) < 0
    if outward.mean() > 0.5:
        nrm = -nrm
        tri = tri[:, ::-1]
    b.add(pos, nrm, tri, mat=1)
    G = 6.0
    b.add(*quad([-G, -1.3, -G], [-G, -1.3, G], [G, -1.3, G], [G, -1.3, -G], [0, 1, 0]), mat=0)
    albedo = [[1, 1, 1], [1, 1, 1]]
    cam = dict(look_from=(0.0, 2.2, 4.6), look_at=(0.0, -0.1, 0.0), up=(0, 1, 0), fov_y=_FOV)
    return b.finish("knob", albedo, cam)


def _terrain_height(x, z):
    return (1.10 * np.sin(0.9 * x) * np.cos(1.1 * z) + 0.45 * np.sin(2.3 * x + 0.7) * np.sin(1.9 * z + 0.3)
            + 0.12 * np.sin(5.1 * x) * np.cos(4.7 * z + 1.1))


def terrain_spheres(grid=601, n_spheres=64, n_phi=34, extent=8.0, offset=(0.0, 0.0, 0.0), seed=SEED):
    """C3 block: (grid-1)^2*2 terrain triangles + n_spheres spheres of 4*n_phi*(n_phi-1) triangles.
    Defaults: 720 000 + 64*4 488 = 1 007 232 triangles."""
    b = _Builder()
    _terrain_block(b, grid, n_spheres, n_phi, extent, np.asarray(offset, np.float64), seed)
    albedo = [[1, 1, 1], [1, 1, 1]]
    cam = dict(look_from=(0.0, 4.2, 8.5), look_at=(0.0, -0.6, 0.0), up=(0, 1, 0), fov_y=_FOV)
    return b.finish("terrain_spheres", albedo, cam)


def _terrain_block(b, grid, n_spheres, n_phi, extent, offset, seed, rot=0.0):
    rng = np.random.default_rng(seed)
    x = np.linspace(-extent, extent, grid)
    z = np.linspace(-extent, extent, grid)
    X, Z = np.meshgrid(x, z, indexing="ij")
    Y = _terrain_height(X, Z)
    P = np.stack([X, Y, Z], -1)
    pos, nrm, tri = grid_surface(P)
    if nrm[:, 1].mean() < 0:
        nrm = -nrm
        tri = tri[:, ::-1]
    c, s = np.cos(rot), np.sin(rot)
    R = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
    b.add(pos @ R.T + offset, nrm @ R.T, tri, mat=0)
    for k in range(n_spheres):
        cx, cz = rng.uniform(-0.85 * extent, 0.85 * extent, 2)
        rad = rng.uniform(0.18, 0.55)
        cy = float(_terrain_height(cx, cz)) + rad * rng.uniform(0.6, 1.6)
        sp, sn, st = sphere([cx, cy, cz], rad, n_phi)
        b.add(sp @ R.T + offset, sn @ R.T, st, mat=1)


def flattened_blocks(n_blocks=10, grid=601, n_spheres=64, n_phi=34):
    """C4: n_blocks rigidly transformed copies of the 1 M block, flattened into ONE vertex/index pool
    (the reference has no instancing: one triangle soup, src/accel/optix_backend.h:283-364)."""
    b = _Builder()
    ext = 8.0
    cols = int(np.ceil(np.sqrt(n_blocks)))
    for k in range(n_blocks):
        gx, gz = k % cols, k // cols
        off = np.array([(gx - (cols - 1) / 2) * 2 * ext, 0.15 * k, (gz - (cols - 1) / 2) * 2 * ext])
        _terrain_block(b, grid, n_spheres, n_phi, ext, off, SEED + k, rot=0.5 * np.pi * (k % 4))
    albedo = [[1, 1, 1], [1, 1, 1]]
    span = cols * ext
    cam = dict(look_from=(0.0, 0.55 * span, 1.35 * span), look_at=(0.0, 0.0, 0.0), up=(0, 1, 0), fov_y=_FOV)
    return b.finish("flattened_blocks", albedo, cam)


def interior(grid=289, n_pillars=8, pillar_grid=(64, 128), skylight=True, grey=0.5):
    """C5: closed room whose six walls are displaced grids (6*(grid-1)^2*2 triangles) plus displaced
    pillars; ~1 M triangles by default.  Under the reference's semantics a sealed room renders black
    (light only enters when a ray escapes, src/main.cpp:407), so by default a 1 x 1 skylight is cut
    into the ceiling."""
    b = _Builder()
    L = 4.0
    t = np.linspace(-L - 0.25, L + 0.25, grid)      # walls overlap at the edges: bumps (< 0.1) cannot open a seam
    A, B = np.meshgrid(t, t, indexing="ij")
    bump = 0.06 * np.sin(3.1 * A) * np.cos(2.7 * B) + 0.03 * np.sin(7.3 * A + 0.4) * np.sin(6.1 * B)

    def wall(axis, sign):
        coord = sign * (L + bump)  # bumps go outward so walls never intersect each other's interior badly
        comps = [None, None, None]
        comps[axis] = coord
        others = [k for k in range(3) if k != axis]
        comps[others[0]] = A
        comps[others[1]] = B
        P = np.stack(comps, -1)
        pos, nrm, tri = grid_surface(P)
        inward = -sign * nrm[:, axis]
        if inward.mean() < 0:
            nrm = -nrm
            tri = tri[:, ::-1]
        return pos, nrm, tri

    for axis in range(3):
        for sign in (-1.0, 1.0):
            pos, nrm, tri = wall(axis, sign)
            if skylight and axis == 1 and sign > 0:
                c = pos[tri].mean(1)
                keep = ~((np.abs(c[:, 0]) < 0.5) & (np.abs(c[:, 2]) < 0.5))
                tri = tri[keep]
            b.add(pos, nrm, tri, mat=0)
    rng = np.random.default_rng(SEED)
    nu, nv = pillar_grid
    for k in range(n_pillars):
        ang = 2 * np.pi * k / n_pillars
        cx, cz = 2.4 * np.cos(ang), 2.4 * np.sin(ang)
        u = np.linspace(0, 2 * np.pi, nu, endpoint=False)
        h = np.linspace(-L - 0.2, L + 0.2, nv)
        U, H = np.meshgrid(u, h, indexing="ij")
        rad = 0.35 * (1 + 0.15 * np.sin(5 * U + k) * np.cos(3 * H))
        P = np.stack([cx + rad * np.cos(U), H, cz + rad * np.sin(U)], -1)
        pos, nrm, tri = grid_surface(P, wrap_u=True)
        out = np.einsum("ij,ij->i", nrm, pos - np.array([cx, 0, cz])[None, :] * np.array([1, 0, 1]))
        if out.mean() < 0:
            nrm = -nrm
            tri = tri[:, ::-1]
        b.add(pos, nrm, tri, mat=1)
    albedo = [[grey] * 3, [grey] * 3]
    cam = dict(look_from=(0.0, 0.0, 3.6), look_at=(0.0, -0.3, 0.0), up=(0, 1, 0), fov_y=60.0 / 180.0 * np.pi)
    return b.finish("interior", albedo, cam)


def spiral(n=80, q=0.75, n_phi=8):
    """test scene with a DEEP tree: n small spheres in a geometric progression of size and position towards the
    origin (the SAH peels one off per split): spiral(80, 0.75) -> 17 920 triangles, 17 node levels; (60, 0.7) -> 13"""
    b = _Builder()
    for i in range(n):
        s = q ** i
        b.add(*sphere([4.0 * s, 0.3 * s, 0.0], 0.15 * s, n_phi), mat=0)
    cam = dict(look_from=(0.0, 0.0, 6.0), look_at=(1.0, 0.0, 0.0), up=(0, 1, 0), fov_y=40.0)
    return b.finish("spiral", [[1, 1, 1]], cam)


CONFIGS = {
    # name: (generator, W, H, spp, depth)   — BASELINE.json configs[0..4]
    "C1": (lambda: cornell(50), 256, 256, 4, 5),
    "C2": (lambda: knob(500, 200), 1024, 1024, 16, 8),
    "C3": (lambda: terrain_spheres(), 1920, 1080, 64, 8),
    "C4": (lambda: flattened_blocks(10), 3840, 2160, 256, 8),
    "C5": (lambda: interior(), 1920, 1080, 128, 16),
}
