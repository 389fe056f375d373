"""ctypes binding of the CPU oracle (oracle/libwpt_oracle.so) and of the real Embree 3.6.1
reference (oracle/_ref/libembree_ref.so).  TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by the product."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


class OrcMesh(C.Structure):
    _fields_ = [("pos", C.c_void_p), ("pos_idx", C.c_void_p), ("nrm", C.c_void_p), ("nrm_idx", C.c_void_p),
                ("mat_id", C.c_void_p), ("albedo", C.c_void_p),
                ("n_tris", C.c_uint32), ("n_pos", C.c_uint32), ("n_nrm", C.c_uint32), ("n_mat", C.c_uint32),
                ("uvs", C.c_void_p), ("uv_idx", C.c_void_p), ("tex", C.c_void_p)]


class OrcTexture(C.Structure):
    _fields_ = [("rgb", C.c_void_p), ("w", C.c_int32), ("h", C.c_int32)]


class OrcCamera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("x", C.c_float * 3), ("y", C.c_float * 3), ("z", C.c_float * 3),
                ("film_size_y", C.c_float), ("focal_dist", C.c_float), ("fov_y", C.c_float),
                ("lens_radius", C.c_float)]


class OrcRenderCfg(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp_begin", C.c_int32), ("spp_count", C.c_int32),
                ("depth", C.c_int32), ("rng_mode", C.c_int32), ("seed", C.c_uint64),
                ("tmin", C.c_float), ("tmax", C.c_float)]


class OrcPcg(C.Structure):
    _fields_ = [("state", C.c_uint64), ("inc", C.c_uint64)]


RNG_PCG32, RNG_PHILOX = 0, 1
_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        p = os.path.join(_HERE, "libwpt_oracle.so")
        if not os.path.exists(p):
            raise RuntimeError("oracle not built: run `make -C oracle` (or __graft_entry__.build())")
        L = C.CDLL(p)
        L.orc_render.restype = C.c_uint64
        L.orc_bvh_build.restype = C.c_void_p
        L.orc_bvh_build.argtypes = [C.POINTER(OrcMesh)]
        L.orc_bvh_free.argtypes = [C.c_void_p]
        L.orc_pcg32_next_u32.restype = C.c_uint32
        L.orc_pcg32_next_f32.restype = C.c_float
        L.orc_u32_to_f32.restype = C.c_float
        L.orc_num_threads.restype = C.c_int
        L.orc_tri_intersect.restype = C.c_int
        _lib = L
    return _lib


def ref_available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libembree_ref.so"))


def ref():
    global _ref
    if _ref is None:
        p = os.path.join(_HERE, "_ref", "libembree_ref.so")
        if not os.path.exists(p):
            raise RuntimeError("oracle/_ref/libembree_ref.so missing (built only where /root/reference exists)")
        L = C.CDLL(p)
        L.eref_scene_create.restype = C.c_void_p
        L.eref_scene_create.argtypes = [_f32p, C.c_uint32, _i32p, C.c_uint32, C.c_int, C.c_int]
        L.eref_scene_destroy.argtypes = [C.c_void_p]
        L.eref_intersect1_fn.restype = C.c_void_p
        L.eref_version.restype = C.c_char_p
        _ref = L
    return _ref


def ref_builder_sah(pos, idx, branching=7, max_leaf=3, c_trav=1.0, c_isect=0.3, quality=1):
    """SAH cost of the tree the reference tree's own builder (rtcBuildBVH, include/embree3/rtcore_builder.h:26-110) makes
    of the same triangles with the given cost constants.  Returns dict(sah, nodes, leaves, depth, prims)."""
    pos = _c(pos, np.float32).reshape(-1, 3)
    idx = _c(idx, np.int32).reshape(-1, 3)
    out = (C.c_double * 5)()
    rc = ref().eref_builder_sah(pos.reshape(-1).ctypes.data_as(C.c_void_p), C.c_uint32(pos.shape[0]),
                                idx.reshape(-1).ctypes.data_as(C.c_void_p), C.c_uint32(idx.shape[0]), C.c_uint(branching),
                                C.c_uint(max_leaf), C.c_float(c_trav), C.c_float(c_isect), C.c_int(quality), out)
    if rc != 0:
        raise RuntimeError(f"rtcBuildBVH failed ({rc})")
    return dict(sah=out[0], nodes=int(out[1]), leaves=int(out[2]), depth=int(out[3]), prims=int(out[4]))


def ref_obj_available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libtinyobj_ref.so"))


def ref_load_obj(path):
    """The reference's own loader chain: its vendored tinyobjloader flattened like load_meshes
    (src/main.cpp:143-223), through oracle/_ref/libtinyobj_ref.so (tinyobj_ref_wrap.cpp)."""
    L = C.CDLL(os.path.join(_HERE, "_ref", "libtinyobj_ref.so"))
    L.tor_load.restype = C.c_void_p
    L.tor_counts.argtypes = [C.c_void_p, C.c_void_p]
    L.tor_copy.argtypes = [C.c_void_p] * 9
    L.tor_free.argtypes = [C.c_void_p]
    h = L.tor_load(str(path).encode())
    if not h:
        raise RuntimeError(f"tinyobjloader failed on {path}")
    cnt = (C.c_uint32 * 5)()
    L.tor_counts(h, cnt)
    V, Vn, Vt, T, M = [int(x) for x in cnt]
    out = {"positions": np.zeros((V, 3), np.float32), "normals": np.zeros((Vn, 3), np.float32),
           "texcoords": np.zeros((Vt, 2), np.float32), "kd": np.zeros((M, 3), np.float32),
           "position_triplets": np.zeros((T, 3), np.int32), "normal_triplets": np.zeros((T, 3), np.int32),
           "texcoord_triplets": np.zeros((T, 3), np.int32), "material_id": np.zeros(T, np.int32)}
    L.tor_copy(h, *[out[k].ctypes.data for k in ("positions", "normals", "texcoords", "kd", "position_triplets",
                                                 "normal_triplets", "texcoord_triplets", "material_id")])
    L.tor_free(h)
    return out


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


class Mesh:
    """Holds the numpy arrays alive and exposes an orc_mesh."""

    def __init__(self, pos, pos_idx, nrm=None, nrm_idx=None, mat_id=None, albedo=None):
        self.pos = _c(pos, np.float32).reshape(-1, 3)
        self.pos_idx = _c(pos_idx, np.int32).reshape(-1, 3)
        self.nrm = _c(nrm if nrm is not None else np.zeros((1, 3)), np.float32).reshape(-1, 3)
        self.nrm_idx = _c(nrm_idx if nrm_idx is not None else np.zeros_like(self.pos_idx), np.int32).reshape(-1, 3)
        T = self.pos_idx.shape[0]
        self.mat_id = _c(mat_id if mat_id is not None else np.zeros(T), np.int32)
        self.albedo = _c(albedo if albedo is not None else np.ones((1, 3)), np.float32).reshape(-1, 3)
        self.c = OrcMesh(self.pos.ctypes.data, self.pos_idx.ctypes.data, self.nrm.ctypes.data,
                         self.nrm_idx.ctypes.data, self.mat_id.ctypes.data, self.albedo.ctypes.data,
                         T, self.pos.shape[0], self.nrm.shape[0], self.albedo.shape[0])
        self._bvh = None
        self._eref = None

    def set_textures(self, uvs, uv_idx, textures):
        """textures: dict material -> float32 array (h, w, 3); uvs (n,2), uv_idx (T,3) with -1 = none"""
        self._uvs = _c(uvs, np.float32).reshape(-1, 2)
        self._uvidx = _c(uv_idx, np.int32).reshape(-1, 3)
        n_mat = self.albedo.shape[0]
        self._tex_arr = (OrcTexture * n_mat)()
        self._tex_keep = {}
        for m, img in textures.items():
            img = _c(img, np.float32)
            self._tex_keep[m] = img
            self._tex_arr[m] = OrcTexture(img.ctypes.data, img.shape[1], img.shape[0])
        self.c.uvs = self._uvs.ctypes.data
        self.c.uv_idx = self._uvidx.ctypes.data
        self.c.tex = C.cast(self._tex_arr, C.c_void_p).value

    @property
    def n_tris(self):
        return self.pos_idx.shape[0]

    # -- intersector back-ends: (fn_ptr, ctx) pairs with the orc_intersect1_fn signature
    def backend(self, kind):
        L = lib()
        if kind == "brute":
            return C.cast(L.orc_intersect1_brute, C.c_void_p).value, C.addressof(self.c)
        if kind == "bvh":
            if self._bvh is None:
                self._bvh = L.orc_bvh_build(C.byref(self.c))
            return C.cast(L.orc_intersect1_bvh, C.c_void_p).value, self._bvh
        if kind in ("embree", "embree_robust"):
            R = ref()
            key = kind
            if self._eref is None:
                self._eref = {}
            if key not in self._eref:
                self._eref[key] = R.eref_scene_create(self.pos.reshape(-1), self.pos.shape[0],
                                                      self.pos_idx.reshape(-1), self.n_tris,
                                                      1 if kind == "embree_robust" else 0, 0)
            return R.eref_intersect1_fn(), self._eref[key]
        raise ValueError(kind)

    def intersect(self, kind, o, d, tmin=1e-3, tmax=1e20, mask=None, threads=0):
        """o,d: (n,3) float32. Returns id,t,u,v (masked-out / missed lanes keep id=-2/-1, t,u,v = nan)."""
        o = _c(o, np.float32).reshape(-1, 3)
        d = _c(d, np.float32).reshape(-1, 3)
        n = o.shape[0]
        cols = [np.ascontiguousarray(o[:, k]) for k in range(3)] + [np.ascontiguousarray(d[:, k]) for k in range(3)]
        hid = np.full(n, -2, np.int32)
        t = np.full(n, np.nan, np.float32)
        u = np.full(n, np.nan, np.float32)
        v = np.full(n, np.nan, np.float32)
        fn, ctx = self.backend(kind)
        if isinstance(tmin, np.ndarray) or isinstance(tmax, np.ndarray):
            # per-ray windows (src/ray.h:20-26): both as float32 arrays of n
            assert mask is None
            tmin = np.ascontiguousarray(np.broadcast_to(np.asarray(tmin, np.float32), (n,)))
            tmax = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, np.float32), (n,)))
            lib().orc_intersect_streams_win(C.c_void_p(fn), C.c_void_p(ctx),
                                            *[c.ctypes.data_as(C.c_void_p) for c in cols],
                                            tmin.ctypes.data_as(C.c_void_p), tmax.ctypes.data_as(C.c_void_p), C.c_uint32(n),
                                            hid.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p),
                                            u.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p), C.c_int(threads))
            return hid, t, u, v
        if mask is None:
            mptr, msz = None, 0
        else:
            mask = _c(mask, np.uint8)
            mptr, msz = mask.ctypes.data_as(C.c_void_p), mask.size
        lib().orc_intersect_streams(C.c_void_p(fn), C.c_void_p(ctx),
                                    *[c.ctypes.data_as(C.c_void_p) for c in cols],
                                    C.c_float(tmin), C.c_float(tmax), mptr, C.c_uint32(msz), C.c_uint32(n),
                                    hid.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p),
                                    u.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p), C.c_int(threads))
        return hid, t, u, v

    def render(self, kind, cam, W, H, spp_begin, spp_count, depth, rng_mode=RNG_PCG32, seed=0,
               tmin=1e-3, tmax=1e20, threads=0):
        """Returns (film[3,H*W] un-normalised sum, segments)."""
        film = np.zeros((3, W * H), np.float32)
        cfg = OrcRenderCfg(W, H, spp_begin, spp_count, depth, rng_mode, seed, tmin, tmax)
        fn, ctx = self.backend(kind)
        seg = lib().orc_render(C.byref(self.c), C.c_void_p(fn), C.c_void_p(ctx), C.byref(cam), C.byref(cfg),
                               film[0].ctypes.data_as(C.c_void_p), film[1].ctypes.data_as(C.c_void_p),
                               film[2].ctypes.data_as(C.c_void_p), C.c_int(threads))
        return film, int(seg)

    def shade_step(self, o, d, tri, t, u, v, s):
        arr = lambda x: (C.c_float * len(x))(*[float(a) for a in x])
        no, nd, w = (C.c_float * 3)(), (C.c_float * 3)(), (C.c_float * 3)()
        lib().orc_shade_step(C.byref(self.c), arr(o), arr(d), C.c_int32(int(tri)), C.c_float(t), C.c_float(u),
                             C.c_float(v), arr(s), no, nd, w)
        return np.array(no, np.float32), np.array(nd, np.float32), np.array(w, np.float32)

    def __del__(self):
        try:
            if self._bvh is not None:
                lib().orc_bvh_free(C.c_void_p(self._bvh))
            if self._eref:
                for h in self._eref.values():
                    ref().eref_scene_destroy(C.c_void_p(h))
        except Exception:
            pass


def camera(look_from, look_at, up=(0, 1, 0), lens_radius=0.0, focal_dist=1.0, fov_y=40.0 / 180.0 * np.pi,
           film_size_y=0.035):
    cam = OrcCamera()
    f3 = lambda x: (C.c_float * 3)(*[float(a) for a in x])
    lib().orc_camera_init(C.byref(cam), f3(look_from), f3(look_at), f3(up), C.c_float(lens_radius),
                          C.c_float(focal_dist), C.c_float(fov_y), C.c_float(film_size_y))
    return cam
