"""Python host-side mirror of the reference's interface for the hot path, over the C ABI of
libwpt.so (include/wpt.h) via ctypes.  Class and method names follow the reference
(src/main.cpp: Scene.add_triangle_mesh / commit / intersect; src/pinhole.h: ThinlensCamera;
src/fimage.h: Fimage.save_pfm) so tests read like the reference's own driver.

PyTorch is plumbing only: device buffers, streams, torch.distributed.  All compute is in libwpt.so;
there is no fallback — a missing library or a missing GPU raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# WPT_LIB: another build of the same library (libwpt_checked.so, tests/test_gpu_checked.py)
LIB_PATH = os.environ.get("WPT_LIB", os.path.join(_HERE, "libwpt.so"))
WPT_MAX_DEPTH = 64


class WptError(RuntimeError):
    """Raised for any non-zero libwpt status (the reference throws std::runtime_error,
    src/accel/optix_backend.h:25-82)."""

    def __init__(self, code, msg):
        super().__init__(f"wpt error {code}: {msg}")
        self.code = code


class MeshDesc(C.Structure):
    _fields_ = [("positions", C.c_void_p), ("position_triplets", C.c_void_p), ("normals", C.c_void_p),
                ("normal_triplets", C.c_void_p), ("material_id", C.c_void_p), ("albedo", C.c_void_p),
                ("n_positions", C.c_uint32), ("n_triangles", C.c_uint32), ("n_normals", C.c_uint32),
                ("n_materials", C.c_uint32)]


class BvhStats(C.Structure):
    _fields_ = [("n_triangles", C.c_uint32), ("n_nodes", C.c_uint32), ("n_leaf_slots", C.c_uint32),
                ("max_depth", C.c_uint32), ("sah_cost", C.c_double), ("build_ms", C.c_double),
                ("upload_ms", C.c_double), ("node_bytes", C.c_uint64), ("tri_bytes", C.c_uint64),
                ("attr_bytes", C.c_uint64)]


class RayStreams(C.Structure):
    _fields_ = [("ox", C.c_void_p), ("oy", C.c_void_p), ("oz", C.c_void_p), ("dx", C.c_void_p),
                ("dy", C.c_void_p), ("dz", C.c_void_p), ("tmin", C.c_void_p), ("tmax", C.c_void_p),
                ("tmin_scalar", C.c_float), ("tmax_scalar", C.c_float)]


class HitStreams(C.Structure):
    _fields_ = [("tri_id", C.c_void_p), ("t", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p)]


class HitAttrStreams(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("px", "py", "pz", "gnx", "gny", "gnz", "snx", "sny", "snz", "tu", "tv",
                                          "material_id")]


class Camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("x", C.c_float * 3), ("y", C.c_float * 3), ("z", C.c_float * 3),
                ("film_size_y", C.c_float), ("focal_dist", C.c_float), ("fov_y", C.c_float),
                ("lens_radius", C.c_float)]


class RenderConfig(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp_begin", C.c_int32), ("spp_count", C.c_int32),
                ("depth", C.c_int32), ("seed", C.c_uint64), ("tmin", C.c_float), ("tmax", C.c_float)]


class Stats(C.Structure):
    _fields_ = [("segments", C.c_uint64), ("live", C.c_uint64 * WPT_MAX_DEPTH), ("passes", C.c_uint32),
                ("kernel_launches", C.c_uint32), ("ms_total", C.c_float), ("ms_raygen", C.c_float),
                ("ms_trace", C.c_float), ("ms_shade", C.c_float), ("retraced", C.c_uint64), ("tail_from", C.c_int32),
                ("reserved", C.c_int32)]


_lib = None


def lib():
    """Load libwpt.so (in-tree).  Fails loudly if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise WptError(-1, f"{LIB_PATH} not built: run __graft_entry__.build() or make -C smallpt-enoki-optix_b200")
        L = C.CDLL(LIB_PATH)
        L.wpt_last_error.restype = C.c_char_p
        L.wpt_obj_free.argtypes = [C.c_void_p]
        L.wpt_scene_refit.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.wpt_scene_refit_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.wpt_obj_mesh.argtypes = [C.c_void_p, C.c_void_p]
        L.wpt_obj_texcoords.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.wpt_obj_material_name.argtypes = [C.c_void_p, C.c_uint32]
        L.wpt_host_alloc.argtypes = [C.c_size_t, C.c_void_p]
        L.wpt_host_free.argtypes = [C.c_void_p]
        L.wpt_host_register.argtypes = [C.c_void_p, C.c_size_t]
        L.wpt_host_unregister.argtypes = [C.c_void_p]
        L.wpt_host_unregister.restype = None
        L.wpt_film_shard_layout.argtypes = [C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.wpt_film_reduce_scatter.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_void_p]
        L.wpt_film_shard_finish.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p]
        L.wpt_nccl_comm_destroy.argtypes = [C.c_void_p]
        L.wpt_film_reduce.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise WptError(rc, lib().wpt_last_error().decode())


def _f3(v):
    return (C.c_float * 3)(*[float(a) for a in v])


def _np(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def _ptr(t):
    """device (or host) pointer of a torch tensor / numpy array / None"""
    if t is None:
        return None
    if isinstance(t, np.ndarray):
        return C.c_void_p(t.ctypes.data)
    return C.c_void_p(t.data_ptr())


def _stream_handle(stream):
    if stream is None:
        import torch
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if isinstance(stream, int):
        return C.c_void_p(stream)
    return C.c_void_p(stream.cuda_stream)


class ThinlensCamera:
    """src/pinhole.h:7-72.  Same constructor argument order as the reference."""

    def __init__(self, look_from, look_at, up, lens_radius, focal_dist, fov_y, film_size_y=0.035):
        self.c = Camera()
        _check(lib().wpt_camera_init(C.byref(self.c), _f3(look_from), _f3(look_at), _f3(up), C.c_float(lens_radius),
                                     C.c_float(focal_dist), C.c_float(fov_y), C.c_float(film_size_y)))

    @classmethod
    def reference_default(cls):
        """the camera main() builds, src/main.cpp:383"""
        return cls((0.0, 3.03, 5.0), (0.0, 0.03, 0.0), (0.0, 1.0, 0.0), 0.0, 1.0, 40.0 / 180.0 * np.pi)


def load_meshes(path, use_kd=False):
    """load_meshes (src/main.cpp:133-251) through wpt_obj_load.  Returns a dict of host numpy arrays:
    position_triplets (T,3) i32, positions (V,3) f32, normal_triplets (T,3), normals (Vn,3),
    texcoord_triplets (T,3), texcoords (Vt,2), material_id (T,) [tinyobj id + 1], albedo (M+1,3)
    [all white unless use_kd], material_names."""
    L = lib()
    h = C.c_void_p()
    _check(L.wpt_obj_load(str(path).encode(), C.c_int(1 if use_kd else 0), C.byref(h)))
    try:
        d = MeshDesc()
        _check(L.wpt_obj_mesh(h, C.byref(d)))
        T, V, Vn, M = d.n_triangles, d.n_positions, d.n_normals, d.n_materials

        def arr(ptr, n, ct, dt):
            if not ptr or n == 0:
                return np.zeros(0, dt)
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(n,)).astype(dt, copy=True)

        uv, uvt, nuv = C.c_void_p(), C.c_void_p(), C.c_uint32()
        _check(L.wpt_obj_texcoords(h, C.byref(uv), C.byref(uvt), C.byref(nuv)))
        L.wpt_obj_material_name.restype = C.c_char_p
        out = {
            "position_triplets": arr(d.position_triplets, 3 * T, C.c_int32, np.int32).reshape(-1, 3),
            "positions": arr(d.positions, 3 * V, C.c_float, np.float32).reshape(-1, 3),
            "normal_triplets": (arr(d.normal_triplets, 3 * T, C.c_int32, np.int32).reshape(-1, 3) if Vn
                                else np.full((T, 3), -1, np.int32)),
            "normals": arr(d.normals, 3 * Vn, C.c_float, np.float32).reshape(-1, 3),
            "texcoord_triplets": arr(uvt.value, 3 * T, C.c_int32, np.int32).reshape(-1, 3),
            "texcoords": arr(uv.value, 2 * nuv.value, C.c_float, np.float32).reshape(-1, 2),
            "material_id": arr(d.material_id, T, C.c_int32, np.int32),
            "albedo": arr(d.albedo, 3 * M, C.c_float, np.float32).reshape(-1, 3),
            "material_names": [L.wpt_obj_material_name(h, C.c_uint32(i)).decode() for i in range(M)],
        }
    finally:
        L.wpt_obj_free(h)
    return out


class _PinnedBlock:
    """One wpt_host_alloc allocation, exposed to numpy through the array interface: every array made from it — and every
    view of those — keeps this object as its base, so the memory is returned (wpt_host_free) only after the last of them
    has gone, never under a live view."""

    def __init__(self, n_floats):
        self._p = C.c_void_p()
        _check(lib().wpt_host_alloc(C.c_size_t(4 * max(n_floats, 1)), C.byref(self._p)))
        self.__array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (self._p.value, False), "version": 3}

    def __del__(self):
        try:
            if self._p:
                lib().wpt_host_free(self._p)
                self._p = C.c_void_p()
        except Exception:
            pass


class PinnedArray:
    """float32 numpy array over pinned host memory from wpt_host_alloc.  The allocation belongs to `array` (and to any
    view taken from it), not to this object: it is freed when the last of them is dropped."""

    def __init__(self, shape):
        self.array = np.asarray(_PinnedBlock(int(np.prod(shape)))).reshape(shape)


def scene_from_obj(path, device=0, use_kd=False):
    """main()'s load_meshes -> Scene::add_triangle_mesh -> commit (src/main.cpp:362-366)"""
    m = load_meshes(path, use_kd)
    s = Scene()
    has_n = m["normals"].size > 0
    has_t = m["texcoords"].size > 0
    s.add_triangle_mesh(m["positions"], m["position_triplets"], m["normals"] if has_n else None,
                        m["normal_triplets"] if has_n else None, m["material_id"], m["albedo"],
                        texcoords=m["texcoords"] if has_t else None,
                        texcoord_triplets=m["texcoord_triplets"] if has_t else None)
    s.commit(device)
    return s


class Scene:
    """src/main.cpp:286-352."""

    def __init__(self):
        self._h = C.c_void_p()
        self._keep = None
        self.device = None
        self.n_triangles = 0

    def add_triangle_mesh(self, positions, position_triplets, normals=None, normal_triplets=None,
                          material_id=None, albedo=None, texcoords=None, texcoord_triplets=None):
        """Host arrays as produced by load_meshes (src/main.cpp:133-251); see also wpt_obj.load_meshes."""
        if self._h:
            raise WptError(3, "Scene already holds a mesh")
        pos = _np(positions, np.float32).reshape(-1)
        idx = _np(position_triplets, np.int32).reshape(-1)
        d = MeshDesc()
        d.positions, d.position_triplets = pos.ctypes.data, idx.ctypes.data
        d.n_positions, d.n_triangles = pos.size // 3, idx.size // 3
        keep = [pos, idx]
        if normals is not None and normal_triplets is not None:
            nrm = _np(normals, np.float32).reshape(-1)
            nidx = _np(normal_triplets, np.int32).reshape(-1)
            d.normals, d.normal_triplets, d.n_normals = nrm.ctypes.data, nidx.ctypes.data, nrm.size // 3
            keep += [nrm, nidx]
        if material_id is not None:
            mat = _np(material_id, np.int32).reshape(-1)
            d.material_id = mat.ctypes.data
            keep.append(mat)
        if albedo is not None:
            alb = _np(albedo, np.float32).reshape(-1)
            d.albedo, d.n_materials = alb.ctypes.data, alb.size // 3
            keep.append(alb)
        _check(lib().wpt_scene_create(C.byref(d), C.byref(self._h)))
        self.n_triangles = d.n_triangles
        if texcoords is not None and texcoord_triplets is not None:
            uv = _np(texcoords, np.float32).reshape(-1)
            uvi = _np(texcoord_triplets, np.int32).reshape(-1)
            assert uvi.size == idx.size
            _check(lib().wpt_scene_set_texcoords(self._h, C.c_void_p(uv.ctypes.data), C.c_void_p(uvi.ctypes.data),
                                                 C.c_uint32(uv.size // 2)))

    def set_texture(self, material, image):
        """ImageTexture of one material (src/main.cpp:35-80): image float32 (h, w, 3); before commit()"""
        img = _np(image, np.float32)
        assert img.ndim == 3 and img.shape[2] == 3
        _check(lib().wpt_scene_set_texture(self._h, C.c_uint32(material), C.c_void_p(img.ctypes.data),
                                           C.c_uint32(img.shape[1]), C.c_uint32(img.shape[0])))

    def commit(self, device=0, builder="sah", cache_path=None):
        """builder "sah": host SAH build (default); "lbvh": Morton-order build on the device (wpt_scene_commit_lbvh)"""
        if not self._h:
            raise WptError(3, "Scene.commit: no mesh")
        self.from_cache = False
        if cache_path is not None:
            f = C.c_int(0)
            _check(lib().wpt_scene_commit_cached(self._h, C.c_int(device), str(cache_path).encode(), C.byref(f)))
            self.from_cache = bool(f.value)
        elif builder == "lbvh":
            _check(lib().wpt_scene_commit_lbvh(self._h, C.c_int(device)))
        else:
            _check(lib().wpt_scene_commit(self._h, C.c_int(device)))
        self.device = device

    def refit(self, positions):
        """wpt_scene_refit: moved vertices, same topology.  positions: host (V,3) float32, or a CUDA float32
        tensor of 3V elements (device path, asynchronous on the current stream)."""
        if isinstance(positions, np.ndarray) or isinstance(positions, (list, tuple)):
            pos = _np(positions, np.float32).reshape(-1)
            _check(lib().wpt_scene_refit(self._h, C.c_void_p(pos.ctypes.data), C.c_uint32(pos.size // 3)))
        else:
            assert positions.is_cuda and positions.is_contiguous()
            _check(lib().wpt_scene_refit_device(self._h, C.c_void_p(positions.data_ptr()), _stream_handle(None)))

    def refit_status(self):
        """True if the last refit saw a non-finite vertex (wpt_scene_refit_status; synchronises)"""
        f = C.c_int(0)
        _check(lib().wpt_scene_refit_status(self._h, C.byref(f), _stream_handle(None)))
        return bool(f.value)

    def stats(self):
        s = BvhStats()
        _check(lib().wpt_scene_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in BvhStats._fields_}

    def intersect(self, origin, direction, active=None, tmin=1e-3, tmax=1e20, out=None, stream=None):
        """OptixBackend::intersect's launch (src/accel/optix_backend.h:422-460).
        origin, direction: 3-tuples of 1-D CUDA float32 tensors (the SoA streams of Ray3C);
        active: None, or a uint8/bool CUDA tensor of size n or 1.  Returns (tri_id, t, u, v) CUDA tensors;
        lanes masked out keep whatever `out` held."""
        import torch
        n = origin[0].numel()
        for t_ in list(origin) + list(direction):
            assert t_.is_cuda and t_.dtype == torch.float32 and t_.is_contiguous() and t_.numel() == n
        if out is None:
            dev = origin[0].device
            out = (torch.full((n,), -2, dtype=torch.int32, device=dev),
                   torch.full((n,), float("nan"), dtype=torch.float32, device=dev),
                   torch.full((n,), float("nan"), dtype=torch.float32, device=dev),
                   torch.full((n,), float("nan"), dtype=torch.float32, device=dev))
        rs = self._ray_streams(origin, direction, tmin, tmax)
        hs = HitStreams(*[t_.data_ptr() for t_ in out])
        mptr, msz = None, 0
        if active is not None:
            if active.dtype == torch.bool:
                active = active.view(torch.uint8)
            assert active.is_cuda and active.dtype == torch.uint8
            mptr, msz = C.c_void_p(active.data_ptr()), active.numel()
        _check(lib().wpt_intersect(self._h, C.byref(rs), mptr, C.c_uint32(msz), C.c_uint32(n), C.byref(hs),
                                   _stream_handle(stream)))
        return out

    @staticmethod
    def _ray_streams(origin, direction, tmin=1e-3, tmax=1e20):
        """struct Params' eight ray streams (src/optixdata.h:11-29): tmin / tmax are per-ray CUDA float32 tensors (the
        form Ray3C carries, src/ray.h:20-26) or scalars"""
        import torch
        n = origin[0].numel()
        rs = RayStreams(*[t_.data_ptr() for t_ in list(origin) + list(direction)], None, None, 0.0, 0.0)
        for name, val in (("tmin", tmin), ("tmax", tmax)):
            if torch.is_tensor(val):
                assert val.is_cuda and val.dtype == torch.float32 and val.is_contiguous() and val.numel() == n
                setattr(rs, name, val.data_ptr())
            else:
                setattr(rs, name + "_scalar", float(val))
        return rs

    def hit_attributes(self, origin, direction, hit, want=("p", "gn", "sn", "uv", "mat"), stream=None):
        """The derived attributes OptixBackend::intersect returns with a hit (src/accel/optix_backend.h:463-486) and
        the material index of Scene::intersect (src/main.cpp:325).  origin / direction: the ray streams; hit: the
        (tri_id, t, u, v) tensors of intersect().  Returns a dict of CUDA tensors: p [3][n], gn [3][n], sn [3][n],
        uv [2][n], mat [n]; lanes that missed keep NaN / -1."""
        import torch
        n = hit[0].numel()
        dev = hit[0].device
        out, hs = {}, HitAttrStreams()
        for key, names in (("p", ("px", "py", "pz")), ("gn", ("gnx", "gny", "gnz")), ("sn", ("snx", "sny", "snz")),
                           ("uv", ("tu", "tv"))):
            if key in want:
                out[key] = torch.full((len(names), n), float("nan"), dtype=torch.float32, device=dev)
                for k, nm in enumerate(names):
                    setattr(hs, nm, out[key][k].data_ptr())
        if "mat" in want:
            out["mat"] = torch.full((n,), -1, dtype=torch.int32, device=dev)
            hs.material_id = out["mat"].data_ptr()
        rs = self._ray_streams(origin, direction)
        h = HitStreams(*[t_.data_ptr() for t_ in hit])
        _check(lib().wpt_hit_attributes(self._h, C.byref(rs), C.byref(h), C.c_uint32(n), C.byref(hs), _stream_handle(stream)))
        return out

    def intersect_host(self, o, d, active=None, tmin=1e-3, tmax=1e20):
        """Host numpy in/out through wpt_intersect_host (copies inside).  o, d: (n,3) float32."""
        o = _np(o, np.float32).reshape(-1, 3)
        d = _np(d, np.float32).reshape(-1, 3)
        n = o.shape[0]
        cols = [np.ascontiguousarray(o[:, k]) for k in range(3)] + [np.ascontiguousarray(d[:, k]) for k in range(3)]
        rs = RayStreams(*[c.ctypes.data for c in cols], None, None, tmin, tmax)
        hid = np.full(n, -2, np.int32)
        t = np.full(n, np.nan, np.float32)
        u = np.full(n, np.nan, np.float32)
        v = np.full(n, np.nan, np.float32)
        hs = HitStreams(hid.ctypes.data, t.ctypes.data, u.ctypes.data, v.ctypes.data)
        mptr, msz = None, 0
        if active is not None:
            active = _np(active, np.uint8)
            mptr, msz = C.c_void_p(active.ctypes.data), active.size
        _check(lib().wpt_intersect_host(self._h, C.byref(rs), mptr, C.c_uint32(msz), C.c_uint32(n), C.byref(hs)))
        return hid, t, u, v

    def shade_step(self, o, d, tri_id, t, u, v, s0, s1):
        """per-function device entry point (wpt_shade_step); all 1-D CUDA tensors, o/d 3-tuples"""
        import torch
        n = tri_id.numel()
        dev = tri_id.device
        no = [torch.zeros(n, device=dev) for _ in range(3)]
        nd = [torch.zeros(n, device=dev) for _ in range(3)]
        w = [torch.zeros(n, device=dev) for _ in range(3)]
        arr = lambda ts: (C.c_void_p * 3)(*[t_.data_ptr() for t_ in ts])
        _check(lib().wpt_shade_step(self._h, C.c_uint32(n), arr(o), arr(d), _ptr(tri_id), _ptr(t), _ptr(u), _ptr(v),
                                    _ptr(s0), _ptr(s1), arr(no), arr(nd), arr(w), _stream_handle(None)))
        return no, nd, w

    def close(self):
        if self._h:
            lib().wpt_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Renderer:
    """The sample/bounce loop of main() (src/main.cpp:385-429) as a wavefront pipeline."""

    def __init__(self, scene, width, height, spp_per_pass=1):
        self.scene = scene
        self.width, self.height = width, height
        self._h = C.c_void_p()
        _check(lib().wpt_renderer_create(scene._h, C.c_int32(width), C.c_int32(height), C.c_int32(spp_per_pass),
                                         C.byref(self._h)))

    def set_profile(self, on):
        _check(lib().wpt_renderer_set_profile(self._h, C.c_int(1 if on else 0)))

    def set_depth_seeding(self, on=True):
        """True / False, or an int > 1 = seed margin in 1/10000 (tests)"""
        _check(lib().wpt_renderer_set_depth_seeding(self._h, C.c_int(int(on))))

    def set_depth_shard(self, comm, rank, n_ranks):
        """wpt_renderer_set_depth_shard: rank `rank` of `n_ranks` traces its share of the depth pre-pass, the hit distances
        are all-gathered on `comm` (a wpt_nccl communicator handle) inside render()"""
        _check(lib().wpt_renderer_set_depth_shard(self._h, comm, C.c_int(rank), C.c_int(n_ranks)))

    def set_tail(self, first_bounce=0):
        """bounces >= first_bounce run in one fused trace+shade launch per pass; 0 = default, negative = off"""
        _check(lib().wpt_renderer_set_tail(self._h, C.c_int(int(first_bounce))))

    def set_tuning(self, trace_variant=0, ctas_per_sm=0, overlap_lanes=0):
        _check(lib().wpt_renderer_set_tuning(self._h, C.c_int(trace_variant), C.c_int(ctas_per_sm), C.c_int(overlap_lanes)))

    def _cfg(self, spp_begin, spp_count, depth, seed, tmin, tmax):
        return RenderConfig(self.width, self.height, spp_begin, spp_count, depth, seed, tmin, tmax)

    def render(self, camera, film, spp_begin, spp_count, depth, seed=0, tmin=1e-3, tmax=1e20, stream=None):
        """film: CUDA float32 tensor [3, H*W] (planar, un-normalised sum is ADDED). Asynchronous."""
        assert film.is_cuda and film.is_contiguous() and film.numel() == 3 * self.width * self.height
        N = self.width * self.height
        base = film.data_ptr()
        cfg = self._cfg(spp_begin, spp_count, depth, seed, tmin, tmax)
        _check(lib().wpt_render(self._h, C.byref(camera.c), C.byref(cfg), C.c_void_p(base), C.c_void_p(base + 4 * N),
                                C.c_void_p(base + 8 * N), _stream_handle(stream)))

    def finish(self, stream=None):
        st = Stats()
        _check(lib().wpt_render_finish(self._h, C.byref(st), _stream_handle(stream)))
        return _stats_dict(st)

    def render_host(self, camera, spp_begin, spp_count, depth, seed=0, tmin=1e-3, tmax=1e20, out=None, copy=True):
        """whole frame to HOST planar film [3, H*W], normalised by spp_count (src/main.cpp:369-439).  Without `out` the
        frame lands in the renderer's own pinned film; by default the caller gets a copy of it (an array of its own, at
        the price of one more pass over 12 bytes per pixel on the host).  copy=False hands out the pinned film itself: no
        extra pass, but the NEXT render_host of this renderer overwrites it.  With `out` the frame is written there and
        `out` is returned."""
        N = self.width * self.height
        if out is None:
            # the renderer's own pinned film (wpt_host_alloc): the D2H copies are plain DMA transfers
            if getattr(self, "_pinned", None) is None:
                self._pinned = PinnedArray((3, N))
            out = self._pinned.array
        film = out
        cfg = self._cfg(spp_begin, spp_count, depth, seed, tmin, tmax)
        st = Stats()
        _check(lib().wpt_render_host(self._h, C.byref(camera.c), C.byref(cfg), C.c_void_p(film[0].ctypes.data),
                                     C.c_void_p(film[1].ctypes.data), C.c_void_p(film[2].ctypes.data), C.byref(st)))
        own = film is getattr(getattr(self, "_pinned", None), "array", None)
        return (film.copy() if own and copy else film), _stats_dict(st)

    def close(self):
        if self._h:
            lib().wpt_renderer_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _stats_dict(st):
    d = {k: getattr(st, k) for k, _ in Stats._fields_ if k != "live"}
    d["live"] = [int(x) for x in st.live]
    return d


def film_scale(film, scale, stream=None):
    N = film.numel() // 3
    base = film.data_ptr()
    _check(lib().wpt_film_scale(C.c_void_p(base), C.c_void_p(base + 4 * N), C.c_void_p(base + 8 * N), C.c_uint32(N),
                                C.c_float(scale), _stream_handle(stream)))


def raygen(camera, width, height, spp_index, seed=0):
    """wpt_raygen: camera rays of one global sample index for all pixels -> (o[3], d[3]) CUDA tensors"""
    import torch
    N = width * height
    o = [torch.empty(N, device="cuda") for _ in range(3)]
    d = [torch.empty(N, device="cuda") for _ in range(3)]
    _check(lib().wpt_raygen(C.byref(camera.c), C.c_int32(width), C.c_int32(height), C.c_int32(spp_index),
                            C.c_uint64(seed), *[_ptr(t) for t in o + d], _stream_handle(None)))
    return o, d


def philox(n, c1, c2, key):
    import torch
    out = torch.empty(4 * n, dtype=torch.int32, device="cuda")
    _check(lib().wpt_philox(C.c_uint32(n), C.c_uint32(c1), C.c_uint32(c2), C.c_uint64(key), _ptr(out),
                            _stream_handle(None)))
    return out.view(n, 4)


class Fimage:
    """src/fimage.h:5-59"""

    @staticmethod
    def save_pfm(r, g, b, width, height, filepath):
        r, g, b = (_np(x, np.float32).reshape(-1) for x in (r, g, b))
        _check(lib().wpt_save_pfm(str(filepath).encode(), C.c_void_p(r.ctypes.data), C.c_void_p(g.ctypes.data),
                                  C.c_void_p(b.ctypes.data), C.c_uint32(width), C.c_uint32(height)))


def film_save(path, film, width, height, spp_done, seed=0):
    """checkpoint of a render in progress (wpt_film_save): film = HOST un-normalised planar [3, H*W] float32"""
    f = _np(film, np.float32).reshape(3, -1)
    assert f.shape[1] == width * height
    _check(lib().wpt_film_save(str(path).encode(), C.c_void_p(f[0].ctypes.data), C.c_void_p(f[1].ctypes.data),
                               C.c_void_p(f[2].ctypes.data), C.c_uint32(width), C.c_uint32(height), C.c_uint64(spp_done),
                               C.c_uint64(seed)))


def film_info(path):
    """(width, height, spp_done, seed) of a checkpoint (wpt_film_info)"""
    w, h, n, s = C.c_uint32(), C.c_uint32(), C.c_uint64(), C.c_uint64()
    _check(lib().wpt_film_info(str(path).encode(), C.byref(w), C.byref(h), C.byref(n), C.byref(s)))
    return w.value, h.value, n.value, s.value


def film_load(path, width, height):
    """-> (film [3, H*W] float32 un-normalised, spp_done, seed) (wpt_film_load)"""
    f = np.empty((3, width * height), np.float32)
    n, s = C.c_uint64(), C.c_uint64()
    _check(lib().wpt_film_load(str(path).encode(), C.c_void_p(f[0].ctypes.data), C.c_void_p(f[1].ctypes.data),
                               C.c_void_p(f[2].ctypes.data), C.c_uint32(width), C.c_uint32(height), C.byref(n), C.byref(s)))
    return f, n.value, s.value


def scene_from(sd, device=0, builder="sah", cache_path=None):
    """SceneData (scenes.py) -> committed Scene"""
    s = Scene()
    s.add_triangle_mesh(sd.pos, sd.idx, sd.nrm, sd.nidx, sd.mat, sd.albedo)
    s.commit(device, builder, cache_path)
    return s


def camera_from(sd):
    c = sd.camera
    return ThinlensCamera(c["look_from"], c["look_at"], c["up"], 0.0, 1.0, c["fov_y"])
