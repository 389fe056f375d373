// wpt_shim.hpp — header-only C++ shim over the C ABI (wpt.h) that keeps the reference's type and
// method names, so that its main() (src/main.cpp:355-445) reads the same after the swap:
//
//   reference (Enoki + OptiX)                           here (libwpt.so)
//   Scene::add_triangle_mesh(path)   main.cpp:288-310   wpt_obj_load + wpt_scene_create
//   Scene::commit()                  main.cpp:312-318   wpt_scene_commit (SAH BVH build + upload)
//   Scene::intersect(ray, active)    main.cpp:320-340   wpt_intersect on the same SoA streams, + wpt_hit_attributes for
//                                                       the Interaction (position, normals, texcoord, material)
//   Frame3<Real>                     coordframe.h:5-54  Frame3 (host scalar form, Frame3S)
//   N GPUs (no counterpart)                             ShardedPathLoop: spp shards + ONE wpt_film_reduce (NCCL)
//   ThinlensCamera(...)              pinhole.h:9-25     wpt_camera_init
//   the sample / bounce loop         main.cpp:385-429   PathLoop::render -> wpt_render_host
//   Fimage::save_pfm                 fimage.h:33-58     wpt_save_pfm
//   StopWatch                        stopwatch.h:6-27   same (std::chrono)
//
// Errors: the C ABI returns codes; like the reference's OPTIX_CHECK / CUDA_CHECK
// (src/accel/optix_backend.h:25-82) the shim throws std::runtime_error with the message.
#pragma once
#include <chrono>
#include <cmath>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "wpt.h"

namespace wpt_shim {

inline void check(int rc)
{
    if (rc != WPT_OK) throw std::runtime_error(std::string("wpt: ") + wpt_last_error());
}

struct Real3 {
    float x, y, z;
    Real3() : x(0), y(0), z(0) {}
    Real3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};

// src/coordframe.h:5-54, scalar instantiation (Frame3S): orthonormal basis about `normal` as the local Y axis;
// m_world_from_local is row-major with the basis vectors as COLUMNS.  Same operation order as the reference
// (and as the device code, csrc/shade_core.h frame_to_world).
struct Frame3 {
    Frame3() : m_world_from_local{1, 0, 0, 0, 1, 0, 0, 0, 1} {}
    explicit Frame3(const Real3 &normal)
    {
        const float sign = std::copysign(1.0f, normal.y);
        const float a = -1.0f / (sign + normal.y);
        const float b = normal.z * normal.x * a;
        set(Real3(sign + normal.x * normal.x * a, -normal.x, b), normal,
            Real3(sign * b, -sign * normal.z, 1.0f + sign * normal.z * normal.z * a));
    }
    Frame3(const Real3 &basis_x, const Real3 &basis_y, const Real3 &basis_z) { set(basis_x, basis_y, basis_z); }
    Real3 to_local(const Real3 &w) const        // transpose(M) * world
    {
        const float *m = m_world_from_local;
        return Real3(std::fma(m[6], w.z, std::fma(m[3], w.y, m[0] * w.x)), std::fma(m[7], w.z, std::fma(m[4], w.y, m[1] * w.x)),
                     std::fma(m[8], w.z, std::fma(m[5], w.y, m[2] * w.x)));
    }
    Real3 to_world(const Real3 &l) const        // M * local
    {
        const float *m = m_world_from_local;
        return Real3(std::fma(m[2], l.z, std::fma(m[1], l.y, m[0] * l.x)), std::fma(m[5], l.z, std::fma(m[4], l.y, m[3] * l.x)),
                     std::fma(m[8], l.z, std::fma(m[7], l.y, m[6] * l.x)));
    }
    float m_world_from_local[9];

private:
    void set(const Real3 &x, const Real3 &y, const Real3 &z)
    {
        const float m[9] = {x.x, y.x, z.x, x.y, y.y, z.y, x.z, y.z, z.z};
        for (int k = 0; k < 9; k++) m_world_from_local[k] = m[k];
    }
};
using Frame3S = Frame3;

// a device array of n elements (the role of Enoki's CUDAArray<T>, enoki/cuda.h:210-220, at this boundary)
template <class T>
struct DeviceArray {
    DeviceArray() = default;
    explicit DeviceArray(size_t n_, int device = 0) : n(n_)
    {
        if (n) check(wpt_device_alloc(device, n * sizeof(T), reinterpret_cast<void **>(&p)));
    }
    DeviceArray(const DeviceArray &) = delete;
    DeviceArray &operator=(const DeviceArray &) = delete;
    DeviceArray(DeviceArray &&o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DeviceArray &operator=(DeviceArray &&o) noexcept
    {
        if (this != &o) { wpt_device_free(p); p = o.p; n = o.n; o.p = nullptr; o.n = 0; }
        return *this;
    }
    ~DeviceArray() { wpt_device_free(p); }
    static DeviceArray copy(const T *host, size_t n_, int device = 0)      // CUDAArray<T>::copy
    {
        DeviceArray a(n_, device);
        if (n_) check(wpt_device_upload(a.p, host, n_ * sizeof(T), nullptr));
        return a;
    }
    std::vector<T> to_host() const
    {
        std::vector<T> h(n);
        if (n) check(wpt_device_download(h.data(), p, n * sizeof(T), nullptr));
        return h;
    }
    T *data() const { return p; }
    size_t size() const { return n; }
    T *p = nullptr;
    size_t n = 0;
};
using RealC = DeviceArray<float>;
using IntC = DeviceArray<int32_t>;

// src/stopwatch.h:6-27
struct StopWatch {
    void reset() { start = std::chrono::high_resolution_clock::now(); }
    long long timeMilliSec() const
    {
        return std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - start).count();
    }
    long long timeNanoSec() const
    {
        return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::high_resolution_clock::now() - start).count();
    }
    std::chrono::high_resolution_clock::time_point start;
};

// src/pinhole.h:7-72
struct ThinlensCamera {
    ThinlensCamera(const Real3 &look_from, const Real3 &look_at, const Real3 &up, float lens_radius, float focal_dist,
                   float fov_y, float film_size_y = 0.035f)
    {
        const float f[3] = {look_from.x, look_from.y, look_from.z}, a[3] = {look_at.x, look_at.y, look_at.z},
                    u[3] = {up.x, up.y, up.z};
        check(wpt_camera_init(&c, f, a, u, lens_radius, focal_dist, fov_y, film_size_y));
    }
    wpt_camera c;
};

// SoA ray / hit streams in DEVICE memory: Ray<Real3C> (src/ray.h:5-33) and TriangleHitInfoC
// (src/accel/optix_backend.h:99-134) reduced to what crosses the boundary (src/optixdata.h:11-29)
struct Ray3C {
    const float *ox, *oy, *oz, *dx, *dy, *dz;
    uint32_t size;
    float tmin = 0.001f, tmax = 1e20f;      // src/ray.h:16
};
struct TriangleHitInfoC {
    int32_t *tri_id;
    float *t, *u, *v;
};
// the complete TriangleHitInfo<IntC, RealC> of OptixBackend::intersect (src/accel/optix_backend.h:99-134): hit record
// + derived attributes, all device arrays of n lanes, owned here
struct TriangleHitInfoFullC {
    explicit TriangleHitInfoFullC(size_t n, int device = 0)
        : m_tri_id(n, device), m_t(n, device), m_barycentric{RealC(n, device), RealC(n, device)},
          m_position{RealC(n, device), RealC(n, device), RealC(n, device)},
          m_geometry_normal{RealC(n, device), RealC(n, device), RealC(n, device)},
          m_shading_normal{RealC(n, device), RealC(n, device), RealC(n, device)},
          m_texcoord{RealC(n, device), RealC(n, device)}
    {
    }
    IntC m_tri_id;
    RealC m_t;
    RealC m_barycentric[2], m_position[3], m_geometry_normal[3], m_shading_normal[3], m_texcoord[2];
};
// src/main.cpp:255-284 as far as the path carries it: the hit's attributes and, instead of the Bsdf POINTER the
// reference gathers (main.cpp:325-326), the material INDEX (m_scatter) — the virtual-call dispatch is an albedo-table
// lookup on the device.  m_is_in_medium / m_has_light / m_light are constant false / null in the reference (:332-336).
struct Interaction {
    explicit Interaction(size_t n, int device = 0) : m_hit(n, device), m_scatter(n, device) {}
    TriangleHitInfoFullC m_hit;      // m_position, m_shading_normal, m_geometry_normal, m_texcoord (z = 0)
    IntC m_scatter;                  // material index per lane
};

// src/main.cpp:286-352
struct Scene {
    Scene() = default;
    Scene(const Scene &) = delete;
    Scene &operator=(const Scene &) = delete;
    ~Scene() { wpt_scene_destroy(handle); }

    // use_kd = false is the reference as shipped (every material white, src/main.cpp:244)
    void add_triangle_mesh(const std::string &obj_filepath, bool use_kd = false)
    {
        if (handle) throw std::runtime_error("wpt: Scene already holds a mesh");
        wpt_obj *o = nullptr;
        check(wpt_obj_load(obj_filepath.c_str(), use_kd ? 1 : 0, &o));
        wpt_mesh_desc d;
        int rc = wpt_obj_mesh(o, &d);
        if (rc == WPT_OK) rc = wpt_scene_create(&d, &handle);
        n_triangles = d.n_triangles;
        wpt_obj_free(o);
        check(rc);
    }
    void add_triangle_mesh(const wpt_mesh_desc &d)
    {
        if (handle) throw std::runtime_error("wpt: Scene already holds a mesh");
        check(wpt_scene_create(&d, &handle));
        n_triangles = d.n_triangles;
    }
    void commit(int device = 0) { check(wpt_scene_commit(handle, device)); }

    // closest hit on device streams; `active` (device bytes, size n or 1) may be null = all live.
    // tri_id = -1 on a miss (src/wavefront_isect.cu:70); masked-out lanes are left untouched.
    void intersect(const Ray3C &ray, const uint8_t *active, uint32_t active_size, const TriangleHitInfoC &hit,
                   void *stream = nullptr) const
    {
        wpt_ray_streams r{ray.ox, ray.oy, ray.oz, ray.dx, ray.dy, ray.dz, nullptr, nullptr, ray.tmin, ray.tmax};
        wpt_hit_streams h{hit.tri_id, hit.t, hit.u, hit.v};
        check(wpt_intersect(handle, &r, active, active_size, ray.size, &h, stream));
    }

    // Scene::intersect as the reference declares it (src/main.cpp:320-340): closest hit + derived attributes + the
    // material of every hit lane.  Lanes that miss keep tri_id = -1 and undefined attributes, as there.
    void intersect(const Ray3C &ray, const uint8_t *active, uint32_t active_size, Interaction &out, void *stream = nullptr) const
    {
        TriangleHitInfoFullC &h = out.m_hit;
        intersect(ray, active, active_size, TriangleHitInfoC{h.m_tri_id.data(), h.m_t.data(), h.m_barycentric[0].data(),
                                                             h.m_barycentric[1].data()}, stream);
        wpt_ray_streams r{ray.ox, ray.oy, ray.oz, ray.dx, ray.dy, ray.dz, nullptr, nullptr, ray.tmin, ray.tmax};
        wpt_hit_streams hs{h.m_tri_id.data(), h.m_t.data(), h.m_barycentric[0].data(), h.m_barycentric[1].data()};
        wpt_hit_attr_streams a{h.m_position[0].data(), h.m_position[1].data(), h.m_position[2].data(),
                               h.m_geometry_normal[0].data(), h.m_geometry_normal[1].data(), h.m_geometry_normal[2].data(),
                               h.m_shading_normal[0].data(), h.m_shading_normal[1].data(), h.m_shading_normal[2].data(),
                               h.m_texcoord[0].data(), h.m_texcoord[1].data(), out.m_scatter.data()};
        check(wpt_hit_attributes(handle, &r, &hs, ray.size, &a, stream));
    }

    wpt_scene *handle = nullptr;
    uint32_t n_triangles = 0;
};

// planar film on the host: the three channels of `SpectrumC film` (src/main.cpp:369) after readback
struct Film {
    Film(int w, int h) : width(w), height(h), r((size_t)w * h), g((size_t)w * h), b((size_t)w * h) {}
    int width, height;
    std::vector<float> r, g, b;
};

// the sample / bounce loop of main() (src/main.cpp:385-429) as one call
struct PathLoop {
    PathLoop(const Scene &scene, int width, int height, int spp_per_pass = 4)
    {
        check(wpt_renderer_create(scene.handle, width, height, spp_per_pass, &handle));
        w = width; h = height;
    }
    PathLoop(const PathLoop &) = delete;
    PathLoop &operator=(const PathLoop &) = delete;
    ~PathLoop() { wpt_renderer_destroy(handle); }

    // film = (1/num_samples) * sum over samples of the escaped throughput; returns the device-counted stats
    wpt_stats render(const ThinlensCamera &camera, int num_samples, int num_bounces, Film &film, uint64_t seed = 0,
                     int first_sample = 0)
    {
        wpt_render_config cfg{w, h, first_sample, num_samples, num_bounces, seed, 0.001f, 1e20f};
        wpt_stats st;
        check(wpt_render_host(handle, &camera.c, &cfg, film.r.data(), film.g.data(), film.b.data(), &st));
        return st;
    }
    wpt_renderer *handle = nullptr;
    int w = 0, h = 0;
};

// The same loop sharded over the GPUs of one box: GPU g renders the global sample indices
// [g*S/G, (g+1)*S/G) of all pixels from its own copy of the scene, the partial films are summed onto GPU 0 with ONE
// wpt_film_reduce (ncclReduce over NVLink, a communicator per GPU inside one process), scaled by 1/S and read back.
struct ShardedPathLoop {
    ShardedPathLoop(const wpt_mesh_desc &mesh, int n_gpus, int width, int height, int spp_per_pass = 4)
        : w(width), h(height), G(n_gpus), comms((size_t)n_gpus, nullptr)
    {
        std::vector<int> devs((size_t)G);
        for (int g = 0; g < G; g++) devs[(size_t)g] = g;
        for (int g = 0; g < G; g++) {
            scenes.emplace_back(new Scene());
            scenes.back()->add_triangle_mesh(mesh);
            scenes.back()->commit(g);
            loops.emplace_back(new PathLoop(*scenes.back(), width, height, spp_per_pass));
            films.emplace_back(3 * (size_t)w * h, g);
        }
        if (G > 1) check(wpt_nccl_comm_create_all(G, devs.data(), comms.data()));
    }
    ShardedPathLoop(const ShardedPathLoop &) = delete;
    ShardedPathLoop &operator=(const ShardedPathLoop &) = delete;
    ~ShardedPathLoop()
    {
        for (void *c : comms) wpt_nccl_comm_destroy(c);
    }
    // returns the total number of ray segments over all GPUs
    uint64_t render(const ThinlensCamera &camera, int num_samples, int num_bounces, Film &film, uint64_t seed = 0)
    {
        const uint32_t n = (uint32_t)w * (uint32_t)h;
        for (int g = 0; g < G; g++) {                       // every GPU starts its shard; nothing here waits
            const int begin = (int)((int64_t)num_samples * g / G), end = (int)((int64_t)num_samples * (g + 1) / G);
            float *f = films[(size_t)g].data();
            check(wpt_device_set(g));
            wpt_render_config cfg{w, h, begin, end - begin, num_bounces, seed, 0.001f, 1e20f};
            check(wpt_device_zero(f, 3 * (size_t)n * sizeof(float), nullptr));            // film = 0 (src/main.cpp:369)
            check(wpt_render(loops[(size_t)g]->handle, &camera.c, &cfg, f, f + n, f + 2 * (size_t)n, nullptr));
        }
        if (G > 1) {
            // one reduce per GPU, queued behind that GPU's frame on its stream
            check(wpt_nccl_group_start());
            for (int g = 0; g < G; g++) {
                check(wpt_device_set(g));
                float *f = films[(size_t)g].data();
                check(wpt_film_reduce(comms[(size_t)g], f, f + n, f + 2 * (size_t)n, n, 0, nullptr));
            }
            check(wpt_nccl_group_end());
        }
        uint64_t segments = 0;
        for (int g = 0; g < G; g++) {
            wpt_stats st;
            check(wpt_device_sync(g));
            check(wpt_render_finish(loops[(size_t)g]->handle, &st, nullptr));
            segments += st.segments;
        }
        check(wpt_device_sync(0));
        float *f0 = films[0].data();
        check(wpt_film_scale(f0, f0 + n, f0 + 2 * (size_t)n, n, 1.0f / (float)num_samples, nullptr));   // src/main.cpp:429
        check(wpt_device_download(film.r.data(), f0, n * sizeof(float), nullptr));
        check(wpt_device_download(film.g.data(), f0 + n, n * sizeof(float), nullptr));
        check(wpt_device_download(film.b.data(), f0 + 2 * (size_t)n, n * sizeof(float), nullptr));
        return segments;
    }
    int w, h, G;
    std::vector<std::unique_ptr<Scene>> scenes;
    std::vector<std::unique_ptr<PathLoop>> loops;
    std::vector<RealC> films;
    std::vector<void *> comms;
};

// src/fimage.h:5-59
struct Fimage {
    static void save_pfm(const float *r, const float *g, const float *b, size_t width, size_t height,
                         const std::string &filepath)
    {
        check(wpt_save_pfm(filepath.c_str(), r, g, b, (uint32_t)width, (uint32_t)height));
    }
    static void save_pfm_mono(const float *image, size_t width, size_t height, const std::string &filepath)
    {
        save_pfm(image, image, image, width, height, filepath);     // fimage.h:7-31 writes the value three times
    }
};

}  // namespace wpt_shim
