// wpt_shim.hpp — header-only C++ shim over the C ABI (wpt.h) that keeps the reference's type and
// method names, so that its main() (src/main.cpp:355-445) reads the same after the swap:
//
//   reference (Enoki + OptiX)                           here (libwpt.so)
//   Scene::add_triangle_mesh(path)   main.cpp:288-310   wpt_obj_load + wpt_scene_create
//   Scene::commit()                  main.cpp:312-318   wpt_scene_commit (SAH BVH build + upload)
//   Scene::intersect(ray, active)    main.cpp:320-340   wpt_intersect on the same SoA streams
//   ThinlensCamera(...)              pinhole.h:9-25     wpt_camera_init
//   the sample / bounce loop         main.cpp:385-429   PathLoop::render -> wpt_render_host
//   Fimage::save_pfm                 fimage.h:33-58     wpt_save_pfm
//   StopWatch                        stopwatch.h:6-27   same (std::chrono)
//
// Errors: the C ABI returns codes; like the reference's OPTIX_CHECK / CUDA_CHECK
// (src/accel/optix_backend.h:25-82) the shim throws std::runtime_error with the message.
#pragma once
#include <chrono>
#include <cstdint>
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
