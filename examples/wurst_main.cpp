// wurst_main.cpp — the reference's main() (src/main.cpp:355-445) on libwpt.so through include/wpt_shim.hpp:
// load an OBJ, commit, render num_samples x num_bounces with the reference's default thin-lens camera,
// normalise, write wurst.pfm.  Same statement order; the per-sample / per-bounce loop body
// (main.cpp:385-426) is the one call PathLoop::render.
//
//   wurst [scene.obj [out.pfm [width [height [num_samples [num_bounces]]]]]] [--kd] [--gpus N]     (flags anywhere)
// --gpus N shards the samples over N GPUs of the box (ShardedPathLoop: one wpt_film_reduce per frame), no Python.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <vector>

#include "wpt_shim.hpp"

using namespace wpt_shim;

int main(int argc, char **argv)
{
    // flags may stand anywhere; what is left are the reference's positional arguments
    bool use_kd = false;
    int n_gpus = 1;
    std::vector<const char *> pos;
    for (int i = 1; i < argc; i++) {
        if (std::strcmp(argv[i], "--kd") == 0) use_kd = true;
        else if (std::strcmp(argv[i], "--gpus") == 0) {
            if (i + 1 >= argc) { std::cerr << "wurst: --gpus needs a count" << std::endl; return 2; }
            n_gpus = std::atoi(argv[++i]);
        } else if (std::strncmp(argv[i], "--", 2) == 0) {
            std::cerr << "usage: wurst [scene.obj [out.pfm [width [height [samples [bounces]]]]]] [--kd] [--gpus N]" << std::endl;
            return 2;
        } else pos.push_back(argv[i]);
    }
    const size_t np = pos.size();
    // all const (src/main.cpp:357-362)
    const char *obj = np > 0 ? pos[0] : "mitsuba.obj";
    const char *out = np > 1 ? pos[1] : "wurst.pfm";
    const int width = np > 2 ? std::atoi(pos[2]) : 512;
    const int height = np > 3 ? std::atoi(pos[3]) : 512;
    const int num_samples = np > 4 ? std::atoi(pos[4]) : 100;
    const int num_bounces = np > 5 ? std::atoi(pos[5]) : 2;
    if (np > 6 || width <= 0 || height <= 0 || num_samples <= 0 || num_bounces <= 0) {
        std::cerr << "wurst: width, height, samples and bounces must be positive integers (at most six positional arguments)" << std::endl;
        return 2;
    }
    if (n_gpus < 1 || (n_gpus > 1 && n_gpus > wpt_device_count())) {
        std::cerr << "wurst: --gpus " << n_gpus << ": this box has " << wpt_device_count() << " CUDA device(s)" << std::endl;
        return 2;
    }
    if (n_gpus > 1) {
        try {
            // the same main(), samples sharded over n_gpus GPUs: every GPU commits its own copy of the scene
            wpt_obj *o = nullptr;
            check(wpt_obj_load(obj, use_kd ? 1 : 0, &o));
            wpt_mesh_desc d;
            check(wpt_obj_mesh(o, &d));
            ShardedPathLoop loop(d, n_gpus, width, height);
            wpt_obj_free(o);
            Film film(width, height);
            StopWatch sw;
            sw.reset();
            const float pi = 3.14159265358979323846f;
            const ThinlensCamera thinlens(Real3(0.0f, 3.03f, 5.0f), Real3(0.0f, 0.03f, 0.0f), Real3(0.0f, 1.0f, 0.0f), 0.00f,
                                          1.0f, 40.0f / 180.0f * pi);
            const uint64_t segments = loop.render(thinlens, num_samples, num_bounces, film);
            std::cout << sw.timeMilliSec() << std::endl;
            std::cerr << segments << " ray segments on " << n_gpus << " GPUs" << std::endl;
            std::cout << "writing image" << std::endl;
            Fimage::save_pfm(film.r.data(), film.g.data(), film.b.data(), width, height, out);
            std::cout << "done" << std::endl;
        } catch (const std::exception &e) {
            std::cerr << e.what() << std::endl;
            return 1;
        }
        return 0;
    }
    try {
        // load meshes and materials (src/main.cpp:364-366)
        Scene scene;
        scene.add_triangle_mesh(obj, use_kd);
        scene.commit();

        // init film (src/main.cpp:369)
        Film film(width, height);

        // init timer (src/main.cpp:372-373)
        StopWatch sw;
        sw.reset();

        // camera (src/main.cpp:383)
        const float pi = 3.14159265358979323846f;
        const ThinlensCamera thinlens(Real3(0.0f, 3.03f, 5.0f), Real3(0.0f, 0.03f, 0.0f), Real3(0.0f, 1.0f, 0.0f), 0.00f,
                                      1.0f, 40.0f / 180.0f * pi);

        // samples x bounces, film /= num_samples (src/main.cpp:385-429)
        PathLoop loop(scene, width, height);
        const wpt_stats st = loop.render(thinlens, num_samples, num_bounces, film);

        std::cout << sw.timeMilliSec() << std::endl;    // src/main.cpp:431
        std::cerr << st.segments << " ray segments, " << st.ms_total << " ms on the device, " << st.kernel_launches
                  << " kernel launches" << std::endl;

        // image write (src/main.cpp:433-442)
        std::cout << "writing image" << std::endl;
        Fimage::save_pfm(film.r.data(), film.g.data(), film.b.data(), width, height, out);
        std::cout << "done" << std::endl;
    } catch (const std::exception &e) {
        std::cerr << e.what() << std::endl;
        return 1;
    }
    return 0;
}
