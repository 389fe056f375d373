// embree_ref_wrap.cpp — OUR thin C wrapper over the PUBLIC API of the reference's vendored
// Embree 3.6.1 (dependencies/embree/include/embree3/rtcore.h).  TEST INFRASTRUCTURE ONLY.
// It is compiled into oracle/_ref/libembree_ref.so together with Embree's own unmodified sources
// (recipe: oracle/ref_embree.mk) and is used (a) to pin oracle/wpt_oracle.c and (b) as the
// "reference" CPU baseline (bench.py --impl reference).  Scene setup follows SURVEY §8c:
// one RTC_GEOMETRY_TYPE_TRIANGLE, FLOAT3 vertices, UINT3 indices in face order so that
// primID == the reference's tri_id; device config "isa=avx2" as tutorials/verify/verify.cpp:2331.
#include <embree3/rtcore.h>
#include <cstdint>
#include <cstring>
#include <cstdio>
#include <limits>
#ifdef _OPENMP
#include <omp.h>
#endif

struct ErefScene { RTCDevice dev; RTCScene scene; };

extern "C" {

__attribute__((visibility("default")))
void *eref_scene_create(const float *pos, uint32_t n_pos, const int32_t *idx, uint32_t n_tris, int robust,
                        int build_threads)
{
    char cfg[128];
    if (build_threads > 0) snprintf(cfg, sizeof cfg, "isa=avx2,threads=%d", build_threads);
    else snprintf(cfg, sizeof cfg, "isa=avx2");
    RTCDevice dev = rtcNewDevice(cfg);
    if (!dev) return nullptr;
    RTCScene scene = rtcNewScene(dev);
    if (robust) rtcSetSceneFlags(scene, RTC_SCENE_FLAG_ROBUST);
    RTCGeometry geom = rtcNewGeometry(dev, RTC_GEOMETRY_TYPE_TRIANGLE);
    float *vb = (float *)rtcSetNewGeometryBuffer(geom, RTC_BUFFER_TYPE_VERTEX, 0, RTC_FORMAT_FLOAT3,
                                                 3 * sizeof(float), n_pos);
    unsigned *ib = (unsigned *)rtcSetNewGeometryBuffer(geom, RTC_BUFFER_TYPE_INDEX, 0, RTC_FORMAT_UINT3,
                                                       3 * sizeof(unsigned), n_tris);
    memcpy(vb, pos, sizeof(float) * 3 * (size_t)n_pos);
    memcpy(ib, idx, sizeof(unsigned) * 3 * (size_t)n_tris);
    rtcCommitGeometry(geom);
    rtcAttachGeometry(scene, geom);
    rtcReleaseGeometry(geom);
    rtcCommitScene(scene);
    ErefScene *s = new ErefScene{dev, scene};
    return s;
}

__attribute__((visibility("default")))
void eref_scene_destroy(void *p)
{
    ErefScene *s = (ErefScene *)p;
    if (!s) return;
    rtcReleaseScene(s->scene);
    rtcReleaseDevice(s->dev);
    delete s;
}

// Same signature as orc_intersect1_fn (oracle/wpt_oracle.h) so the oracle's loop can run on Embree.
// Ray setup per doc/src/api/rtcIntersect1.md:24-36.
__attribute__((visibility("default")))
void eref_intersect1(void *p, const float o[3], const float d[3], float tmin, float tmax,
                     int32_t *id, float *t, float *u, float *v)
{
    ErefScene *s = (ErefScene *)p;
    RTCIntersectContext ctx;
    rtcInitIntersectContext(&ctx);
    RTCRayHit rh;
    rh.ray.org_x = o[0]; rh.ray.org_y = o[1]; rh.ray.org_z = o[2];
    rh.ray.dir_x = d[0]; rh.ray.dir_y = d[1]; rh.ray.dir_z = d[2];
    rh.ray.tnear = tmin; rh.ray.tfar = tmax; rh.ray.time = 0.0f;
    rh.ray.mask = 0xFFFFFFFFu; rh.ray.id = 0; rh.ray.flags = 0;
    rh.hit.geomID = RTC_INVALID_GEOMETRY_ID;
    rh.hit.primID = RTC_INVALID_GEOMETRY_ID;
    rh.hit.instID[0] = RTC_INVALID_GEOMETRY_ID;
    rtcIntersect1(s->scene, &ctx, &rh);
    if (rh.hit.geomID == RTC_INVALID_GEOMETRY_ID) { *id = -1; return; }
    *id = (int32_t)rh.hit.primID; *t = rh.ray.tfar; *u = rh.hit.u; *v = rh.hit.v;
}

__attribute__((visibility("default")))
void *eref_intersect1_fn(void) { return (void *)&eref_intersect1; }

__attribute__((visibility("default")))
void eref_intersect_streams(void *p, const float *ox, const float *oy, const float *oz,
                            const float *dx, const float *dy, const float *dz, float tmin, float tmax,
                            const uint8_t *mask, uint32_t mask_size, uint32_t n,
                            int32_t *id, float *t, float *u, float *v, int threads)
{
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 1024) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        const int live = mask ? (mask_size == 1 ? mask[0] : mask[i]) : 1;
        if (!live) continue;
        const float o[3] = {ox[i], oy[i], oz[i]}, d[3] = {dx[i], dy[i], dz[i]};
        int32_t hid = -1; float ht = 0, hu = 0, hv = 0;
        eref_intersect1(p, o, d, tmin, tmax, &hid, &ht, &hu, &hv);
        id[i] = hid;
        if (hid >= 0) { t[i] = ht; u[i] = hu; v[i] = hv; }
    }
}

__attribute__((visibility("default")))
const char *eref_version(void) { return RTC_VERSION_STRING; }
}
