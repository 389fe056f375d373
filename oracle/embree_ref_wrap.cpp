// embree_ref_wrap.cpp — OUR thin C wrapper over the PUBLIC API of the reference's vendored
// Embree 3.6.1 (dependencies/embree/include/embree3/rtcore.h).  TEST INFRASTRUCTURE ONLY.
// It is compiled into oracle/_ref/libembree_ref.so together with Embree's own unmodified sources
// (recipe: oracle/ref_embree.mk) and is used (a) to pin oracle/wpt_oracle.c and (b) as the
// "reference" CPU baseline (bench.py --impl reference).  Scene setup follows SURVEY §8c:
// one RTC_GEOMETRY_TYPE_TRIANGLE, FLOAT3 vertices, UINT3 indices in face order so that
// primID == the reference's tri_id; device config "isa=avx2" as tutorials/verify/verify.cpp:2331.
#include <embree3/rtcore.h>
#include <embree3/rtcore_builder.h>
#include <vector>
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

// ---- SAH yardstick: the reference tree's own BVH builder, rtcBuildBVH
// (include/embree3/rtcore_builder.h:26-110), on the same primitives with the product's cost constants; the tree is
// only built to be measured: cost = sum over inner nodes A(node)/A(root) * c_trav + sum over leaves A(leaf)/A(root) *
// c_isect * prims — the formula the product reports as wpt_bvh_stats.sah_cost.
namespace {
struct YNode { float clo[8][3], chi[8][3]; unsigned n_children; YNode *child[8]; unsigned n_prims; };   // bounds of the CHILDREN, as Embree hands them out
struct YStats { double inner = 0, leaf = 0; size_t nodes = 0, leaves = 0, prims = 0; unsigned depth = 0; };
void *y_create_node(RTCThreadLocalAllocator a, unsigned n, void *)
{
    YNode *nd = (YNode *)rtcThreadLocalAlloc(a, sizeof(YNode), 16);
    nd->n_children = n; nd->n_prims = 0;
    for (unsigned i = 0; i < 8; i++) nd->child[i] = nullptr;
    return nd;
}
void y_set_children(void *p, void **c, unsigned n, void *) { for (unsigned i = 0; i < n; i++) ((YNode *)p)->child[i] = (YNode *)c[i]; }
void y_set_bounds(void *p, const RTCBounds **b, unsigned n, void *)
{
    YNode *nd = (YNode *)p;      // called BEFORE the children exist (kernels/common/rtcore_builder.cpp): keep them here
    for (unsigned i = 0; i < n && i < 8; i++) {
        nd->clo[i][0] = b[i]->lower_x; nd->clo[i][1] = b[i]->lower_y; nd->clo[i][2] = b[i]->lower_z;
        nd->chi[i][0] = b[i]->upper_x; nd->chi[i][1] = b[i]->upper_y; nd->chi[i][2] = b[i]->upper_z;
    }
}
void *y_create_leaf(RTCThreadLocalAllocator a, const RTCBuildPrimitive *, size_t n, void *)
{
    YNode *nd = (YNode *)rtcThreadLocalAlloc(a, sizeof(YNode), 16);
    nd->n_children = 0; nd->n_prims = (unsigned)n;
    for (unsigned i = 0; i < 8; i++) nd->child[i] = nullptr;
    return nd;
}
double y_half_area(const float lo[3], const float hi[3])
{
    const double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}
void y_walk(const YNode *nd, const float lo[3], const float hi[3], double a_root, double c_trav, double c_isect,
            unsigned depth, YStats &st)
{
    const double a = y_half_area(lo, hi) / a_root;
    if (depth > st.depth) st.depth = depth;
    if (nd->n_children == 0) { st.leaf += a * c_isect * nd->n_prims; st.leaves++; st.prims += nd->n_prims; return; }
    st.inner += a * c_trav; st.nodes++;
    for (unsigned i = 0; i < nd->n_children; i++)
        if (nd->child[i]) y_walk(nd->child[i], nd->clo[i], nd->chi[i], a_root, c_trav, c_isect, depth + 1, st);
}
}  // namespace

// out[0] = SAH cost, out[1] = inner nodes, out[2] = leaves, out[3] = depth, out[4] = primitives found in leaves
// quality: 0 = RTC_BUILD_QUALITY_LOW (Morton), 1 = MEDIUM (binned SAH), 2 = HIGH (SAH + spatial splits are NOT enabled
// without a splitPrimitive callback: HIGH then equals MEDIUM with more bins)
__attribute__((visibility("default")))
int eref_builder_sah(const float *pos, uint32_t n_pos, const int32_t *idx, uint32_t n_tris, unsigned branching,
                     unsigned max_leaf, float c_trav, float c_isect, int quality, double *out)
{
    (void)n_pos;
    RTCDevice dev = rtcNewDevice("isa=avx2");
    if (!dev) return 1;
    RTCBVH bvh = rtcNewBVH(dev);
    std::vector<RTCBuildPrimitive> prims(n_tris);
    float rlo[3] = {1e30f, 1e30f, 1e30f}, rhi[3] = {-1e30f, -1e30f, -1e30f};
    for (uint32_t t = 0; t < n_tris; t++) {
        float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
        for (int v = 0; v < 3; v++) {
            const float *p = pos + 3 * (size_t)idx[3 * t + v];
            for (int k = 0; k < 3; k++) { lo[k] = p[k] < lo[k] ? p[k] : lo[k]; hi[k] = p[k] > hi[k] ? p[k] : hi[k]; }
        }
        RTCBuildPrimitive &bp = prims[t];
        bp.lower_x = lo[0]; bp.lower_y = lo[1]; bp.lower_z = lo[2]; bp.geomID = 0;
        bp.upper_x = hi[0]; bp.upper_y = hi[1]; bp.upper_z = hi[2]; bp.primID = t;
        for (int k = 0; k < 3; k++) { rlo[k] = lo[k] < rlo[k] ? lo[k] : rlo[k]; rhi[k] = hi[k] > rhi[k] ? hi[k] : rhi[k]; }
    }
    RTCBuildArguments a = rtcDefaultBuildArguments();
    a.buildQuality = quality == 0 ? RTC_BUILD_QUALITY_LOW : quality == 2 ? RTC_BUILD_QUALITY_HIGH : RTC_BUILD_QUALITY_MEDIUM;
    a.maxBranchingFactor = branching;
    a.maxDepth = 64;
    a.sahBlockSize = 1;
    a.minLeafSize = 1;
    a.maxLeafSize = max_leaf;
    a.traversalCost = c_trav;
    a.intersectionCost = c_isect;
    a.bvh = bvh;
    a.primitives = prims.data();
    a.primitiveCount = prims.size();
    a.primitiveArrayCapacity = prims.size();
    a.createNode = y_create_node;
    a.setNodeChildren = y_set_children;
    a.setNodeBounds = y_set_bounds;
    a.createLeaf = y_create_leaf;
    a.userPtr = nullptr;
    YNode *root = (YNode *)rtcBuildBVH(&a);
    int rc = 0;
    if (!root) rc = 2;
    else {
        YStats st;
        y_walk(root, rlo, rhi, y_half_area(rlo, rhi), c_trav, c_isect, 1, st);
        out[0] = st.inner + st.leaf; out[1] = (double)st.nodes; out[2] = (double)st.leaves; out[3] = st.depth; out[4] = (double)st.prims;
    }
    rtcReleaseBVH(bvh);
    rtcReleaseDevice(dev);
    return rc;
}

__attribute__((visibility("default")))
const char *eref_version(void) { return RTC_VERSION_STRING; }
}
