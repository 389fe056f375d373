/*
 * wpt_oracle.h — CPU restatement of the reference's hot path.   TEST INFRASTRUCTURE ONLY.
 *
 * This is the parity oracle for the B200 path tracer.  It is NOT part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * compile, link, load or call anything in oracle/.  libwpt.so never does.
 *
 * What it restates (all paths relative to /root/reference):
 *   PCG32                      dependencies/enoki/include/enoki/random.h:62-77,116-118
 *   ThinlensCamera             src/pinhole.h:9-56
 *   warps                      src/mapping.h:5-27
 *   Frame3                     src/coordframe.h:17-48   (normal = local Y axis)
 *   bary interp / geo normal   src/add_math.h:3-16
 *   Ray defaults               src/ray.h:15-26          (tmin 1e-3, tmax 1e20)
 *   derived hit attributes     src/accel/optix_backend.h:463-486
 *   material fetch             src/main.cpp:325-326, 109-117 (Lambert, weight = albedo)
 *   sample/bounce loop + film  src/main.cpp:376-429
 *   closest-hit contract       src/wavefront_isect.cu:48-112 (id -1 on miss; t,u,v; u,v = weights
 *                              of vertex 1 and 2; masked-out lanes untouched)
 *
 * The intersector the reference calls is OptiX 7.0 (driver resident, not in the tree: the SDK
 * path is hard-coded at CMakeLists.txt:41; call sites src/accel/optix_backend.h:460,
 * src/wavefront_isect.cu:103-105).  Its algorithm is therefore restated from the published
 * contract (closest hit, tmin < t <= tmax, no culling) using the accept inequalities of the
 * reference's vendored CPU intersector Embree 3.6.1
 * (dependencies/embree/kernels/geometry/triangle_intersector_moeller.h:75-113).
 *
 * Pinning: the reference has no tests or golden vectors of its own for this path.
 * The oracle is pinned against (i) the real Embree 3.6.1 built from the reference tree
 * (oracle/_ref/libembree_ref.so, recipe oracle/ref_embree.mk) on shared ray buffers,
 * (ii) Embree's own known-answer tests (tutorials/verify/verify.cpp:2312-2399 TriangleHitTest,
 * tutorials/minimal/minimal.cpp:72-213), (iii) the canonical PCG32 and Philox4x32-10 known-answer
 * vectors.  See tests/test_oracle_*.py and tests/golden/.
 */
#ifndef WPT_ORACLE_H
#define WPT_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---------------- RNG ---------------- */
typedef struct { uint64_t state, inc; } orc_pcg32;
void     orc_pcg32_seed(orc_pcg32 *r, uint64_t initstate, uint64_t initseq);
uint32_t orc_pcg32_next_u32(orc_pcg32 *r);
float    orc_pcg32_next_f32(orc_pcg32 *r);
/* Philox4x32-10 (Salmon et al. 2011, Random123) — the product's counter-based generator. */
void     orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
float    orc_u32_to_f32(uint32_t u); /* (u>>9 | 0x3f800000) - 1, random.h:116-118 */

/* ---------------- camera / warps / frame ---------------- */
typedef struct {
    float origin[3], x[3], y[3], z[3];
    float film_size_y, focal_dist, fov_y, lens_radius;
} orc_camera;
void orc_camera_init(orc_camera *c, const float from[3], const float at[3], const float up[3],
                     float lens_radius, float focal_dist, float fov_y, float film_size_y);
void orc_camera_sample_pos(const orc_camera *c, const float s[2], float out[3]);
void orc_camera_sample_dir(const orc_camera *c, int px, int py, const float pos[3], int W, int H,
                           const float s[2], float out[3]);
void orc_disk_from_square(const float s[2], float out[2]);
void orc_cosine_hemisphere(const float s[2], float out[3]);
/* m = world_from_local, row-major 3x3; columns are basis_x, basis_y(=normal), basis_z */
void orc_frame_from_normal(const float n[3], float m[9]);
void orc_frame_to_world(const float m[9], const float v[3], float out[3]);
void orc_frame_to_local(const float m[9], const float v[3], float out[3]);
void orc_barycentric3(const float a[3], const float b[3], const float c[3], float u, float v, float out[3]);
void orc_geometry_normal(const float p1[3], const float p2[3], const float p3[3], float out[3]);

/* ---------------- mesh + intersectors ---------------- */
typedef struct {
    const float   *pos;      /* 3*n_pos, AoS-flattened like Scene::m_positions (main.cpp:297) */
    const int32_t *pos_idx;  /* 3*n_tris */
    const float   *nrm;      /* 3*n_nrm */
    const int32_t *nrm_idx;  /* 3*n_tris */
    const int32_t *mat_id;   /* n_tris */
    const float   *albedo;   /* 3*n_mat */
    uint32_t n_tris, n_pos, n_nrm, n_mat;
    /* optional (may all be NULL): texture coordinates (main.cpp:215-223) and one ImageTexture per material
     * (main.cpp:35-80); a material whose texture is NULL / 1x1 uses its constant albedo */
    const float   *uvs;      /* 2*n_uv */
    const int32_t *uv_idx;   /* 3*n_tris, -1 = none */
    const struct orc_texture *tex;   /* n_mat entries */
} orc_mesh;
typedef struct orc_texture { const float *rgb; int32_t w, h; } orc_texture;
/* ImageTexture::eval (src/main.cpp:62-76): bilinear, clamped; texel (x,y) at y*w + x (the reference's :52 uses
 * y*size.y + x, a bug for non-square images, fixed here and in the product) */
void orc_texture_eval(const orc_texture *t, float u, float v, float rgb[3]);

/* single ray / single triangle, Embree accept rules; on accept updates *tfar,*u,*v, returns 1 */
int orc_tri_intersect(const float o[3], const float d[3], float tnear, float *tfar,
                      const float v0[3], const float v1[3], const float v2[3], float *u, float *v);

typedef void (*orc_intersect1_fn)(void *ctx, const float o[3], const float d[3], float tmin, float tmax,
                                  int32_t *id, float *t, float *u, float *v);

void orc_intersect1_brute(void *mesh /* orc_mesh* */, const float o[3], const float d[3], float tmin,
                          float tmax, int32_t *id, float *t, float *u, float *v);

typedef struct orc_bvh orc_bvh;
orc_bvh *orc_bvh_build(const orc_mesh *m);
void     orc_bvh_free(orc_bvh *b);
void     orc_intersect1_bvh(void *bvh /* orc_bvh* */, const float o[3], const float d[3], float tmin,
                            float tmax, int32_t *id, float *t, float *u, float *v);

/* SoA stream form of OptixBackend::intersect's launch (optixdata.h:11-29): mask_size 1 = broadcast;
 * masked-out lanes are left untouched. threads<=0 -> all cores. */
void orc_intersect_streams(orc_intersect1_fn fn, void *ctx,
                           const float *ox, const float *oy, const float *oz,
                           const float *dx, const float *dy, const float *dz,
                           float tmin, float tmax, const uint8_t *mask, uint32_t mask_size, uint32_t n,
                           int32_t *id, float *t, float *u, float *v, int threads);

/* per-ray windows (src/optixdata.h:20-21, src/ray.h:20-26) */
void orc_intersect_streams_win(orc_intersect1_fn fn, void *ctx,
                               const float *ox, const float *oy, const float *oz,
                               const float *dx, const float *dy, const float *dz,
                               const float *tmin, const float *tmax, uint32_t n,
                               int32_t *id, float *t, float *u, float *v, int threads);

/* ---------------- the loop ---------------- */
enum { ORC_RNG_PCG32 = 0, ORC_RNG_PHILOX = 1 };
typedef struct {
    int32_t  width, height;
    int32_t  spp_begin, spp_count, depth;
    int32_t  rng_mode;     /* ORC_RNG_PCG32 = the reference's generator; ORC_RNG_PHILOX = the product's */
    uint64_t seed;         /* philox key (pcg32 ignores it: reference seeds with the default state) */
    float    tmin, tmax;
} orc_render_cfg;

/* Adds the UN-normalised radiance sum of samples [spp_begin, spp_begin+spp_count) into film_r/g/b
 * (planar, pixel_index = y*W + x, main.cpp:379-381).  Returns ray segments actually traced. */
uint64_t orc_render(const orc_mesh *m, orc_intersect1_fn fn, void *ctx, const orc_camera *cam,
                    const orc_render_cfg *cfg, float *film_r, float *film_g, float *film_b, int threads);

/* One path vertex step, exposed for per-function parity tests: given a hit (tri id,u,v,t) on ray (o,d)
 * and two uniforms, return next origin, next direction, and the bsdf weight. */
void orc_shade_step(const orc_mesh *m, const float o[3], const float d[3], int32_t tri, float t, float u,
                    float v, const float s[2], float next_o[3], float next_d[3], float weight[3]);

/* The derived hit attributes OptixBackend::intersect returns with a hit (src/accel/optix_backend.h:463-486):
 * position o + t d (:469), geometry normal (:472-476, add_math.h:9-16), barycentric-interpolated shading normal
 * (:483-484) and texcoord (:479-480; uvs f32[2*n], uv_idx i32[3*n_tris], either may be NULL -> (0,0)), and the
 * material index Scene::intersect gathers (src/main.cpp:325). */
void orc_hit_attributes(const orc_mesh *m, const float *uvs, const int32_t *uv_idx, const float o[3], const float d[3],
                        int32_t tri, float t, float u, float v, float p[3], float gn[3], float sn[3], float uv[2],
                        int32_t *mat);

int orc_num_threads(void);
#ifdef __cplusplus
}
#endif
#endif
