/*
 * wpt_oracle.c — CPU restatement of the reference hot path (see wpt_oracle.h).
 * TEST INFRASTRUCTURE ONLY — never linked into or called from the product.
 * Build: gcc -O2 -ffp-contract=off -fopenmp -fPIC -shared (oracle/Makefile).
 * -ffp-contract=off because the reference only fuses where Enoki calls fmadd explicitly
 * (dependencies/enoki/include/enoki/cuda.h:323-376: every other op is a separate .rn PTX instruction).
 */
#include "wpt_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_PI 3.14159265358979323846f

/* ------------------------------------------------------------------ RNG */
/* enoki/random.h:36-38 */
#define PCG32_MULT 0x5851f42d4c957f2dULL

/* enoki/random.h:71-77 (XSH-RR output, LCG step) */
uint32_t orc_pcg32_next_u32(orc_pcg32 *r)
{
    uint64_t old = r->state;
    r->state = old * PCG32_MULT + r->inc;
    uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xorshifted >> rot) | (xorshifted << ((32u - rot) & 31u));
}

/* enoki/random.h:62-68 */
void orc_pcg32_seed(orc_pcg32 *r, uint64_t initstate, uint64_t initseq)
{
    r->state = 0u;
    r->inc = (initseq << 1u) | 1u;
    orc_pcg32_next_u32(r);
    r->state += initstate;
    orc_pcg32_next_u32(r);
}

/* enoki/random.h:116-118 */
float orc_u32_to_f32(uint32_t u)
{
    uint32_t bits = (u >> 9) | 0x3f800000u;
    float f;
    memcpy(&f, &bits, 4);
    return f - 1.0f;
}

float orc_pcg32_next_f32(orc_pcg32 *r) { return orc_u32_to_f32(orc_pcg32_next_u32(r)); }

/* Philox4x32-10, Salmon/Moraes/Dror/Shaw SC'11; constants from Random123 philox.h */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* ------------------------------------------------------------------ small vector helpers */
/* enoki/array_static.h:924-935: dot = x*x', then fmadd chain */
static inline float dot3(const float a[3], const float b[3])
{
    return fmaf(a[2], b[2], fmaf(a[1], b[1], a[0] * b[0]));
}
/* enoki/array_router.h:674-684: cross = fmsub(a.yzx, b.zxy, a.zxy * b.yzx) */
static inline void cross3(const float a[3], const float b[3], float o[3])
{
    float x = fmaf(a[1], b[2], -(a[2] * b[1]));
    float y = fmaf(a[2], b[0], -(a[0] * b[2]));
    float z = fmaf(a[0], b[1], -(a[1] * b[0]));
    o[0] = x; o[1] = y; o[2] = z;
}
/* enoki/array_router.h:663-665: v * rsqrt(dot(v,v)) (rsqrt.approx on device; exact here) */
static inline void normalize3(const float a[3], float o[3])
{
    float r = 1.0f / sqrtf(dot3(a, a));
    o[0] = a[0] * r; o[1] = a[1] * r; o[2] = a[2] * r;
}

/* ------------------------------------------------------------------ frame (src/coordframe.h) */
/* coordframe.h:17-30 */
void orc_frame_from_normal(const float n[3], float m[9])
{
    const float sign = copysignf(1.0f, n[1]);
    const float a = -1.0f / (sign + n[1]);
    const float b = n[2] * n[0] * a;
    const float bx[3] = { sign + n[0] * n[0] * a, -n[0], b };
    const float by[3] = { n[0], n[1], n[2] };
    const float bz[3] = { sign * b, -sign * n[2], 1.0f + sign * n[2] * n[2] * a };
    m[0] = bx[0]; m[1] = by[0]; m[2] = bz[0];
    m[3] = bx[1]; m[4] = by[1]; m[5] = bz[1];
    m[6] = bx[2]; m[7] = by[2]; m[8] = bz[2];
}
/* coordframe.h:45-48 with enoki/matrix.h:177-184: col0*v0, then fmadd(col_i, v_i, sum) */
void orc_frame_to_world(const float m[9], const float v[3], float out[3])
{
    float r[3];
    for (int i = 0; i < 3; i++)
        r[i] = fmaf(m[3 * i + 2], v[2], fmaf(m[3 * i + 1], v[1], m[3 * i + 0] * v[0]));
    out[0] = r[0]; out[1] = r[1]; out[2] = r[2];
}
/* coordframe.h:40-43: transpose(M) * v */
void orc_frame_to_local(const float m[9], const float v[3], float out[3])
{
    float r[3];
    for (int i = 0; i < 3; i++)
        r[i] = fmaf(m[6 + i], v[2], fmaf(m[3 + i], v[1], m[i] * v[0]));
    out[0] = r[0]; out[1] = r[1]; out[2] = r[2];
}

/* ------------------------------------------------------------------ warps (src/mapping.h) */
/* mapping.h:15-27 (Shirley concentric map) */
void orc_disk_from_square(const float s[2], float out[2])
{
    const float a = s[0] * 2.0f - 1.0f, b = s[1] * 2.0f - 1.0f;
    const float a2 = a * a, b2 = b * b;
    const int sel = a2 > b2;
    const float r = sel ? a : b;
    const float phi = sel ? (ORC_PI / 4.0f) * (b / a) : (ORC_PI / 2.0f) - (ORC_PI / 4.0f) * (a / b);
    out[0] = r * cosf(phi);
    out[1] = r * sinf(phi);
}
/* mapping.h:5-11 (Y-up cosine-weighted hemisphere) */
void orc_cosine_hemisphere(const float s[2], float out[3])
{
    const float sin_phi = sqrtf(1.0f - s[0]);
    const float theta = ORC_PI * 2.0f * s[1];
    out[0] = cosf(theta) * sin_phi;
    out[1] = sqrtf(s[0]);
    out[2] = sinf(theta) * sin_phi;
}

/* ------------------------------------------------------------------ camera (src/pinhole.h) */
/* pinhole.h:9-25 */
void orc_camera_init(orc_camera *c, const float from[3], const float at[3], const float up[3],
                     float lens_radius, float focal_dist, float fov_y, float film_size_y)
{
    float d[3] = { at[0] - from[0], at[1] - from[1], at[2] - from[2] };
    memcpy(c->origin, from, sizeof(float) * 3);
    normalize3(d, c->z);
    float t[3];
    cross3(up, c->z, t);
    normalize3(t, c->x);
    cross3(c->z, c->x, t);
    normalize3(t, c->y);
    c->film_size_y = film_size_y;
    c->focal_dist = focal_dist;
    c->fov_y = fov_y;
    c->lens_radius = lens_radius;
}
static void cam_matrix(const orc_camera *c, float m[9])
{
    /* Frame3C(m_x, m_y, m_z): columns x,y,z (coordframe.h:32-38) */
    for (int i = 0; i < 3; i++) { m[3 * i] = c->x[i]; m[3 * i + 1] = c->y[i]; m[3 * i + 2] = c->z[i]; }
}
/* pinhole.h:27-32 — lens lies in the local x–z plane (sic) */
void orc_camera_sample_pos(const orc_camera *c, const float s[2], float out[3])
{
    float m[9], disk[2], w[3];
    cam_matrix(c, m);
    if (c->lens_radius == 0.0f) {
        /* disk*0: identical to the reference except for its NaN at s==(0.5,0.5) (0/0 in mapping.h:22) */
        disk[0] = disk[1] = 0.0f;
    } else {
        orc_disk_from_square(s, disk);
        disk[0] *= c->lens_radius; disk[1] *= c->lens_radius;
    }
    const float lens[3] = { disk[0], 0.0f, disk[1] };
    orc_frame_to_world(m, lens, w);
    out[0] = w[0] + c->origin[0]; out[1] = w[1] + c->origin[1]; out[2] = w[2] + c->origin[2];
}
/* pinhole.h:34-56 */
void orc_camera_sample_dir(const orc_camera *c, int px, int py, const float pos[3], int W, int H,
                           const float s[2], float out[3])
{
    float m[9], lens_local[3];
    cam_matrix(c, m);
    const float rel[3] = { pos[0] - c->origin[0], pos[1] - c->origin[1], pos[2] - c->origin[2] };
    orc_frame_to_local(m, rel, lens_local);
    const float dist = c->film_size_y * 0.5f / tanf(c->fov_y * 0.5f);
    const float ratio = (float)W / (float)H;
    const float ndc_x = ((float)px + s[0]) / (float)W;
    const float ndc_y = ((float)py + s[1]) / (float)H;
    const float fsx = ratio * c->film_size_y, fsy = c->film_size_y;
    const float film[3] = { (0.5f - ndc_x) * fsx, (0.5f - ndc_y) * fsy, dist };
    const float focal[3] = { c->focal_dist * film[0] / film[2], c->focal_dist * film[1] / film[2],
                             c->focal_dist * film[2] / film[2] };
    const float dl[3] = { focal[0] - lens_local[0], focal[1] - lens_local[1], focal[2] - lens_local[2] };
    float dn[3];
    normalize3(dl, dn);
    orc_frame_to_world(m, dn, out);
}

/* ------------------------------------------------------------------ add_math.h */
/* add_math.h:3-7 */
void orc_barycentric3(const float a[3], const float b[3], const float c[3], float u, float v, float out[3])
{
    const float w = 1.0f - u - v;
    for (int i = 0; i < 3; i++) out[i] = w * a[i] + u * b[i] + v * c[i];
}
/* add_math.h:9-16 */
void orc_geometry_normal(const float p1[3], const float p2[3], const float p3[3], float out[3])
{
    const float e1[3] = { p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2] };
    const float e2[3] = { p3[0] - p1[0], p3[1] - p1[1], p3[2] - p1[2] };
    float un[3];
    cross3(e1, e2, un);
    normalize3(un, out);
}

/* ------------------------------------------------------------------ ray / triangle */
/* Accept rules and op order of embree/kernels/geometry/triangle_intersector_moeller.h:75-113
 * (e1 = v0-v1, e2 = v2-v0, Ng = cross(e2,e1): kernels/geometry/triangle.h:54, moeller.h:122;
 *  cross/dot as msub/madd chains: common/math/vec3.h:221,215).  t = T/|den| (Embree: T*rcp(|den|)). */
static inline void e_cross(const float a[3], const float b[3], float o[3])
{
    o[0] = fmaf(a[1], b[2], -(a[2] * b[1]));
    o[1] = fmaf(a[2], b[0], -(a[0] * b[2]));
    o[2] = fmaf(a[0], b[1], -(a[1] * b[0]));
}
static inline float e_dot(const float a[3], const float b[3])
{
    return fmaf(a[0], b[0], fmaf(a[1], b[1], a[2] * b[2]));
}
int orc_tri_intersect(const float o[3], const float d[3], float tnear, float *tfar,
                      const float v0[3], const float v1[3], const float v2[3], float *u, float *v)
{
    const float e1[3] = { v0[0] - v1[0], v0[1] - v1[1], v0[2] - v1[2] };
    const float e2[3] = { v2[0] - v0[0], v2[1] - v0[1], v2[2] - v0[2] };
    float Ng[3], R[3];
    e_cross(e2, e1, Ng);
    const float C[3] = { v0[0] - o[0], v0[1] - o[1], v0[2] - o[2] };
    e_cross(C, d, R);
    const float den = e_dot(Ng, d);
    const float absDen = fabsf(den);
    const float sgn = den < 0.0f || (den == 0.0f && signbit(den)) ? -1.0f : 1.0f;
    const float U = e_dot(R, e2) * sgn;
    const float V = e_dot(R, e1) * sgn;
    if (!(den != 0.0f && U >= 0.0f && V >= 0.0f && U + V <= absDen)) return 0;
    const float T = e_dot(Ng, C) * sgn;
    if (!(absDen * tnear < T && T <= absDen * (*tfar))) return 0;
    *tfar = T / absDen;
    *u = U / absDen;
    *v = V / absDen;
    return 1;
}

static inline void tri_verts(const orc_mesh *m, uint32_t i, const float **a, const float **b, const float **c)
{
    *a = m->pos + 3 * (size_t)m->pos_idx[3 * i + 0];
    *b = m->pos + 3 * (size_t)m->pos_idx[3 * i + 1];
    *c = m->pos + 3 * (size_t)m->pos_idx[3 * i + 2];
}

/* closest hit; tie policy: smaller t wins, equal t -> lower primitive id (order independent) */
static inline void consider(const orc_mesh *m, uint32_t i, const float o[3], const float d[3], float tmin,
                            int32_t *id, float *t, float *u, float *v)
{
    const float *a, *b, *c;
    tri_verts(m, i, &a, &b, &c);
    float tt = *t, uu, vv;
    if (orc_tri_intersect(o, d, tmin, &tt, a, b, c, &uu, &vv)) {
        if (*id < 0 || tt < *t || (tt == *t && (int32_t)i < *id)) { *t = tt; *u = uu; *v = vv; *id = (int32_t)i; }
    }
}

void orc_intersect1_brute(void *mesh, const float o[3], const float d[3], float tmin, float tmax,
                          int32_t *id, float *t, float *u, float *v)
{
    const orc_mesh *m = (const orc_mesh *)mesh;
    int32_t bid = -1; float bt = tmax, bu = 0, bv = 0;
    for (uint32_t i = 0; i < m->n_tris; i++) consider(m, i, o, d, tmin, &bid, &bt, &bu, &bv);
    *id = bid;
    if (bid >= 0) { *t = bt; *u = bu; *v = bv; }
}

/* ------------------------------------------------------------------ a plain median-split BVH2
 * Only there so the oracle finishes in seconds at test sizes; deliberately shares nothing with
 * the product's SAH/wide-node builder. */
typedef struct { float lo[3], hi[3]; uint32_t left, count; /* count>0: leaf, left=first */ } orc_node;
struct orc_bvh { const orc_mesh *m; orc_node *nodes; uint32_t *prims; uint32_t n_nodes; };

typedef struct { float c[3]; float lo[3], hi[3]; } orc_pref;

static void bounds_of(const orc_pref *pr, const uint32_t *prims, uint32_t first, uint32_t count,
                      float lo[3], float hi[3], float clo[3], float chi[3])
{
    for (int k = 0; k < 3; k++) { lo[k] = clo[k] = INFINITY; hi[k] = chi[k] = -INFINITY; }
    for (uint32_t i = first; i < first + count; i++) {
        const orc_pref *p = pr + prims[i];
        for (int k = 0; k < 3; k++) {
            lo[k] = fminf(lo[k], p->lo[k]); hi[k] = fmaxf(hi[k], p->hi[k]);
            clo[k] = fminf(clo[k], p->c[k]); chi[k] = fmaxf(chi[k], p->c[k]);
        }
    }
}

static uint32_t build_rec(orc_bvh *b, const orc_pref *pr, uint32_t first, uint32_t count)
{
    uint32_t me = b->n_nodes++;
    float clo[3], chi[3];
    orc_node nd;
    bounds_of(pr, b->prims, first, count, nd.lo, nd.hi, clo, chi);
    int axis = 0;
    float ext = chi[0] - clo[0];
    for (int k = 1; k < 3; k++) if (chi[k] - clo[k] > ext) { ext = chi[k] - clo[k]; axis = k; }
    if (count <= 4 || !(ext > 0.0f)) {
        nd.left = first; nd.count = count;
        b->nodes[me] = nd;
        return me;
    }
    const float mid = 0.5f * (clo[axis] + chi[axis]);
    uint32_t i = first, j = first + count;
    while (i < j) {
        if (pr[b->prims[i]].c[axis] < mid) i++;
        else { uint32_t t = b->prims[i]; b->prims[i] = b->prims[--j]; b->prims[j] = t; }
    }
    if (i == first || i == first + count) i = first + count / 2;
    nd.count = 0;
    b->nodes[me] = nd;
    uint32_t l = build_rec(b, pr, first, i - first);
    (void)l; /* left child is always me+1 */
    uint32_t r = build_rec(b, pr, i, first + count - i);
    b->nodes[me].left = r; /* store right child; left is me+1 */
    return me;
}

orc_bvh *orc_bvh_build(const orc_mesh *m)
{
    orc_bvh *b = (orc_bvh *)calloc(1, sizeof(orc_bvh));
    b->m = m;
    b->nodes = (orc_node *)malloc(sizeof(orc_node) * (2 * (size_t)m->n_tris + 1));
    b->prims = (uint32_t *)malloc(sizeof(uint32_t) * ((size_t)m->n_tris + 1));
    orc_pref *pr = (orc_pref *)malloc(sizeof(orc_pref) * ((size_t)m->n_tris + 1));
    for (uint32_t i = 0; i < m->n_tris; i++) {
        const float *a, *bb, *c;
        tri_verts(m, i, &a, &bb, &c);
        for (int k = 0; k < 3; k++) {
            pr[i].lo[k] = fminf(a[k], fminf(bb[k], c[k]));
            pr[i].hi[k] = fmaxf(a[k], fmaxf(bb[k], c[k]));
            pr[i].c[k] = 0.5f * (pr[i].lo[k] + pr[i].hi[k]);
        }
        b->prims[i] = i;
    }
    if (m->n_tris) build_rec(b, pr, 0, m->n_tris);
    free(pr);
    return b;
}
void orc_bvh_free(orc_bvh *b)
{
    if (!b) return;
    free(b->nodes); free(b->prims); free(b);
}

static inline int slab(const orc_node *n, const float o[3], const float inv[3], float tmin, float tmax)
{
    /* conservative: widened by 2 ulp-ish so grazing rays that Moeller-Trumbore accepts are never culled */
    float t0 = tmin, t1 = tmax;
    for (int k = 0; k < 3; k++) {
        float a = (n->lo[k] - o[k]) * inv[k], b = (n->hi[k] - o[k]) * inv[k];
        float lo = fminf(a, b), hi = fmaxf(a, b);
        /* NaN (0*inf) -> fminf/fmaxf drop it and keep the other operand: treated as no constraint */
        if (lo == lo) t0 = fmaxf(t0, lo - fabsf(lo) * 4e-7f - 1e-30f);
        if (hi == hi) t1 = fminf(t1, hi + fabsf(hi) * 4e-7f + 1e-30f);
    }
    return t0 <= t1;
}

void orc_intersect1_bvh(void *bvh, const float o[3], const float d[3], float tmin, float tmax,
                        int32_t *id, float *t, float *u, float *v)
{
    const orc_bvh *b = (const orc_bvh *)bvh;
    int32_t bid = -1; float bt = tmax, bu = 0, bv = 0;
    if (b->n_nodes) {
        const float inv[3] = { 1.0f / d[0], 1.0f / d[1], 1.0f / d[2] };
        uint32_t stack[128]; int sp = 0;
        stack[sp++] = 0;
        while (sp) {
            uint32_t ni = stack[--sp];
            const orc_node *n = b->nodes + ni;
            /* bt is inclusive upper bound: equal-t candidates must still be visited for the tie rule */
            if (!slab(n, o, inv, tmin, bt)) continue;
            if (n->count) {
                for (uint32_t i = n->left; i < n->left + n->count; i++)
                    consider(b->m, b->prims[i], o, d, tmin, &bid, &bt, &bu, &bv);
            } else {
                stack[sp++] = n->left;
                stack[sp++] = ni + 1;
            }
        }
    }
    *id = bid;
    if (bid >= 0) { *t = bt; *u = bu; *v = bv; }
}

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* wavefront_isect.cu:80-112 as a host loop */
void orc_intersect_streams(orc_intersect1_fn fn, void *ctx,
                           const float *ox, const float *oy, const float *oz,
                           const float *dx, const float *dy, const float *dz,
                           float tmin, float tmax, const uint8_t *mask, uint32_t mask_size, uint32_t n,
                           int32_t *id, float *t, float *u, float *v, int threads)
{
    if (threads <= 0) threads = orc_num_threads();
#pragma omp parallel for schedule(dynamic, 1024) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        const int live = mask ? (mask_size == 1 ? mask[0] : mask[i]) : 1;
        if (!live) continue;
        const float o[3] = { ox[i], oy[i], oz[i] }, d[3] = { dx[i], dy[i], dz[i] };
        int32_t hid = -1; float ht = 0, hu = 0, hv = 0;
        fn(ctx, o, d, tmin, tmax, &hid, &ht, &hu, &hv);
        id[i] = hid;
        if (hid >= 0) { t[i] = ht; u[i] = hu; v[i] = hv; }
    }
}

/* the same with per-ray windows: struct Params carries tmin / tmax as ARRAYS (src/optixdata.h:20-21, filled from
 * Ray3C::m_tmin / m_tmax, src/ray.h:20-26; read per lane at wavefront_isect.cu:98-99) */
void orc_intersect_streams_win(orc_intersect1_fn fn, void *ctx,
                               const float *ox, const float *oy, const float *oz,
                               const float *dx, const float *dy, const float *dz,
                               const float *tmin, const float *tmax, uint32_t n,
                               int32_t *id, float *t, float *u, float *v, int threads)
{
    if (threads <= 0) threads = orc_num_threads();
#pragma omp parallel for schedule(dynamic, 1024) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        const float o[3] = { ox[i], oy[i], oz[i] }, d[3] = { dx[i], dy[i], dz[i] };
        int32_t hid = -1; float ht = 0, hu = 0, hv = 0;
        fn(ctx, o, d, tmin[i], tmax[i], &hid, &ht, &hu, &hv);
        id[i] = hid;
        if (hid >= 0) { t[i] = ht; u[i] = hu; v[i] = hv; }
    }
}

/* ImageTexture::eval, src/main.cpp:62-76 (texel_fetch_wrap :56-60 clamps; index bug of :52 fixed) */
void orc_texture_eval(const orc_texture *t, float u, float v, float rgb[3])
{
    const float su = u * (float)t->w - 0.5f, sv = v * (float)t->h - 0.5f;
    const float fx = floorf(su), fy = floorf(sv);
    const int x = (int)fx, y = (int)fy;
    const float dx1 = su - fx, dy1 = sv - fy, dx2 = 1.0f - dx1, dy2 = 1.0f - dy1;
    const int W = t->w, H = t->h;
#define CL(a, n) ((a) < 0 ? 0 : ((a) > (n) - 1 ? (n) - 1 : (a)))
    const int x0 = CL(x, W), x1 = CL(x + 1, W), y0 = CL(y, H), y1 = CL(y + 1, H);
#undef CL
    for (int k = 0; k < 3; k++) {
        const float t00 = t->rgb[3 * (y0 * W + x0) + k], t01 = t->rgb[3 * (y1 * W + x0) + k];
        const float t10 = t->rgb[3 * (y0 * W + x1) + k], t11 = t->rgb[3 * (y1 * W + x1) + k];
        rgb[k] = t00 * dx2 * dy2 + t01 * dx2 * dy1 + t10 * dx1 * dy2 + t11 * dx1 * dy1;
    }
}

/* ------------------------------------------------------------------ shading step
 * optix_backend.h:469 (p = o + t d), :483-484 (interpolated, un-normalised shading normal),
 * main.cpp:325-326 (material), :414-424 (frame, Lambert sample, next ray). */
void orc_shade_step(const orc_mesh *m, const float o[3], const float d[3], int32_t tri, float t, float u,
                    float v, const float s[2], float next_o[3], float next_d[3], float weight[3])
{
    for (int k = 0; k < 3; k++) next_o[k] = o[k] + t * d[k];
    float n[3], fr[9], local[3], gn[3];
    const float *n0 = gn, *n1 = gn, *n2 = gn;
    if (m->nrm && m->nrm_idx && m->nrm_idx[3 * tri] >= 0 && m->nrm_idx[3 * tri + 1] >= 0 && m->nrm_idx[3 * tri + 2] >= 0) {
        n0 = m->nrm + 3 * (size_t)m->nrm_idx[3 * tri + 0];
        n1 = m->nrm + 3 * (size_t)m->nrm_idx[3 * tri + 1];
        n2 = m->nrm + 3 * (size_t)m->nrm_idx[3 * tri + 2];
    } else {
        /* DOCUMENTED DEVIATION (include/wpt.h): a face without normals — where the reference
         * gathers out of bounds (optix_backend.h:412-420 with index -1) — shades with its geometric normal,
         * normalize(cross(p1-p0, p2-p0)) (add_math.h:9-16), at all three corners. */
        const float *p0 = m->pos + 3 * (size_t)m->pos_idx[3 * tri], *p1 = m->pos + 3 * (size_t)m->pos_idx[3 * tri + 1],
                    *p2 = m->pos + 3 * (size_t)m->pos_idx[3 * tri + 2];
        const float e1[3] = { p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2] };
        const float e2[3] = { p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2] };
        gn[0] = e1[1] * e2[2] - e1[2] * e2[1]; gn[1] = e1[2] * e2[0] - e1[0] * e2[2]; gn[2] = e1[0] * e2[1] - e1[1] * e2[0];
        const float l = sqrtf(gn[0] * gn[0] + gn[1] * gn[1] + gn[2] * gn[2]);
        if (l > 0) { gn[0] /= l; gn[1] /= l; gn[2] /= l; } else { gn[0] = 0; gn[1] = 1; gn[2] = 0; }
    }
    orc_barycentric3(n0, n1, n2, u, v, n);
    orc_frame_from_normal(n, fr);
    orc_cosine_hemisphere(s, local);
    orc_frame_to_world(fr, local, next_d);
    const int32_t mat = m->mat_id ? m->mat_id[tri] : 0;
    for (int k = 0; k < 3; k++) weight[k] = m->albedo ? m->albedo[3 * mat + k] : 1.0f;
    if (m->tex && m->tex[mat].rgb && m->tex[mat].w * m->tex[mat].h > 1) {
        /* LambertBsdf::sample -> m_reflectance->eval(texcoord) (main.cpp:116) at the interpolated texcoord
         * (optix_backend.h:479-480); faces without texcoords evaluate at (0,0) */
        float tu = 0.0f, tv = 0.0f;
        if (m->uvs && m->uv_idx && m->uv_idx[3 * tri] >= 0 && m->uv_idx[3 * tri + 1] >= 0 && m->uv_idx[3 * tri + 2] >= 0) {
            const float w = 1.0f - u - v;
            const float *a = m->uvs + 2 * (size_t)m->uv_idx[3 * tri], *b = m->uvs + 2 * (size_t)m->uv_idx[3 * tri + 1],
                        *c = m->uvs + 2 * (size_t)m->uv_idx[3 * tri + 2];
            tu = w * a[0] + u * b[0] + v * c[0];
            tv = w * a[1] + u * b[1] + v * c[1];
        }
        orc_texture_eval(&m->tex[mat], tu, tv, weight);
    }
}

/* optix_backend.h:463-486 + main.cpp:325; faces without normals: the documented deviation of orc_shade_step */
void orc_hit_attributes(const orc_mesh *m, const float *uvs, const int32_t *uv_idx, const float o[3], const float d[3],
                        int32_t tri, float t, float u, float v, float p[3], float gn[3], float sn[3], float uv[2],
                        int32_t *mat)
{
    for (int k = 0; k < 3; k++) p[k] = o[k] + t * d[k];                                     /* :469 */
    const float *p0 = m->pos + 3 * (size_t)m->pos_idx[3 * tri], *p1 = m->pos + 3 * (size_t)m->pos_idx[3 * tri + 1],
                *p2 = m->pos + 3 * (size_t)m->pos_idx[3 * tri + 2];
    orc_geometry_normal(p0, p1, p2, gn);                                                    /* :472-476 */
    if (m->nrm && m->nrm_idx && m->nrm_idx[3 * tri] >= 0 && m->nrm_idx[3 * tri + 1] >= 0 && m->nrm_idx[3 * tri + 2] >= 0) {
        orc_barycentric3(m->nrm + 3 * (size_t)m->nrm_idx[3 * tri], m->nrm + 3 * (size_t)m->nrm_idx[3 * tri + 1],
                         m->nrm + 3 * (size_t)m->nrm_idx[3 * tri + 2], u, v, sn);           /* :483-484 */
    } else {
        const float e1[3] = { p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2] };
        const float e2[3] = { p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2] };
        float g[3] = { e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0] };
        const float l = sqrtf(g[0] * g[0] + g[1] * g[1] + g[2] * g[2]);
        if (l > 0) { g[0] /= l; g[1] /= l; g[2] /= l; } else { g[0] = 0; g[1] = 1; g[2] = 0; }
        orc_barycentric3(g, g, g, u, v, sn);
    }
    uv[0] = uv[1] = 0.0f;
    if (uvs && uv_idx && uv_idx[3 * tri] >= 0 && uv_idx[3 * tri + 1] >= 0 && uv_idx[3 * tri + 2] >= 0) {    /* :479-480 */
        const float w = 1.0f - u - v;
        for (int k = 0; k < 2; k++)
            uv[k] = w * uvs[2 * (size_t)uv_idx[3 * tri] + k] + u * uvs[2 * (size_t)uv_idx[3 * tri + 1] + k] +
                    v * uvs[2 * (size_t)uv_idx[3 * tri + 2] + k];
    }
    *mat = m->mat_id ? m->mat_id[tri] : 0;                                                  /* main.cpp:325 */
}

/* ------------------------------------------------------------------ the loop (main.cpp:376-429)
 * Pixels never interact, so the reference's "all pixels per sample" array ops are run pixel-outer.
 * RNG draw order per sample: lens(2), jitter(2), then 2 per bounce for EVERY lane, live or not
 * (main.cpp:395,396,413) — kept so the PCG32 stream position matches the reference's. */
typedef struct {
    int mode; orc_pcg32 pcg; uint32_t key[2]; uint32_t pixel, spp;
} orc_rng;

static void rng_draw(orc_rng *r, uint32_t slot, float out[4], int count)
{
    if (r->mode == ORC_RNG_PCG32) {
        for (int i = 0; i < count; i++) out[i] = orc_pcg32_next_f32(&r->pcg);
    } else {
        const uint32_t ctr[4] = { r->pixel, r->spp, slot, 0u };
        uint32_t o[4];
        orc_philox4x32_10(ctr, r->key, o);
        for (int i = 0; i < 4; i++) out[i] = orc_u32_to_f32(o[i]);
    }
}

uint64_t orc_render(const orc_mesh *m, orc_intersect1_fn fn, void *ctx, const orc_camera *cam,
                    const orc_render_cfg *cfg, float *film_r, float *film_g, float *film_b, int threads)
{
    const int W = cfg->width, H = cfg->height;
    const int64_t N = (int64_t)W * H;
    uint64_t segments = 0;
    if (threads <= 0) threads = orc_num_threads();
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : segments) num_threads(threads)
    for (int64_t pix = 0; pix < N; pix++) {
        const int py = (int)(pix / W), px = (int)(pix % W); /* main.cpp:379-381 */
        orc_rng rng;
        rng.mode = cfg->rng_mode;
        rng.pixel = (uint32_t)pix;
        rng.key[0] = (uint32_t)cfg->seed; rng.key[1] = (uint32_t)(cfg->seed >> 32);
        if (rng.mode == ORC_RNG_PCG32) {
            orc_pcg32_seed(&rng.pcg, 0x853c49e6748fea9bULL, (uint64_t)pix); /* main.cpp:376 */
            const int64_t skip = (int64_t)cfg->spp_begin * (4 + 2 * cfg->depth);
            for (int64_t i = 0; i < skip; i++) orc_pcg32_next_u32(&rng.pcg);
        }
        float acc[3] = { 0, 0, 0 };
        for (int s = cfg->spp_begin; s < cfg->spp_begin + cfg->spp_count; s++) {
            rng.spp = (uint32_t)s;
            float contrib[3] = { 1.0f, 1.0f, 1.0f }; /* main.cpp:391 */
            int active = 1;                          /* main.cpp:392 */
            float xi[4], org[3], dir[3];
            rng_draw(&rng, 0, xi, 4);
            orc_camera_sample_pos(cam, xi, org);                         /* main.cpp:395 */
            orc_camera_sample_dir(cam, px, py, org, W, H, xi + 2, dir); /* main.cpp:396 */
            for (int j = 0; j < cfg->depth; j++) {                       /* main.cpp:399 */
                float sc[4];
                rng_draw(&rng, 1u + (uint32_t)j, sc, 2);                 /* main.cpp:413 */
                if (!active) continue;
                int32_t id = -1; float t = 0, u = 0, v = 0;
                fn(ctx, org, dir, cfg->tmin, cfg->tmax, &id, &t, &u, &v); /* main.cpp:404 */
                segments++;
                if (id < 0) {                                            /* main.cpp:407,410 */
                    acc[0] += contrib[0]; acc[1] += contrib[1]; acc[2] += contrib[2];
                    active = 0;
                    continue;
                }
                float no[3], nd[3], w[3];
                orc_shade_step(m, org, dir, id, t, u, v, sc, no, nd, w);
                for (int k = 0; k < 3; k++) { contrib[k] *= w[k]; org[k] = no[k]; dir[k] = nd[k]; } /* :422-424 */
            }
        }
        film_r[pix] += acc[0]; film_g[pix] += acc[1]; film_b[pix] += acc[2];
    }
    return segments;
}
