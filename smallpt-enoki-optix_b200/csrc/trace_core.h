// trace_core.h — the ray/node and ray/triangle tests and the single-ray traversal loop.
// __host__ __device__ so the exact same logic that runs in the sm_100a kernels can be exercised by
// the CPU-only unit tests (tests build it into a separate test-only library; the product never
// executes it on the host).
//
// Replaces optixTrace + __closesthit__ch + __miss__ms (src/wavefront_isect.cu:48-71,103-105).
// Triangle test: accept rules and operation order of Embree's Moeller-Trumbore intersector
// (dependencies/embree/kernels/geometry/triangle_intersector_moeller.h:75-113): den != 0, U >= 0,
// V >= 0, U+V <= |den|, |den|*tnear < T <= |den|*tfar.
//
// Instruction budget of the node test (sm_100a: the ALU pipe — PRMT/LOP3/SHF/SEL/FMNMX/FSETP —
// issues at half the rate of the FMA pipe — FFMA/FMUL/IMAD — so ALU-pipe instructions are what is
// minimised): per child 6 FFMA, 2 FMNMX3, 2 FMNMX, 1 FSET, 1 FFMA (mask accumulate); per child PAIR
// 6 SEL (near/far by ray sign) and 6 IMAD.SHL (even child's planes); the odd child's planes are used
// as floats as loaded (their low 16 bits are the sibling's plane: the builder rounds so that the
// perturbation is conservative).
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>
#include "bvh.h"

#if defined(__CUDACC__)
#define WPT_HD __host__ __device__ __forceinline__
#else
#define WPT_HD inline
#endif

namespace wpt {

// ---------------------------------------------------------------- bit helpers
WPT_HD float u2f(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
WPT_HD uint32_t f2u(float f)
{
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
WPT_HD int high_bit(uint32_t x)  // index of the highest set bit, x != 0
{
#if defined(__CUDA_ARCH__)
    return 31 - __clz((int)x);
#else
    return 31 - __builtin_clz(x);
#endif
}
WPT_HD int pop_count(uint32_t x)
{
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
// byte `i` (0..7) of the 64-bit value {hi,lo}
WPT_HD uint32_t byte_of(uint32_t lo, uint32_t hi, uint32_t i)
{
#if defined(__CUDA_ARCH__)
    return __byte_perm(lo, hi, i) & 0xffu;
#else
    return (uint32_t)((((uint64_t)hi << 32) | lo) >> (8 * i)) & 0xffu;
#endif
}
// permutation of the bit positions of an 8-bit mask by XOR with o: result bit (s ^ o) = m bit s.
// Internal children are visited in order of decreasing (slot ^ oct_inv) — slot s holds the child
// lying furthest towards corner s of the node, so that is front to back for the ray's octant.
WPT_HD uint32_t perm8(uint32_t m, uint32_t o)
{
    if (o & 1u) m = ((m & 0x55u) << 1) | ((m >> 1) & 0x55u);
    if (o & 2u) m = ((m & 0x33u) << 2) | ((m >> 2) & 0x33u);
    if (o & 4u) m = ((m & 0x0fu) << 4) | ((m >> 4) & 0x0fu);
    return m;
}

// ---------------------------------------------------------------- ray
struct RayCtx {
    float ox, oy, oz, dx, dy, dz;
    float ix, iy, iz;     // reciprocal direction (zero components replaced by +-1e-30 first)
    float tmin;
    uint32_t oct_inv;     // dx>=0 | dy>=0<<1 | dz>=0<<2
};

WPT_HD RayCtx make_ray(float ox, float oy, float oz, float dx, float dy, float dz, float tmin)
{
    RayCtx r;
    r.ox = ox; r.oy = oy; r.oz = oz; r.dx = dx; r.dy = dy; r.dz = dz; r.tmin = tmin;
    const float eps = 1e-30f;
    const float sx = fabsf(dx) > eps ? dx : copysignf(eps, dx);
    const float sy = fabsf(dy) > eps ? dy : copysignf(eps, dy);
    const float sz = fabsf(dz) > eps ? dz : copysignf(eps, dz);
    r.ix = 1.0f / sx; r.iy = 1.0f / sy; r.iz = 1.0f / sz;
    r.oct_inv = (sx >= 0.0f ? 1u : 0u) | (sy >= 0.0f ? 2u : 0u) | (sz >= 0.0f ? 4u : 0u);
    return r;
}

// ---------------------------------------------------------------- node test
// n0 = (p.x, p.y, p.z, e|imask<<24); q[2*axis + side] = the four plane words of that axis/side.
// Returns the 8-bit mask of children (slot order) whose box the ray segment [tmin, tmax] touches.
WPT_HD uint32_t node_hits8(const uint4 n0, const uint4 qxl, const uint4 qxh, const uint4 qyl, const uint4 qyh,
                           const uint4 qzl, const uint4 qzh, const RayCtx &r, float tmax)
{
    const uint32_t e = n0.w;
    const float ax = u2f((e & 0xffu) << 23) * r.ix;
    const float ay = u2f(((e >> 8) & 0xffu) << 23) * r.iy;
    const float az = u2f(((e >> 16) & 0xffu) << 23) * r.iz;
    const float bx = (u2f(n0.x) - r.ox) * r.ix;
    const float by = (u2f(n0.y) - r.oy) * r.iy;
    const float bz = (u2f(n0.z) - r.oz) * r.iz;
    const bool px = r.ix >= 0.0f, py = r.iy >= 0.0f, pz = r.iz >= 0.0f;
    float acc = 8388608.0f;   // 2^23: the low mantissa bits collect the hit mask
#define WPT_PAIR(J, W)                                                                             \
    {                                                                                              \
        const uint32_t xn = px ? qxl.W : qxh.W, xf = px ? qxh.W : qxl.W;                           \
        const uint32_t yn = py ? qyl.W : qyh.W, yf = py ? qyh.W : qyl.W;                           \
        const uint32_t zn = pz ? qzl.W : qzh.W, zf = pz ? qzh.W : qzl.W;                           \
        {   /* even child: low halves */                                                           \
            const float tx0 = fmaf(u2f(xn << 16), ax, bx), tx1 = fmaf(u2f(xf << 16), ax, bx);      \
            const float ty0 = fmaf(u2f(yn << 16), ay, by), ty1 = fmaf(u2f(yf << 16), ay, by);      \
            const float tz0 = fmaf(u2f(zn << 16), az, bz), tz1 = fmaf(u2f(zf << 16), az, bz);      \
            const float tn = fmaxf(fmaxf(tx0, ty0), fmaxf(tz0, r.tmin));                           \
            const float tf = fminf(fminf(tx1, ty1), fminf(tz1, tmax));                             \
            acc = fmaf(tn <= tf ? 1.0f : 0.0f, (float)(1 << (2 * J)), acc);                        \
        }                                                                                          \
        {   /* odd child: the words as they are */                                                 \
            const float tx0 = fmaf(u2f(xn), ax, bx), tx1 = fmaf(u2f(xf), ax, bx);        \
            const float ty0 = fmaf(u2f(yn), ay, by), ty1 = fmaf(u2f(yf), ay, by);        \
            const float tz0 = fmaf(u2f(zn), az, bz), tz1 = fmaf(u2f(zf), az, bz);        \
            const float tn = fmaxf(fmaxf(tx0, ty0), fmaxf(tz0, r.tmin));                           \
            const float tf = fminf(fminf(tx1, ty1), fminf(tz1, tmax));                             \
            acc = fmaf(tn <= tf ? 1.0f : 0.0f, (float)(2 << (2 * J)), acc);                        \
        }                                                                                          \
    }
    WPT_PAIR(0, x) WPT_PAIR(1, y) WPT_PAIR(2, z) WPT_PAIR(3, w)
#undef WPT_PAIR
    return f2u(acc) & 0xffu;
}

// ---------------------------------------------------------------- triangle test
// t0 = (v0.xyz, e1.x) t1 = (e1.yz, e2.xy) t2 = (e2.z, Ng.xyz).  On accept updates tmax,u,v.
// Like Embree (triangle_intersector_moeller.h:109-113, rcp = approximation + one Newton step) the
// accepted hit is scaled by ONE reciprocal of |den| instead of three divisions.
WPT_HD float rcp_nr(float a)
{
#if defined(__CUDA_ARCH__)
    float r;
    a = fmaxf(a, 1e-37f);     // accept path only: keeps a denormal |den| from turning into inf
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r * fmaf(-a, r, 2.0f);
#else
    return 1.0f / a;
#endif
}
WPT_HD bool tri_hit(const float4 t0, const float4 t1, const float4 t2, const RayCtx &r, float &tmax,
                    float &u, float &v)
{
    const float cx = t0.x - r.ox, cy = t0.y - r.oy, cz = t0.z - r.oz;
    // R = cross(C, D)
    const float rx = fmaf(cy, r.dz, -(cz * r.dy));
    const float ry = fmaf(cz, r.dx, -(cx * r.dz));
    const float rz = fmaf(cx, r.dy, -(cy * r.dx));
    const float ngx = t2.y, ngy = t2.z, ngz = t2.w;
    const float den = fmaf(ngx, r.dx, fmaf(ngy, r.dy, ngz * r.dz));
    const float absden = fabsf(den);
    const uint32_t sgn = f2u(den) & 0x80000000u;
    const float U = u2f(f2u(fmaf(rx, t1.z, fmaf(ry, t1.w, rz * t2.x))) ^ sgn);   // dot(R, e2)
    const float V = u2f(f2u(fmaf(rx, t0.w, fmaf(ry, t1.x, rz * t1.y))) ^ sgn);   // dot(R, e1)
    if (!(den != 0.0f && U >= 0.0f && V >= 0.0f && U + V <= absden)) return false;
    const float T = u2f(f2u(fmaf(ngx, cx, fmaf(ngy, cy, ngz * cz))) ^ sgn);      // dot(Ng, C)
    if (!(absden * r.tmin < T && T <= absden * tmax)) return false;
    const float rc = rcp_nr(absden);
    tmax = T * rc;
    u = U * rc;
    v = V * rc;
    return true;
}

// ---------------------------------------------------------------- one ray, start to finish
struct Hit {
    int32_t slot;   // leaf slot (index into tris / prim_id), -1 = miss
    float t, u, v;
};

WPT_HD uint4 ld_node(const uint4 *p)
{
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}
// one 256-bit load (sm_100a LDG.E.256): two consecutive 16-byte quads in a single L1 wavefront per
// lane — divergent traversal is bound by L1 wavefronts per node visit, not by bytes
WPT_HD void ld_node2(const uint4 *p, uint4 &a, uint4 &b)
{
#if defined(__CUDA_ARCH__)
    asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
        : "l"(p));
#else
    a = p[0]; b = p[1];
#endif
}
WPT_HD float4 ld_tri(const float4 *p)
{
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}

// triangle bit mask (bit i = triangle tri_base + i) of the leaf children in `leaf` (slot mask)
WPT_HD uint32_t leaf_tri_mask(uint32_t leaf, uint32_t meta_lo, uint32_t meta_hi)
{
    uint32_t T = 0;
    while (leaf) {
        const int s = high_bit(leaf);
        leaf ^= 1u << s;
        const uint32_t m = byte_of(meta_lo, meta_hi, (uint32_t)s);
        T |= (m >> 5) << (m & 31u);     // unary count << offset
    }
    return T;
}

// Traversal entry: x = index of the first internal child of a node, y = (pending internal children,
// permuted to visiting priority) << 24 | imask.  STACK must provide push(uint2) / pop() / empty().
template <class STACK>
WPT_HD Hit trace_ray(const BvhView &bvh, const RayCtx &r, float tmax, STACK &stack, uint64_t *counters = nullptr)
{
    Hit hit;
    hit.slot = -1; hit.t = tmax; hit.u = 0.0f; hit.v = 0.0f;
    uint2 G; G.x = 0u; G.y = 0x01000000u;  // root: base 0, one pending child, imask 0 -> rel index 0
    for (;;) {
        const int bit = high_bit(G.y);
        G.y ^= 1u << bit;
        if (G.y >> 24) stack.push(G);
        const uint32_t slot = ((uint32_t)bit - 24u) ^ r.oct_inv;
        const uint32_t rel = (uint32_t)pop_count(G.y & ((1u << slot) - 1u) & 0xffu);
        const uint4 *np = bvh.nodes + (size_t)(G.x + rel) * 8;
        uint4 n0, n1, qxl, qxh, qyl, qyh, qzl, qzh;
        ld_node2(np, n0, n1); ld_node2(np + 2, qxl, qxh); ld_node2(np + 4, qyl, qyh); ld_node2(np + 6, qzl, qzh);
        const uint32_t h8 = node_hits8(n0, qxl, qxh, qyl, qyh, qzl, qzh, r, hit.t);
        const uint32_t imask = n0.w >> 24;
        G.x = n1.x;
        G.y = (perm8(h8 & imask, r.oct_inv) << 24) | imask;
        uint32_t T = leaf_tri_mask(h8 & ~imask, n1.z, n1.w);
        if (counters) { counters[0]++; counters[1] += (uint64_t)pop_count(T); }
        while (T) {
            const int tb = high_bit(T);
            T ^= 1u << tb;
            const uint32_t ti = n1.y + (uint32_t)tb;
            const float4 *tp = bvh.tris + (size_t)ti * 3;
            if (tri_hit(ld_tri(tp), ld_tri(tp + 1), ld_tri(tp + 2), r, hit.t, hit.u, hit.v)) hit.slot = (int32_t)ti;
        }
        if (!(G.y >> 24)) {
            if (stack.empty()) break;
            G = stack.pop();
        }
    }
    return hit;
}

// thread-private stack (used by the CPU tests and the simple kernel)
struct LocalStack {
    uint2 e[kStackSize];
    int n = 0;
    WPT_HD void push(uint2 v) { e[n++] = v; }
    WPT_HD uint2 pop() { return e[--n]; }
    WPT_HD bool empty() const { return n == 0; }
};

}  // namespace wpt
