// trace_core.h — the ray/node and ray/triangle tests and the single-ray traversal loop.
// __host__ __device__ so the exact same logic that runs in the sm_100a kernels can be exercised by
// the CPU-only unit tests (tests build it into a separate test-only library; the product never
// executes it on the host).
//
// Replaces optixTrace + __closesthit__ch + __miss__ms (src/wavefront_isect.cu:48-71,103-105).
// Triangle test: accept rules and operation order of Embree's Moeller-Trumbore intersector
// (dependencies/embree/kernels/geometry/triangle_intersector_moeller.h:75-113): den != 0, U >= 0,
// V >= 0, U+V <= |den|, |den|*tnear < T <= |den|*tfar.
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
// PTX prmt.b32, default mode: selector nibble bits 2:0 pick a byte of {b,a}, bit 3 replicates that
// byte's sign bit instead.  (__byte_perm ignores bit 3, so the sign mode needs the PTX form.)
WPT_HD uint32_t byte_perm(uint32_t a, uint32_t b, uint32_t sel)
{
#if defined(__CUDA_ARCH__)
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
#else
    const uint64_t src = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) {
        const uint32_t n = (sel >> (4 * i)) & 0xf;
        uint32_t byte = (uint32_t)(src >> (8 * (n & 7))) & 0xff;
        if (n & 8) byte = (byte & 0x80) ? 0xff : 0x00;
        r |= byte << (8 * i);
    }
    return r;
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

// float with bits 0x4380qq00 = 256 + q/128: one PRMT turns a quantised byte into an FMA operand
// (no I2F).  t = f*A + B with A = 128*scale/d and B = (p-o)/d - 256*A gives (p + q*scale - o)/d.
template <int J>
WPT_HD float q2f(uint32_t word)
{
    return u2f(byte_perm(word, 0x43800000u, 0x7604u | (J << 4)));
}

// ---------------------------------------------------------------- ray
struct RayCtx {
    float ox, oy, oz, dx, dy, dz;
    float ix, iy, iz;     // reciprocal direction (zero components replaced by +-1e-30 first)
    float tmin;
    uint32_t oct_inv4;    // (dx>=0 | dy>=0<<1 | dz>=0<<2) replicated into four bytes
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
    const uint32_t oi = (sx >= 0.0f ? 1u : 0u) | (sy >= 0.0f ? 2u : 0u) | (sz >= 0.0f ? 4u : 0u);
    r.oct_inv4 = oi * 0x01010101u;
    return r;
}

// ---------------------------------------------------------------- node test
// Returns the 32-bit hit mask: bits 24..31 internal children at priority (slot ^ oct_inv), bits
// 0..23 triangles (offset from the node's tri_base).
WPT_HD uint32_t node_hits(const uint4 n0, const uint4 n1, const uint4 n2, const uint4 n3, const uint4 n4,
                          const RayCtx &r, float tmax)
{
    const uint32_t e = n0.w;
    const float ax = u2f(((e & 0xffu) + 7u) << 23) * r.ix;
    const float ay = u2f((((e >> 8) & 0xffu) + 7u) << 23) * r.iy;
    const float az = u2f((((e >> 16) & 0xffu) + 7u) << 23) * r.iz;
    const float bx = fmaf(-256.0f, ax, (u2f(n0.x) - r.ox) * r.ix);
    const float by = fmaf(-256.0f, ay, (u2f(n0.y) - r.oy) * r.iy);
    const float bz = fmaf(-256.0f, az, (u2f(n0.z) - r.oz) * r.iz);

    // near/far plane bytes per axis, by ray direction sign
    const bool px = r.ix >= 0.0f, py = r.iy >= 0.0f, pz = r.iz >= 0.0f;
    const uint32_t xn[2] = { px ? n2.x : n3.z, px ? n2.y : n3.w };
    const uint32_t xf[2] = { px ? n3.z : n2.x, px ? n3.w : n2.y };
    const uint32_t yn[2] = { py ? n2.z : n4.x, py ? n2.w : n4.y };
    const uint32_t yf[2] = { py ? n4.x : n2.z, py ? n4.y : n2.w };
    const uint32_t zn[2] = { pz ? n3.x : n4.z, pz ? n3.y : n4.w };
    const uint32_t zf[2] = { pz ? n4.z : n3.x, pz ? n4.w : n3.y };

    uint32_t hitmask = 0;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint32_t meta4 = h ? n1.w : n1.z;
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = byte_perm(is_inner4 << 3, 0u, 0xba98u);  // 0xff per inner byte
        const uint32_t bit_index4 = (meta4 ^ (r.oct_inv4 & inner_mask4)) & 0x1f1f1f1fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
#define WPT_CHILD(J)                                                                              \
        {                                                                                         \
            const float tx0 = fmaf(q2f<J>(xn[h]), ax, bx), tx1 = fmaf(q2f<J>(xf[h]), ax, bx);     \
            const float ty0 = fmaf(q2f<J>(yn[h]), ay, by), ty1 = fmaf(q2f<J>(yf[h]), ay, by);     \
            const float tz0 = fmaf(q2f<J>(zn[h]), az, bz), tz1 = fmaf(q2f<J>(zf[h]), az, bz);     \
            const float tn = fmaxf(fmaxf(tx0, ty0), fmaxf(tz0, r.tmin));                          \
            const float tf = fminf(fminf(tx1, ty1), fminf(tz1, tmax));                            \
            if (tn <= tf)                                                                         \
                hitmask |= ((child_bits4 >> (8 * J)) & 0xffu) << ((bit_index4 >> (8 * J)) & 0xffu); \
        }
        WPT_CHILD(0) WPT_CHILD(1) WPT_CHILD(2) WPT_CHILD(3)
#undef WPT_CHILD
    }
    return hitmask;
}

// ---------------------------------------------------------------- triangle test
// t0 = (v0.xyz, e1.x) t1 = (e1.yz, e2.xy) t2 = (e2.z, Ng.xyz).  On accept updates tmax,u,v.
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
    tmax = T / absden;
    u = U / absden;
    v = V / absden;
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
WPT_HD float4 ld_tri(const float4 *p)
{
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}

// STACK must provide push(uint2) / pop() / empty()
template <class STACK>
WPT_HD Hit trace_ray(const BvhView &bvh, const RayCtx &r, float tmax, STACK &stack)
{
    Hit hit;
    hit.slot = -1; hit.t = tmax; hit.u = 0.0f; hit.v = 0.0f;
    const uint32_t oct_inv = r.oct_inv4 & 7u;
    uint2 G; G.x = 0u; G.y = 0x80000000u;  // root: base 0, one pending child, imask 0 -> rel index 0
    for (;;) {
        uint2 T; T.x = 0u; T.y = 0u;
        if (G.y & 0xff000000u) {
            const int bit = high_bit(G.y);
            G.y &= ~(1u << bit);
            if (G.y & 0xff000000u) stack.push(G);
            const uint32_t slot = ((uint32_t)bit - 24u) ^ oct_inv;
            const uint32_t rel = (uint32_t)pop_count(G.y & ~(0xffffffffu << slot) & 0xffu);
            const uint4 *np = bvh.nodes + (size_t)(G.x + rel) * 5;
            const uint4 n0 = ld_node(np), n1 = ld_node(np + 1), n2 = ld_node(np + 2), n3 = ld_node(np + 3),
                        n4 = ld_node(np + 4);
            const uint32_t hm = node_hits(n0, n1, n2, n3, n4, r, hit.t);
            G.x = n1.x; G.y = (hm & 0xff000000u) | (n0.w >> 24);
            T.x = n1.y; T.y = hm & 0x00ffffffu;
        } else {
            T = G; G.x = 0u; G.y = 0u;
        }
        while (T.y) {
            const int bit = high_bit(T.y);
            T.y &= ~(1u << bit);
            const uint32_t ti = T.x + (uint32_t)bit;
            const float4 *tp = bvh.tris + (size_t)ti * 3;
            if (tri_hit(ld_tri(tp), ld_tri(tp + 1), ld_tri(tp + 2), r, hit.t, hit.u, hit.v)) hit.slot = (int32_t)ti;
        }
        if (!(G.y & 0xff000000u)) {
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
