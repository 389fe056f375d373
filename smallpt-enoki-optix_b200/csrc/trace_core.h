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
// Instruction budget of the node test on sm_100a (measured, tools/mb/pipes.cu): PRMT/LOP3/SHF/SEL/
// FSET/FMNMX3 share the ALU pipe at 0.5 warp-instr/clk/SMSP; FFMA (and 2-input FMNMX) issue at
// 1/clk; IDP.4A and IMAD run on the heavy FMA pipe at 0.5/clk, beside the ALU pipe.  Per child:
// 6 IDP.4A (plane byte -> float operand), 6 FFMA, 2 FMNMX3 + 2 FMNMX, 1 FSET, 1 FFMA (mask
// accumulate); per node 12 SEL (near/far words by ray sign).
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
// d = c + sum of the four byte products of a and b (unsigned)
WPT_HD uint32_t dp4a_u(uint32_t a, uint32_t b, uint32_t c)
{
#if defined(__CUDA_ARCH__)
    return __dp4a(a, b, c);
#else
    for (int i = 0; i < 4; i++) c += ((a >> (8 * i)) & 0xffu) * ((b >> (8 * i)) & 0xffu);
    return c;
#endif
}
// plane byte J of `word` as the float 256 + q/256 (bits 0x43800000 + 128 q): one IDP.4A, no ALU-pipe
// instruction.  With A = 256*scale/d and B = (p-o)/d - 256*A:  fma(f, A, B) = (p + q*scale - o)/d.
template <int J>
WPT_HD float q2f(uint32_t word)
{
    return u2f(dp4a_u(word, 0x80u << (8 * J), 0x43800000u));
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
// q0 = (p.x, p.y, p.z, base); q1 = (xlo[0..3], xlo[4..7], xhi[0..3], xhi[4..7]); q2 likewise for y,
// q3 for z.  Byte 7 of each array is header: xlo e[0]+8, xhi e[1]+8, ylo e[2]+8, yhi imask,
// zlo / zhi triangle counts.  Returns the 7-bit mask of children (slot order) whose box the ray
// segment [tmin, tmax] touches (empty slots hold an inverted box and never hit).
// node_hits7_raw: the same mask in bits 0..6 of the float bits of 2^23 + mask (bits 7..31 are not zero)
WPT_HD uint32_t node_hits7_raw(const uint4 q0, const uint4 q1, const uint4 q2, const uint4 q3, const RayCtx &r,
                               float tmax)
{
    const float ax = u2f((q1.y >> 24) << 23) * r.ix;     // 256 * scale / d
    const float ay = u2f((q1.w >> 24) << 23) * r.iy;
    const float az = u2f((q2.y >> 24) << 23) * r.iz;
    const float bx = fmaf(-256.0f, ax, (u2f(q0.x) - r.ox) * r.ix);
    const float by = fmaf(-256.0f, ay, (u2f(q0.y) - r.oy) * r.iy);
    const float bz = fmaf(-256.0f, az, (u2f(q0.z) - r.oz) * r.iz);
    const bool px = r.ix >= 0.0f, py = r.iy >= 0.0f, pz = r.iz >= 0.0f;
    const uint32_t xn[2] = { px ? q1.x : q1.z, px ? q1.y : q1.w }, xf[2] = { px ? q1.z : q1.x, px ? q1.w : q1.y };
    const uint32_t yn[2] = { py ? q2.x : q2.z, py ? q2.y : q2.w }, yf[2] = { py ? q2.z : q2.x, py ? q2.w : q2.y };
    const uint32_t zn[2] = { pz ? q3.x : q3.z, pz ? q3.y : q3.w }, zf[2] = { pz ? q3.z : q3.x, pz ? q3.w : q3.y };
    float acc = 8388608.0f;   // 2^23: the low mantissa bits collect the hit mask
#define WPT_CHILD(H, J)                                                                           \
    {                                                                                             \
        const float tx0 = fmaf(q2f<J>(xn[H]), ax, bx), tx1 = fmaf(q2f<J>(xf[H]), ax, bx);         \
        const float ty0 = fmaf(q2f<J>(yn[H]), ay, by), ty1 = fmaf(q2f<J>(yf[H]), ay, by);         \
        const float tz0 = fmaf(q2f<J>(zn[H]), az, bz), tz1 = fmaf(q2f<J>(zf[H]), az, bz);         \
        const float tn = fmaxf(fmaxf(tx0, ty0), fmaxf(tz0, r.tmin));                              \
        const float tf = fminf(fminf(tx1, ty1), fminf(tz1, tmax));                                \
        acc = fmaf(tn <= tf ? 1.0f : 0.0f, (float)(1 << (4 * H + J)), acc);                       \
    }
    WPT_CHILD(0, 0) WPT_CHILD(0, 1) WPT_CHILD(0, 2) WPT_CHILD(0, 3) WPT_CHILD(1, 0) WPT_CHILD(1, 1) WPT_CHILD(1, 2)
#undef WPT_CHILD
    return f2u(acc);
}
WPT_HD uint32_t node_hits7(const uint4 q0, const uint4 q1, const uint4 q2, const uint4 q3, const RayCtx &r,
                           float tmax)
{
    return node_hits7_raw(q0, q1, q2, q3, r, tmax) & 0x7fu;
}

#if defined(__CUDACC__)
// ---------------------------------------------------------------- node test, packed fp32 (device only)
// sm_100a FFMA2 (fma.rn.f32x2): two IEEE fp32 FMAs per issued instruction — bit-identical results to
// node_hits7, 21 issue slots instead of 42 for the plane distances (measured, tools/mb/pipes2.cu:
// FFMA2 issues at half rate on the pipe IDP.4A uses, so it trades pipe cycles for issue slots, and
// the traversal is issue bound).  Pairs: (x,y) of the near planes, (x,y) of the far planes, and
// (z near, z far) of one child.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 p, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ uint32_t node_hits7_p(const uint4 q0, const uint4 q1, const uint4 q2, const uint4 q3,
                                                 const RayCtx &r, float tmax)
{
    const float ax = u2f((q1.y >> 24) << 23) * r.ix;     // 256 * scale / d
    const float ay = u2f((q1.w >> 24) << 23) * r.iy;
    const float az = u2f((q2.y >> 24) << 23) * r.iz;
    const float bx = fmaf(-256.0f, ax, (u2f(q0.x) - r.ox) * r.ix);
    const float by = fmaf(-256.0f, ay, (u2f(q0.y) - r.oy) * r.iy);
    const float bz = fmaf(-256.0f, az, (u2f(q0.z) - r.oz) * r.iz);
    const f32x2 Axy = pack2(ax, ay), Bxy = pack2(bx, by), Azz = pack2(az, az), Bzz = pack2(bz, bz);
    const bool px = r.ix >= 0.0f, py = r.iy >= 0.0f, pz = r.iz >= 0.0f;
    const uint32_t xn[2] = { px ? q1.x : q1.z, px ? q1.y : q1.w }, xf[2] = { px ? q1.z : q1.x, px ? q1.w : q1.y };
    const uint32_t yn[2] = { py ? q2.x : q2.z, py ? q2.y : q2.w }, yf[2] = { py ? q2.z : q2.x, py ? q2.w : q2.y };
    const uint32_t zn[2] = { pz ? q3.x : q3.z, pz ? q3.y : q3.w }, zf[2] = { pz ? q3.z : q3.x, pz ? q3.w : q3.y };
    float acc = 8388608.0f;
#define WPT_CHILD(H, J)                                                                           \
    {                                                                                             \
        float tx0, ty0, tx1, ty1, tz0, tz1;                                                       \
        unpack2(fma2(pack2(q2f<J>(xn[H]), q2f<J>(yn[H])), Axy, Bxy), tx0, ty0);                   \
        unpack2(fma2(pack2(q2f<J>(xf[H]), q2f<J>(yf[H])), Axy, Bxy), tx1, ty1);                   \
        unpack2(fma2(pack2(q2f<J>(zn[H]), q2f<J>(zf[H])), Azz, Bzz), tz0, tz1);                   \
        const float tn = fmaxf(fmaxf(tx0, ty0), fmaxf(tz0, r.tmin));                              \
        const float tf = fminf(fminf(tx1, ty1), fminf(tz1, tmax));                                \
        acc = fmaf(tn <= tf ? 1.0f : 0.0f, (float)(1 << (4 * H + J)), acc);                       \
    }
    WPT_CHILD(0, 0) WPT_CHILD(0, 1) WPT_CHILD(0, 2) WPT_CHILD(0, 3) WPT_CHILD(1, 0) WPT_CHILD(1, 1) WPT_CHILD(1, 2)
#undef WPT_CHILD
    return f2u(acc) & 0x7fu;
}
#endif

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
    int32_t prim;   // original triangle index, -1 = miss
    float t, u, v;
};

// one 256-bit load (sm_100a LDG.E.256): two consecutive 16-byte quads for ONE L1 wavefront per lane
WPT_HD void ld_block_half(const uint4 *p, uint4 &a, uint4 &b)
{
#if defined(__CUDA_ARCH__)
    asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
        : "l"(p));
#else
    a = p[0]; b = p[1];
#endif
}
WPT_HD float4 as_f4(const uint4 q) { float4 f; f.x = u2f(q.x); f.y = u2f(q.y); f.z = u2f(q.z); f.w = u2f(q.w); return f; }

// triangle bit mask (bit i = i-th triangle of the node) of the leaf children in `leaf` (slot mask);
// cnt2 = the 2-bit triangle counts of slots 0..6
WPT_HD uint32_t leaf_tri_mask(uint32_t leaf, uint32_t cnt2)
{
    uint32_t T = 0;
    while (leaf) {
        const int s = high_bit(leaf);
        leaf ^= 1u << s;
        const uint32_t below = cnt2 & ((1u << (2 * s)) - 1u);
        const uint32_t off = (uint32_t)(pop_count(below & 0x1555u) + 2 * pop_count(below & 0x2aaau));
        const uint32_t cnt = (cnt2 >> (2 * s)) & 3u;
        T |= ((1u << cnt) - 1u) << off;
    }
    return T;
}

// Traversal entry: x = first child block of a node, y = (pending internal children, permuted to
// visiting priority) << 24 | imask.  STACK must provide push(uint2) / pop() / empty().
template <class STACK>
WPT_HD Hit trace_ray(const BvhView &bvh, const RayCtx &r, float tmax, STACK &stack, uint64_t *counters = nullptr)
{
    Hit hit;
    hit.prim = -1; hit.t = tmax; hit.u = 0.0f; hit.v = 0.0f;
    uint2 G; G.x = 0u; G.y = 0x01000000u;  // root: block 0, one pending child, imask 0 -> rel index 0
    for (;;) {
        const int bit = high_bit(G.y);
        G.y ^= 1u << bit;
        if (G.y >> 24) stack.push(G);
        const uint32_t slot = ((uint32_t)bit - 24u) ^ r.oct_inv;
        const uint32_t rel = (uint32_t)pop_count(G.y & ((1u << slot) - 1u) & 0xffu);
        const uint4 *np = bvh.blocks + (size_t)(G.x + rel) * 4;
        uint4 q0, q1, q2, q3;
        ld_block_half(np, q0, q1); ld_block_half(np + 2, q2, q3);
        const uint32_t h7 = node_hits7(q0, q1, q2, q3, r, hit.t);
        const uint32_t imask = q2.w >> 24;
        G.x = q0.w;
        G.y = (perm8(h7 & imask, r.oct_inv) << 24) | imask;
        uint32_t T = leaf_tri_mask(h7 & ~imask, (q3.y >> 24) | ((q3.w >> 24) << 8));
        if (counters) { counters[0]++; counters[1] += (uint64_t)pop_count(T); }
        const uint32_t tri0 = q0.w + (uint32_t)pop_count(imask);
        while (T) {
            const int tb = high_bit(T);
            T ^= 1u << tb;
            const uint4 *tp = bvh.blocks + (size_t)(tri0 + (uint32_t)tb) * 4;
            uint4 a, b, c, d;
            ld_block_half(tp, a, b); ld_block_half(tp + 2, c, d);
            if (tri_hit(as_f4(a), as_f4(b), as_f4(c), r, hit.t, hit.u, hit.v)) hit.prim = (int32_t)d.x;
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
