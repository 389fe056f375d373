// trace_kernels.cu — hand-written sm_100a closest-hit BVH traversal (replaces optixLaunch,
// src/accel/optix_backend.h:460, and the OptiX programs src/wavefront_isect.cu:48-112):
//   trace_wq          persistent warps, dynamic ray fetch, warp-shared triangle queue (the default)
//   trace_wq<TailIO>  the same kernel with the shading step fused in: a path whose segment ends inside the
//                     kernel is shaded by its own lane (src/main.cpp:407-424) and continues with its next
//                     segment in place — one launch covers every remaining bounce of a pass
//   trace_persistent  per-lane triangle loop (comparison variant)
//   trace_simple      one thread per ray (comparison variant)
// Nothing here is a dense contraction: no tensor cores.  The traversal is bound by the L1 data stage
// (one wavefront per lane of every divergent 256-bit load) together with instruction issue
// (profiles/r1_trace_wq_ncu_64spp.txt, profiles/r2_*).
#include "kernels.cuh"
#include "shade_core.h"
#include "trace_core.h"
#include <type_traits>

namespace wpt {

// WPT_CHECKS build (libwpt_checked.so, tests/test_gpu_checked.py): compute-sanitizer is closed on this pool, so the
// invariants its memcheck / racecheck would watch in the warp-queue code are asserted by the kernel itself and counted
// per kind; wpt_debug_check_failures() reads the counters.  The product build compiles the checks away.
#ifdef WPT_CHECKS
__device__ unsigned int g_check_fail[8];
#define WPT_CHECK(cond, code) do { if (!(cond)) atomicAdd(&g_check_fail[code], 1u); } while (0)
int read_check_failures(unsigned int out[8], int reset)
{
    if (cudaMemcpyFromSymbol(out, g_check_fail, sizeof(unsigned int) * 8) != cudaSuccess) return 1;
    if (reset) { unsigned int z[8] = {0}; if (cudaMemcpyToSymbol(g_check_fail, z, sizeof z) != cudaSuccess) return 1; }
    return 0;
}
#else
#define WPT_CHECK(cond, code) ((void)0)
int read_check_failures(unsigned int *, int) { return -1; }
#endif
// check kinds: 0 queue slot out of range | 1 item slot out of range | 2 block index beyond the tree | 3 traversal
// stack out of range | 4 fetched ray index beyond the queue | 5 queue not empty after a drain | 6 item count differs
// from the scanned total | 7 owner lane of a queued entry out of range

// ------------------------------------------------------------------------------------------------
// ray sources / hit sinks
// ------------------------------------------------------------------------------------------------
struct NoPathState {};

// MODE 0: plain queue records, scalar tmax.  MODE 1 (depth-seeded camera rays, api.cu wpt_render): the ray's own tmax
// travels in b.w (bounce 0: the throughput there is 1 by construction); a ray that finds nothing inside its seed is
// appended to retry_list.  MODE 2: the second trace of those rays — indirect = that list, full tmax.
template <int MODE>
struct QueueIOT {
    static constexpr bool kContinue = false;
    using PathState = NoPathState;
    RayQueue q;
    int4 *hits;
    float tmin, tmax;
    uint32_t *retry_list, *retry_count;   // MODE 1
    const uint32_t *indirect;             // MODE 2
    __device__ __forceinline__ bool load(uint32_t i, float &ox, float &oy, float &oz, float &dx, float &dy,
                                         float &dz, float &t0, float &t1) const
    {
        const uint32_t j = MODE == 2 ? indirect[i] : i;
        const float4 a = __ldcs(q.a + j);   // streaming: do not displace the BVH in L2
        const float4 b = __ldcs(q.b + j);
        ox = a.x; oy = a.y; oz = a.z; dx = b.x; dy = b.y; dz = b.z; t0 = tmin; t1 = MODE == 1 ? b.w : tmax;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, int32_t prim, float t, float u, float v) const
    {
        const uint32_t j = MODE == 2 ? indirect[i] : i;
        __stcs(hits + j, make_int4(prim, __float_as_int(t), __float_as_int(u), __float_as_int(v)));
        if (MODE == 1 && prim < 0 && t < 1e19f) retry_list[atomicAdd(retry_count, 1u)] = j;
    }
};
using QueueIO = QueueIOT<0>;

struct StreamIODev {
    static constexpr bool kContinue = false;
    using PathState = NoPathState;
    StreamIO s;
    __device__ __forceinline__ bool load(uint32_t i, float &ox, float &oy, float &oz, float &dx, float &dy,
                                         float &dz, float &t0, float &t1) const
    {
        // wavefront_isect.cu:86-87: size-1 mask broadcasts; masked-out lanes are skipped entirely
        if (s.mask) {
            const uint8_t m = s.mask_size == 1 ? s.mask[0] : s.mask[i];
            if (!m) return false;
        }
        ox = __ldcs(s.ox + i); oy = __ldcs(s.oy + i); oz = __ldcs(s.oz + i);
        dx = __ldcs(s.dx + i); dy = __ldcs(s.dy + i); dz = __ldcs(s.dz + i);
        // per-ray windows (src/ray.h:20-26, src/optixdata.h:20-21).  The closest hit is folded with an integer
        // min on the float bits of t, which needs t >= 0: a negative tmin is clamped to 0 (a hit behind the
        // origin is not a hit of the reference's rays either: src/ray.h:16 starts them at 1e-3)
        t0 = fmaxf(s.tmin ? __ldcs(s.tmin + i) : s.tmin_s, 0.0f);
        t1 = s.tmax ? __ldcs(s.tmax + i) : s.tmax_s;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, int32_t prim, float t, float u, float v) const
    {
        s.tri_id[i] = prim;                 // -1 on miss (wavefront_isect.cu:70)
        if (prim >= 0) { s.t[i] = t; s.u[i] = u; s.v[i] = v; }
    }
};

// Fused tail (api.cu wpt_render, bounces >= tail_from of a pass): rays are fetched from the queue of bounce
// `bounce0`; a finished segment is shaded in place — miss: film += contrib (src/main.cpp:407); hit and not the last
// segment: attribute gather, frame, cosine sample, contrib *= albedo (src/main.cpp:413-424) — and the lane goes
// straight on with the next segment.  Same Philox counters (pixel, global sample, 1 + bounce) as shade_kernel, so the
// paths are the wavefront pipeline's paths.
struct TailIO {
    static constexpr bool kContinue = true;
    struct PathState {           // per warp, in shared memory: what shade_kernel reads from the queue record
        uint32_t pid[32];        // pixel | sample-in-pass << 24
        float    tr[32], tg[32], tb[32];
        uint32_t bounce[32];
    };
    RayQueue q;
    float tmin, tmax;
    const float4 *tri_attr;
    const float *albedo;
    float *film_r, *film_g, *film_b;
    int bounce0, depth;
    uint32_t spp_first, k0, k1;
    __device__ __forceinline__ bool load(uint32_t i, float &ox, float &oy, float &oz, float &dx, float &dy,
                                         float &dz, float &t0, float &t1, PathState &ps, unsigned lane) const
    {
        const float4 a = __ldcs(q.a + i);
        const float4 b = __ldcs(q.b + i);
        const float2 c = __ldcs(q.c + i);
        ox = a.x; oy = a.y; oz = a.z; dx = b.x; dy = b.y; dz = b.z; t0 = tmin; t1 = tmax;
        ps.pid[lane] = __float_as_uint(a.w);
        ps.tr[lane] = b.w; ps.tg[lane] = c.x; ps.tb[lane] = c.y;
        ps.bounce[lane] = (uint32_t)bounce0;
        return true;
    }
    // the segment (o, d) of this lane ended with (prim, t, u, v): returns true and the next segment if the path goes on
    __device__ __forceinline__ bool next_segment(PathState &ps, unsigned lane, uint32_t *live_cta, float &ox, float &oy,
                                                 float &oz, float &dx, float &dy, float &dz, int32_t prim, float t,
                                                 float u, float v) const
    {
        const uint32_t pid = ps.pid[lane], pixel = pid & 0x00ffffffu, j = ps.bounce[lane];
        const float tr = ps.tr[lane], tg = ps.tg[lane], tb = ps.tb[lane];
        if (prim < 0) {
            atomicAdd(film_r + pixel, tr);
            atomicAdd(film_g + pixel, tg);
            atomicAdd(film_b + pixel, tb);
            return false;
        }
        if ((int)j + 1 >= depth) return false;   // path is cut after `depth` segments (src/main.cpp:399)
        const float4 *ap = tri_attr + (size_t)prim * 3;
        const float4 q0 = __ldg(ap), q1 = __ldg(ap + 1), q2 = __ldg(ap + 2);
        const int mat = __float_as_int(q2.y);
        uint32_t rnd[4];
        philox4x32_10(pixel, spp_first + (pid >> 24), 1u + j, 0u, k0, k1, rnd);
        V3 no, nd;
        shade_step(v3(ox, oy, oz), v3(dx, dy, dz), t, u, v, v3(q0.x, q0.y, q0.z), v3(q0.w, q1.x, q1.y),
                   v3(q1.z, q1.w, q2.x), u32_to_unit(rnd[0]), u32_to_unit(rnd[1]), no, nd);
        ps.tr[lane] = tr * __ldg(albedo + 3 * mat);            // contrib *= bsdf weight (src/main.cpp:422)
        ps.tg[lane] = tg * __ldg(albedo + 3 * mat + 1);
        ps.tb[lane] = tb * __ldg(albedo + 3 * mat + 2);
        ps.bounce[lane] = j + 1u;
        atomicAdd(live_cta + j + 1u, 1u);
        ox = no.x; oy = no.y; oz = no.z; dx = nd.x; dy = nd.y; dz = nd.z;
        return true;
    }
};

template <class IO>
__device__ __forceinline__ bool io_load(const IO &io, uint32_t i, float &ox, float &oy, float &oz, float &dx, float &dy,
                                        float &dz, float &t0, float &t1, typename IO::PathState &ps, unsigned lane)
{
    if constexpr (IO::kContinue) return io.load(i, ox, oy, oz, dx, dy, dz, t0, t1, ps, lane);
    else return io.load(i, ox, oy, oz, dx, dy, dz, t0, t1);
}

// ------------------------------------------------------------------------------------------------
// persistent-thread traversal with dynamic ray fetch, per-lane triangle loop (comparison variant 2)
// ------------------------------------------------------------------------------------------------
constexpr int kTraceBlock = 128;
constexpr int kSmemStack = 10;     // entries kept in shared memory per thread; the rest spills to local

template <class IO, int FETCH_THRESHOLD>
__global__ void __launch_bounds__(kTraceBlock, 5)
trace_persistent(const BvhView bvh, const IO io, const uint32_t *__restrict__ n_ptr, uint32_t n_static,
                 uint32_t *__restrict__ cursor, unsigned long long *__restrict__ live_acc)
{
    __shared__ uint2 sstack[kSmemStack * kTraceBlock];
    __shared__ uint8_t perm_lut[8 * 128];   // perm_lut[o*128 + m] = perm8(m, o): visiting priority of slot mask m
    for (uint32_t i = threadIdx.x; i < 8u * 128u; i += kTraceBlock) perm_lut[i] = (uint8_t)perm8(i & 127u, i >> 7);
    __syncthreads();
    const uint32_t n = n_ptr ? *n_ptr : n_static;
    if (live_acc && blockIdx.x == 0 && threadIdx.x == 0) *live_acc += n;
    const unsigned lane = threadIdx.x & 31u;
    uint2 *const my = sstack + threadIdx.x;
    uint2 lstack[kStackSize - kSmemStack];
    int sp = 0;

    bool active = false, exhausted = false;
    uint32_t ray_i = 0;
    const uint8_t *lut = perm_lut;
    RayCtx r;
    Hit hit;
    uint2 G;
    G.x = G.y = 0u;
    hit.prim = -1; hit.t = hit.u = hit.v = 0.0f;

    for (;;) {
        // ---- refill idle lanes once too few of the warp are working: one warp-aggregated atomic on
        // the global cursor.  Single flat loop: every iteration is one node step of every active lane.
        const unsigned act = __ballot_sync(0xffffffffu, active);
        if (__popc(act) < (FETCH_THRESHOLD > 0 ? FETCH_THRESHOLD : 1)) {
            if (!exhausted) {
                const unsigned idle = ~act;
                const int cnt = __popc(idle);
                const int leader = __ffs(idle) - 1;
                uint32_t base = 0;
                if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)cnt);
                base = __shfl_sync(0xffffffffu, base, leader);
                if (base + (uint32_t)cnt >= n) exhausted = true;
                if (!active) {
                    const uint32_t i = base + (uint32_t)__popc(idle & ((1u << lane) - 1u));
                    float ox, oy, oz, dx, dy, dz, t0, t1;
                    if (i < n && io.load(i, ox, oy, oz, dx, dy, dz, t0, t1)) {
                        r = make_ray(ox, oy, oz, dx, dy, dz, t0);
                        lut = perm_lut + r.oct_inv * 128u;
                        hit.prim = -1; hit.t = t1; hit.u = hit.v = 0.0f;
                        G.x = 0u; G.y = 0x01000000u;
                        sp = 0;
                        ray_i = i;
                        active = true;
                    }
                }
                if (!__any_sync(0xffffffffu, active)) {
                    if (exhausted) break;
                    continue;
                }
            } else if (act == 0u) {
                break;
            }
        }
        if (active) {
            // next internal child, front to back: highest pending priority bit
            const int bit = 31 - __clz((int)G.y);
            G.y ^= 1u << bit;
            if (G.y >> 24) {
                if (sp < kSmemStack) my[sp * kTraceBlock] = G; else lstack[sp - kSmemStack] = G;
                sp++;
            }
            const uint32_t slot = ((uint32_t)bit - 24u) ^ r.oct_inv;
            const uint32_t rel = (uint32_t)__popc(G.y & ((1u << slot) - 1u) & 0xffu);
            const uint4 *np = bvh.blocks + (size_t)(G.x + rel) * 4;
            uint4 q0, q1, q2, q3;
            ld_block_half(np, q0, q1); ld_block_half(np + 2, q2, q3);
            const uint32_t h7 = node_hits7(q0, q1, q2, q3, r, hit.t);
            const uint32_t imask = q2.w >> 24;
            G.x = q0.w;
            G.y = ((uint32_t)lut[h7 & imask] << 24) | imask;
            uint32_t T = leaf_tri_mask(h7 & ~imask, (q3.y >> 24) | ((q3.w >> 24) << 8));
            const uint32_t tri0 = q0.w + (uint32_t)__popc(imask);
            while (T) {
                const int tb = 31 - __clz((int)T);
                T ^= 1u << tb;
                const uint4 *tp = bvh.blocks + (size_t)(tri0 + (uint32_t)tb) * 4;
                uint4 ta, tb4, tc, td;
                ld_block_half(tp, ta, tb4); ld_block_half(tp + 2, tc, td);
                if (tri_hit(as_f4(ta), as_f4(tb4), as_f4(tc), r, hit.t, hit.u, hit.v)) hit.prim = (int32_t)td.x;
            }
            if (!(G.y >> 24)) {
                if (sp == 0) {
                    io.store(ray_i, hit.prim, hit.t, hit.u, hit.v);
                    active = false;
                } else {
                    sp--;
                    G = sp < kSmemStack ? my[sp * kTraceBlock] : lstack[sp - kSmemStack];
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// persistent traversal with a WARP-SHARED TRIANGLE QUEUE (the default trace kernel)
// ------------------------------------------------------------------------------------------------
// ncu of the per-lane loop above (profiles/): on incoherent rays the triangle loop ran with 2.8 of
// 32 lanes and took 40 % of all issued instructions.  Here a node step only APPENDS the leaves it hit to a
// queue shared by the warp; whenever 32 entries are queued (and before every refill) the queued triangles are
// tested by whichever lane is free — for whichever ray of the warp they belong to — and the result is folded
// into that ray's closest hit with a shared-memory 64-bit atomicMin on (float bits of t, primitive): t > 0, so
// the integer order is the float order, and among equal t the lowest primitive index wins whatever the order
// of the tests (the record of a ray is written by exactly one lane; hits are reproducible run to run).
// Ray records and closest hits therefore live in shared memory.  A ray whose stack runs empty parks until the
// queue has been drained (before the next refill), then stores its hit.
constexpr int kQueueCap = 64;
constexpr int kItemCap = 96;

struct WarpShared {
    float4   ray_a[32];      // ox, oy, oz, dx
    float4   ray_b[32];      // dy, dz, tmin, -
    unsigned long long hit_tp[32];   // closest hit so far: float bits of t << 32 | primitive (0xffffffff = none)
    float    hit_u[32], hit_v[32];
    uint2    queue[kQueueCap];   // one entry per node step that hit leaves: x = owner lane << 27 | first
                                 // triangle block of the node, y = hit leaf slots << 16 | 2-bit triangle counts
    uint32_t items[kItemCap];    // balanced flush: owner lane << 27 | triangle block, one per queued triangle
};
__device__ __forceinline__ float ws_tmax(const WarpShared &ws, unsigned i)
{
    return __uint_as_float(reinterpret_cast<const uint32_t *>(ws.hit_tp)[2u * i + 1u]);
}
__device__ __forceinline__ int32_t ws_prim(const WarpShared &ws, unsigned i)
{
    return (int32_t)reinterpret_cast<const uint32_t *>(ws.hit_tp)[2u * i];
}
__device__ __forceinline__ unsigned long long hit_key(float t, uint32_t prim)
{
    return ((unsigned long long)__float_as_uint(t) << 32) | prim;
}

// 256-bit block-half load with a selectable L1 hint.  Measured on C3 (profiles/r1_microbench.md): triangle blocks
// are hardly ever re-used by the same SM, so they are loaded L1::no_allocate (+0.8 %); node loads must allocate —
// the top three levels of the tree live in L1 (no_allocate on nodes: -20 %).
template <int M>
__device__ __forceinline__ void ld_half_m(const uint4 *p, uint4 &a, uint4 &b)
{
#define WPT_LD(S) asm(S " {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(p))
    if (M == 0) WPT_LD("ld.global.nc.v8.b32");
    if (M == 2) WPT_LD("ld.global.nc.L1::no_allocate.v8.b32");
#undef WPT_LD
}

__device__ __forceinline__ uint32_t bfind_u32(uint32_t x)   // index of the highest set bit (FLO.U32)
{
    uint32_t r;
    asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(x));
    return r;
}

// FLUSH 0 — every lane < k takes one queued entry (the leaves one node step hit, 1..3 triangles each, up to 21), tests
// them one after the other against the owner's ray and folds the closest accepted hit into the owner's record
__device__ __forceinline__ void flush_queue_entries(const BvhView &bvh, WarpShared &ws, unsigned lane, uint32_t &qn, uint32_t k)
{
    __syncwarp();
    bool cand = false;
    uint32_t owner = 0, prim = 0;
    float t = 0.0f, u = 0.0f, v = 0.0f;
    if (lane < k) {
        const uint2 e = ws.queue[qn - k + lane];
        owner = e.x >> 27;
        const float4 ra = ws.ray_a[owner], rb = ws.ray_b[owner];
        RayCtx r;
        r.ox = ra.x; r.oy = ra.y; r.oz = ra.z; r.dx = ra.w; r.dy = rb.x; r.dz = rb.y; r.tmin = rb.z;
        t = ws_tmax(ws, owner);
        uint32_t T = leaf_tri_mask(e.y >> 16, e.y & 0x3fffu);   // bit i = i-th triangle of the node
        const uint4 *t0p = bvh.blocks + (size_t)(e.x & 0x07ffffffu) * 4;
        while (T) {
            const uint32_t tb = bfind_u32(T);
            T ^= 1u << tb;
            const uint4 *tp = t0p + (size_t)tb * 4;
            uint4 ta, tb4, tc, td;
            ld_half_m<2>(tp, ta, tb4); ld_half_m<2>(tp + 2, tc, td);
            if (tri_hit(as_f4(ta), as_f4(tb4), as_f4(tc), r, t, u, v)) { cand = true; prim = td.x; }
        }
        if (cand) atomicMin(&ws.hit_tp[owner], hit_key(t, prim));
    }
    __syncwarp();
    if (cand && ws.hit_tp[owner] == hit_key(t, prim)) { ws.hit_u[owner] = u; ws.hit_v[owner] = v; }
    __syncwarp();
    qn -= k;
}

// FLUSH 1 — balanced: the k entries are first expanded into one item per TRIANGLE (exclusive warp scan of the triangle
// counts), then every lane tests one item per round: ceil(triangles / 32) rounds at full width instead of
// max(triangles per entry) rounds at the width of the entry count (ncu r1: 13.1 of 32 lanes at the 32-entry flush,
// 7.9 at the drain before a refill)
__device__ __forceinline__ void flush_queue_balanced(const BvhView &bvh, WarpShared &ws, unsigned lane, uint32_t &qn, uint32_t k)
{
    __syncwarp();
    uint32_t T = 0, ex = 0;
    if (lane < k) {
        const uint2 e = ws.queue[qn - k + lane];
        ex = e.x;
        T = leaf_tri_mask(e.y >> 16, e.y & 0x3fffu);
    }
    qn -= k;
    for (;;) {
        const uint32_t c = (uint32_t)__popc(T);
        uint32_t incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
            if ((int)lane >= d) incl += up;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        uint32_t pos = incl - c;
        WPT_CHECK(__reduce_add_sync(0xffffffffu, c) == total, 6);
        while (T && pos < (uint32_t)kItemCap) {
            const uint32_t tb = bfind_u32(T);
            T ^= 1u << tb;
            WPT_CHECK(pos < (uint32_t)kItemCap, 1);
            WPT_CHECK((ex & 0x07ffffffu) + tb < bvh.n_blocks, 2);
            ws.items[pos++] = ex + tb;      // owner << 27 | (first block + tb): blocks < 2^27 (wpt_scene_commit)
        }
        __syncwarp();
        const uint32_t m = total < (uint32_t)kItemCap ? total : (uint32_t)kItemCap;
        for (uint32_t base = 0; base < m; base += 32u) {
            bool cand = false;
            uint32_t owner = 0, prim = 0;
            float t = 0.0f, u = 0.0f, v = 0.0f;
            if (base + lane < m) {
                WPT_CHECK(base + lane < (uint32_t)kItemCap, 1);
                const uint32_t it = ws.items[base + lane];
                owner = it >> 27;
                WPT_CHECK(owner < 32u, 7);
                WPT_CHECK((it & 0x07ffffffu) < bvh.n_blocks, 2);
                const float4 ra = ws.ray_a[owner], rb = ws.ray_b[owner];
                RayCtx r;
                r.ox = ra.x; r.oy = ra.y; r.oz = ra.z; r.dx = ra.w; r.dy = rb.x; r.dz = rb.y; r.tmin = rb.z;
                t = ws_tmax(ws, owner);
                const uint4 *tp = bvh.blocks + (size_t)(it & 0x07ffffffu) * 4;
                uint4 ta, tb4, tc, td;
                ld_half_m<2>(tp, ta, tb4); ld_half_m<2>(tp + 2, tc, td);
                if (tri_hit(as_f4(ta), as_f4(tb4), as_f4(tc), r, t, u, v)) {
                    cand = true; prim = td.x;
                    atomicMin(&ws.hit_tp[owner], hit_key(t, prim));
                }
            }
            __syncwarp();
            if (cand && ws.hit_tp[owner] == hit_key(t, prim)) { ws.hit_u[owner] = u; ws.hit_v[owner] = v; }
        }
        if (total <= (uint32_t)kItemCap) break;   // warp-uniform; otherwise the lanes still hold the rest in T
        __syncwarp();
    }
    __syncwarp();
}

template <int FLUSH>
__device__ __forceinline__ void flush_queue(const BvhView &bvh, WarpShared &ws, unsigned lane, uint32_t &qn, uint32_t k)
{
    if (FLUSH == 0) flush_queue_entries(bvh, ws, lane, qn, k);
    else flush_queue_balanced(bvh, ws, lane, qn, k);
}

__device__ __forceinline__ void sts64(uint32_t addr, uint2 v)
{
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ uint2 lds64(uint32_t addr)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
    return v;
}

template <class IO, int FETCH_THRESHOLD, int MIN_CTAS, int FLUSH = 1, bool DEEP = true, int SST = kSmemStack>
__global__ void __launch_bounds__(kTraceBlock, MIN_CTAS)
trace_wq(const BvhView bvh, const IO io, const uint32_t *__restrict__ n_ptr, uint32_t n_static,
         uint32_t *__restrict__ cursor, unsigned long long *__restrict__ live_acc)
{
    __shared__ uint2 sstack[SST * kTraceBlock];
    __shared__ uint8_t perm_lut[8 * 128];
    __shared__ WarpShared wsh[kTraceBlock / 32];
    __shared__ typename IO::PathState psh[kTraceBlock / 32];
    __shared__ uint32_t live_cta[IO::kContinue ? WPT_MAX_DEPTH : 1];
    for (uint32_t i = threadIdx.x; i < 8u * 128u; i += kTraceBlock) perm_lut[i] = (uint8_t)perm8(i & 127u, i >> 7);
    if (IO::kContinue && threadIdx.x < WPT_MAX_DEPTH) live_cta[threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t n = n_ptr ? *n_ptr : n_static;
    if (!IO::kContinue && live_acc && blockIdx.x == 0 && threadIdx.x == 0) *live_acc += n;
    const unsigned lane = threadIdx.x & 31u;
    WarpShared &ws = wsh[threadIdx.x >> 5];
    typename IO::PathState &ps = psh[threadIdx.x >> 5];
    // traversal stack: a shared-memory byte address that moves by one row (kTraceBlock entries) per
    // push / pop.  DEEP: entries beyond kSmemStack spill to local memory; the launcher picks the
    // non-DEEP instantiation (no overflow code at all) when max_depth - 1 <= SST
    constexpr uint32_t kRow = kTraceBlock * (uint32_t)sizeof(uint2);
    const uint32_t s_base = (uint32_t)__cvta_generic_to_shared(sstack + threadIdx.x);
    const uint32_t s_limit = s_base + SST * kRow;
    uint32_t s_top = s_base;
    uint2 lstack[DEEP ? kStackSize - SST : 1];

    // state: 0 = idle, 1 = traversing, 2 = parked (stack empty, waiting for the queue to drain)
    int state = 0;
    bool exhausted = false;
    uint32_t ray_i = 0, qn = 0;
    const uint8_t *lut = perm_lut;
    RayCtx r;
    uint2 G;
    float tmax_r = 0.0f;   // this lane's closest t so far: changes only inside flush_queue, re-read after every flush
    G.x = G.y = 0u;
    r = make_ray(0.f, 0.f, 0.f, 0.f, 0.f, 1.f, 0.f);

    // (re)arms this lane with a ray: record and empty closest hit to shared memory, traversal starts at the root
    auto arm = [&](float ox, float oy, float oz, float dx, float dy, float dz, float t0, float t1) {
        r = make_ray(ox, oy, oz, dx, dy, dz, t0);
        lut = perm_lut + r.oct_inv * 128u;
        ws.ray_a[lane] = make_float4(ox, oy, oz, dx);
        ws.ray_b[lane] = make_float4(dy, dz, t0, 0.0f);
        ws.hit_tp[lane] = hit_key(t1, 0xffffffffu);
        tmax_r = t1;
        G.x = 0u; G.y = 0x01000000u;
        s_top = s_base;
        state = 1;
    };

    for (;;) {
        unsigned act = __ballot_sync(0xffffffffu, state == 1);
        // while rays are left to fetch: refill once too few lanes traverse.  Afterwards only parked rays are left to
        // finish (or, in the fused tail, to continue with their next segment)
        const bool refill = __popc(act) < (FETCH_THRESHOLD > 0 ? FETCH_THRESHOLD : 1) &&
                            (!exhausted || act == 0u || __any_sync(0xffffffffu, state == 2));
        if (refill) {
            // drain the queue so that parked rays are final, finish them, then refill
            if (qn) {
                do flush_queue<FLUSH>(bvh, ws, lane, qn, qn < 32u ? qn : 32u); while (qn);
                WPT_CHECK(qn == 0u, 5);
                tmax_r = ws_tmax(ws, lane);
            }
            if (state == 2) {
                state = 0;
                const int32_t prim = ws_prim(ws, lane);
                const float t = ws_tmax(ws, lane), u = ws.hit_u[lane], v = ws.hit_v[lane];
                if constexpr (IO::kContinue) {
                    const float4 ra = ws.ray_a[lane], rb = ws.ray_b[lane];
                    float ox = ra.x, oy = ra.y, oz = ra.z, dx = ra.w, dy = rb.x, dz = rb.y;
                    if (io.next_segment(ps, lane, live_cta, ox, oy, oz, dx, dy, dz, prim, t, u, v))
                        arm(ox, oy, oz, dx, dy, dz, io.tmin, io.tmax);
                } else {
                    io.store(ray_i, prim, t, u, v);
                }
            }
            if (IO::kContinue) act = __ballot_sync(0xffffffffu, state == 1);
            if (!exhausted) {
                const unsigned idle = ~act;
                if (idle) {
                    const int cnt = __popc(idle);
                    const int leader = __ffs(idle) - 1;
                    uint32_t base = 0;
                    if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)cnt);
                    base = __shfl_sync(0xffffffffu, base, leader);
                    if (base + (uint32_t)cnt >= n) exhausted = true;
                    if (state == 0) {
                        const uint32_t i = base + (uint32_t)__popc(idle & ((1u << lane) - 1u));
                        float ox, oy, oz, dx, dy, dz, t0, t1;
                        WPT_CHECK(i < n || exhausted, 4);
                        if (i < n && io_load(io, i, ox, oy, oz, dx, dy, dz, t0, t1, ps, lane)) {
                            arm(ox, oy, oz, dx, dy, dz, t0, t1);
                            ray_i = i;
                        }
                    }
                }
                if (!__any_sync(0xffffffffu, state == 1)) {
                    if (exhausted) break;
                    continue;
                }
            } else if (!__any_sync(0xffffffffu, state == 1)) {
                break;
            }
        }
        uint32_t L = 0, tri0, cnt2;   // hit leaf slots of this step (tri0, cnt2 are defined where L != 0)
        if (state == 1) {
            // next internal child, front to back: highest pending priority bit (24..31)
            const uint32_t bit = bfind_u32(G.y);
            G.y ^= 1u << bit;
            if (G.y >> 24) {
                WPT_CHECK(DEEP ? (s_top < s_limit || (s_top - s_limit) / kRow < (uint32_t)(kStackSize - SST)) : s_top < s_limit, 3);
                if (!DEEP || s_top < s_limit) sts64(s_top, G); else lstack[(s_top - s_limit) / kRow] = G;
                s_top += kRow;
            }
            const uint32_t slot = (bit ^ r.oct_inv) & 7u;       // = (bit - 24) ^ oct_inv
            const uint32_t rel = (uint32_t)__popc(G.y & ((1u << slot) - 1u) & 0xffu);
            WPT_CHECK(G.x + rel < bvh.n_blocks, 2);
            const uint4 *np = bvh.blocks + (size_t)(G.x + rel) * 4;
            uint4 q0, q1, q2, q3;
            ld_half_m<0>(np, q0, q1); ld_half_m<0>(np + 2, q2, q3);
            // bits 0..6 = hit mask; the bits above are masked off by imask (<= 0x7f) / the 0x7f below
            const uint32_t h7 = node_hits7_raw(q0, q1, q2, q3, r, tmax_r);
            const uint32_t imask = q2.w >> 24;
            G.x = q0.w;
            G.y = ((uint32_t)lut[h7 & imask] << 24) | imask;
            L = h7 & ~imask & 0x7fu;
            cnt2 = __byte_perm(q3.y, q3.w, 0x3373);   // low half = the 2-bit triangle counts, high half unused
            tri0 = q0.w + (uint32_t)__popc(imask);
            if (!(G.y >> 24)) {
                if (s_top == s_base) {
                    state = 2;
                } else {
                    s_top -= kRow;
                    G = (!DEEP || s_top < s_limit) ? lds64(s_top) : lstack[(s_top - s_limit) / kRow];
                }
            }
        }
        // ---- append this step's hit leaves to the warp queue: ONE entry per lane (all its hit leaves)
        {
            const bool has = L != 0u;
            const unsigned b = __ballot_sync(0xffffffffu, has);
            if (b) {
                // qn < 32 here (every iteration ends with qn < 32), so at most 63 entries are ever queued
                WPT_CHECK(qn + (uint32_t)__popc(b) <= (uint32_t)kQueueCap, 0);
                if (has)
                    ws.queue[qn + (uint32_t)__popc(b & ((1u << lane) - 1u))] = make_uint2((lane << 27) | tri0, (L << 16) | (cnt2 & 0xffffu));
                qn += (uint32_t)__popc(b);
                if (qn >= 32u) {
                    flush_queue<FLUSH>(bvh, ws, lane, qn, 32u);
                    tmax_r = ws_tmax(ws, lane);
                }
            }
        }
    }
    if constexpr (IO::kContinue) {
        // rays per bounce of this CTA -> wpt_stats.live (rays entering bounce j)
        __syncthreads();
        if (live_acc) {
            if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(live_acc + io.bounce0, (unsigned long long)n);
            if (threadIdx.x < WPT_MAX_DEPTH && live_cta[threadIdx.x])
                atomicAdd(live_acc + threadIdx.x, (unsigned long long)live_cta[threadIdx.x]);
        }
    }
}

// one thread per ray, whole traversal in one go (comparison variant / small launches)
template <class IO>
__global__ void __launch_bounds__(kTraceBlock)
trace_simple(const BvhView bvh, const IO io, const uint32_t *__restrict__ n_ptr, uint32_t n_static,
             unsigned long long *__restrict__ live_acc)
{
    const uint32_t n = n_ptr ? *n_ptr : n_static;
    if (live_acc && blockIdx.x == 0 && threadIdx.x == 0) *live_acc += n;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float ox, oy, oz, dx, dy, dz, t0, t1;
        if (!io.load(i, ox, oy, oz, dx, dy, dz, t0, t1)) continue;
        const RayCtx r = make_ray(ox, oy, oz, dx, dy, dz, t0);
        LocalStack st;
        const Hit h = trace_ray(bvh, r, t1, st);
        io.store(i, h.prim, h.t, h.u, h.v);
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
// CTAs per SM of a persistent kernel: the occupancy calculator's answer, cached per (kernel instantiation, device) —
// instantiations differ in shared memory (stack rows) and so in occupancy, hence the key is the function pointer
template <class K>
static int ctas_per_sm(K kern, const TraceTuning &tune)
{
    if (tune.ctas_per_sm > 0) return tune.ctas_per_sm;
    struct Entry { const void *fn; int dev, ctas; };
    static thread_local Entry cache[64];
    static thread_local int n_cached = 0;
    int dev = 0;
    cudaGetDevice(&dev);
    const void *fn = reinterpret_cast<const void *>(kern);
    for (int i = 0; i < n_cached; i++)
        if (cache[i].fn == fn && cache[i].dev == dev) return cache[i].ctas;
    int c = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c, kern, kTraceBlock, 0) != cudaSuccess || c <= 0) c = 4;
    if (n_cached < 64) cache[n_cached++] = Entry{fn, dev, c};
    return c;
}

// the warp-queue kernel: 64 registers, 8 CTAs (32 warps) per SM; the stack lives in shared memory whenever the tree
// allows (a traversal needs at most max_depth - 1 entries: C1/C2 8 levels, C3 11, C4 13)
template <class IO, int THRESH, int MIN_CTAS, int FLUSH>
static void launch_wq(const BvhView &bvh, const IO &io, const uint32_t *n_ptr, uint32_t n_static, uint32_t *cursor,
                      unsigned long long *live_acc, int num_sms, const TraceTuning &tune, cudaStream_t st, int force_deep = -1)
{
    auto launch = [&](auto kern) {
        kern<<<(uint32_t)(num_sms * ctas_per_sm(kern, tune)), kTraceBlock, 0, st>>>(bvh, io, n_ptr, n_static, cursor, live_acc);
    };
    if (force_deep == 2) launch(trace_wq<IO, THRESH, MIN_CTAS, FLUSH, true, 2>);   // test: two shared rows, the rest in local memory
    else if (bvh.max_depth <= 11u) launch(trace_wq<IO, THRESH, MIN_CTAS, FLUSH, false, 10>);
    else if (bvh.max_depth <= 15u) launch(trace_wq<IO, THRESH, MIN_CTAS, FLUSH, false, 14>);
    else launch(trace_wq<IO, THRESH, MIN_CTAS, FLUSH, true, kSmemStack>);
}

// tune.variant: 0 default | 1 one thread per ray | 2 per-lane triangle loop | 12 default with a two-row shared stack
// (exercises the local-memory overflow) | 20 per-entry flush (the round-1 kernel) | 21..26 balanced flush with refill
// thresholds 20 / 24 / 28 / 16 / 18 / 22.  Variants 1 and 2 exist for plain ray sources only; every other source runs the default.
template <class IO>
static void launch_trace_any(const BvhView &bvh, const IO &io, const uint32_t *n_ptr, uint32_t n_static,
                             uint32_t *cursor, unsigned long long *live_acc, int num_sms,
                             const TraceTuning &tune, cudaStream_t st)
{
    if constexpr (!IO::kContinue) {
        if (tune.variant == 1) {
            trace_simple<IO><<<(uint32_t)num_sms * 16u, kTraceBlock, 0, st>>>(bvh, io, n_ptr, n_static, live_acc);
            return;
        }
        if (tune.variant == 2) {
            auto kern = trace_persistent<IO, 20>;
            kern<<<(uint32_t)(num_sms * ctas_per_sm(kern, tune)), kTraceBlock, 0, st>>>(bvh, io, n_ptr, n_static, cursor, live_acc);
            return;
        }
    }
    // 30..33: experiments on the seeded camera rays only (QueueIOT<1>); every other source runs the default
    if (tune.variant >= 30 && tune.variant <= 33) {
        if constexpr (std::is_same<IO, QueueIOT<1>>::value) {
            switch (tune.variant) {
            case 30: launch_wq<IO, 24, 8, 1>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); return;
            case 31: launch_wq<IO, 28, 8, 1>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); return;
            case 32: launch_wq<IO, 20, 8, 0>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); return;
            default: launch_wq<IO, 24, 8, 0>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); return;
            }
        } else {
            launch_wq<IO, kDefaultThresh, 8, kDefaultFlush>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st);
            return;
        }
    }
    switch (tune.variant) {
    case 12: launch_wq<IO, kDefaultThresh, 8, kDefaultFlush>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st, 2); break;
    case 20: launch_wq<IO, 20, 8, 0>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); break;
    case 21: launch_wq<IO, 20, 8, 1>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); break;
    case 22: launch_wq<IO, 24, 8, 1>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); break;
    case 23: launch_wq<IO, 28, 8, 1>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); break;
    case 24: launch_wq<IO, 16, 8, 1>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); break;
    case 25: launch_wq<IO, 18, 8, 1>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); break;
    case 26: launch_wq<IO, 22, 8, 1>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); break;
    default: launch_wq<IO, kDefaultThresh, 8, kDefaultFlush>(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); break;
    }
}

void launch_trace_queue(const BvhView &bvh, const RayQueue &q, int4 *hits, const uint32_t *n_ptr,
                        uint32_t *cursor, unsigned long long *live_acc, float tmin, float tmax,
                        uint32_t n_max, int num_sms, const TraceTuning &tune, cudaStream_t st,
                        int seeded, uint32_t *retry_list, uint32_t *retry_count, const uint32_t *indirect)
{
    if (indirect) {
        QueueIOT<2> io;
        io.q = q; io.hits = hits; io.tmin = tmin; io.tmax = tmax; io.retry_list = nullptr; io.retry_count = nullptr; io.indirect = indirect;
        launch_trace_any(bvh, io, n_ptr, n_max, cursor, live_acc, num_sms, tune, st);
    } else if (seeded) {
        QueueIOT<1> io;
        io.q = q; io.hits = hits; io.tmin = tmin; io.tmax = tmax; io.retry_list = retry_list; io.retry_count = retry_count; io.indirect = nullptr;
        launch_trace_any(bvh, io, n_ptr, n_max, cursor, live_acc, num_sms, tune, st);
    } else {
        QueueIO io;
        io.q = q; io.hits = hits; io.tmin = tmin; io.tmax = tmax; io.retry_list = nullptr; io.retry_count = nullptr; io.indirect = nullptr;
        launch_trace_any(bvh, io, n_ptr, n_max, cursor, live_acc, num_sms, tune, st);
    }
}

void launch_trace_streams(const BvhView &bvh, const StreamIO &s, uint32_t n, uint32_t *cursor, int num_sms,
                          const TraceTuning &tune, cudaStream_t st)
{
    StreamIODev io;
    io.s = s;
    launch_trace_any(bvh, io, nullptr, n, cursor, nullptr, num_sms, tune, st);
}

void launch_trace_tail(const BvhView &bvh, const RayQueue &q, const uint32_t *n_ptr, uint32_t *cursor,
                       unsigned long long *live_acc, const TailArgs &a, uint32_t n_max, int num_sms,
                       const TraceTuning &tune, cudaStream_t st)
{
    TailIO io;
    io.q = q; io.tmin = a.tmin; io.tmax = a.tmax; io.tri_attr = a.tri_attr; io.albedo = a.albedo;
    io.film_r = a.film_r; io.film_g = a.film_g; io.film_b = a.film_b; io.bounce0 = a.bounce0; io.depth = a.depth;
    io.spp_first = (uint32_t)a.spp_first; io.k0 = (uint32_t)a.seed; io.k1 = (uint32_t)(a.seed >> 32);
    launch_trace_any(bvh, io, n_ptr, n_max, cursor, live_acc, num_sms, tune, st);
}

}  // namespace wpt
