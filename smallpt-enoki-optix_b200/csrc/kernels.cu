// kernels.cu — hand-written sm_100a kernels of the wavefront path tracer:
//   trace_persistent  closest-hit BVH traversal (replaces optixLaunch, src/accel/optix_backend.h:460,
//                     and the OptiX programs src/wavefront_isect.cu:48-112)
//   raygen_kernel     camera rays (src/main.cpp:395-396 -> src/pinhole.h:27-56)
//   shade_kernel      fused miss-accumulate / attribute interpolation / frame / Lambert sample /
//                     next-ray write with block-aggregated warp-ballot compaction
//                     (src/main.cpp:407-424, src/accel/optix_backend.h:463-486)
// Nothing here is a dense contraction: no tensor cores.  The traversal is bound by the L1 data stage
// (one wavefront per lane of every divergent 256-bit load) together with instruction issue
// (profiles/r1_trace_wq_ncu_64spp.txt); the stream kernels are HBM bound.
#include "kernels.cuh"
#include "shade_core.h"
#include "trace_core.h"

namespace wpt {

// ------------------------------------------------------------------------------------------------
// ray sources / hit sinks
// ------------------------------------------------------------------------------------------------
// MODE 0: plain queue records, scalar tmax.  MODE 1 (depth-seeded camera rays, api.cu wpt_render): the ray's own tmax
// travels in b.w (bounce 0: the throughput there is 1 by construction); a ray that finds nothing inside its seed is
// appended to retry_list.  MODE 2: the second trace of those rays — indirect = that list, full tmax.
template <int MODE>
struct QueueIOT {
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
        t0 = s.tmin ? __ldcs(s.tmin + i) : s.tmin_s;
        t1 = s.tmax ? __ldcs(s.tmax + i) : s.tmax_s;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, int32_t prim, float t, float u, float v) const
    {
        s.tri_id[i] = prim;                 // -1 on miss (wavefront_isect.cu:70)
        if (prim >= 0) { s.t[i] = t; s.u[i] = u; s.v[i] = v; }
    }
};

// ------------------------------------------------------------------------------------------------
// persistent-thread traversal with dynamic ray fetch
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
// 32 lanes and took 40 % of all issued instructions.  Here a node step only APPENDS the triangles
// of the leaves it hit to a queue shared by the warp; whenever 32 are queued, every lane tests one
// of them — for whichever ray of the warp it belongs to — and folds the result into that ray's
// closest hit with a shared-memory atomicMin on the (positive) float bits of t.  Ray records and
// closest hits therefore live in shared memory.  A ray whose stack runs empty parks until the queue
// has been drained (before the next refill), then stores its hit.
constexpr int kQueueCap = 64;

struct WarpShared {
    float4   ray_a[32];      // ox, oy, oz, dx
    float4   ray_b[32];      // dy, dz, tmin, -
    uint32_t hit_t[32];      // float bits of the closest t so far (atomicMin target)
    float    hit_u[32], hit_v[32];
    int32_t  hit_prim[32];
    uint2    queue[kQueueCap];   // one entry per node step that hit leaves: x = owner lane << 27 | first
                                 // triangle block of the node, y = hit leaf slots << 16 | 2-bit triangle counts
};

// 256-bit block-half load with a selectable L1 hint.  Measured on C3 (profiles/r1_microbench.md): triangle blocks
// are hardly ever re-used by the same SM, so they are loaded L1::no_allocate (+0.8 %); node loads must allocate —
// the top three levels of the tree live in L1 (no_allocate on nodes: -20 %).
template <int M>
__device__ __forceinline__ void ld_half_m(const uint4 *p, uint4 &a, uint4 &b)
{
#define WPT_LD(S) asm(S " {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(p))
    if (M == 0) WPT_LD("ld.global.nc.v8.b32");
    if (M == 1) WPT_LD("ld.global.v8.b32");
    if (M == 2) WPT_LD("ld.global.nc.L1::no_allocate.v8.b32");
    if (M == 3) WPT_LD("ld.global.nc.L1::evict_last.v8.b32");
    if (M == 4) WPT_LD("ld.global.nc.L1::evict_first.v8.b32");
#undef WPT_LD
}

// every lane < k takes one queued entry (the leaves one node step hit, 1..3 triangles each), tests
// them against the owner's ray and folds the closest accepted hit into the owner's record
template <int LDT>
__device__ __forceinline__ void flush_queue(const BvhView &bvh, WarpShared &ws, unsigned lane, uint32_t &qn, uint32_t k)
{
    __syncwarp();
    bool cand = false;
    uint32_t owner = 0, prim = 0;
    float t = 0.0f, u = 0.0f, v = 0.0f;
    if (lane < k) {
        const uint2 e = ws.queue[qn - k + lane];
        owner = e.x >> 27;
        const uint32_t cnt2 = e.y & 0x3fffu;
        uint32_t L = e.y >> 16;
        const float4 ra = ws.ray_a[owner], rb = ws.ray_b[owner];
        RayCtx r;
        r.ox = ra.x; r.oy = ra.y; r.oz = ra.z; r.dx = ra.w; r.dy = rb.x; r.dz = rb.y; r.tmin = rb.z;
        t = __uint_as_float(ws.hit_t[owner]);
        // triangle bit mask of the hit leaves (bit i = i-th triangle of the node)
        uint32_t T = leaf_tri_mask(L, cnt2);
        const uint4 *t0p = bvh.blocks + (size_t)(e.x & 0x07ffffffu) * 4;
        while (T) {
            const int tb = 31 - __clz((int)T);
            T ^= 1u << tb;
            const uint4 *tp = t0p + (size_t)tb * 4;
            uint4 ta, tb4, tc, td;
            ld_half_m<LDT>(tp, ta, tb4); ld_half_m<LDT>(tp + 2, tc, td);
            if (tri_hit(as_f4(ta), as_f4(tb4), as_f4(tc), r, t, u, v)) { cand = true; prim = td.x; }
        }
        if (cand) atomicMin(&ws.hit_t[owner], __float_as_uint(t));
    }
    __syncwarp();
    if (cand && ws.hit_t[owner] == __float_as_uint(t)) { ws.hit_u[owner] = u; ws.hit_v[owner] = v; ws.hit_prim[owner] = (int32_t)prim; }
    __syncwarp();
    qn -= k;
}

__device__ __forceinline__ uint32_t bfind_u32(uint32_t x)   // index of the highest set bit (FLO.U32)
{
    uint32_t r;
    asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(x));
    return r;
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

template <class IO, int FETCH_THRESHOLD, int MIN_CTAS, bool PACKED = false, int LDN = 0, int LDT = 2, bool DEEP = true, int SST = kSmemStack>
__global__ void __launch_bounds__(kTraceBlock, MIN_CTAS)
trace_wq(const BvhView bvh, const IO io, const uint32_t *__restrict__ n_ptr, uint32_t n_static,
         uint32_t *__restrict__ cursor, unsigned long long *__restrict__ live_acc)
{
    __shared__ uint2 sstack[SST * kTraceBlock];
    __shared__ uint8_t perm_lut[8 * 128];
    __shared__ WarpShared wsh[kTraceBlock / 32];
    for (uint32_t i = threadIdx.x; i < 8u * 128u; i += kTraceBlock) perm_lut[i] = (uint8_t)perm8(i & 127u, i >> 7);
    __syncthreads();
    const uint32_t n = n_ptr ? *n_ptr : n_static;
    if (live_acc && blockIdx.x == 0 && threadIdx.x == 0) *live_acc += n;
    const unsigned lane = threadIdx.x & 31u;
    WarpShared &ws = wsh[threadIdx.x >> 5];
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

    for (;;) {
        const unsigned act = __ballot_sync(0xffffffffu, state == 1);
        if (__popc(act) < (FETCH_THRESHOLD > 0 ? FETCH_THRESHOLD : 1)) {
            // drain the queue so that parked rays are final, store them, then refill
            if (qn) {
                do flush_queue<LDT>(bvh, ws, lane, qn, qn < 32u ? qn : 32u); while (qn);
                tmax_r = __uint_as_float(ws.hit_t[lane]);
            }
            if (state == 2) {
                io.store(ray_i, ws.hit_prim[lane], __uint_as_float(ws.hit_t[lane]), ws.hit_u[lane], ws.hit_v[lane]);
                state = 0;
            }
            if (!exhausted) {
                const unsigned idle = ~act;
                const int cnt = __popc(idle);
                const int leader = __ffs(idle) - 1;
                uint32_t base = 0;
                if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)cnt);
                base = __shfl_sync(0xffffffffu, base, leader);
                if (base + (uint32_t)cnt >= n) exhausted = true;
                if (state == 0) {
                    const uint32_t i = base + (uint32_t)__popc(idle & ((1u << lane) - 1u));
                    float ox, oy, oz, dx, dy, dz, t0, t1;
                    if (i < n && io.load(i, ox, oy, oz, dx, dy, dz, t0, t1)) {
                        r = make_ray(ox, oy, oz, dx, dy, dz, t0);
                        lut = perm_lut + r.oct_inv * 128u;
                        ws.ray_a[lane] = make_float4(ox, oy, oz, dx);
                        ws.ray_b[lane] = make_float4(dy, dz, t0, 0.0f);
                        ws.hit_t[lane] = __float_as_uint(t1);
                        tmax_r = t1;
                        ws.hit_prim[lane] = -1;
                        G.x = 0u; G.y = 0x01000000u;
                        s_top = s_base;
                        ray_i = i;
                        state = 1;
                    }
                }
                if (!__any_sync(0xffffffffu, state == 1)) {
                    if (exhausted) break;
                    continue;
                }
            } else if (act == 0u) {
                break;
            }
        }
        uint32_t L = 0, tri0, cnt2;   // hit leaf slots of this step (tri0, cnt2 are defined where L != 0)
        if (state == 1) {
            // next internal child, front to back: highest pending priority bit (24..31)
            const uint32_t bit = bfind_u32(G.y);
            G.y ^= 1u << bit;
            if (G.y >> 24) {
                if (!DEEP || s_top < s_limit) sts64(s_top, G); else lstack[(s_top - s_limit) / kRow] = G;
                s_top += kRow;
            }
            const uint32_t slot = (bit ^ r.oct_inv) & 7u;       // = (bit - 24) ^ oct_inv
            const uint32_t rel = (uint32_t)__popc(G.y & ((1u << slot) - 1u) & 0xffu);
            const uint4 *np = bvh.blocks + (size_t)(G.x + rel) * 4;
            uint4 q0, q1, q2, q3;
            ld_half_m<LDN>(np, q0, q1); ld_half_m<LDN>(np + 2, q2, q3);
            // bits 0..6 = hit mask; the bits above are masked off by imask (<= 0x7f) / the 0x7f below
            const uint32_t h7 = PACKED ? node_hits7_p(q0, q1, q2, q3, r, tmax_r)
                                       : node_hits7_raw(q0, q1, q2, q3, r, tmax_r);
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
                if (has)
                    ws.queue[qn + (uint32_t)__popc(b & ((1u << lane) - 1u))] = make_uint2((lane << 27) | tri0, (L << 16) | (cnt2 & 0xffffu));
                qn += (uint32_t)__popc(b);
                if (qn >= 32u) {
                    flush_queue<LDT>(bvh, ws, lane, qn, 32u);
                    tmax_r = __uint_as_float(ws.hit_t[lane]);
                }
            }
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

// the default kernel: 64 registers, 8 CTAs (32 warps) per SM; the stack lives in shared memory whenever the tree
// allows (a traversal needs at most max_depth - 1 entries: C1/C2 8 levels, C3 11, C4 13)
template <class IO>
static void launch_trace_default(const BvhView &bvh, const IO &io, const uint32_t *n_ptr, uint32_t n_static,
                                 uint32_t *cursor, unsigned long long *live_acc, int num_sms,
                                 const TraceTuning &tune, cudaStream_t st)
{
    auto launch = [&](auto kern) {
        int occ = tune.ctas_per_sm;
        if (occ <= 0) {
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kTraceBlock, 0);
            if (occ <= 0) occ = 4;
        }
        kern<<<(uint32_t)(num_sms * occ), kTraceBlock, 0, st>>>(bvh, io, n_ptr, n_static, cursor, live_acc);
    };
    if (bvh.max_depth <= 11u) launch(trace_wq<IO, 20, 8, false, 0, 2, false, 10>);
    else if (bvh.max_depth <= 15u) launch(trace_wq<IO, 20, 8, false, 0, 2, false, 14>);
    else launch(trace_wq<IO, 20, 8>);
}

template <class IO>
static void launch_trace_any(const BvhView &bvh, const IO &io, const uint32_t *n_ptr, uint32_t n_static,
                             uint32_t *cursor, unsigned long long *live_acc, int num_sms,
                             const TraceTuning &tune, cudaStream_t st)
{
    if (tune.variant == 1) {
        const uint32_t blocks = (uint32_t)num_sms * 16u;
        trace_simple<IO><<<blocks, kTraceBlock, 0, st>>>(bvh, io, n_ptr, n_static, live_acc);
        return;
    }
    int per_sm = tune.ctas_per_sm;
    if (per_sm <= 0) {
        static int cached = 0;
        if (!cached) {
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cached, trace_persistent<IO, 20>, kTraceBlock, 0);
            if (cached <= 0) cached = 4;
        }
        per_sm = cached;
    }
    if (tune.variant == 2) {
        const uint32_t blocks = (uint32_t)(num_sms * per_sm);
        trace_persistent<IO, 20><<<blocks, kTraceBlock, 0, st>>>(bvh, io, n_ptr, n_static, cursor, live_acc);
        return;
    }
    auto launch = [&](auto kern) {
        int occ = tune.ctas_per_sm;
        if (occ <= 0) {
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kTraceBlock, 0);
            if (occ <= 0) occ = 4;
        }
        kern<<<(uint32_t)(num_sms * occ), kTraceBlock, 0, st>>>(bvh, io, n_ptr, n_static, cursor, live_acc);
    };
    switch (tune.variant) {
    case 3: launch(trace_wq<IO, 26, 6>); break;
    case 4: launch(trace_wq<IO, 14, 6>); break;
    case 5: launch(trace_wq<IO, 20, 8>); break;
    case 6: launch(trace_wq<IO, 24, 8>); break;
    case 7: launch(trace_wq<IO, 20, 7>); break;
    case 8: launch(trace_wq<IO, 20, 6>); break;
    case 9: launch(trace_wq<IO, 20, 7, true>); break;
    case 10: launch(trace_wq<IO, 20, 6, true>); break;
    case 11: launch(trace_wq<IO, 20, 7, false, 0, 0>); break;   // triangle loads allocate in L1 (the old default)
    case 12: launch(trace_wq<IO, 20, 8, false, 0, 2, true, 2>); break;   // test variant: two shared-memory stack rows, the rest overflows to local memory

    default: launch_trace_default(bvh, io, n_ptr, n_static, cursor, live_acc, num_sms, tune, st); break;
    }
}

void launch_trace_queue(const BvhView &bvh, const RayQueue &q, int4 *hits, const uint32_t *n_ptr,
                        uint32_t *cursor, unsigned long long *live_acc, float tmin, float tmax,
                        uint32_t n_max, int num_sms, const TraceTuning &tune, cudaStream_t st,
                        int seeded, uint32_t *retry_list, uint32_t *retry_count, const uint32_t *indirect)
{
    // the seeded first trace and its second trace always run the default kernel (tune.variant applies to plain launches)
    if (indirect) {
        QueueIOT<2> io;
        io.q = q; io.hits = hits; io.tmin = tmin; io.tmax = tmax; io.retry_list = nullptr; io.retry_count = nullptr; io.indirect = indirect;
        launch_trace_default(bvh, io, n_ptr, n_max, cursor, live_acc, num_sms, tune, st);
    } else if (seeded) {
        QueueIOT<1> io;
        io.q = q; io.hits = hits; io.tmin = tmin; io.tmax = tmax; io.retry_list = retry_list; io.retry_count = retry_count; io.indirect = nullptr;
        launch_trace_default(bvh, io, n_ptr, n_max, cursor, live_acc, num_sms, tune, st);
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

// ------------------------------------------------------------------------------------------------
// camera rays
// ------------------------------------------------------------------------------------------------
// Lane order inside a frame: 8x4 pixel tiles per warp when the resolution allows it (more coherent
// primary rays than scanline rows); the pixel index travels in the path id, so order is free.
__device__ __forceinline__ void lane_to_pixel(uint32_t p, int W, int H, int &x, int &y)
{
    if (((W & 7) | (H & 3)) == 0) {
        const uint32_t tile = p >> 5, in = p & 31u, tw = (uint32_t)W >> 3;
        x = (int)((tile % tw) * 8u + (in & 7u));
        y = (int)((tile / tw) * 4u + (in >> 3));
    } else {
        x = (int)(p % (uint32_t)W);
        y = (int)(p / (uint32_t)W);
    }
}

// seed != nullptr: b.w carries the ray's tmax seed (seed[pixel]) instead of the constant throughput 1 — the trace and
// shade kernels of bounce 0 are told (QueueIO::seeded, shade_kernel's seeded0).  centre != 0: one ray per pixel
// through the pixel centre and the lens centre (the depth pre-pass; not a sample of the estimator).
__global__ void __launch_bounds__(256)
raygen_kernel(const wpt_camera cam, int W, int H, int spp_first, int spp_in_pass, uint32_t k0, uint32_t k1,
              RayQueue q, uint32_t *__restrict__ qcount0, const float *__restrict__ seed, int centre)
{
    const uint32_t N = (uint32_t)W * (uint32_t)H;
    const uint32_t total = N * (uint32_t)spp_in_pass;
    if (blockIdx.x == 0 && threadIdx.x == 0) *qcount0 = total;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const uint32_t s_local = i / N, p = i - s_local * N;
        int x, y;
        lane_to_pixel(p, W, H, x, y);
        const uint32_t pixel = (uint32_t)y * (uint32_t)W + (uint32_t)x;   // src/main.cpp:379-381
        uint32_t rnd[4];
        philox4x32_10(pixel, (uint32_t)spp_first + s_local, 0u, 0u, k0, k1, rnd);
        float xi[4] = { u32_to_unit(rnd[0]), u32_to_unit(rnd[1]), u32_to_unit(rnd[2]), u32_to_unit(rnd[3]) };
        if (centre) xi[0] = xi[1] = xi[2] = xi[3] = 0.5f;
        V3 o, d;
        camera_ray(cam, x, y, W, H, xi, o, d);
        __stcs(q.a + i, make_float4(o.x, o.y, o.z, __uint_as_float(pixel | (s_local << 24))));
        __stcs(q.b + i, make_float4(d.x, d.y, d.z, seed ? __ldg(seed + pixel) : 1.0f));   // contrib = 1 (src/main.cpp:391)
        __stcs(q.c + i, make_float2(1.0f, 1.0f));
    }
}

void launch_raygen(const wpt_camera &cam, int W, int H, int spp_first, int spp_in_pass, uint64_t seed,
                   const RayQueue &q, uint32_t *qcount0, cudaStream_t st, const float *tmax_seed, int centre)
{
    const uint64_t total = (uint64_t)W * H * spp_in_pass;
    const uint32_t blocks = (uint32_t)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    raygen_kernel<<<blocks ? blocks : 1, 256, 0, st>>>(cam, W, H, spp_first, spp_in_pass, (uint32_t)seed,
                                                       (uint32_t)(seed >> 32), q, qcount0, tmax_seed, centre);
}

// ------------------------------------------------------------------------------------------------
// depth pre-pass -> per-pixel tmax seeds for the camera rays of a frame
// ------------------------------------------------------------------------------------------------
// depth[pixel] = hit distance of the pixel-centre ray (inf on a miss); the queue is in 8x4-tile order, the pixel
// travels in a.w
__global__ void __launch_bounds__(256)
depth_scatter_kernel(const float4 *__restrict__ qa, const int4 *__restrict__ hits, uint32_t n, float *__restrict__ depth)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int4 h = hits[i];
        depth[__float_as_uint(qa[i].w) & 0x00ffffffu] = h.x >= 0 ? __int_as_float(h.y) : __int_as_float(0x7f800000);
    }
}
// seed[pixel] = margin * max of the 3x3 neighbourhood of depth (the jittered rays of a pixel lie inside the hull
// of those nine centres); no seed (1e20) where any of them missed.  A wrong seed costs a second trace, never a result.
__global__ void __launch_bounds__(256)
depth_seed_kernel(const float *__restrict__ depth, int W, int H, float margin, float *__restrict__ seed)
{
    const uint32_t N = (uint32_t)W * (uint32_t)H;
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < N; p += gridDim.x * blockDim.x) {
        const int y = (int)(p / (uint32_t)W), x = (int)(p % (uint32_t)W);
        float m = 0.0f;
        for (int dy = -1; dy <= 1; dy++)
            for (int dx = -1; dx <= 1; dx++) {
                const int xx = min(max(x + dx, 0), W - 1), yy = min(max(y + dy, 0), H - 1);
                m = fmaxf(m, depth[(uint32_t)yy * (uint32_t)W + (uint32_t)xx]);
            }
        seed[p] = m < 1e19f ? m * margin : 1e20f;
    }
}
void launch_depth_seeds(const float4 *qa, const int4 *hits, int W, int H, float margin, float *depth, float *seed,
                        cudaStream_t st)
{
    const uint32_t n = (uint32_t)W * (uint32_t)H, blocks = (n + 255) / 256 < 148u * 16u ? (n + 255) / 256 : 148u * 16u;
    depth_scatter_kernel<<<blocks, 256, 0, st>>>(qa, hits, n, depth);
    depth_seed_kernel<<<blocks, 256, 0, st>>>(depth, W, H, margin, seed);
}

__global__ void __launch_bounds__(256)
raygen_streams_kernel(const wpt_camera cam, int W, int H, int spp, uint32_t k0, uint32_t k1, float *ox,
                      float *oy, float *oz, float *dx, float *dy, float *dz)
{
    const uint32_t N = (uint32_t)W * (uint32_t)H;
    for (uint32_t pixel = blockIdx.x * blockDim.x + threadIdx.x; pixel < N; pixel += gridDim.x * blockDim.x) {
        const int y = (int)(pixel / (uint32_t)W), x = (int)(pixel % (uint32_t)W);
        uint32_t rnd[4];
        philox4x32_10(pixel, (uint32_t)spp, 0u, 0u, k0, k1, rnd);
        const float xi[4] = { u32_to_unit(rnd[0]), u32_to_unit(rnd[1]), u32_to_unit(rnd[2]), u32_to_unit(rnd[3]) };
        V3 o, d;
        camera_ray(cam, x, y, W, H, xi, o, d);
        ox[pixel] = o.x; oy[pixel] = o.y; oz[pixel] = o.z;
        dx[pixel] = d.x; dy[pixel] = d.y; dz[pixel] = d.z;
    }
}

void launch_raygen_streams(const wpt_camera &cam, int W, int H, int spp_index, uint64_t seed, float *ox,
                           float *oy, float *oz, float *dx, float *dy, float *dz, cudaStream_t st)
{
    const uint32_t N = (uint32_t)W * H;
    const uint32_t blocks = (N + 255) / 256;
    raygen_streams_kernel<<<blocks ? blocks : 1, 256, 0, st>>>(cam, W, H, spp_index, (uint32_t)seed,
                                                               (uint32_t)(seed >> 32), ox, oy, oz, dx, dy, dz);
}

// ------------------------------------------------------------------------------------------------
// fused shade + sample + accumulate + compact
// ------------------------------------------------------------------------------------------------
constexpr int kShadeBlock = 256;

__global__ void __launch_bounds__(kShadeBlock)
shade_kernel(const RayQueue in, const int4 *__restrict__ hits, const uint32_t *__restrict__ n_ptr,
             const RayQueue out, uint32_t *__restrict__ n_out, const float4 *__restrict__ tri_attr,
             const float *__restrict__ albedo, int bounce, int depth, int spp_first, uint32_t k0, uint32_t k1,
             float *__restrict__ film_r, float *__restrict__ film_g, float *__restrict__ film_b, int seeded0)
{
    __shared__ uint32_t warp_cnt[kShadeBlock / 32];
    __shared__ uint32_t block_base;
    const uint32_t n = *n_ptr;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const bool last = bounce + 1 >= depth;   // path is cut after `depth` segments (src/main.cpp:399)
    const uint32_t stride = gridDim.x * blockDim.x;
    // all threads of a block run the same number of iterations (block-uniform loop bound)
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += stride) {
        const uint32_t i = base + threadIdx.x;
        bool survive = false;
        float4 na, nb;
        float2 nc;
        if (i < n) {
            const int4 h = __ldcs(hits + i);
            const float4 a = __ldcs(in.a + i);
            float4 b = __ldcs(in.b + i);
            if (seeded0) b.w = 1.0f;   // bounce 0 of a seeded pass: b.w held the ray's tmax seed, the throughput is 1
            const float2 c = __ldcs(in.c + i);
            const uint32_t pid = __float_as_uint(a.w);
            const uint32_t pixel = pid & 0x00ffffffu;
            if (h.x < 0) {
                // escaped: film += contrib (src/main.cpp:407); fire-and-forget RED.ADD.F32
                atomicAdd(film_r + pixel, b.w);
                atomicAdd(film_g + pixel, c.x);
                atomicAdd(film_b + pixel, c.y);
            } else if (!last) {
                const float4 *ap = tri_attr + (size_t)h.x * 3;
                const float4 q0 = __ldg(ap), q1 = __ldg(ap + 1), q2 = __ldg(ap + 2);
                const int mat = __float_as_int(q2.y);
                uint32_t rnd[4];
                philox4x32_10(pixel, (uint32_t)spp_first + (pid >> 24), 1u + (uint32_t)bounce, 0u, k0, k1, rnd);
                V3 no, nd;
                shade_step(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), __int_as_float(h.y), __int_as_float(h.z),
                           __int_as_float(h.w), v3(q0.x, q0.y, q0.z), v3(q0.w, q1.x, q1.y), v3(q1.z, q1.w, q2.x),
                           u32_to_unit(rnd[0]), u32_to_unit(rnd[1]), no, nd);
                const float wr = __ldg(albedo + 3 * mat), wg = __ldg(albedo + 3 * mat + 1),
                            wb = __ldg(albedo + 3 * mat + 2);
                na = make_float4(no.x, no.y, no.z, a.w);
                nb = make_float4(nd.x, nd.y, nd.z, b.w * wr);     // contrib *= bsdf weight (src/main.cpp:422)
                nc = make_float2(c.x * wg, c.y * wb);
                survive = true;
            }
        }
        if (last) continue;   // block-uniform
        // ---- compaction: warp ballot -> per-block offsets -> one atomic per block on the queue tail
        const unsigned bal = __ballot_sync(0xffffffffu, survive);
        if (lane == 0) warp_cnt[warp] = (uint32_t)__popc(bal);
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t tot = 0;
#pragma unroll
            for (int w = 0; w < kShadeBlock / 32; w++) { const uint32_t c = warp_cnt[w]; warp_cnt[w] = tot; tot += c; }
            block_base = tot ? atomicAdd(n_out, tot) : 0u;
        }
        __syncthreads();
        if (survive) {
            const uint32_t o = block_base + warp_cnt[warp] + (uint32_t)__popc(bal & ((1u << lane) - 1u));
            __stcs(out.a + o, na);
            __stcs(out.b + o, nb);
            __stcs(out.c + o, nc);
        }
        __syncthreads();   // warp_cnt / block_base are reused by the next iteration
    }
}

void launch_shade(const RayQueue &in, const int4 *hits, const uint32_t *n_ptr, const RayQueue &out,
                  uint32_t *n_out, const float4 *tri_attr, const float *albedo, int bounce, int depth,
                  int spp_first, uint64_t seed, float *film_r, float *film_g, float *film_b, uint32_t n_max,
                  int num_sms, cudaStream_t st, int seeded0)
{
    uint32_t blocks = (n_max + kShadeBlock - 1) / kShadeBlock;
    const uint32_t cap = (uint32_t)num_sms * 8u;
    if (blocks > cap) blocks = cap;
    if (blocks == 0) blocks = 1;
    shade_kernel<<<blocks, kShadeBlock, 0, st>>>(in, hits, n_ptr, out, n_out, tri_attr, albedo, bounce, depth,
                                                 spp_first, (uint32_t)seed, (uint32_t)(seed >> 32), film_r,
                                                 film_g, film_b, seeded0);
}

// per-function entry point: one shading step on SoA streams (parity tests)
struct ShadeStreams {
    const float *o[3], *d[3];
    const int32_t *tri_id;
    const float *t, *u, *v, *s0, *s1;
    float *no[3], *nd[3], *w[3];
};
__global__ void shade_streams_kernel(uint32_t n, ShadeStreams p, const float4 *__restrict__ tri_attr,
                                     const float *__restrict__ albedo)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t tri = p.tri_id[i];
    if (tri < 0) return;
    const float4 *ap = tri_attr + (size_t)tri * 3;
    const float4 q0 = __ldg(ap), q1 = __ldg(ap + 1), q2 = __ldg(ap + 2);
    const int mat = __float_as_int(q2.y);
    V3 no, nd;
    shade_step(v3(p.o[0][i], p.o[1][i], p.o[2][i]), v3(p.d[0][i], p.d[1][i], p.d[2][i]), p.t[i], p.u[i], p.v[i],
               v3(q0.x, q0.y, q0.z), v3(q0.w, q1.x, q1.y), v3(q1.z, q1.w, q2.x), p.s0[i], p.s1[i], no, nd);
    p.no[0][i] = no.x; p.no[1][i] = no.y; p.no[2][i] = no.z;
    p.nd[0][i] = nd.x; p.nd[1][i] = nd.y; p.nd[2][i] = nd.z;
    p.w[0][i] = albedo[3 * mat]; p.w[1][i] = albedo[3 * mat + 1]; p.w[2][i] = albedo[3 * mat + 2];
}
void launch_shade_streams(uint32_t n, const float *const o[3], const float *const d[3], const int32_t *tri_id,
                          const float *t, const float *u, const float *v, const float *s0, const float *s1,
                          const float4 *tri_attr, const float *albedo, float *const no[3], float *const nd[3],
                          float *const w[3], cudaStream_t st)
{
    ShadeStreams p;
    for (int k = 0; k < 3; k++) { p.o[k] = o[k]; p.d[k] = d[k]; p.no[k] = no[k]; p.nd[k] = nd[k]; p.w[k] = w[k]; }
    p.tri_id = tri_id; p.t = t; p.u = u; p.v = v; p.s0 = s0; p.s1 = s1;
    if (n) shade_streams_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, p, tri_attr, albedo);
}

// ------------------------------------------------------------------------------------------------
__global__ void film_scale_kernel(float *r, float *g, float *b, uint32_t n, float scale)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        r[i] *= scale; g[i] *= scale; b[i] *= scale;     // film /= num_samples (src/main.cpp:429)
    }
}
void launch_film_scale(float *r, float *g, float *b, uint32_t n, float scale, cudaStream_t st)
{
    uint32_t blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (n) film_scale_kernel<<<blocks, 256, 0, st>>>(r, g, b, n, scale);
}

__global__ void philox_kernel(uint32_t n, uint32_t c1, uint32_t c2, uint32_t k0, uint32_t k1, uint32_t *out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t r[4];
    philox4x32_10(i, c1, c2, 0u, k0, k1, r);
    out[4 * i] = r[0]; out[4 * i + 1] = r[1]; out[4 * i + 2] = r[2]; out[4 * i + 3] = r[3];
}
void launch_philox(uint32_t n, uint32_t c1, uint32_t c2, uint64_t key, uint32_t *out, cudaStream_t st)
{
    if (n) philox_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, c1, c2, (uint32_t)key, (uint32_t)(key >> 32), out);
}

}  // namespace wpt
