// kernels.cu — hand-written sm_100a stream kernels of the wavefront path tracer (the traversal kernels live in
// trace_kernels.cu):
//   raygen_kernel     camera rays (src/main.cpp:395-396 -> src/pinhole.h:27-56)
//   shade_kernel      fused miss-accumulate / attribute interpolation / frame / Lambert sample /
//                     next-ray write with block-aggregated warp-ballot compaction
//                     (src/main.cpp:407-424, src/accel/optix_backend.h:463-486)
//   hit_attributes_kernel   the derived hit attributes OptixBackend::intersect returns with a hit
//                     (src/accel/optix_backend.h:463-486)
// All HBM bound: coalesced SoA streams, one 48-byte gather per hit.
#include "kernels.cuh"
#include "shade_core.h"
#include "trace_core.h"
#include <cstdlib>

namespace wpt {
constexpr int kDefaultShadeBlocks = 6;   // measured (profiles/r2_grid_c3.txt): 6 blocks at 40 registers 7.2 ms of shade per C3 frame, 8 at 32 registers 7.5 ms; a grid of 8 per SM on 6 resident (round 1) 8.05 ms

// SM count of the current device (grid sizing of the grid-stride stream kernels), cached per device
static int device_sms()
{
    static int cached[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    int &c = cached[dev & 63];
    if (c <= 0) {
        if (cudaDeviceGetAttribute(&c, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || c <= 0) c = 148;
    }
    return c;
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
raygen_kernel(const wpt_camera cam, const CameraFrame cf, int W, int H, int spp_first, int spp_in_pass, uint32_t k0, uint32_t k1,
              RayQueue q, uint32_t *__restrict__ qcount0, const float *__restrict__ seed, int centre,
              uint32_t lane_first, uint32_t lane_count)
{
    // lanes [lane_first, lane_first + lane_count) of every sample: the whole frame, or one rank's share of the depth pre-pass.
    // A thread owns a lane and walks its samples: the pixel decode and the seed load are paid once per lane, and the
    // warp's stores of one sample are three coalesced runs at sample * N + lane, as before.
    const uint32_t N = lane_count;
    if (blockIdx.x == 0 && threadIdx.x == 0) *qcount0 = N * (uint32_t)spp_in_pass;
    for (uint32_t l = blockIdx.x * blockDim.x + threadIdx.x; l < N; l += gridDim.x * blockDim.x) {
        int x, y;
        lane_to_pixel(lane_first + l, W, H, x, y);
        const uint32_t pixel = (uint32_t)y * (uint32_t)W + (uint32_t)x;   // src/main.cpp:379-381
        const float bw = seed ? __ldg(seed + pixel) : 1.0f;                // contrib = 1 (src/main.cpp:391)
        for (uint32_t s_local = 0; s_local < (uint32_t)spp_in_pass; s_local++) {
            uint32_t rnd[4];
            philox4x32_10(pixel, (uint32_t)spp_first + s_local, 0u, 0u, k0, k1, rnd);
            float xi[4] = { u32_to_unit(rnd[0]), u32_to_unit(rnd[1]), u32_to_unit(rnd[2]), u32_to_unit(rnd[3]) };
            if (centre) xi[0] = xi[1] = xi[2] = xi[3] = 0.5f;
            V3 o, d;
            camera_ray(cam, cf, x, y, W, H, xi, o, d);
            const uint32_t i = s_local * N + l;
            __stcs(q.a + i, make_float4(o.x, o.y, o.z, __uint_as_float(pixel | (s_local << 24))));
            __stcs(q.b + i, make_float4(d.x, d.y, d.z, bw));
            __stcs(q.c + i, make_float2(1.0f, 1.0f));
        }
    }
}

void launch_raygen(const wpt_camera &cam, int W, int H, int spp_first, int spp_in_pass, uint64_t seed,
                   const RayQueue &q, uint32_t *qcount0, cudaStream_t st, const float *tmax_seed, int centre,
                   uint32_t lane_first, uint32_t lane_count)
{
    if (lane_count == 0) { lane_first = 0; lane_count = (uint32_t)W * (uint32_t)H; }
    const uint64_t cap = (uint64_t)device_sms() * 16u;
    const uint64_t want = ((uint64_t)lane_count + 255) / 256;
    const uint32_t blocks = (uint32_t)(want < cap ? want : cap);
    raygen_kernel<<<blocks ? blocks : 1, 256, 0, st>>>(cam, camera_frame_constants(cam, W, H), W, H, spp_first, spp_in_pass,
                                                       (uint32_t)seed, (uint32_t)(seed >> 32), q, qcount0, tmax_seed, centre,
                                                       lane_first, lane_count);
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
// sharded depth pre-pass (wpt_renderer_set_depth_shard): every rank traces the pixel-centre rays of its own lane window
// and writes their hit distances, in lane order, into its chunk of one buffer; after the all-gather every rank scatters
// the whole buffer to pixels and derives the same seeds
__global__ void __launch_bounds__(256)
depth_pack_kernel(const int4 *__restrict__ hits, uint32_t n, float *__restrict__ tlane)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int4 h = hits[i];
        tlane[i] = h.x >= 0 ? __int_as_float(h.y) : __int_as_float(0x7f800000);
    }
}
__global__ void __launch_bounds__(256)
depth_scatter_lanes_kernel(const float *__restrict__ tlane, int W, int H, float *__restrict__ depth)
{
    const uint32_t n = (uint32_t)W * (uint32_t)H;
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        int x, y;
        lane_to_pixel(p, W, H, x, y);
        depth[(uint32_t)y * (uint32_t)W + (uint32_t)x] = tlane[p];
    }
}
void launch_depth_pack(const int4 *hits, uint32_t n, float *tlane_chunk, cudaStream_t st)
{
    const uint32_t cap = (uint32_t)device_sms() * 16u, blocks = (n + 255) / 256 < cap ? (n + 255) / 256 : cap;
    if (n) depth_pack_kernel<<<blocks, 256, 0, st>>>(hits, n, tlane_chunk);
}
void launch_depth_seeds_from_lanes(const float *tlane, int W, int H, float margin, float *depth, float *seed, cudaStream_t st)
{
    const uint32_t n = (uint32_t)W * (uint32_t)H, cap = (uint32_t)device_sms() * 16u;
    const uint32_t blocks = (n + 255) / 256 < cap ? (n + 255) / 256 : cap;
    depth_scatter_lanes_kernel<<<blocks, 256, 0, st>>>(tlane, W, H, depth);
    depth_seed_kernel<<<blocks, 256, 0, st>>>(depth, W, H, margin, seed);
}

void launch_depth_seeds(const float4 *qa, const int4 *hits, int W, int H, float margin, float *depth, float *seed,
                        cudaStream_t st)
{
    const uint32_t n = (uint32_t)W * (uint32_t)H, cap = (uint32_t)device_sms() * 16u;
    const uint32_t blocks = (n + 255) / 256 < cap ? (n + 255) / 256 : cap;
    depth_scatter_kernel<<<blocks, 256, 0, st>>>(qa, hits, n, depth);
    depth_seed_kernel<<<blocks, 256, 0, st>>>(depth, W, H, margin, seed);
}

__global__ void __launch_bounds__(256)
raygen_streams_kernel(const wpt_camera cam, const CameraFrame cf, int W, int H, int spp, uint32_t k0, uint32_t k1, float *ox,
                      float *oy, float *oz, float *dx, float *dy, float *dz)
{
    const uint32_t N = (uint32_t)W * (uint32_t)H;
    for (uint32_t pixel = blockIdx.x * blockDim.x + threadIdx.x; pixel < N; pixel += gridDim.x * blockDim.x) {
        const int y = (int)(pixel / (uint32_t)W), x = (int)(pixel % (uint32_t)W);
        uint32_t rnd[4];
        philox4x32_10(pixel, (uint32_t)spp, 0u, 0u, k0, k1, rnd);
        const float xi[4] = { u32_to_unit(rnd[0]), u32_to_unit(rnd[1]), u32_to_unit(rnd[2]), u32_to_unit(rnd[3]) };
        V3 o, d;
        camera_ray(cam, cf, x, y, W, H, xi, o, d);
        ox[pixel] = o.x; oy[pixel] = o.y; oz[pixel] = o.z;
        dx[pixel] = d.x; dy[pixel] = d.y; dz[pixel] = d.z;
    }
}

void launch_raygen_streams(const wpt_camera &cam, int W, int H, int spp_index, uint64_t seed, float *ox,
                           float *oy, float *oz, float *dx, float *dy, float *dz, cudaStream_t st)
{
    const uint32_t N = (uint32_t)W * H;
    const uint32_t blocks = (N + 255) / 256;
    raygen_streams_kernel<<<blocks ? blocks : 1, 256, 0, st>>>(cam, camera_frame_constants(cam, W, H), W, H, spp_index, (uint32_t)seed,
                                                               (uint32_t)(seed >> 32), ox, oy, oz, dx, dy, dz);
}

// ------------------------------------------------------------------------------------------------
// fused shade + sample + accumulate + compact
// ------------------------------------------------------------------------------------------------
constexpr int kShadeBlock = 256;

// TEX: some material of the scene carries an image texture (wpt_scene_set_texture): the bsdf weight of such a material
// is ImageTexture::eval at the interpolated texcoord (src/main.cpp:62-76,116) instead of the constant albedo
template <bool TEX, int MINB>
__global__ void __launch_bounds__(kShadeBlock, MINB)
shade_kernel(const RayQueue in, const int4 *__restrict__ hits, const uint32_t *__restrict__ n_ptr,
             const RayQueue out, uint32_t *__restrict__ n_out, const float4 *__restrict__ tri_attr,
             const float *__restrict__ albedo, int bounce, int depth, int spp_first, uint32_t k0, uint32_t k1,
             float *__restrict__ film_r, float *__restrict__ film_g, float *__restrict__ film_b, int seeded0,
             const TexArgs tex)
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
                float wr = __ldg(albedo + 3 * mat), wg = __ldg(albedo + 3 * mat + 1), wb = __ldg(albedo + 3 * mat + 2);
                if (TEX) {
                    const MatTex mt = tex.mattex[mat];
                    if (mt.w * mt.h > 1u) {
                        const float2 *uv = tex.triuv + (size_t)h.x * 3;
                        float tu, tv, rgb[3];
                        interp_uv(__ldg(uv), __ldg(uv + 1), __ldg(uv + 2), __int_as_float(h.z), __int_as_float(h.w), tu, tv);
                        texture_eval(tex.texels, mt, tu, tv, rgb);
                        wr = rgb[0]; wg = rgb[1]; wb = rgb[2];
                    }
                }
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
                  int num_sms, cudaStream_t st, int seeded0, const TexArgs *tex)
{
    // resident blocks per SM: 6 at the compiler's 40 registers, 8 when held to 32 (16 bytes of spill); WPT_SHADE_BLOCKS
    // selects (experiments), default = measured winner (profiles/r2_stream_kernels_ncu.txt)
    static const int minb = [] { const char *e = getenv("WPT_SHADE_BLOCKS"); return e ? atoi(e) : kDefaultShadeBlocks; }();
    uint32_t blocks = (n_max + kShadeBlock - 1) / kShadeBlock;
    const uint32_t cap = (uint32_t)num_sms * (uint32_t)(minb >= 8 ? 8 : 6);
    if (blocks > cap) blocks = cap;
    if (blocks == 0) blocks = 1;
    const TexArgs none{nullptr, nullptr, nullptr};
#define WPT_SHADE_ARGS in, hits, n_ptr, out, n_out, tri_attr, albedo, bounce, depth, spp_first, (uint32_t)seed, (uint32_t)(seed >> 32), film_r, film_g, film_b, seeded0
    if (tex && tex->texels) shade_kernel<true, 4><<<blocks, kShadeBlock, 0, st>>>(WPT_SHADE_ARGS, *tex);
    else if (minb >= 8) shade_kernel<false, 8><<<blocks, kShadeBlock, 0, st>>>(WPT_SHADE_ARGS, none);
    else shade_kernel<false, 6><<<blocks, kShadeBlock, 0, st>>>(WPT_SHADE_ARGS, none);
#undef WPT_SHADE_ARGS
}

// per-function entry point: one shading step on SoA streams (parity tests)
struct ShadeStreams {
    const float *o[3], *d[3];
    const int32_t *tri_id;
    const float *t, *u, *v, *s0, *s1;
    float *no[3], *nd[3], *w[3];
};
__global__ void shade_streams_kernel(uint32_t n, ShadeStreams p, const float4 *__restrict__ tri_attr,
                                     const float *__restrict__ albedo, const TexArgs tex)
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
    float w[3] = {albedo[3 * mat], albedo[3 * mat + 1], albedo[3 * mat + 2]};
    if (tex.texels) {
        const MatTex mt = tex.mattex[mat];
        if (mt.w * mt.h > 1u) {
            const float2 *uv = tex.triuv + (size_t)tri * 3;
            float tu, tv;
            interp_uv(uv[0], uv[1], uv[2], p.u[i], p.v[i], tu, tv);
            texture_eval(tex.texels, mt, tu, tv, w);
        }
    }
    p.w[0][i] = w[0]; p.w[1][i] = w[1]; p.w[2][i] = w[2];
}
void launch_shade_streams(uint32_t n, const float *const o[3], const float *const d[3], const int32_t *tri_id,
                          const float *t, const float *u, const float *v, const float *s0, const float *s1,
                          const float4 *tri_attr, const float *albedo, float *const no[3], float *const nd[3],
                          float *const w[3], cudaStream_t st, const TexArgs *tex)
{
    ShadeStreams p;
    for (int k = 0; k < 3; k++) { p.o[k] = o[k]; p.d[k] = d[k]; p.no[k] = no[k]; p.nd[k] = nd[k]; p.w[k] = w[k]; }
    p.tri_id = tri_id; p.t = t; p.u = u; p.v = v; p.s0 = s0; p.s1 = s1;
    if (n) shade_streams_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, p, tri_attr, albedo, tex ? *tex : TexArgs{nullptr, nullptr, nullptr});
}

// ------------------------------------------------------------------------------------------------
// derived hit attributes (src/accel/optix_backend.h:463-486): what OptixBackend::intersect returns beside
// (tri_id, t, u, v) — position o + t d (:469), geometry normal normalize(cross(p1-p0, p2-p0)) (:472-476,
// src/add_math.h:9-16), interpolated texcoord (:479-480) and shading normal (:483-484, src/add_math.h:3-7) — plus the
// material index Scene::intersect gathers next (src/main.cpp:325).  Lanes with tri_id < 0 are left untouched.
__global__ void __launch_bounds__(256)
hit_attributes_kernel(uint32_t n, const HitAttrArgs a)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t tri = a.tri_id[i];
    if (tri < 0) return;
    const float u = a.u[i], v = a.v[i];
    if (a.p[0]) {
        const float t = a.t[i];
        a.p[0][i] = __fadd_rn(a.ox[i], __fmul_rn(t, a.dx[i]));
        a.p[1][i] = __fadd_rn(a.oy[i], __fmul_rn(t, a.dy[i]));
        a.p[2][i] = __fadd_rn(a.oz[i], __fmul_rn(t, a.dz[i]));
    }
    if (a.ng[0]) {
        const int32_t i0 = a.pidx[3 * (size_t)tri], i1 = a.pidx[3 * (size_t)tri + 1], i2 = a.pidx[3 * (size_t)tri + 2];
        const V3 p0 = v3(a.pos[3 * (size_t)i0], a.pos[3 * (size_t)i0 + 1], a.pos[3 * (size_t)i0 + 2]);
        const V3 p1 = v3(a.pos[3 * (size_t)i1], a.pos[3 * (size_t)i1 + 1], a.pos[3 * (size_t)i1 + 2]);
        const V3 p2 = v3(a.pos[3 * (size_t)i2], a.pos[3 * (size_t)i2 + 1], a.pos[3 * (size_t)i2 + 2]);
        const V3 e1 = v3(p1.x - p0.x, p1.y - p0.y, p1.z - p0.z), e2 = v3(p2.x - p0.x, p2.y - p0.y, p2.z - p0.z);
        // enoki/array_router.h:674-684: cross = fmsub(a.yzx, b.zxy, a.zxy * b.yzx); normalize = v * rsqrt(dot)
        const V3 un = v3(fmaf(e1.y, e2.z, -(e1.z * e2.y)), fmaf(e1.z, e2.x, -(e1.x * e2.z)), fmaf(e1.x, e2.y, -(e1.y * e2.x)));
        const V3 g = normalize3(un);
        a.ng[0][i] = g.x; a.ng[1][i] = g.y; a.ng[2][i] = g.z;
    }
    const float w = __fsub_rn(__fsub_rn(1.0f, u), v);
    if (a.ns[0] || a.mat) {
        const float4 *ap = a.tri_attr + (size_t)tri * 3;
        const float4 q0 = __ldg(ap), q1 = __ldg(ap + 1), q2 = __ldg(ap + 2);
        if (a.ns[0]) {
            a.ns[0][i] = __fadd_rn(__fadd_rn(__fmul_rn(w, q0.x), __fmul_rn(u, q0.w)), __fmul_rn(v, q1.z));
            a.ns[1][i] = __fadd_rn(__fadd_rn(__fmul_rn(w, q0.y), __fmul_rn(u, q1.x)), __fmul_rn(v, q1.w));
            a.ns[2][i] = __fadd_rn(__fadd_rn(__fmul_rn(w, q0.z), __fmul_rn(u, q1.y)), __fmul_rn(v, q2.x));
        }
        if (a.mat) a.mat[i] = __float_as_int(q2.y);
    }
    if (a.uv[0]) {
        float tu = 0.0f, tv = 0.0f;
        if (a.uvs) {
            const int32_t j0 = a.uvidx[3 * (size_t)tri], j1 = a.uvidx[3 * (size_t)tri + 1], j2 = a.uvidx[3 * (size_t)tri + 2];
            if (j0 >= 0 && j1 >= 0 && j2 >= 0) {
                tu = __fadd_rn(__fadd_rn(__fmul_rn(w, a.uvs[2 * (size_t)j0]), __fmul_rn(u, a.uvs[2 * (size_t)j1])), __fmul_rn(v, a.uvs[2 * (size_t)j2]));
                tv = __fadd_rn(__fadd_rn(__fmul_rn(w, a.uvs[2 * (size_t)j0 + 1]), __fmul_rn(u, a.uvs[2 * (size_t)j1 + 1])), __fmul_rn(v, a.uvs[2 * (size_t)j2 + 1]));
            }
        }
        a.uv[0][i] = tu; a.uv[1][i] = tv;
    }
}
void launch_hit_attributes(uint32_t n, const HitAttrArgs &a, cudaStream_t st)
{
    if (n) hit_attributes_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, a);
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
    const uint32_t cap = (uint32_t)device_sms() * 8u;
    if (blocks > cap) blocks = cap;
    if (n) film_scale_kernel<<<blocks, 256, 0, st>>>(r, g, b, n, scale);
}

// one contiguous run of film floats (a rank's slice after wpt_film_reduce_scatter)
__global__ void film_scale_flat_kernel(float *p, uint32_t n, float scale)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] *= scale;
}
void launch_film_scale_flat(float *p, uint32_t n, float scale, cudaStream_t st)
{
    uint32_t blocks = (n + 255) / 256;
    const uint32_t cap = (uint32_t)device_sms() * 8u;
    if (blocks > cap) blocks = cap;
    if (n) film_scale_flat_kernel<<<blocks, 256, 0, st>>>(p, n, scale);
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
