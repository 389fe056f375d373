// kernels.cuh — launch-side declarations shared by kernels.cu and api.cu
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>
#include "bvh.h"
#include "../../include/wpt.h"

namespace wpt {

// In-flight path record, 40 B in three coalesced streams (16 B + 16 B + 8 B per lane):
//   a = (o.x, o.y, o.z, path id bits)   path id = pixel | (sample-in-pass << 24)
//   b = (d.x, d.y, d.z, throughput.r)
//   c = (throughput.g, throughput.b)
struct RayQueue {
    float4 *a;
    float4 *b;
    float2 *c;
};

// Per-pass device counters.  qcount[j] = rays entering bounce j; cursor[j] = dynamic-fetch cursor of
// the trace kernel of bounce j.
struct PassCounters {
    uint32_t qcount[WPT_MAX_DEPTH + 1];
    uint32_t cursor[WPT_MAX_DEPTH + 1];
    uint32_t retry_count, retry_cursor;   // seeded camera rays that found nothing inside their interval
};

struct TraceTuning {
    int variant;      // 0 = warp-queue kernel (default); the others are comparison / experiment variants, trace_kernels.cu
    int ctas_per_sm;  // persistent grid = num_sms * ctas_per_sm (0 = the occupancy calculator's answer)
};
// defaults of the warp-queue kernel (measured, profiles/r2_*): refill threshold (lanes), flush scheme (1 = balanced)
constexpr int kDefaultThresh = 20;
constexpr int kDefaultFlush = 1;

// the fused tail (trace_wq<TailIO>): everything shade_kernel needs, for all remaining bounces of a pass
struct TailArgs {
    float tmin, tmax;
    const float4 *tri_attr;
    const float *albedo;
    float *film_r, *film_g, *film_b;
    int bounce0, depth, spp_first;
    uint64_t seed;
};

// image textures (wpt_scene_set_texture): per-material window into one texel pool, per-triangle texcoords
struct MatTex;
struct TexArgs {
    const float4 *texels;     // null: no material of the scene has an image texture
    const MatTex *mattex;     // per material
    const float2 *triuv;      // 3 per ORIGINAL triangle
};

// wpt_hit_attributes (src/accel/optix_backend.h:463-486): device streams; any output group may be null
struct HitAttrArgs {
    const float *ox, *oy, *oz, *dx, *dy, *dz;
    const int32_t *tri_id;
    const float *t, *u, *v;
    const float *pos;
    const int32_t *pidx;
    const float4 *tri_attr;
    const float *uvs;
    const int32_t *uvidx;
    float *p[3], *ng[3], *ns[3], *uv[2];
    int32_t *mat;
};

// SoA stream front-end of wpt_intersect (mirror of struct Params, src/optixdata.h:11-29)
struct StreamIO {
    const float *ox, *oy, *oz, *dx, *dy, *dz, *tmin, *tmax;
    float tmin_s, tmax_s;
    const uint8_t *mask;
    uint32_t mask_size;
    int32_t *tri_id;
    float *t, *u, *v;
};

void launch_trace_queue(const BvhView &bvh, const RayQueue &q, int4 *hits, const uint32_t *n_ptr,
                        uint32_t *cursor, unsigned long long *live_acc, float tmin, float tmax,
                        uint32_t n_max, int num_sms, const TraceTuning &tune, cudaStream_t st,
                        int seeded = 0, uint32_t *retry_list = nullptr, uint32_t *retry_count = nullptr,
                        const uint32_t *indirect = nullptr);
void launch_trace_streams(const BvhView &bvh, const StreamIO &io, uint32_t n, uint32_t *cursor, int num_sms,
                          const TraceTuning &tune, cudaStream_t st);
void launch_trace_tail(const BvhView &bvh, const RayQueue &q, const uint32_t *n_ptr, uint32_t *cursor,
                       unsigned long long *live_acc, const TailArgs &a, uint32_t n_max, int num_sms,
                       const TraceTuning &tune, cudaStream_t st);
// WPT_CHECKS build only: per-kind counters of violated invariants (-1 in the product build)
int read_check_failures(unsigned int out[8], int reset);
void launch_hit_attributes(uint32_t n, const HitAttrArgs &a, cudaStream_t st);
void launch_raygen(const wpt_camera &cam, int W, int H, int spp_first, int spp_in_pass, uint64_t seed,
                   const RayQueue &q, uint32_t *qcount0, cudaStream_t st, const float *tmax_seed = nullptr,
                   int centre = 0, uint32_t lane_first = 0, uint32_t lane_count = 0);
void launch_depth_pack(const int4 *hits, uint32_t n, float *tlane_chunk, cudaStream_t st);
void launch_depth_seeds_from_lanes(const float *tlane, int W, int H, float margin, float *depth, float *seed, cudaStream_t st);
void launch_depth_seeds(const float4 *qa, const int4 *hits, int W, int H, float margin, float *depth, float *seed,
                        cudaStream_t st);
void launch_raygen_streams(const wpt_camera &cam, int W, int H, int spp_index, uint64_t seed, float *ox,
                           float *oy, float *oz, float *dx, float *dy, float *dz, cudaStream_t st);
void launch_shade(const RayQueue &in, const int4 *hits, const uint32_t *n_ptr, const RayQueue &out,
                  uint32_t *n_out, const float4 *tri_attr, const float *albedo, int bounce, int depth,
                  int spp_first, uint64_t seed, float *film_r, float *film_g, float *film_b, uint32_t n_max,
                  int num_sms, cudaStream_t st, int seeded0 = 0, const TexArgs *tex = nullptr);
void launch_shade_streams(uint32_t n, const float *const o[3], const float *const d[3], const int32_t *tri_id,
                          const float *t, const float *u, const float *v, const float *s0, const float *s1,
                          const float4 *tri_attr, const float *albedo, float *const no[3], float *const nd[3],
                          float *const w[3], cudaStream_t st, const TexArgs *tex = nullptr);
void launch_film_scale(float *r, float *g, float *b, uint32_t n, float scale, cudaStream_t st);
void launch_film_scale_flat(float *p, uint32_t n, float scale, cudaStream_t st);
void launch_refit(Block64 *blocks, const uint32_t *level_nodes, const uint32_t *level_begin_host, uint32_t n_levels,
                  const float *pos, const int32_t *pidx, float *boxes, TriAttr48 *attr, const uint8_t *geo_normal,
                  cudaStream_t st, uint32_t *bad_flag = nullptr);
// lbvh_build.cu: topology of a 7-wide tree built on the device (geometry is then written by launch_refit)
bool lbvh_build_topology(const float *d_pos, const int32_t *d_pidx, uint32_t n_tris, Block64 **d_blocks_out,
                         uint32_t *n_blocks_out, uint32_t **d_level_nodes_out, std::vector<uint32_t> &level_begin,
                         uint32_t *n_nodes_out, uint32_t *max_depth_out, const char **err, cudaStream_t st);
void launch_philox(uint32_t n, uint32_t c1, uint32_t c2, uint64_t key, uint32_t *out, cudaStream_t st);

}  // namespace wpt
