/*
 * wpt.h — C ABI of libwpt.so, the B200-native replacement for the reference's per-sample /
 * per-bounce hot path.  Plain pointers and sizes only; no C++/torch types cross this boundary.
 * Every entry point returns 0 on success or a WPT_E_* code; wpt_last_error() gives the text
 * (thread-local).  Nothing throws across the ABI (the reference throws std::runtime_error from
 * OPTIX_CHECK/CUDA_CHECK, src/accel/optix_backend.h:25-82 — the C++ shim include/wpt_shim.hpp
 * re-raises our codes the same way).
 *
 * All file:line citations are relative to the reference tree (/root/reference).
 *
 * There is NO CPU fallback: every compute entry point needs a CUDA device and fails with
 * WPT_E_CUDA otherwise.
 */
#ifndef WPT_H
#define WPT_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define WPT_VERSION 200

enum {
    WPT_OK = 0,
    WPT_E_INVALID = 1,  /* bad argument */
    WPT_E_CUDA = 2,     /* CUDA runtime error (text in wpt_last_error) */
    WPT_E_STATE = 3,    /* call order (e.g. intersect before commit) */
    WPT_E_LIMIT = 4,    /* size limit exceeded (W*H > 2^24, BVH depth, ...) */
    WPT_E_IO = 5
};

const char *wpt_last_error(void);
int wpt_version(void);
/* Self-checking build (make libwpt_checked.so, -DWPT_CHECKS): the traversal kernels count violations of their own
 * invariants — warp-queue / item / stack slots in range, block indices inside the tree, ray indices inside the queue,
 * queue empty after a drain — per kind in out[8]; reset != 0 zeroes them.  WPT_E_STATE in the product build, whose
 * kernels carry no checks. */
int wpt_debug_check_failures(uint32_t out[8], int reset);

/* ------------------------------------------------------------------ scene
 * Replaces Scene::add_triangle_mesh's device upload (src/main.cpp:288-310) and
 * Scene::commit -> OptixBackend::init + set_triangles_soup (src/main.cpp:312-318,
 * src/accel/optix_backend.h:178-186, 283-364).  Array shapes are exactly the reference's
 * AoS-flattened host vectors (src/main.cpp:160-223):
 *   positions f32[3*n_positions], position_triplets i32[3*n_triangles] (global face order ==
 *   tri_id), normals f32[3*n_normals], normal_triplets i32[3*n_triangles],
 *   material_id i32[n_triangles] (src/main.cpp:185), albedo f32[3*n_materials] (the constant
 *   1x1 reflectance texture of each LambertBsdf, src/main.cpp:40-44,104-106).
 * Deviation: normal_triplets == NULL, or a negative index in a face, makes that
 * face use its geometric normal instead of the reference's out-of-bounds gather.
 * material_id == NULL -> all 0; albedo == NULL -> one white material (src/main.cpp:234,244).
 * Host arrays are copied at create; the caller may free them afterwards. */
typedef struct wpt_scene wpt_scene;
typedef struct {
    const float   *positions;
    const int32_t *position_triplets;
    const float   *normals;
    const int32_t *normal_triplets;
    const int32_t *material_id;
    const float   *albedo;
    uint32_t n_positions, n_triangles, n_normals, n_materials;
} wpt_mesh_desc;

typedef struct {
    uint32_t n_triangles, n_nodes, n_leaf_slots, max_depth;
    double   sah_cost;      /* of the wide tree, c_node=1, c_tri=0.3, relative to root area */
    double   build_ms, upload_ms;
    uint64_t node_bytes, tri_bytes, attr_bytes;
} wpt_bvh_stats;

int  wpt_scene_create(const wpt_mesh_desc *mesh, wpt_scene **out);
/* Optional texture coordinates (src/main.cpp:215-223: texcoords f32[2*n_uvs], texcoord_triplets i32[3*n_triangles],
 * -1 = none): only wpt_hit_attributes reads them.  Call before wpt_scene_commit. */
int  wpt_scene_set_texcoords(wpt_scene *s, const float *uvs, const int32_t *uv_triplets, uint32_t n_uvs);
/* Image texture of one material: rgb f32[3*width*height], texel (x, y) at y*width + x.  The bsdf weight of that
 * material becomes ImageTexture::eval (src/main.cpp:62-76) at the hit's interpolated texcoord (wpt_scene_set_texcoords):
 * bilinear, clamped, same operation order as the reference.  Its texel_fetch indexes y * size.y + x (src/main.cpp:52),
 * wrong for non-square images; this library (and the oracle) use the row length, y * size.x + x.  A 1x1 texture is the
 * constant albedo the reference builds for every material (src/main.cpp:40-44,104-106).  Call before wpt_scene_commit. */
int  wpt_scene_set_texture(wpt_scene *s, uint32_t material, const float *rgb, uint32_t width, uint32_t height);
/* SAH BVH build on the host, flatten to 64-byte 7-wide quantised nodes + 64-byte triangle blocks (one buffer),
 * one upload to `device`.  WPT_E_LIMIT if the tree is deeper than 32 levels or holds more than 2^27 blocks
 * (nodes + triangles; about 115 M triangles).  If the vertices were last moved by wpt_scene_refit_device, a
 * re-commit rebuilds from those device-side positions (read back first). */
int  wpt_scene_commit(wpt_scene *s, int device);
/* The same commit with the tree built ON THE DEVICE (the reference's own build is a device build
 * too: optixAccelBuild, src/accel/optix_backend.h:309-358): 63-bit Morton codes, radix sort, binary radix tree
 * (Karras 2012), bottom-up boxes, level-by-level collapse to 7-wide nodes with <= 3-triangle leaves, then the block
 * geometry by the refit kernels.  10 ms per million triangles instead of 0.3 s, for scenes whose topology
 * changes between frames; the trees are of lower quality than the host SAH builder's (traversal ~1.2x slower, measured), hits
 * are the same closest hits.  sah_cost is reported as 0 (not evaluated).  WPT_E_LIMIT if the tree is deeper than 32. */
int  wpt_scene_commit_lbvh(wpt_scene *s, int device);
/* wpt_scene_commit with the host-built tree kept in a file: if `cache_path` holds the tree of exactly this mesh (hash of
 * positions and indices, library version) it is loaded instead of built, otherwise the tree is built and the file written
 * (atomically: temp file + rename).  For the G ranks of a multi-GPU job on one box (one process per GPU, same mesh): one
 * rank commits first, the others then load — one SAH build (C4: 3.5-4.7 s) instead of G.  *from_cache (optional) = 1 if
 * the tree was loaded.  A cache that cannot be read or written is not an error: the tree is built.  The file is outside
 * input and is treated as such: besides the mesh hash it must carry this build's block layout and builder revision, a
 * payload checksum, a depth the traversal stack holds, a level table that partitions the nodes, child runs inside the
 * buffer and face indices inside the mesh — anything else is ignored and rebuilt.  It is written with mode 0600. */
int  wpt_scene_commit_cached(wpt_scene *s, int device, const char *cache_path, int *from_cache);
int  wpt_scene_stats(const wpt_scene *s, wpt_bvh_stats *out);
/* Moved vertices, same topology: bottom-up refit of the node boxes and rewrite of the triangle blocks ON THE
 * DEVICE, one launch per tree level (the reference rebuilds its GAS from scratch,
 * src/accel/optix_backend.h:283-364).  Hits stay exact — boxes are re-quantised conservatively — but tree
 * quality decays with the deformation; commit again to rebuild.  Vertex normals are kept; faces that shade
 * with the geometric normal get the new one.  wpt_scene_refit takes HOST positions f32[3*n_positions]
 * (count must match; synchronous); wpt_scene_refit_device takes a DEVICE array and is asynchronous on
 * `stream`; the caller orders it against renders in flight on other streams (the refit rewrites the blocks the
 * traversal reads).  Device-side positions are not validated: a NaN/Inf vertex yields an undefined tree until the
 * next wpt_scene_commit, which reads the positions back and rejects non-finite ones. */
int  wpt_scene_refit(wpt_scene *s, const float *positions, uint32_t n_positions);
int  wpt_scene_refit_device(wpt_scene *s, const float *d_positions, void *stream);
/* After a refit: *non_finite = 1 if a NaN / Inf vertex reached the tree (the refit kernels raise a flag word; the tree
 * is then undefined until the next commit).  Waits for `stream`. */
int  wpt_scene_refit_status(const wpt_scene *s, int *non_finite, void *stream);
void wpt_scene_destroy(wpt_scene *s);

/* ------------------------------------------------------------------ OBJ / MTL ingestion
 * Replaces load_meshes (src/main.cpp:133-251): tinyobjloader with triangulate = true, all shapes
 * flattened into one pool, tri_id = global face index in file order (src/main.cpp:165-188),
 * material_id = tinyobj id + 1 with id 0 the default white material (src/main.cpp:185,229-234).
 * use_kd == 0 reproduces the reference as shipped (every material white, src/main.cpp:244);
 * use_kd != 0 gives each material its MTL Kd colour (the commented-out intent, src/main.cpp:239-242).
 * wpt_obj_mesh fills a wpt_mesh_desc whose pointers stay valid until wpt_obj_free; faces without
 * normals carry normal index -1 (geometric-normal fallback, see above).  Texture coordinates are
 * loaded like the reference does (src/main.cpp:215-223) and exposed, but nothing on the path reads
 * them (the reference's textures are constant 1x1, src/main.cpp:40-44). */
typedef struct wpt_obj wpt_obj;
int  wpt_obj_load(const char *path, int use_kd, wpt_obj **out);
int  wpt_obj_mesh(const wpt_obj *o, wpt_mesh_desc *out);
int  wpt_obj_texcoords(const wpt_obj *o, const float **uvs, const int32_t **uv_triplets, uint32_t *n_uvs);
/* name of material i in [0, n_materials): "" for the default material 0 */
const char *wpt_obj_material_name(const wpt_obj *o, uint32_t i);
void wpt_obj_free(wpt_obj *o);

/* ------------------------------------------------------------------ closest-hit on ray streams
 * Replaces OptixBackend::intersect's launch (src/accel/optix_backend.h:422-460) and the device
 * programs src/wavefront_isect.cu:48-112.  Field-for-field mirror of struct Params
 * (src/optixdata.h:11-29): eight DEVICE f32 streams, a bool mask of size n or 1 (size 1 =
 * broadcast, wavefront_isect.cu:86; NULL = all live), four DEVICE result streams.
 * tmin/tmax stream pointers may be NULL, in which case the scalars are used (the reference
 * materialises two constant arrays per bounce, src/ray.h:20-26).
 * Contract: tri_id = -1 on miss (wavefront_isect.cu:70); t,u,v written only on hit; lanes masked
 * out are left untouched (wavefront_isect.cu:87); u,v are the weights of vertex 1 and 2
 * (src/add_math.h:6); no culling; tmin < t <= tmax.  `stream` is a cudaStream_t (NULL = default
 * stream).  Asynchronous. */
typedef struct {
    const float *ox, *oy, *oz, *dx, *dy, *dz;
    const float *tmin, *tmax; /* optional streams */
    float tmin_scalar, tmax_scalar;
} wpt_ray_streams;
typedef struct {
    int32_t *tri_id;
    float   *t, *u, *v;
} wpt_hit_streams;

/* Per-ray tmin below 0 is clamped to 0 (the closest hit is folded with an integer min on the float bits of t).
 * At most 4096 wpt_intersect calls of one scene may be in flight at the same time (ring of fetch cursors).
 * Equal-t hits (shared edges) resolve to the LOWEST primitive index, whatever the order of the tests. */
int wpt_intersect(const wpt_scene *s, const wpt_ray_streams *rays, const uint8_t *mask, uint32_t mask_size,
                  uint32_t n, const wpt_hit_streams *hits, void *stream);
/* Same, HOST buffers in and out (copies inside; synchronous).  A convenience for tests and one-off queries: it allocates
 * and frees its device staging on every call (cudaMalloc / cudaFree wait for the whole device).  A host that calls once
 * per bounce keeps its rays on the device and uses wpt_intersect, or uses wpt_intersect_host_static below. */
int wpt_intersect_host(const wpt_scene *s, const wpt_ray_streams *rays, const uint8_t *mask, uint32_t mask_size,
                       uint32_t n, const wpt_hit_streams *hits);

/* wpt_intersect_host without a single allocation.  No counterpart in the reference (its Enoki arrays allocate on every
 * bounce, src/ray.h:20-26); it exists for hosts that have to keep cudaMalloc out of their per-bounce loop:
 *   - a plugin inside an application whose own pool allocator has already claimed the whole GPU (Enoki's and
 *     PyTorch's caching allocators do): a raw cudaMalloc for staging then fails with out-of-memory although the
 *     application has plenty of room in its pool;
 *   - a frame loop that cannot afford cudaMalloc / cudaFree, which wait for the whole device and so stall every
 *     other stream of the process once per call;
 *   - a thread working while another thread of the process records a CUDA graph in the default (global) capture
 *     mode, where cudaMalloc from any thread breaks the recording.
 * Same arguments, contract and results as wpt_intersect_host (HOST buffers in and out, bit-identical hits), plus the
 * stream the copies and the kernel are queued on (NULL = default stream); it returns after that stream has drained.
 * The device-side staging wpt_intersect_host allocates per call (eight ray streams, four result streams, the mask and
 * the kernel's ray-fetch counter) lives in a fixed area inside the library image instead:
 *   - it holds WPT_STATIC_MAX_RAYS rays; a larger n is refused with WPT_E_LIMIT and nothing is done — split the batch;
 *   - it costs 12.3 MiB of device memory on every GPU that loads libwpt.so, from the load to the end of the process,
 *     whether this entry point is ever called or not;
 *   - there is ONE area per GPU and process, and it is not locked: two calls on the same GPU whose work overlaps there
 *     (two threads, two streams) overwrite each other's rays and results.  Run them one after the other;
 *   - the area keeps what the last call left in it.  The call zeroes the counter the kernel starts from and uploads
 *     the caller's result arrays first, so masked-out and missed lanes come back exactly as the caller passed them;
 *   - the first call on each GPU asks the runtime where that GPU's copy of the area is (cudaGetSymbolAddress) and
 *     remembers it; make that first call at set-up, before any stream capture begins in the process.
 * The scene itself is still allocated by wpt_scene_commit at load time; this call, like wpt_intersect, allocates
 * nothing, on the device or through any library. */
#define WPT_STATIC_MAX_RAYS (1u << 18)
int wpt_intersect_host_static(const wpt_scene *s, const wpt_ray_streams *rays, const uint8_t *mask,
                              uint32_t mask_size, uint32_t n, const wpt_hit_streams *hits, void *stream);

/* ------------------------------------------------------------------ derived hit attributes
 * What OptixBackend::intersect returns WITH the hit (src/accel/optix_backend.h:463-486, TriangleHitInfo :99-134) and
 * Scene::intersect turns into an Interaction (src/main.cpp:320-340): position o + t d, geometry normal
 * normalize(cross(p1-p0, p2-p0)) (src/add_math.h:9-16), barycentric-interpolated UN-normalised shading normal and
 * texcoord (src/add_math.h:3-7), and the material index (src/main.cpp:325).  DEVICE streams; an output group
 * (p*, gn*, sn*, t*) is either given completely or left NULL; lanes with tri_id < 0 are left untouched.
 * Faces without normals get the geometric normal as shading normal (the deviation stated above); scenes without
 * texcoords yield (0, 0). */
typedef struct {
    float *px, *py, *pz;
    float *gnx, *gny, *gnz;
    float *snx, *sny, *snz;
    float *tu, *tv;
    int32_t *material_id;
} wpt_hit_attr_streams;
int wpt_hit_attributes(const wpt_scene *s, const wpt_ray_streams *rays, const wpt_hit_streams *hits, uint32_t n,
                       const wpt_hit_attr_streams *out, void *stream);

/* ------------------------------------------------------------------ camera (src/pinhole.h:7-72) */
typedef struct {
    float origin[3], x[3], y[3], z[3];
    float film_size_y, focal_dist, fov_y, lens_radius;
} wpt_camera;
/* ThinlensCamera ctor, src/pinhole.h:9-25 (host scalar math there too). */
int wpt_camera_init(wpt_camera *c, const float look_from[3], const float look_at[3], const float up[3],
                    float lens_radius, float focal_dist, float fov_y, float film_size_y);

/* ------------------------------------------------------------------ the sample / bounce loop
 * Replaces src/main.cpp:385-426: per sample, camera ray (pinhole.h:27-56), then `depth` segments of
 * intersect -> film += contrib on miss -> Lambert bounce about the interpolated shading normal.
 * RNG: counter-based Philox4x32-10, counter = (pixel, global spp index, slot, 0), key = seed;
 * slot 0 -> lens(2)+jitter(2), slot 1+j -> scatter(2) of bounce j.  Keying on the GLOBAL spp index
 * makes a [spp_begin, spp_begin+spp_count) shard on another GPU a disjoint, decorrelated part of
 * the same estimator. */
typedef struct {
    int32_t  width, height;
    int32_t  spp_begin, spp_count, depth;
    uint64_t seed;
    float    tmin, tmax;       /* src/ray.h:16 defaults: 1e-3, 1e20 */
} wpt_render_config;

#define WPT_MAX_DEPTH 64
typedef struct {
    uint64_t segments;                    /* ray segments actually traced (device counted) */
    uint64_t live[WPT_MAX_DEPTH];         /* rays entering bounce j, summed over passes */
    uint32_t passes, kernel_launches;
    float    ms_total, ms_raygen, ms_trace, ms_shade; /* CUDA-event times; per-kernel only if profile=1 */
    uint64_t retraced;                    /* depth-seeded camera rays that had to be traced a second time (not in segments) */
    int32_t  tail_from;                   /* first bounce the fused tail kernel covered in this render, -1 = none */
    int32_t  reserved;
} wpt_stats;

typedef struct wpt_renderer wpt_renderer;
/* Owns the double-buffered ray/hit queues for width*height*spp_per_pass paths in flight, all
 * allocated here: wpt_render allocates nothing. profile != 0 brackets every kernel with events. */
int  wpt_renderer_create(const wpt_scene *s, int32_t width, int32_t height, int32_t spp_per_pass,
                         wpt_renderer **out);
void wpt_renderer_destroy(wpt_renderer *r);
int  wpt_renderer_set_profile(wpt_renderer *r, int profile);
/* Depth seeding of the camera rays (default on): for pinhole cameras and spp_count >= 8, wpt_render first traces one
 * pixel-centre ray per pixel and starts the samples' camera rays with tmax = 1.0005 x the 3x3-neighbourhood maximum of
 * those hit distances; rays that find nothing inside their seed are traced again with the full interval, so hits,
 * counters and film are those of the unseeded render.  on = 0 switches it off; on > 1 sets the margin to on/10000
 * (a margin below 1 makes most seeds too short: only useful to exercise the second trace). */
int  wpt_renderer_set_depth_seeding(wpt_renderer *r, int on);
/* Sharded depth pre-pass for a frame split over n_ranks GPUs (wpt_film_reduce below): every rank needs the same seeds,
 * so instead of all of them repeating the whole one-ray-per-pixel pre-pass, rank g traces 1/n_ranks of it and the hit
 * distances are all-gathered on `nccl_comm` (ncclAllGather, W*H floats in total) inside wpt_render, on its stream.
 * EVERY rank of the communicator must then call wpt_render for every frame (it is a collective), with the same camera
 * kind, tmax and depth-seeding setting; the pre-pass then runs whatever the rank's own sample count (no >= 8 rule).  The seeds, hence
 * the frame, are bit-identical to the unsharded pre-pass.  nccl_comm = NULL or n_ranks <= 1 switches it off.
 * Ranks that disagree on any of those (one with a thin lens, a finite tmax or seeding switched off) do not all enter
 * the all-gather: the others then wait in it for ever, with no diagnostic — NCCL has no timeout of its own.
 * WPT_E_INVALID for more ranks than pixels. */
int  wpt_renderer_set_depth_shard(wpt_renderer *r, void *nccl_comm, int rank, int n_ranks);
/* Fused tail: bounces >= first_bounce of every pass run in ONE launch of the traversal kernel with the shading step
 * fused in (a path whose segment ends is shaded by its own lane and continues in place) instead of one trace + one
 * shade launch per bounce — late bounces hold few paths and a persistent launch has a fixed ramp-up/drain cost.
 * Same Philox counters, same paths, same film up to fp32 atomic order.  first_bounce >= 1; negative = off (a wavefront
 * launch pair for every bounce); 0 = automatic (default): from bounce 2 on when a pass holds at most 2^24 paths, off
 * for larger passes, where it does not pay (profiles/r2_grid_c3.txt). */
int  wpt_renderer_set_tail(wpt_renderer *r, int first_bounce);
/* Tuning knobs (0 = default): which trace kernel; CTAs per SM of a trace launch; overlap_lanes != 0 renders
 * alternate passes on two internal streams (each lane holds half of spp_per_pass in its own queues) so that
 * the HBM-bound raygen/shade kernels of one lane overlap the issue-bound trace kernel of the other.  Same
 * paths, same film up to fp32 atomic order.  Not allowed while a render is in flight (WPT_E_STATE). */
int  wpt_renderer_set_tuning(wpt_renderer *r, int trace_variant, int ctas_per_sm, int overlap_lanes);

/* Adds the UN-normalised sum of the requested samples into DEVICE planar film_r/g/b
 * (pixel_index = y*W + x, src/main.cpp:379-381,407).  Asynchronous on `stream`; `stats` (optional,
 * host) is filled by wpt_render_finish, which synchronises. */
int wpt_render(wpt_renderer *r, const wpt_camera *cam, const wpt_render_config *cfg,
               float *film_r, float *film_g, float *film_b, void *stream);
/* wpt_render may be recorded into a CUDA graph (call it between cudaStreamBeginCapture and cudaStreamEndCapture on
 * `stream`): it only enqueues kernels, memsets and event fork/joins.  After a replay, wpt_render_finish(stream of the
 * replay) returns the counters of that replay (ms_total = 0); it may be called after every replay. */
int wpt_render_finish(wpt_renderer *r, wpt_stats *stats, void *stream);
/* film *= 1/spp (src/main.cpp:429), device, async */
int wpt_film_scale(float *film_r, float *film_g, float *film_b, uint32_t n, float scale, void *stream);
/* Whole frame, HOST film out (zero -> render -> scale 1/spp_count -> D2H), src/main.cpp:369-439.  The three
 * channel copies are queued asynchronously behind the frame and waited for once; give it a film from wpt_host_alloc
 * (pinned) and they are plain DMA transfers. */
int wpt_render_host(wpt_renderer *r, const wpt_camera *cam, const wpt_render_config *cfg,
                    float *film_r, float *film_g, float *film_b, wpt_stats *stats);

/* pinned host memory (cudaMallocHost) for films and ray buffers that cross the boundary every frame */
int  wpt_host_alloc(size_t bytes, void **out);
void wpt_host_free(void *p);

/* device buffers for hosts without CUDA headers of their own (include/wpt_shim.hpp); the reference's Enoki arrays
 * do this for it (enoki/cuda.h:210-220).  upload is asynchronous on `stream`, download synchronises it. */
int  wpt_device_count(void);
int  wpt_device_alloc(int device, size_t bytes, void **out);
void wpt_device_free(void *p);
int  wpt_device_upload(void *dst_device, const void *src_host, size_t bytes, void *stream);
int  wpt_device_download(void *dst_host, const void *src_device, size_t bytes, void *stream);
int  wpt_device_sync(int device);            /* makes `device` current and waits for all its work */
int  wpt_device_set(int device);             /* makes `device` current (the stream-level calls act on the current device) */
int  wpt_device_zero(void *dst_device, size_t bytes, void *stream);

/* ------------------------------------------------------------------ multi-GPU: one reduce of the partial films
 * GPU g renders the global sample indices [g*spp/G, (g+1)*spp/G) of ALL pixels (spp_begin/spp_count of
 * wpt_render_config; the RNG is keyed on the global index), the scene is replicated, and the only exchange is ONE
 * sum of the partial films onto `root` — ncclReduce(sum, f32) over NVLink — followed by wpt_film_scale(1/spp) there.
 * The reference is single-GPU (src/main.cpp:354-446); this is what its main() calls once per frame when sharded
 * (examples/wurst_main.cpp --gpus N).  NCCL is bound at run time (dlopen libnccl.so.2); WPT_E_STATE if absent.
 *   one process per GPU:  rank 0: wpt_nccl_unique_id(id) -> ship the 128 bytes to the other ranks by any means
 *                         every rank: wpt_nccl_comm_create(id, n_ranks, rank, device, &comm)
 *   one process, N GPUs:  wpt_nccl_comm_create_all(N, devices, comms); the N wpt_film_reduce calls of a frame go
 *                         between wpt_nccl_group_start() and wpt_nccl_group_end()
 * wpt_film_reduce is asynchronous on `stream` and in place: afterwards the root's film holds the sum (one collective
 * of 3n floats when the three planes are contiguous, a group of three otherwise). */
int  wpt_nccl_unique_id(void *id128);
int  wpt_nccl_comm_create(const void *id128, int n_ranks, int rank, int device, void **comm);
int  wpt_nccl_comm_create_all(int n_devices, const int *devices, void **comms);
void wpt_nccl_comm_destroy(void *comm);
int  wpt_nccl_group_start(void);
int  wpt_nccl_group_end(void);
int  wpt_film_reduce(void *nccl_comm, float *film_r, float *film_g, float *film_b, uint32_t n, int root, void *stream);

/* The same exchange without a root (no counterpart in the reference either): ncclReduceScatter leaves on rank g the
 * sum of floats [first, first + count) of the FLATTENED planar film (r plane, then g, then b: 3n floats), in place, where
 * wpt_film_shard_layout(n, g, n_ranks, &padded, &first, &count) says which.  `film` must be one device buffer of at least
 * `padded` floats (3n rounded up to a multiple of n_ranks; the padding is summed too, zero it once).  Afterwards the rest
 * of the buffer is undefined.  wpt_film_shard_finish then multiplies this rank's slice by `scale` (the 1/spp of
 * src/main.cpp:429) and, if host_film != NULL, queues the copy of the slice to host_film + first — host_film is the
 * WHOLE film of 3n floats; when it is one pinned buffer shared by the ranks (one process: wpt_host_alloc; one process per
 * GPU: a shared mapping that every rank passes to wpt_host_register) the n_ranks copies run in parallel and assemble the
 * frame on the host.  Each rank moves and copies 1/n_ranks of what the root of wpt_film_reduce does.  All asynchronous
 * on `stream`. */
int  wpt_film_shard_layout(uint32_t n, int rank, int n_ranks, uint64_t *padded_floats, uint64_t *first, uint64_t *count);
int  wpt_film_reduce_scatter(void *nccl_comm, float *film, uint32_t n, int rank, int n_ranks, void *stream);
int  wpt_film_shard_finish(float *film, uint32_t n, int rank, int n_ranks, float scale, float *host_film, void *stream);
/* page-locks an existing host range for DMA (cudaHostRegister, portable), e.g. a shared-memory mapping of the film */
int  wpt_host_register(void *p, size_t bytes);
void wpt_host_unregister(void *p);

/* ------------------------------------------------------------------ per-function device entry points
 * (parity tests of the fused kernels' pieces against the oracle; all DEVICE buffers) */
/* camera rays for pixels [0,W*H) of one global sample index: o,d as 6 planar streams */
int wpt_raygen(const wpt_camera *cam, int32_t width, int32_t height, int32_t spp_index, uint64_t seed,
               float *ox, float *oy, float *oz, float *dx, float *dy, float *dz, void *stream);
/* one shading step on given hits (tri_id>=0 lanes): next origin/direction/weight, planar streams */
int wpt_shade_step(const wpt_scene *s, uint32_t n, const float *o_xyz[3], const float *d_xyz[3],
                   const int32_t *tri_id, const float *t, const float *u, const float *v,
                   const float *s0, const float *s1, float *next_o[3], float *next_d[3], float *weight[3],
                   void *stream);
/* Philox4x32-10 on device: out[4*i..] = philox(ctr=(i, c1, c2, 0), key) */
int wpt_philox(uint32_t n, uint32_t c1, uint32_t c2, uint64_t key, uint32_t *out, void *stream);

/* ------------------------------------------------------------------ output (src/fimage.h:33-58)
 * PFM: "PF\nW H\n-1\n", rows bottom-up, interleaved RGB float32 LE.  HOST planar inputs. */
int wpt_save_pfm(const char *path, const float *r, const float *g, const float *b, uint32_t width,
                 uint32_t height);

/* ------------------------------------------------------------------ checkpoint / resume
 * No counterpart in the reference, whose film lives on the device until main() ends (src/main.cpp:434-442).  A frame is
 * a sum over samples and the RNG is keyed on the global sample index, so the UN-normalised film plus the number of samples
 * in it is a complete checkpoint: wpt_film_save writes HOST planar r, g, b (f32[width*height] each, before the 1/spp of
 * src/main.cpp:429), spp_done and the seed to one file, atomically (temp file + rename) and with a checksum; a later
 * process loads it (wpt_film_load: dimensions must match, WPT_E_IO for a truncated or damaged file), uploads the planes
 * and goes on with wpt_render(spp_begin = spp_done, same seed) into them.  wpt_film_info reads the header only. */
int wpt_film_save(const char *path, const float *r, const float *g, const float *b, uint32_t width, uint32_t height,
                  uint64_t spp_done, uint64_t seed);
int wpt_film_info(const char *path, uint32_t *width, uint32_t *height, uint64_t *spp_done, uint64_t *seed);
int wpt_film_load(const char *path, float *r, float *g, float *b, uint32_t width, uint32_t height, uint64_t *spp_done,
                  uint64_t *seed);

#ifdef __cplusplus
}
#endif
#endif
