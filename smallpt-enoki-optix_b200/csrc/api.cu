// api.cu — the C ABI of libwpt.so (include/wpt.h).  Host C++; owns device memory, never throws.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>   // types and prototypes only: the library is bound with dlopen (wpt_film_reduce)
#include <nvtx3/nvToolsExt.h>

#include <fcntl.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/wpt.h"
#include "bvh_build.h"
#include "kernels.cuh"
#include "obj_load.h"
#include "shade_core.h"   // MatTex

using namespace wpt;

static thread_local std::string g_err;
static int fail(int code, const std::string &msg)
{
    g_err = msg;
    return code;
}
#define WPT_CUDA(call)                                                                              \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess)                                                                      \
            return fail(WPT_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));            \
    } while (0)

// NVTX ranges around the host-side phases (visible in Nsight Systems / ncu --nvtx; no-ops without a tool attached)
struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

extern "C" const char *wpt_last_error(void) { return g_err.c_str(); }
extern "C" int wpt_version(void) { return WPT_VERSION; }

// libwpt_checked.so (built with -DWPT_CHECKS): the traversal kernels count violations of their own invariants (queue /
// item / stack slots, block indices, ray indices); out[8] receives the counters per kind.  WPT_E_STATE in the product build.
extern "C" int wpt_debug_check_failures(uint32_t out[8], int reset)
{
    if (!out) return fail(WPT_E_INVALID, "wpt_debug_check_failures: null argument");
    const int rc = read_check_failures(out, reset);
    if (rc < 0) return fail(WPT_E_STATE, "wpt_debug_check_failures: this libwpt was built without WPT_CHECKS");
    if (rc > 0) return fail(WPT_E_CUDA, "wpt_debug_check_failures: cannot read the counters");
    return WPT_OK;
}

// ------------------------------------------------------------------------------------------------
constexpr uint32_t kCursorRing = 4096;
struct wpt_scene {
    // host copies (src/main.cpp:296-309 keeps them in Enoki arrays)
    std::vector<float> pos, nrm, albedo, uvs;
    std::vector<int32_t> pidx, nidx, mat, uvidx;
    uint32_t n_tris = 0;
    // image textures per material (wpt_scene_set_texture): rgb f32[3*w*h]; none = the constant albedo
    std::vector<std::vector<float>> tex_rgb;
    std::vector<uint32_t> tex_w, tex_h;
    float4 *d_texels = nullptr;
    MatTex *d_mattex = nullptr;
    float2 *d_triuv = nullptr;
    TexArgs tex{nullptr, nullptr, nullptr};
    bool tree_from_cache = false;
    bool pos_on_device_newer = false;   // wpt_scene_refit_device moved the vertices: the next commit reads them back
    // device
    int device = -1, num_sms = 0;
    bool committed = false;
    Block64 *d_blocks = nullptr;   // nodes and triangles, bvh.h
    TriAttr48 *d_attr = nullptr;
    float *d_albedo = nullptr;
    // ray-fetch cursors of wpt_intersect: a ring of kCursorRing, one per call, so that calls on different streams
    // (the BVH is read-only) do not share a counter; more than kCursorRing calls in flight at once would alias
    uint32_t *d_cursor = nullptr;
    mutable std::atomic<uint32_t> cursor_rr{0};
    // refit (wpt_scene_refit): BFS levels of the node blocks, geometric-normal flags; vertex/index/box
    // buffers are allocated by the first refit
    uint32_t *d_level_nodes = nullptr;
    uint8_t *d_geo = nullptr;
    float *d_pos = nullptr, *d_boxes = nullptr, *d_uvs = nullptr;
    int32_t *d_pidx = nullptr, *d_uvidx = nullptr;
    std::vector<uint32_t> level_begin;
    BvhView view{};
    wpt_bvh_stats stats{};
};

extern "C" int wpt_scene_create(const wpt_mesh_desc *m, wpt_scene **out)
{
    if (!m || !out) return fail(WPT_E_INVALID, "wpt_scene_create: null argument");
    if (!m->positions || !m->position_triplets || m->n_triangles == 0 || m->n_positions == 0)
        return fail(WPT_E_INVALID, "wpt_scene_create: positions/position_triplets required, n_triangles > 0");
    if (m->n_triangles > 0x7fffffffu / 3) return fail(WPT_E_LIMIT, "wpt_scene_create: too many triangles");
    wpt_scene *s = new (std::nothrow) wpt_scene();
    if (!s) return fail(WPT_E_INVALID, "out of host memory");
    s->n_tris = m->n_triangles;
    s->pos.assign(m->positions, m->positions + 3 * (size_t)m->n_positions);
    s->pidx.assign(m->position_triplets, m->position_triplets + 3 * (size_t)m->n_triangles);
    if (m->normals && m->normal_triplets && m->n_normals) {
        s->nrm.assign(m->normals, m->normals + 3 * (size_t)m->n_normals);
        s->nidx.assign(m->normal_triplets, m->normal_triplets + 3 * (size_t)m->n_triangles);
        for (int32_t v : s->nidx)
            if (v >= (int32_t)m->n_normals) { delete s; return fail(WPT_E_INVALID, "normal index out of range"); }
    }
    if (m->albedo && m->n_materials) s->albedo.assign(m->albedo, m->albedo + 3 * (size_t)m->n_materials);
    else s->albedo.assign(3, 1.0f);   // src/main.cpp:234: default material Spectrum(1.0)
    const int32_t n_mat = (int32_t)(s->albedo.size() / 3);
    if (m->material_id) {
        s->mat.assign(m->material_id, m->material_id + m->n_triangles);
        for (int32_t v : s->mat)
            if (v < 0 || v >= n_mat) { delete s; return fail(WPT_E_INVALID, "material id out of range"); }
    } else {
        s->mat.assign(m->n_triangles, 0);
    }
    *out = s;
    return WPT_OK;
}

static void scene_free_device(wpt_scene *s)
{
    if (s->device >= 0) cudaSetDevice(s->device);
    cudaFree(s->d_blocks); cudaFree(s->d_attr);
    cudaFree(s->d_albedo); cudaFree(s->d_cursor);
    cudaFree(s->d_level_nodes); cudaFree(s->d_geo); cudaFree(s->d_pos); cudaFree(s->d_boxes); cudaFree(s->d_pidx);
    cudaFree(s->d_uvs); cudaFree(s->d_uvidx); cudaFree(s->d_texels); cudaFree(s->d_mattex); cudaFree(s->d_triuv);
    s->d_uvs = nullptr; s->d_uvidx = nullptr; s->d_texels = nullptr; s->d_mattex = nullptr; s->d_triuv = nullptr;
    s->tex = TexArgs{nullptr, nullptr, nullptr};
    s->d_blocks = nullptr; s->d_attr = nullptr;
    s->d_albedo = nullptr; s->d_cursor = nullptr;
    s->d_level_nodes = nullptr; s->d_geo = nullptr; s->d_pos = nullptr; s->d_boxes = nullptr; s->d_pidx = nullptr;
    s->committed = false;
}

extern "C" void wpt_scene_destroy(wpt_scene *s)
{
    if (!s) return;
    scene_free_device(s);
    delete s;
}

// ---- tree cache (wpt_scene_commit_cached): the host-built tree of a mesh as one file, so that the G ranks of a
// multi-GPU job (one process per GPU, same mesh) build it once instead of G times
namespace {
struct TreeFileHeader {
    uint64_t magic, mesh_hash;
    uint32_t version, n_tris, n_nodes, max_depth;
    uint64_t n_blocks, n_level_nodes, n_level_begin;
    double sah_cost, build_ms;
    uint64_t layout_tag, payload_hash;
};
constexpr uint64_t kTreeMagic = 0x3254504257ull;   // "WBPT2"
// bump when bvh_build.cpp starts to produce different trees or block contents for the same mesh: WPT_VERSION alone
// does not change with every builder edit, and a file left behind by an older build must not be taken for current
constexpr uint64_t kBuilderRevision = 3;
constexpr uint64_t kTreeLayoutTag = (uint64_t)sizeof(Block64) | (uint64_t)kWidth << 8 | (uint64_t)kMaxLeafTris << 16 |
                                    (uint64_t)kStackSize << 24 | kBuilderRevision << 32;
uint64_t fnv1a(const void *p, size_t n, uint64_t h)
{
    const uint8_t *b = static_cast<const uint8_t *>(p);
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 0x100000001b3ull; }
    return h;
}
// FNV-1a over 64-bit words (n a multiple of 4 bytes; a trailing half word is folded in as is): the payload of a 10 M
// triangle tree is 0.7 GB, byte-wise hashing would cost more than the disk read
uint64_t fnv1a_words(const void *p, size_t n, uint64_t h)
{
    const uint8_t *b = static_cast<const uint8_t *>(p);
    size_t i = 0;
    for (; i + 8 <= n; i += 8) { uint64_t w; std::memcpy(&w, b + i, 8); h ^= w; h *= 0x100000001b3ull; h ^= h >> 29; }
    return fnv1a(b + i, n - i, h);
}
uint64_t tree_payload_hash(const BuiltBvh &b)
{
    uint64_t h = 0xcbf29ce484222325ull;
    h = fnv1a_words(b.blocks.data(), b.blocks.size() * sizeof(Block64), h);
    h = fnv1a_words(b.level_nodes.data(), b.level_nodes.size() * sizeof(uint32_t), h);
    return fnv1a_words(b.level_begin.data(), b.level_begin.size() * sizeof(uint32_t), h);
}
uint64_t mesh_hash(const wpt_scene *s)
{
    uint64_t h = 0xcbf29ce484222325ull;
    h = fnv1a(s->pos.data(), s->pos.size() * sizeof(float), h);
    return fnv1a(s->pidx.data(), s->pidx.size() * sizeof(int32_t), h);
}
// What the kernels take on trust from a tree: max_depth picks the traversal stack (a fixed shared-memory stack for
// shallow trees), `base` + child counts are followed without a range check, level_begin drives the refit sweep.  A
// cache file is outside input, so all of that is checked here and a file that fails any check is ignored (the tree is
// then built): header consistent, payload checksum, every node's children inside the buffer, every leaf triangle's
// face index inside the mesh, the level list a partition of the nodes.
bool tree_consistent(const TreeFileHeader &h, const BuiltBvh &b)
{
    if (h.max_depth == 0 || h.max_depth > (uint32_t)kStackSize - 1) return false;
    if (h.n_level_begin < 2 || h.n_level_begin != (uint64_t)h.max_depth + 1) return false;
    if (h.n_level_nodes != h.n_nodes || h.n_nodes == 0) return false;
    if (b.level_begin.front() != 0 || b.level_begin.back() != h.n_level_nodes) return false;
    for (size_t k = 1; k < b.level_begin.size(); k++)
        if (b.level_begin[k] < b.level_begin[k - 1]) return false;
    if (b.level_begin[1] != 1 || b.level_nodes[0] != 0) return false;              // level 0 is the root, block 0
    uint64_t n_tri_blocks = 0;
    for (uint32_t idx : b.level_nodes) {
        if (idx >= h.n_blocks) return false;
        const Node7 &n = b.blocks[idx].node;
        const uint32_t imask = n.yhi[7], cnt2 = (uint32_t)n.zlo[7] | ((uint32_t)n.zhi[7] << 8);
        if (imask & 0x80u) return false;
        uint32_t n_int = 0, n_leaf = 0;
        for (int s = 0; s < kWidth; s++) {
            const uint32_t cnt = (cnt2 >> (2 * s)) & 3u;
            if ((imask >> s) & 1u) { n_int++; if (cnt) return false; }
            else n_leaf += cnt;
        }
        if (n.base == 0 || (uint64_t)n.base + n_int + n_leaf > h.n_blocks) return false;
        for (uint32_t k = 0; k < n_leaf; k++) {
            const int32_t prim = b.blocks[(size_t)n.base + n_int + k].tri.prim;
            if (prim < 0 || (uint32_t)prim >= h.n_tris) return false;
        }
        n_tri_blocks += n_leaf;
    }
    return n_tri_blocks + h.n_nodes == h.n_blocks;
}
bool tree_load(const char *path, uint64_t hash, uint32_t n_tris, BuiltBvh &b)
{
    FILE *f = std::fopen(path, "rb");
    if (!f) return false;
    TreeFileHeader h;
    bool ok = std::fread(&h, sizeof h, 1, f) == 1 && h.magic == kTreeMagic && h.version == WPT_VERSION &&
              h.layout_tag == kTreeLayoutTag && h.mesh_hash == hash && h.n_tris == n_tris && h.n_blocks >= 1 &&
              h.n_blocks <= kMaxBlocks && h.n_level_nodes <= h.n_blocks && h.n_level_begin <= 64;
    if (ok) {
        // the sizes must be those of the file before anything is allocated for them
        const long at = std::ftell(f);
        ok = at >= 0 && std::fseek(f, 0, SEEK_END) == 0;
        const long end = ok ? std::ftell(f) : -1;
        ok = ok && end >= 0 && (uint64_t)(end - at) == h.n_blocks * sizeof(Block64) + (h.n_level_nodes + h.n_level_begin) * sizeof(uint32_t) &&
             std::fseek(f, at, SEEK_SET) == 0;
    }
    if (ok) {
        b.blocks.resize(h.n_blocks); b.level_nodes.resize(h.n_level_nodes); b.level_begin.resize(h.n_level_begin);
        ok = std::fread(b.blocks.data(), sizeof(Block64), h.n_blocks, f) == h.n_blocks &&
             std::fread(b.level_nodes.data(), sizeof(uint32_t), h.n_level_nodes, f) == h.n_level_nodes &&
             std::fread(b.level_begin.data(), sizeof(uint32_t), h.n_level_begin, f) == h.n_level_begin;
        ok = ok && tree_payload_hash(b) == h.payload_hash && tree_consistent(h, b);
        b.n_nodes = h.n_nodes; b.n_tris = h.n_tris; b.max_depth = h.max_depth; b.sah_cost = h.sah_cost; b.build_ms = h.build_ms;
    }
    std::fclose(f);
    if (!ok) b = BuiltBvh();
    return ok;
}
void tree_store(const char *path, uint64_t hash, const BuiltBvh &b)
{
    const std::string tmp = std::string(path) + ".tmp" + std::to_string((unsigned long long)hash & 0xffff) + std::to_string((long)getpid());
    // 0600, and never through a file someone else put there (O_EXCL): caches tend to live in shared directories
    ::unlink(tmp.c_str());                            // left by a run of the same pid that died half-way
    const int fd = ::open(tmp.c_str(), O_WRONLY | O_CREAT | O_EXCL | O_NOFOLLOW, 0600);
    FILE *f = fd >= 0 ? ::fdopen(fd, "wb") : nullptr;
    if (!f) { if (fd >= 0) ::close(fd); return; }     // a cache that cannot be written is not an error
    TreeFileHeader h{kTreeMagic, hash, WPT_VERSION, b.n_tris, b.n_nodes, b.max_depth, b.blocks.size(), b.level_nodes.size(),
                     b.level_begin.size(), b.sah_cost, b.build_ms, kTreeLayoutTag, tree_payload_hash(b)};
    const bool ok = std::fwrite(&h, sizeof h, 1, f) == 1 && std::fwrite(b.blocks.data(), sizeof(Block64), b.blocks.size(), f) == b.blocks.size() &&
                    std::fwrite(b.level_nodes.data(), sizeof(uint32_t), b.level_nodes.size(), f) == b.level_nodes.size() &&
                    std::fwrite(b.level_begin.data(), sizeof(uint32_t), b.level_begin.size(), f) == b.level_begin.size();
    std::fclose(f);
    if (ok) std::rename(tmp.c_str(), path); else std::remove(tmp.c_str());     // atomic: readers see a whole file or none
}
}  // namespace

// builder 0: SAH build on the host (bvh_build.cpp); 1: Morton-order LBVH on the device (lbvh_build.cu)
static int scene_commit(wpt_scene *s, int device, int builder, const char *cache_path = nullptr)
{
    if (!s) return fail(WPT_E_INVALID, "wpt_scene_commit: null scene");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(WPT_E_CUDA, "wpt_scene_commit: no CUDA device (libwpt has no CPU fallback)");
    if (device < 0 || device >= ndev) return fail(WPT_E_INVALID, "wpt_scene_commit: bad device index");
    if (s->committed) {
        if (s->pos_on_device_newer && s->d_pos) {
            // the vertices were last moved by wpt_scene_refit_device: rebuild from those, not from the stale host copy
            // the refit and its copy into d_pos were queued on the caller's stream; a non-blocking stream (every PyTorch
            // side stream is one) is not ordered against the copy below, so wait for the whole device first
            WPT_CUDA(cudaSetDevice(s->device));
            WPT_CUDA(cudaDeviceSynchronize());
            WPT_CUDA(cudaMemcpy(s->pos.data(), s->d_pos, s->pos.size() * sizeof(float), cudaMemcpyDeviceToHost));
            for (float x : s->pos)
                if (!std::isfinite(x)) return fail(WPT_E_INVALID, "wpt_scene_commit: non-finite vertex in the device-side positions");
        }
        scene_free_device(s);
    }
    s->pos_on_device_newer = false;
    NvtxRange nvtx_commit("wpt_scene_commit");

    BuiltBvh b;
    const char *err = nullptr;
    if (builder == 0) {
        const uint64_t hash = cache_path ? mesh_hash(s) : 0;
        const bool cached = cache_path && tree_load(cache_path, hash, s->n_tris, b);
        s->tree_from_cache = cached;
        if (!cached) {
            if (!build_bvh7(s->pos.data(), (uint32_t)(s->pos.size() / 3), s->pidx.data(), s->n_tris, b, &err))
                return fail(std::strstr(err, "deeper") ? WPT_E_LIMIT : WPT_E_INVALID, std::string("BVH build: ") + err);
            if (cache_path) tree_store(cache_path, hash, b);
        }
        // the warp queue of trace_wq packs (owner lane << 27 | block index) into one word
        if (!blocks_within_limit(b.blocks.size()))
            return fail(WPT_E_LIMIT, "wpt_scene_commit: more than 2^27 BVH blocks (about 115 M triangles)");
    } else {
        // the checks build_bvh7 makes on its input
        const uint32_t n_pos = (uint32_t)(s->pos.size() / 3);
        for (int32_t v : s->pidx)
            if (v < 0 || (uint32_t)v >= n_pos) return fail(WPT_E_INVALID, "BVH build: position index out of range");
        for (float x : s->pos)
            if (!std::isfinite(x)) return fail(WPT_E_INVALID, "BVH build: non-finite vertex position");
        if (!blocks_within_limit(2ull * s->n_tris + 2ull))
            return fail(WPT_E_LIMIT, "wpt_scene_commit_lbvh: more than 2^27 BVH blocks");
    }

    // per-ORIGINAL-triangle shading attributes: 3 vertex normals + material, one 48 B gather per hit
    std::vector<TriAttr48> attr(s->n_tris);
    std::vector<uint8_t> geo(s->n_tris, 0);
    const bool has_n = !s->nidx.empty();
    for (uint32_t t = 0; t < s->n_tris; t++) {
        TriAttr48 &a = attr[t];
        bool ok = has_n;
        if (ok) for (int k = 0; k < 3; k++) ok = ok && s->nidx[3 * t + k] >= 0;
        geo[t] = ok ? 0 : 1;
        if (ok) {
            for (int k = 0; k < 3; k++) {
                a.n0[k] = s->nrm[3 * (size_t)s->nidx[3 * t + 0] + k];
                a.n1[k] = s->nrm[3 * (size_t)s->nidx[3 * t + 1] + k];
                a.n2[k] = s->nrm[3 * (size_t)s->nidx[3 * t + 2] + k];
            }
        } else {
            // documented deviation: geometric normal (src/add_math.h:9-16) instead of an OOB gather
            const float *p0 = &s->pos[3 * (size_t)s->pidx[3 * t]], *p1 = &s->pos[3 * (size_t)s->pidx[3 * t + 1]],
                        *p2 = &s->pos[3 * (size_t)s->pidx[3 * t + 2]];
            const float e1[3] = {p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2]};
            const float e2[3] = {p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2]};
            float n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
            const float l = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
            if (l > 0) { n[0] /= l; n[1] /= l; n[2] /= l; } else { n[0] = 0; n[1] = 1; n[2] = 0; }
            for (int k = 0; k < 3; k++) a.n0[k] = a.n1[k] = a.n2[k] = n[k];
        }
        a.mat = s->mat[t];
        a.pad[0] = a.pad[1] = 0;
    }

    const auto t0 = std::chrono::steady_clock::now();
    WPT_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    WPT_CUDA(cudaGetDeviceProperties(&prop, device));
    s->device = device;
    s->num_sms = prop.multiProcessorCount;
    if (builder == 0) WPT_CUDA(cudaMalloc(&s->d_blocks, b.blocks.size() * sizeof(Block64)));
    WPT_CUDA(cudaMalloc(&s->d_attr, attr.size() * sizeof(TriAttr48)));
    WPT_CUDA(cudaMalloc(&s->d_albedo, s->albedo.size() * sizeof(float)));
    WPT_CUDA(cudaMalloc(&s->d_cursor, (kCursorRing + 1) * sizeof(uint32_t)));   // + the refit flag word
    if (builder == 0) WPT_CUDA(cudaMemcpy(s->d_blocks, b.blocks.data(), b.blocks.size() * sizeof(Block64), cudaMemcpyHostToDevice));
    WPT_CUDA(cudaMemcpy(s->d_attr, attr.data(), attr.size() * sizeof(TriAttr48), cudaMemcpyHostToDevice));
    WPT_CUDA(cudaMemcpy(s->d_albedo, s->albedo.data(), s->albedo.size() * sizeof(float), cudaMemcpyHostToDevice));
    if (builder == 0) WPT_CUDA(cudaMalloc(&s->d_level_nodes, b.level_nodes.size() * sizeof(uint32_t)));
    WPT_CUDA(cudaMalloc(&s->d_geo, geo.size()));
    if (builder == 0) WPT_CUDA(cudaMemcpy(s->d_level_nodes, b.level_nodes.data(), b.level_nodes.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    WPT_CUDA(cudaMemcpy(s->d_geo, geo.data(), geo.size(), cudaMemcpyHostToDevice));
    if (builder == 0) s->level_begin = b.level_begin;
    // vertex positions and index triplets: refit (csrc/refit.cu) and wpt_hit_attributes (geometry normal)
    WPT_CUDA(cudaMalloc(&s->d_pos, s->pos.size() * sizeof(float)));
    WPT_CUDA(cudaMalloc(&s->d_pidx, s->pidx.size() * sizeof(int32_t)));
    WPT_CUDA(cudaMemcpy(s->d_pos, s->pos.data(), s->pos.size() * sizeof(float), cudaMemcpyHostToDevice));
    WPT_CUDA(cudaMemcpy(s->d_pidx, s->pidx.data(), s->pidx.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    uint32_t n_blocks = (uint32_t)b.blocks.size();
    if (builder != 0) {
        // device build: topology from the sorted Morton codes, then the geometry of every block (triangle operands,
        // float boxes, quantised planes) by the refit kernels, deepest level first
        const auto tb0 = std::chrono::steady_clock::now();
        NvtxRange nvtx_lbvh("lbvh_build");
        if (!lbvh_build_topology(s->d_pos, s->d_pidx, s->n_tris, &s->d_blocks, &n_blocks, &s->d_level_nodes, s->level_begin,
                                 &b.n_nodes, &b.max_depth, &err, nullptr))
        {
            scene_free_device(s);
            return fail(std::strstr(err, "deeper") ? WPT_E_LIMIT : WPT_E_CUDA, std::string("LBVH build: ") + err);
        }
        WPT_CUDA(cudaMalloc(&s->d_boxes, (size_t)n_blocks * 6 * sizeof(float)));
        launch_refit(s->d_blocks, s->d_level_nodes, s->level_begin.data(), (uint32_t)s->level_begin.size() - 1, s->d_pos,
                     s->d_pidx, s->d_boxes, s->d_attr, s->d_geo, nullptr);
        WPT_CUDA(cudaDeviceSynchronize());
        b.n_tris = s->n_tris;
        b.sah_cost = 0.0;       // not evaluated for the device build
        b.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tb0).count();
    }
    if (!s->uvs.empty()) {
        WPT_CUDA(cudaMalloc(&s->d_uvs, s->uvs.size() * sizeof(float)));
        WPT_CUDA(cudaMalloc(&s->d_uvidx, s->uvidx.size() * sizeof(int32_t)));
        WPT_CUDA(cudaMemcpy(s->d_uvs, s->uvs.data(), s->uvs.size() * sizeof(float), cudaMemcpyHostToDevice));
        WPT_CUDA(cudaMemcpy(s->d_uvidx, s->uvidx.data(), s->uvidx.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    }
    // image textures: one float4 texel pool, a window per material, three texcoords per original triangle
    bool any_tex = false;
    for (size_t m = 0; m < s->tex_rgb.size(); m++) any_tex = any_tex || (s->tex_w[m] * s->tex_h[m] > 1u);
    if (any_tex) {
        const size_t n_mat = s->albedo.size() / 3;
        std::vector<MatTex> mt(n_mat, MatTex{0u, 1u, 1u, 0u});
        std::vector<float4> texels;
        for (size_t m = 0; m < n_mat && m < s->tex_rgb.size(); m++) {
            if (s->tex_w[m] * s->tex_h[m] <= 1u) continue;
            mt[m] = MatTex{(uint32_t)texels.size(), s->tex_w[m], s->tex_h[m], 0u};
            for (size_t k = 0; k < (size_t)s->tex_w[m] * s->tex_h[m]; k++)
                texels.push_back(make_float4(s->tex_rgb[m][3 * k], s->tex_rgb[m][3 * k + 1], s->tex_rgb[m][3 * k + 2], 0.0f));
        }
        std::vector<float2> triuv(3 * (size_t)s->n_tris, make_float2(0.0f, 0.0f));
        if (!s->uvs.empty())
            for (size_t k = 0; k < 3 * (size_t)s->n_tris; k++) {
                // a face without texcoords (any index < 0) evaluates at (0, 0) in all three corners
                const size_t f = k / 3;
                const bool ok = s->uvidx[3 * f] >= 0 && s->uvidx[3 * f + 1] >= 0 && s->uvidx[3 * f + 2] >= 0;
                if (ok) triuv[k] = make_float2(s->uvs[2 * (size_t)s->uvidx[k]], s->uvs[2 * (size_t)s->uvidx[k] + 1]);
            }
        WPT_CUDA(cudaMalloc(&s->d_texels, texels.size() * sizeof(float4)));
        WPT_CUDA(cudaMalloc(&s->d_mattex, mt.size() * sizeof(MatTex)));
        WPT_CUDA(cudaMalloc(&s->d_triuv, triuv.size() * sizeof(float2)));
        WPT_CUDA(cudaMemcpy(s->d_texels, texels.data(), texels.size() * sizeof(float4), cudaMemcpyHostToDevice));
        WPT_CUDA(cudaMemcpy(s->d_mattex, mt.data(), mt.size() * sizeof(MatTex), cudaMemcpyHostToDevice));
        WPT_CUDA(cudaMemcpy(s->d_triuv, triuv.data(), triuv.size() * sizeof(float2), cudaMemcpyHostToDevice));
        s->tex = TexArgs{s->d_texels, s->d_mattex, s->d_triuv};
    }
    WPT_CUDA(cudaDeviceSynchronize());
    const double up_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();

    s->view.blocks = reinterpret_cast<const uint4 *>(s->d_blocks);
    s->view.n_blocks = n_blocks;
    s->view.n_nodes = b.n_nodes;
    s->view.n_tris = b.n_tris;
    s->view.max_depth = b.max_depth;
    s->stats.n_triangles = s->n_tris;
    s->stats.n_nodes = b.n_nodes;
    s->stats.n_leaf_slots = b.n_tris;
    s->stats.max_depth = b.max_depth;
    s->stats.sah_cost = b.sah_cost;
    s->stats.build_ms = b.build_ms;
    s->stats.upload_ms = up_ms;
    s->stats.node_bytes = (uint64_t)b.n_nodes * sizeof(Block64);
    s->stats.tri_bytes = (uint64_t)b.n_tris * sizeof(Block64);
    s->stats.attr_bytes = attr.size() * sizeof(TriAttr48);
    s->committed = true;
    return WPT_OK;
}

extern "C" int wpt_scene_commit(wpt_scene *s, int device) { return scene_commit(s, device, 0); }
extern "C" int wpt_scene_commit_lbvh(wpt_scene *s, int device) { return scene_commit(s, device, 1); }
extern "C" int wpt_scene_commit_cached(wpt_scene *s, int device, const char *cache_path, int *from_cache)
{
    if (!cache_path) return fail(WPT_E_INVALID, "wpt_scene_commit_cached: null path");
    const int rc = scene_commit(s, device, 0, cache_path);
    if (rc == WPT_OK && from_cache) *from_cache = s->tree_from_cache ? 1 : 0;
    return rc;
}

extern "C" int wpt_scene_stats(const wpt_scene *s, wpt_bvh_stats *out)
{
    if (!s || !out) return fail(WPT_E_INVALID, "wpt_scene_stats: null argument");
    if (!s->committed) return fail(WPT_E_STATE, "wpt_scene_stats: scene not committed");
    *out = s->stats;
    return WPT_OK;
}

// ------------------------------------------------------------------------------------------------
// refit for moved vertices (same topology): csrc/refit.cu
static int refit_common(wpt_scene *s, const float *d_positions, cudaStream_t st)
{
    if (!s->d_boxes) WPT_CUDA(cudaMalloc(&s->d_boxes, (size_t)s->view.n_blocks * 6 * sizeof(float)));
    // flag word: cursor slot kCursorRing (one past the ring) — set by the refit kernels on a non-finite vertex
    uint32_t *flag = s->d_cursor + kCursorRing;
    WPT_CUDA(cudaMemsetAsync(flag, 0, sizeof(uint32_t), st));
    launch_refit(s->d_blocks, s->d_level_nodes, s->level_begin.data(), (uint32_t)s->level_begin.size() - 1, d_positions,
                 s->d_pidx, s->d_boxes, s->d_attr, s->d_geo, st, flag);
    WPT_CUDA(cudaGetLastError());
    return WPT_OK;
}

extern "C" int wpt_scene_refit_device(wpt_scene *s, const float *d_positions, void *stream)
{
    if (!s || !d_positions) return fail(WPT_E_INVALID, "wpt_scene_refit_device: null argument");
    if (!s->committed) return fail(WPT_E_STATE, "wpt_scene_refit_device: scene not committed");
    WPT_CUDA(cudaSetDevice(s->device));
    // keep the scene's own device copy current (wpt_hit_attributes reads it; the next wpt_scene_commit reads it back)
    if (d_positions != s->d_pos)
        WPT_CUDA(cudaMemcpyAsync(s->d_pos, d_positions, s->pos.size() * sizeof(float), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    s->pos_on_device_newer = true;
    return refit_common(s, s->d_pos, (cudaStream_t)stream);
}

// Result of the last refit: 0 = fine, 1 = a non-finite vertex reached the tree (the device path cannot validate its
// input beforehand).  Waits for `stream`.
extern "C" int wpt_scene_refit_status(const wpt_scene *s, int *non_finite, void *stream)
{
    if (!s || !non_finite) return fail(WPT_E_INVALID, "wpt_scene_refit_status: null argument");
    if (!s->committed) return fail(WPT_E_STATE, "wpt_scene_refit_status: scene not committed");
    WPT_CUDA(cudaSetDevice(s->device));
    uint32_t f = 0;
    WPT_CUDA(cudaMemcpyAsync(&f, s->d_cursor + kCursorRing, sizeof f, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    WPT_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    *non_finite = f ? 1 : 0;
    return WPT_OK;
}

extern "C" int wpt_scene_refit(wpt_scene *s, const float *positions, uint32_t n_positions)
{
    if (!s || !positions) return fail(WPT_E_INVALID, "wpt_scene_refit: null argument");
    if (!s->committed) return fail(WPT_E_STATE, "wpt_scene_refit: scene not committed");
    if ((size_t)n_positions * 3 != s->pos.size()) return fail(WPT_E_INVALID, "wpt_scene_refit: vertex count differs from the committed mesh");
    for (size_t k = 0; k < s->pos.size(); k++)
        if (!std::isfinite(positions[k])) return fail(WPT_E_INVALID, "wpt_scene_refit: non-finite vertex");
    WPT_CUDA(cudaSetDevice(s->device));
    s->pos.assign(positions, positions + s->pos.size());
    s->pos_on_device_newer = false;
    WPT_CUDA(cudaMemcpy(s->d_pos, positions, s->pos.size() * sizeof(float), cudaMemcpyHostToDevice));
    if (int rc = refit_common(s, s->d_pos, nullptr)) return rc;
    WPT_CUDA(cudaDeviceSynchronize());
    return WPT_OK;
}

// ------------------------------------------------------------------------------------------------
// OBJ / MTL ingestion (src/main.cpp:133-251)
struct wpt_obj {
    ObjMesh mesh;
    std::vector<float> albedo;   // 3 * (1 + M): default material first (src/main.cpp:229-234)
};

extern "C" int wpt_obj_load(const char *path, int use_kd, wpt_obj **out)
{
    if (!path || !out) return fail(WPT_E_INVALID, "wpt_obj_load: null argument");
    wpt_obj *o = new (std::nothrow) wpt_obj();
    if (!o) return fail(WPT_E_INVALID, "out of host memory");
    std::string err;
    if (!load_obj(path, o->mesh, err)) {
        delete o;
        return fail(err.rfind("cannot open", 0) == 0 ? WPT_E_IO : WPT_E_INVALID, "wpt_obj_load: " + err);
    }
    if (o->mesh.material_id.empty()) {
        delete o;
        return fail(WPT_E_INVALID, std::string("wpt_obj_load: no faces in ") + path);
    }
    const size_t M = o->mesh.material_names.size();
    o->albedo.assign(3 * (M + 1), 1.0f);                       // src/main.cpp:234,244: Spectrum(1.0)
    if (use_kd)
        for (size_t k = 0; k < 3 * M; k++) o->albedo[3 + k] = o->mesh.kd[k];   // src/main.cpp:239-242
    *out = o;
    return WPT_OK;
}

extern "C" int wpt_obj_mesh(const wpt_obj *o, wpt_mesh_desc *d)
{
    if (!o || !d) return fail(WPT_E_INVALID, "wpt_obj_mesh: null argument");
    const ObjMesh &m = o->mesh;
    std::memset(d, 0, sizeof *d);
    d->positions = m.positions.data();
    d->position_triplets = m.position_triplets.data();
    d->n_positions = (uint32_t)(m.positions.size() / 3);
    d->n_triangles = (uint32_t)m.material_id.size();
    if (!m.normals.empty()) {
        d->normals = m.normals.data();
        d->normal_triplets = m.normal_triplets.data();
        d->n_normals = (uint32_t)(m.normals.size() / 3);
    }
    d->material_id = m.material_id.data();
    d->albedo = o->albedo.data();
    d->n_materials = (uint32_t)(o->albedo.size() / 3);
    return WPT_OK;
}

extern "C" int wpt_obj_texcoords(const wpt_obj *o, const float **uvs, const int32_t **uv_triplets, uint32_t *n_uvs)
{
    if (!o || !uvs || !uv_triplets || !n_uvs) return fail(WPT_E_INVALID, "wpt_obj_texcoords: null argument");
    *uvs = o->mesh.texcoords.data();
    *uv_triplets = o->mesh.texcoord_triplets.data();
    *n_uvs = (uint32_t)(o->mesh.texcoords.size() / 2);
    return WPT_OK;
}

extern "C" const char *wpt_obj_material_name(const wpt_obj *o, uint32_t i)
{
    if (!o || i == 0 || i > o->mesh.material_names.size()) return "";
    return o->mesh.material_names[i - 1].c_str();
}

extern "C" void wpt_obj_free(wpt_obj *o) { delete o; }

// ------------------------------------------------------------------------------------------------
// WPT_INTERSECT_CTAS (experiments only): CTAs per SM of the persistent trace kernel behind wpt_intersect
static TraceTuning default_tune_init()
{
    const char *e = std::getenv("WPT_INTERSECT_CTAS");
    return TraceTuning{0, e ? std::atoi(e) : 0};
}
static TraceTuning g_default_tune = default_tune_init();

extern "C" int wpt_intersect(const wpt_scene *s, const wpt_ray_streams *rays, const uint8_t *mask,
                             uint32_t mask_size, uint32_t n, const wpt_hit_streams *hits, void *stream)
{
    if (!s || !rays || !hits) return fail(WPT_E_INVALID, "wpt_intersect: null argument");
    if (!s->committed) return fail(WPT_E_STATE, "wpt_intersect: scene not committed");
    if (n == 0) return WPT_OK;
    if (!rays->ox || !rays->oy || !rays->oz || !rays->dx || !rays->dy || !rays->dz || !hits->tri_id || !hits->t ||
        !hits->u || !hits->v)
        return fail(WPT_E_INVALID, "wpt_intersect: null stream");
    if (mask && mask_size != 1 && mask_size != n) return fail(WPT_E_INVALID, "wpt_intersect: mask_size must be 1 or n");
    cudaStream_t st = (cudaStream_t)stream;
    WPT_CUDA(cudaSetDevice(s->device));
    StreamIO io;
    io.ox = rays->ox; io.oy = rays->oy; io.oz = rays->oz; io.dx = rays->dx; io.dy = rays->dy; io.dz = rays->dz;
    io.tmin = rays->tmin; io.tmax = rays->tmax; io.tmin_s = rays->tmin_scalar; io.tmax_s = rays->tmax_scalar;
    io.mask = mask; io.mask_size = mask_size;
    io.tri_id = hits->tri_id; io.t = hits->t; io.u = hits->u; io.v = hits->v;
    uint32_t *cursor = s->d_cursor + (s->cursor_rr.fetch_add(1u, std::memory_order_relaxed) & (kCursorRing - 1u));
    WPT_CUDA(cudaMemsetAsync(cursor, 0, sizeof(uint32_t), st));
    launch_trace_streams(s->view, io, n, cursor, s->num_sms, g_default_tune, st);
    WPT_CUDA(cudaGetLastError());
    return WPT_OK;
}

extern "C" int wpt_intersect_host(const wpt_scene *s, const wpt_ray_streams *rays, const uint8_t *mask,
                                  uint32_t mask_size, uint32_t n, const wpt_hit_streams *hits)
{
    if (!s || !rays || !hits) return fail(WPT_E_INVALID, "wpt_intersect_host: null argument");
    if (!s->committed) return fail(WPT_E_STATE, "wpt_intersect_host: scene not committed");
    if (n == 0) return WPT_OK;
    WPT_CUDA(cudaSetDevice(s->device));
    const size_t fb = sizeof(float) * (size_t)n;
    float *d = nullptr;          // 8 ray streams + 3 float result streams + 1 int stream
    uint8_t *dm = nullptr;
    WPT_CUDA(cudaMalloc(&d, fb * 12));
    const float *src[8] = {rays->ox, rays->oy, rays->oz, rays->dx, rays->dy, rays->dz, rays->tmin, rays->tmax};
    wpt_ray_streams dr = *rays;
    const float **dst[8] = {&dr.ox, &dr.oy, &dr.oz, &dr.dx, &dr.dy, &dr.dz, &dr.tmin, &dr.tmax};
    int rc = WPT_OK;
    for (int k = 0; k < 8 && rc == WPT_OK; k++) {
        if (!src[k]) { if (k < 6) rc = fail(WPT_E_INVALID, "wpt_intersect_host: null stream"); continue; }
        if (cudaMemcpy(d + (size_t)k * n, src[k], fb, cudaMemcpyHostToDevice) != cudaSuccess) rc = fail(WPT_E_CUDA, "H2D copy failed");
        *dst[k] = d + (size_t)k * n;
    }
    if (rc == WPT_OK && mask) {
        const size_t mb = mask_size == 1 ? 1 : n;
        if (cudaMalloc(&dm, mb) != cudaSuccess || cudaMemcpy(dm, mask, mb, cudaMemcpyHostToDevice) != cudaSuccess)
            rc = fail(WPT_E_CUDA, "mask copy failed");
    }
    wpt_hit_streams dh;
    dh.tri_id = reinterpret_cast<int32_t *>(d + (size_t)8 * n);
    dh.t = d + (size_t)9 * n; dh.u = d + (size_t)10 * n; dh.v = d + (size_t)11 * n;
    if (rc == WPT_OK) {
        // lanes the kernel leaves untouched must keep the caller's values
        cudaMemcpy(dh.tri_id, hits->tri_id, fb, cudaMemcpyHostToDevice);
        cudaMemcpy(dh.t, hits->t, fb, cudaMemcpyHostToDevice);
        cudaMemcpy(dh.u, hits->u, fb, cudaMemcpyHostToDevice);
        cudaMemcpy(dh.v, hits->v, fb, cudaMemcpyHostToDevice);
        rc = wpt_intersect(s, &dr, dm, mask_size, n, &dh, nullptr);
    }
    if (rc == WPT_OK) {
        if (cudaMemcpy(hits->tri_id, dh.tri_id, fb, cudaMemcpyDeviceToHost) != cudaSuccess ||
            cudaMemcpy(hits->t, dh.t, fb, cudaMemcpyDeviceToHost) != cudaSuccess ||
            cudaMemcpy(hits->u, dh.u, fb, cudaMemcpyDeviceToHost) != cudaSuccess ||
            cudaMemcpy(hits->v, dh.v, fb, cudaMemcpyDeviceToHost) != cudaSuccess)
            rc = fail(WPT_E_CUDA, std::string("D2H copy failed: ") + cudaGetErrorString(cudaGetLastError()));
    }
    cudaFree(d);
    cudaFree(dm);
    return rc;
}

// ------------------------------------------------------------------------------------------------
// wpt_intersect_host_static: wpt_intersect_host for callers that must not allocate (include/wpt.h).  The device-side
// staging that wpt_intersect_host cudaMallocs per call — 8 ray streams, 4 hit streams, the mask, the kernel's ray-fetch
// cursor — is carved out of ONE array that is part of the library image.  Every device that loads libwpt.so holds a
// copy of it (about 12.3 MiB) until the process exits.
constexpr uint32_t kStaticMaxRays = WPT_STATIC_MAX_RAYS;
constexpr size_t kStaticCursorBytes = 256;                                    // keeps the streams behind it 256-aligned
constexpr size_t kStaticStreamBytes = sizeof(float) * (size_t)kStaticMaxRays; // one f32 / i32 stream at full size
constexpr size_t kStaticBytes = kStaticCursorBytes + 12 * kStaticStreamBytes + kStaticMaxRays /* mask */;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

// In host code `g_static_scratch` is only a handle for the runtime, not an address the GPU understands: the device
// address has to be asked for, per device (each one has its own copy), with that device current.
constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static unsigned char *g_static_base[kStaticMaxDevices] = {};

// `device` must be current.  Looks the address up on the first call for that device and remembers it.
static int static_scratch_base(int device, unsigned char **out)
{
    if (device < 0 || device >= kStaticMaxDevices)
        return fail(WPT_E_LIMIT, "wpt_intersect_host_static: device ordinal beyond " + std::to_string(kStaticMaxDevices - 1));
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_base[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p)
            return fail(WPT_E_CUDA, std::string("wpt_intersect_host_static: cudaGetSymbolAddress(g_static_scratch): ") +
                                        cudaGetErrorString(e));
        g_static_base[device] = static_cast<unsigned char *>(p);
    }
    *out = g_static_base[device];
    return WPT_OK;
}

extern "C" int wpt_intersect_host_static(const wpt_scene *s, const wpt_ray_streams *rays, const uint8_t *mask,
                                         uint32_t mask_size, uint32_t n, const wpt_hit_streams *hits, void *stream)
{
    if (!s || !rays || !hits) return fail(WPT_E_INVALID, "wpt_intersect_host_static: null argument");
    if (!s->committed) return fail(WPT_E_STATE, "wpt_intersect_host_static: scene not committed");
    if (n > kStaticMaxRays)
        return fail(WPT_E_LIMIT, "wpt_intersect_host_static: " + std::to_string(n) + " rays in one call, the built-in staging area holds " +
                                     std::to_string(kStaticMaxRays) + " (WPT_STATIC_MAX_RAYS): split the batch or use wpt_intersect_host");
    if (n == 0) return WPT_OK;
    if (!rays->ox || !rays->oy || !rays->oz || !rays->dx || !rays->dy || !rays->dz || !hits->tri_id || !hits->t ||
        !hits->u || !hits->v)
        return fail(WPT_E_INVALID, "wpt_intersect_host_static: null stream");
    if (mask && mask_size != 1 && mask_size != n)
        return fail(WPT_E_INVALID, "wpt_intersect_host_static: mask_size must be 1 or n");
    cudaStream_t st = (cudaStream_t)stream;
    WPT_CUDA(cudaSetDevice(s->device));
    unsigned char *base = nullptr;
    if (int rc = static_scratch_base(s->device, &base)) return rc;

    // layout: [cursor | 12 streams of kStaticMaxRays 4-byte items | mask]
    uint32_t *cursor = reinterpret_cast<uint32_t *>(base);
    float *d = reinterpret_cast<float *>(base + kStaticCursorBytes);
    uint8_t *dm = base + kStaticCursorBytes + 12 * kStaticStreamBytes;
    auto slot = [&](int k) { return d + (size_t)k * kStaticMaxRays; };
    const size_t fb = sizeof(float) * (size_t)n;

    StreamIO io;
    const float *src[8] = {rays->ox, rays->oy, rays->oz, rays->dx, rays->dy, rays->dz, rays->tmin, rays->tmax};
    const float **dst[8] = {&io.ox, &io.oy, &io.oz, &io.dx, &io.dy, &io.dz, &io.tmin, &io.tmax};
    for (int k = 0; k < 8; k++) {
        *dst[k] = nullptr;
        if (!src[k]) continue;                  // optional tmin / tmax stream: the scalar is used
        WPT_CUDA(cudaMemcpyAsync(slot(k), src[k], fb, cudaMemcpyHostToDevice, st));
        *dst[k] = slot(k);
    }
    io.tmin_s = rays->tmin_scalar; io.tmax_s = rays->tmax_scalar;
    io.mask = nullptr; io.mask_size = mask_size;
    if (mask) {
        WPT_CUDA(cudaMemcpyAsync(dm, mask, mask_size == 1 ? 1 : n, cudaMemcpyHostToDevice, st));
        io.mask = dm;
    }
    // The area still holds the previous call's results.  The kernel writes t, u, v only where it finds a hit and
    // nothing at all on masked lanes, and those lanes must come back with the caller's values: upload them first.
    io.tri_id = reinterpret_cast<int32_t *>(slot(8)); io.t = slot(9); io.u = slot(10); io.v = slot(11);
    WPT_CUDA(cudaMemcpyAsync(io.tri_id, hits->tri_id, fb, cudaMemcpyHostToDevice, st));
    WPT_CUDA(cudaMemcpyAsync(io.t, hits->t, fb, cudaMemcpyHostToDevice, st));
    WPT_CUDA(cudaMemcpyAsync(io.u, hits->u, fb, cudaMemcpyHostToDevice, st));
    WPT_CUDA(cudaMemcpyAsync(io.v, hits->v, fb, cudaMemcpyHostToDevice, st));
    WPT_CUDA(cudaMemsetAsync(cursor, 0, sizeof(uint32_t), st));   // the kernel's ray counter has to start at zero
    launch_trace_streams(s->view, io, n, cursor, s->num_sms, g_default_tune, st);
    WPT_CUDA(cudaGetLastError());
    WPT_CUDA(cudaMemcpyAsync(hits->tri_id, io.tri_id, fb, cudaMemcpyDeviceToHost, st));
    WPT_CUDA(cudaMemcpyAsync(hits->t, io.t, fb, cudaMemcpyDeviceToHost, st));
    WPT_CUDA(cudaMemcpyAsync(hits->u, io.u, fb, cudaMemcpyDeviceToHost, st));
    WPT_CUDA(cudaMemcpyAsync(hits->v, io.v, fb, cudaMemcpyDeviceToHost, st));
    WPT_CUDA(cudaStreamSynchronize(st));
    return WPT_OK;
}

// ------------------------------------------------------------------------------------------------
// ThinlensCamera ctor, src/pinhole.h:9-25
extern "C" int wpt_camera_init(wpt_camera *c, const float from[3], const float at[3], const float up[3],
                               float lens_radius, float focal_dist, float fov_y, float film_size_y)
{
    if (!c || !from || !at || !up) return fail(WPT_E_INVALID, "wpt_camera_init: null argument");
    auto norm = [](const float a[3], float o[3]) {
        const float d = std::fmaf(a[2], a[2], std::fmaf(a[1], a[1], a[0] * a[0]));
        const float r = 1.0f / std::sqrt(d);
        o[0] = a[0] * r; o[1] = a[1] * r; o[2] = a[2] * r;
    };
    auto cross = [](const float a[3], const float b[3], float o[3]) {
        o[0] = std::fmaf(a[1], b[2], -(a[2] * b[1]));
        o[1] = std::fmaf(a[2], b[0], -(a[0] * b[2]));
        o[2] = std::fmaf(a[0], b[1], -(a[1] * b[0]));
    };
    const float d[3] = {at[0] - from[0], at[1] - from[1], at[2] - from[2]};
    float t[3];
    std::memcpy(c->origin, from, sizeof(float) * 3);
    norm(d, c->z);
    cross(up, c->z, t);
    norm(t, c->x);
    cross(c->z, c->x, t);
    norm(t, c->y);
    c->film_size_y = film_size_y; c->focal_dist = focal_dist; c->fov_y = fov_y; c->lens_radius = lens_radius;
    for (int k = 0; k < 3; k++)
        if (!std::isfinite(c->x[k]) || !std::isfinite(c->y[k]) || !std::isfinite(c->z[k]))
            return fail(WPT_E_INVALID, "wpt_camera_init: degenerate look_from/look_at/up");
    return WPT_OK;
}

// ------------------------------------------------------------------------------------------------
// renderer: the sample/bounce loop (src/main.cpp:385-426) as a wavefront pipeline
// ------------------------------------------------------------------------------------------------
// ---- NCCL, bound at run time (see wpt_film_reduce below)
namespace {
struct NcclApi {
    void *lib = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommInitAll) CommInitAll = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclReduce) Reduce = nullptr;
    decltype(&ncclReduceScatter) ReduceScatter = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
};
}  // namespace
static NcclApi g_nccl;
static std::mutex g_nccl_mutex;
static int nccl_load()
{
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    if (g_nccl.lib) return WPT_OK;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(WPT_E_STATE, std::string("NCCL not available: ") + dlerror());
    auto sym = [&](const char *n) { return dlsym(h, n); };
    g_nccl.GetUniqueId = reinterpret_cast<decltype(g_nccl.GetUniqueId)>(sym("ncclGetUniqueId"));
    g_nccl.CommInitRank = reinterpret_cast<decltype(g_nccl.CommInitRank)>(sym("ncclCommInitRank"));
    g_nccl.CommInitAll = reinterpret_cast<decltype(g_nccl.CommInitAll)>(sym("ncclCommInitAll"));
    g_nccl.CommDestroy = reinterpret_cast<decltype(g_nccl.CommDestroy)>(sym("ncclCommDestroy"));
    g_nccl.Reduce = reinterpret_cast<decltype(g_nccl.Reduce)>(sym("ncclReduce"));
    g_nccl.AllGather = reinterpret_cast<decltype(g_nccl.AllGather)>(sym("ncclAllGather"));
    g_nccl.ReduceScatter = reinterpret_cast<decltype(g_nccl.ReduceScatter)>(sym("ncclReduceScatter"));
    g_nccl.GroupStart = reinterpret_cast<decltype(g_nccl.GroupStart)>(sym("ncclGroupStart"));
    g_nccl.GroupEnd = reinterpret_cast<decltype(g_nccl.GroupEnd)>(sym("ncclGroupEnd"));
    g_nccl.GetErrorString = reinterpret_cast<decltype(g_nccl.GetErrorString)>(sym("ncclGetErrorString"));
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommInitAll || !g_nccl.CommDestroy || !g_nccl.Reduce || !g_nccl.AllGather || !g_nccl.ReduceScatter ||
        !g_nccl.GroupStart || !g_nccl.GroupEnd || !g_nccl.GetErrorString)
        return fail(WPT_E_STATE, "NCCL library lacks a required symbol");
    g_nccl.lib = h;
    return WPT_OK;
}
#define WPT_NCCL(call)                                                                              \
    do {                                                                                            \
        ncclResult_t e_ = (call);                                                                   \
        if (e_ != ncclSuccess) return fail(WPT_E_CUDA, std::string(#call) + ": " + g_nccl.GetErrorString(e_)); \
    } while (0)

// fused tail: 0 = automatic — from bounce 2 on when a pass holds at most 2^24 paths (late bounces of small passes are
// where a launch's fixed ramp-up/drain cost shows: C3 at 8 spp per GPU +0.9 %), off for larger passes (64 spp: -1 %);
// measured in profiles/r2_grid_c3.txt
constexpr int kDefaultTailFrom = 0;
constexpr uint64_t kAutoTailMaxPaths = 1ull << 24;
struct wpt_renderer {
    const wpt_scene *scene = nullptr;
    int W = 0, H = 0, spp_per_pass = 1;
    uint32_t capacity = 0;            // paths in flight = W*H*spp_per_pass
    RayQueue q[2]{};
    int4 *hits = nullptr;
    PassCounters *counters = nullptr; // one per pass slot, zeroed per pass
    unsigned long long *live_acc = nullptr;   // [WPT_MAX_DEPTH + 1]: rays per bounce, then the seeded rays traced twice
    float *film = nullptr;            // 3*W*H, used by wpt_render_host only
    // depth seeding (wpt_renderer_set_depth_seeding): hit distance of every pixel-centre ray and the per-pixel tmax
    // seed derived from it, W*H floats each
    float *depth = nullptr, *seed = nullptr;
    int seeding = 1;
    float seed_margin = 1.0005f;
    // sharded depth pre-pass (wpt_renderer_set_depth_shard): this rank traces lanes [rank*chunk, (rank+1)*chunk) of the
    // pixel-centre rays; the hit distances of all ranks are all-gathered (NCCL) into tlane
    void *shard_comm = nullptr;
    int shard_rank = 0, shard_n = 1;
    float *tlane = nullptr;
    uint64_t tlane_floats = 0;
    // fused tail (wpt_renderer_set_tail): bounces >= tail_from of a pass run in ONE launch of trace_wq<TailIO>
    int tail_from = kDefaultTailFrom, tail_used = -1;
    // overlap mode (wpt_renderer_set_tuning, third knob): a second set of queues on its own stream; passes
    // alternate between the two lanes so that the HBM-bound shade/raygen kernels of one overlap the
    // issue-bound trace kernel of the other.  Each lane then holds half of spp_per_pass.
    bool overlap = false;
    RayQueue q1[2]{};
    int4 *hits1 = nullptr;
    PassCounters *counters1 = nullptr;
    unsigned long long *live_acc1 = nullptr;
    cudaStream_t lane_stream[2] = {nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join[2] = {nullptr, nullptr};
    bool profile = false;
    TraceTuning tune{0, 0};
    // bookkeeping of the render in flight
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    std::vector<cudaEvent_t> prof_events;     // pairs around each kernel when profile is on
    std::vector<int> prof_kind;               // 0 raygen, 1 trace, 2 shade
    uint32_t passes = 0, launches = 0;
    int depth_in_flight = 0;
    bool in_flight = false, two_lanes_in_flight = false;
    bool captured = false;   // the last wpt_render was recorded into a CUDA graph (stream capture)
};

extern "C" void wpt_renderer_destroy(wpt_renderer *r)
{
    if (!r) return;
    if (r->scene && r->scene->device >= 0) cudaSetDevice(r->scene->device);
    for (int k = 0; k < 2; k++) { cudaFree(r->q[k].a); cudaFree(r->q[k].b); cudaFree(r->q[k].c); }
    cudaFree(r->hits); cudaFree(r->counters); cudaFree(r->live_acc); cudaFree(r->film);
    cudaFree(r->depth); cudaFree(r->seed); cudaFree(r->tlane);
    for (int k = 0; k < 2; k++) { cudaFree(r->q1[k].a); cudaFree(r->q1[k].b); cudaFree(r->q1[k].c); }
    cudaFree(r->hits1); cudaFree(r->counters1); cudaFree(r->live_acc1);
    for (int k = 0; k < 2; k++) {
        if (r->lane_stream[k]) cudaStreamDestroy(r->lane_stream[k]);
        if (r->ev_join[k]) cudaEventDestroy(r->ev_join[k]);
    }
    if (r->ev_fork) cudaEventDestroy(r->ev_fork);
    if (r->ev_begin) cudaEventDestroy(r->ev_begin);
    if (r->ev_end) cudaEventDestroy(r->ev_end);
    for (cudaEvent_t e : r->prof_events) cudaEventDestroy(e);
    delete r;
}

extern "C" int wpt_renderer_create(const wpt_scene *s, int32_t width, int32_t height, int32_t spp_per_pass,
                                   wpt_renderer **out)
{
    if (!s || !out) return fail(WPT_E_INVALID, "wpt_renderer_create: null argument");
    if (!s->committed) return fail(WPT_E_STATE, "wpt_renderer_create: scene not committed");
    if (width <= 0 || height <= 0) return fail(WPT_E_INVALID, "wpt_renderer_create: bad resolution");
    if ((uint64_t)width * height > (1u << 24)) return fail(WPT_E_LIMIT, "wpt_renderer_create: W*H > 2^24");
    if (spp_per_pass <= 0) spp_per_pass = 1;
    if (spp_per_pass > 256) return fail(WPT_E_LIMIT, "wpt_renderer_create: spp_per_pass > 256");
    const uint64_t cap = (uint64_t)width * height * spp_per_pass;
    if (cap > 0x7fffffffull) return fail(WPT_E_LIMIT, "wpt_renderer_create: too many paths in flight");
    wpt_renderer *r = new (std::nothrow) wpt_renderer();
    if (!r) return fail(WPT_E_INVALID, "out of host memory");
    r->scene = s; r->W = width; r->H = height; r->spp_per_pass = spp_per_pass; r->capacity = (uint32_t)cap;
    cudaError_t e = cudaSetDevice(s->device);
    for (int k = 0; k < 2 && e == cudaSuccess; k++) {
        e = cudaMalloc(&r->q[k].a, cap * sizeof(float4));
        if (e == cudaSuccess) e = cudaMalloc(&r->q[k].b, cap * sizeof(float4));
        if (e == cudaSuccess) e = cudaMalloc(&r->q[k].c, cap * sizeof(float2));
    }
    if (e == cudaSuccess) e = cudaMalloc(&r->hits, cap * sizeof(int4));
    if (e == cudaSuccess) e = cudaMalloc(&r->counters, sizeof(PassCounters));
    if (e == cudaSuccess) e = cudaMalloc(&r->live_acc, sizeof(unsigned long long) * (WPT_MAX_DEPTH + 1));
    if (e == cudaSuccess) e = cudaMalloc(&r->film, sizeof(float) * 3 * (size_t)width * height);
    if (e == cudaSuccess) e = cudaMalloc(&r->depth, sizeof(float) * (size_t)width * height);
    if (e == cudaSuccess) e = cudaMalloc(&r->seed, sizeof(float) * (size_t)width * height);
    if (e == cudaSuccess) e = cudaEventCreate(&r->ev_begin);
    if (e == cudaSuccess) e = cudaEventCreate(&r->ev_end);
    if (e != cudaSuccess) {
        const std::string msg = std::string("wpt_renderer_create: ") + cudaGetErrorString(e);
        wpt_renderer_destroy(r);
        return fail(WPT_E_CUDA, msg);
    }
    *out = r;
    return WPT_OK;
}

extern "C" int wpt_renderer_set_profile(wpt_renderer *r, int profile)
{
    if (!r) return fail(WPT_E_INVALID, "null renderer");
    r->profile = profile != 0;
    return WPT_OK;
}

extern "C" int wpt_renderer_set_depth_seeding(wpt_renderer *r, int on)
{
    if (!r) return fail(WPT_E_INVALID, "null renderer");
    if (r->in_flight && !r->captured) return fail(WPT_E_STATE, "wpt_renderer_set_depth_seeding: render in flight");
    r->seeding = on != 0;
    r->seed_margin = on > 1 ? (float)on * 1e-4f : 1.0005f;   // on > 1: margin in 1/10000 (tests force retries with < 10000)
    return WPT_OK;
}

extern "C" int wpt_renderer_set_depth_shard(wpt_renderer *r, void *nccl_comm, int rank, int n_ranks)
{
    if (!r) return fail(WPT_E_INVALID, "null renderer");
    if (r->in_flight && !r->captured) return fail(WPT_E_STATE, "wpt_renderer_set_depth_shard: render in flight");
    if (!nccl_comm || n_ranks <= 1) { r->shard_comm = nullptr; r->shard_rank = 0; r->shard_n = 1; return WPT_OK; }
    if (rank < 0 || rank >= n_ranks) return fail(WPT_E_INVALID, "wpt_renderer_set_depth_shard: bad rank");
    if ((uint64_t)n_ranks > (uint64_t)r->W * r->H) return fail(WPT_E_INVALID, "wpt_renderer_set_depth_shard: more ranks than pixels");
    if (int rc = nccl_load()) return rc;
    const uint64_t N = (uint64_t)r->W * r->H, chunk = (N + (uint64_t)n_ranks - 1) / (uint64_t)n_ranks;
    if (!r->tlane || r->tlane_floats < chunk * (uint64_t)n_ranks) {
        WPT_CUDA(cudaSetDevice(r->scene->device));
        cudaFree(r->tlane);
        r->tlane = nullptr;
        WPT_CUDA(cudaMalloc(&r->tlane, sizeof(float) * chunk * (uint64_t)n_ranks));
        r->tlane_floats = chunk * (uint64_t)n_ranks;
    }
    r->shard_comm = nccl_comm; r->shard_rank = rank; r->shard_n = n_ranks;
    return WPT_OK;
}

extern "C" int wpt_renderer_set_tail(wpt_renderer *r, int first_bounce)
{
    if (!r) return fail(WPT_E_INVALID, "null renderer");
    if (r->in_flight && !r->captured) return fail(WPT_E_STATE, "wpt_renderer_set_tail: render in flight");
    r->tail_from = first_bounce;
    return WPT_OK;
}

extern "C" int wpt_renderer_set_tuning(wpt_renderer *r, int trace_variant, int ctas_per_sm, int overlap_lanes)
{
    if (!r) return fail(WPT_E_INVALID, "null renderer");
    if (r->in_flight) return fail(WPT_E_STATE, "wpt_renderer_set_tuning: render in flight");
    r->tune.variant = trace_variant;
    r->tune.ctas_per_sm = ctas_per_sm;
    const bool want = overlap_lanes != 0;
    if (want && !r->hits1) {
        const uint64_t cap1 = (uint64_t)r->W * r->H * (uint64_t)((r->spp_per_pass + 1) / 2);
        cudaError_t e = cudaSetDevice(r->scene->device);
        for (int k = 0; k < 2 && e == cudaSuccess; k++) {
            e = cudaMalloc(&r->q1[k].a, cap1 * sizeof(float4));
            if (e == cudaSuccess) e = cudaMalloc(&r->q1[k].b, cap1 * sizeof(float4));
            if (e == cudaSuccess) e = cudaMalloc(&r->q1[k].c, cap1 * sizeof(float2));
        }
        if (e == cudaSuccess) e = cudaMalloc(&r->hits1, cap1 * sizeof(int4));
        if (e == cudaSuccess) e = cudaMalloc(&r->counters1, sizeof(PassCounters));
        if (e == cudaSuccess) e = cudaMalloc(&r->live_acc1, sizeof(unsigned long long) * (WPT_MAX_DEPTH + 1));
        for (int k = 0; k < 2 && e == cudaSuccess; k++) {
            e = cudaStreamCreateWithFlags(&r->lane_stream[k], cudaStreamNonBlocking);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ev_join[k], cudaEventDisableTiming);
        }
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ev_fork, cudaEventDisableTiming);
        if (e != cudaSuccess) return fail(WPT_E_CUDA, std::string("wpt_renderer_set_tuning: ") + cudaGetErrorString(e));
    }
    r->overlap = want;
    return WPT_OK;
}

static int check_cfg(const wpt_renderer *r, const wpt_render_config *cfg)
{
    if (cfg->width != r->W || cfg->height != r->H) return fail(WPT_E_INVALID, "render config resolution != renderer resolution");
    if (cfg->depth < 1 || cfg->depth > WPT_MAX_DEPTH) return fail(WPT_E_LIMIT, "depth must be in [1, 64]");
    if (cfg->spp_count < 0 || cfg->spp_begin < 0) return fail(WPT_E_INVALID, "negative spp range");
    return WPT_OK;
}

extern "C" int wpt_render(wpt_renderer *r, const wpt_camera *cam, const wpt_render_config *cfg, float *film_r,
                          float *film_g, float *film_b, void *stream)
{
    if (!r || !cam || !cfg || !film_r || !film_g || !film_b) return fail(WPT_E_INVALID, "wpt_render: null argument");
    if (int rc = check_cfg(r, cfg)) return rc;
    if (r->in_flight && !r->captured) return fail(WPT_E_STATE, "wpt_render: previous render not finished (call wpt_render_finish)");
    const wpt_scene *s = r->scene;
    cudaStream_t st = (cudaStream_t)stream;
    NvtxRange nvtx_render("wpt_render");
    WPT_CUDA(cudaSetDevice(s->device));
    // CUDA-graph capture (cudaStreamBeginCapture / torch.cuda.graph around this call): everything below is
    // capturable — kernel launches, memsets, and the fork/join of the two lane streams through events; only the
    // timing events are left out (ms_total is 0 for a captured render)
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    WPT_CUDA(cudaStreamIsCapturing(st, &cap));
    const bool capturing = cap == cudaStreamCaptureStatusActive;
    if (capturing && r->profile) return fail(WPT_E_STATE, "wpt_render: per-kernel profiling cannot be captured into a graph");
    WPT_CUDA(cudaMemsetAsync(r->live_acc, 0, sizeof(unsigned long long) * (WPT_MAX_DEPTH + 1), st));
    r->passes = 0; r->launches = 0; r->depth_in_flight = cfg->depth;
    size_t pe = 0;
    r->prof_kind.clear();
    auto mark = [&](int kind) -> cudaError_t {
        if (!r->profile) return cudaSuccess;
        for (int k = 0; k < 2; k++) {
            if (pe + k >= r->prof_events.size()) {
                cudaEvent_t e;
                cudaError_t ce = cudaEventCreate(&e);
                if (ce != cudaSuccess) return ce;
                r->prof_events.push_back(e);
            }
        }
        r->prof_kind.push_back(kind);
        return cudaEventRecord(r->prof_events[pe], st);
    };
    auto done = [&]() -> cudaError_t {
        if (!r->profile) return cudaSuccess;
        cudaError_t ce = cudaEventRecord(r->prof_events[pe + 1], st);
        pe += 2;
        return ce;
    };
    if (!capturing) WPT_CUDA(cudaEventRecord(r->ev_begin, st));
    const uint32_t N = (uint32_t)r->W * r->H;
    // overlap mode: two lanes on their own streams, forked from and joined back into `st`; each lane takes
    // every other pass of half the samples.  Per-kernel profiling
    // needs serial kernels, so it runs single-lane.
    const bool two = r->overlap && !r->profile && cfg->spp_count > 1;
    const int lane_spp = two ? (r->spp_per_pass + 1) / 2 : r->spp_per_pass;
    TraceTuning tune = r->tune;
    // depth seeding of the camera rays: one pixel-centre ray per pixel first (1/spp of the bounce-0 work), its hit
    // distance -> a tmax seed per pixel (margin x the 3x3 neighbourhood maximum, which already brackets the pixel);
    // the samples' camera rays then start with that tmax instead of cfg->tmax, which spares them the node visits
    // behind the surface (simulated on C3: -22..-41 %).  Exact: the closest hit inside the seed is the closest hit;
    // the few rays that find nothing inside it are traced a second time with the full interval.  Pinhole cameras only
    // (every ray of a pixel leaves the same point), the reference's unbounded tmax (src/ray.h:16) only, and only
    // where the pre-pass is a small fraction of the frame.
    // With a sharded pre-pass the decision must be the same on every rank (the all-gather is a collective): it then does
    // not depend on this rank's own sample count (ranks of an uneven split may hold 7 and 8 samples).
    const bool shard_on = r->shard_comm && r->shard_n > 1 && r->tlane;
    const bool seeded = r->seeding && cam->lens_radius == 0.0f && cfg->tmax >= 1e20f && (cfg->spp_count >= 8 || shard_on);
    if (seeded) {
        // every rank of a sharded frame needs the same seeds: with wpt_renderer_set_depth_shard each rank traces 1/n of the
        // pixel-centre rays and the hit distances are all-gathered, instead of every rank repeating the whole pre-pass
        const bool shard = shard_on;
        const uint32_t chunk = shard ? (N + (uint32_t)r->shard_n - 1) / (uint32_t)r->shard_n : N;
        const uint32_t lane0 = shard ? std::min(N, (uint32_t)r->shard_rank * chunk) : 0u;
        const uint32_t lanes = shard ? std::min(N - lane0, chunk) : N;
        WPT_CUDA(cudaMemsetAsync(r->counters, 0, sizeof(PassCounters), st));
        WPT_CUDA(mark(0));
        // a trailing rank of an uneven split may own no pixel at all (10 pixels over 8 ranks: chunk 2, ranks 5..7
        // empty): it launches nothing but still takes part in the all-gather
        if (lanes) launch_raygen(*cam, r->W, r->H, 0, 1, cfg->seed, r->q[0], &r->counters->qcount[0], st, nullptr, 1, lane0, lanes);
        WPT_CUDA(done());
        WPT_CUDA(mark(1));
        if (lanes)
            launch_trace_queue(s->view, r->q[0], r->hits, &r->counters->qcount[0], &r->counters->cursor[0], nullptr,
                               cfg->tmin, cfg->tmax, lanes, s->num_sms, tune, st);
        WPT_CUDA(done());
        WPT_CUDA(mark(0));
        if (shard) {
            if (lanes) launch_depth_pack(r->hits, lanes, r->tlane + (size_t)r->shard_rank * chunk, st);
            WPT_NCCL(g_nccl.AllGather(r->tlane + (size_t)r->shard_rank * chunk, r->tlane, chunk, ncclFloat32,
                                      reinterpret_cast<ncclComm_t>(r->shard_comm), st));
            launch_depth_seeds_from_lanes(r->tlane, r->W, r->H, r->seed_margin, r->depth, r->seed, st);
        } else {
            launch_depth_seeds(r->q[0].a, r->hits, r->W, r->H, r->seed_margin, r->depth, r->seed, st);
        }
        WPT_CUDA(done());
        r->launches += 4;
    }
    if (two) {
        WPT_CUDA(cudaMemsetAsync(r->live_acc1, 0, sizeof(unsigned long long) * (WPT_MAX_DEPTH + 1), st));
        WPT_CUDA(cudaEventRecord(r->ev_fork, st));
        for (int k = 0; k < 2; k++) WPT_CUDA(cudaStreamWaitEvent(r->lane_stream[k], r->ev_fork, 0));
    }
    int pass_idx = 0;
    for (int s0 = 0; s0 < cfg->spp_count; s0 += lane_spp, pass_idx++) {
        const int ns = std::min(lane_spp, cfg->spp_count - s0);
        const int spp_first = cfg->spp_begin + s0;
        const uint32_t n_max = N * (uint32_t)ns;
        const int lane = two ? (pass_idx & 1) : 0;
        cudaStream_t ls = two ? r->lane_stream[lane] : st;
        RayQueue *q = lane ? r->q1 : r->q;
        int4 *hits = lane ? r->hits1 : r->hits;
        PassCounters *counters = lane ? r->counters1 : r->counters;
        unsigned long long *live_acc = lane ? r->live_acc1 : r->live_acc;
        WPT_CUDA(cudaMemsetAsync(counters, 0, sizeof(PassCounters), ls));
        WPT_CUDA(mark(0));
        launch_raygen(*cam, r->W, r->H, spp_first, ns, cfg->seed, q[0], &counters->qcount[0], ls,
                      seeded ? r->seed : nullptr, 0);
        WPT_CUDA(done());
        r->launches++;
        // (the fused tail shades with constant albedos only: scenes with image textures run the wavefront pair)
        const int tail_from = s->tex.texels ? -1 : r->tail_from != 0 ? r->tail_from
                            : ((uint64_t)N * (uint64_t)lane_spp <= kAutoTailMaxPaths ? 2 : -1);
        const int wave_end = (tail_from >= 1 && tail_from < cfg->depth) ? tail_from : cfg->depth;
        r->tail_used = wave_end < cfg->depth ? wave_end : -1;
        for (int j = 0; j < wave_end; j++) {
            const RayQueue &qin = q[j & 1], &qout = q[(j + 1) & 1];
            const bool sj = seeded && j == 0;
            // the retry list of bounce 0 borrows the c stream of the queue that shade 0 fills afterwards
            uint32_t *retry = reinterpret_cast<uint32_t *>(qout.c);
            WPT_CUDA(mark(1));
            launch_trace_queue(s->view, qin, hits, &counters->qcount[j], &counters->cursor[j],
                               live_acc + j, cfg->tmin, cfg->tmax, n_max, s->num_sms, tune, ls,
                               sj, sj ? retry : nullptr, sj ? &counters->retry_count : nullptr, nullptr);
            if (sj) {
                launch_trace_queue(s->view, qin, hits, &counters->retry_count, &counters->retry_cursor, live_acc + WPT_MAX_DEPTH,
                                   cfg->tmin, cfg->tmax, n_max, s->num_sms, tune, ls, 0, nullptr, nullptr, retry);
                r->launches++;
            }
            WPT_CUDA(done());
            WPT_CUDA(mark(2));
            launch_shade(qin, hits, &counters->qcount[j], qout, &counters->qcount[j + 1],
                         reinterpret_cast<const float4 *>(s->d_attr), s->d_albedo, j, cfg->depth, spp_first,
                         cfg->seed, film_r, film_g, film_b, n_max, s->num_sms, ls, sj, &s->tex);
            WPT_CUDA(done());
            r->launches += 2;
        }
        if (wave_end < cfg->depth) {
            // every remaining bounce of this pass in one launch: paths are shaded and continued in place
            TailArgs ta;
            ta.tmin = cfg->tmin; ta.tmax = cfg->tmax;
            ta.tri_attr = reinterpret_cast<const float4 *>(s->d_attr); ta.albedo = s->d_albedo;
            ta.film_r = film_r; ta.film_g = film_g; ta.film_b = film_b;
            ta.bounce0 = wave_end; ta.depth = cfg->depth; ta.spp_first = spp_first; ta.seed = cfg->seed;
            WPT_CUDA(mark(1));
            launch_trace_tail(s->view, q[wave_end & 1], &counters->qcount[wave_end], &counters->cursor[wave_end], live_acc,
                              ta, n_max, s->num_sms, tune, ls);
            WPT_CUDA(done());
            r->launches++;
        }
        r->passes++;
    }
    if (two) {
        for (int k = 0; k < 2; k++) {
            WPT_CUDA(cudaEventRecord(r->ev_join[k], r->lane_stream[k]));
            WPT_CUDA(cudaStreamWaitEvent(st, r->ev_join[k], 0));
        }
    }
    r->two_lanes_in_flight = two;
    if (!capturing) WPT_CUDA(cudaEventRecord(r->ev_end, st));
    r->captured = capturing;
    WPT_CUDA(cudaGetLastError());
    r->in_flight = true;
    return WPT_OK;
}

extern "C" int wpt_render_finish(wpt_renderer *r, wpt_stats *stats, void *stream)
{
    if (!r) return fail(WPT_E_INVALID, "wpt_render_finish: null renderer");
    if (!r->in_flight) return fail(WPT_E_STATE, "wpt_render_finish: nothing in flight");
    cudaStream_t st = (cudaStream_t)stream;
    WPT_CUDA(cudaSetDevice(r->scene->device));
    WPT_CUDA(cudaStreamSynchronize(st));
    if (!r->captured) r->in_flight = false;   // a captured render may be replayed and finished any number of times
    if (!stats) return WPT_OK;
    std::memset(stats, 0, sizeof *stats);
    unsigned long long live[WPT_MAX_DEPTH + 1];
    WPT_CUDA(cudaMemcpy(live, r->live_acc, sizeof live, cudaMemcpyDeviceToHost));
    for (int j = 0; j < r->depth_in_flight; j++) { stats->live[j] = live[j]; stats->segments += live[j]; }
    stats->retraced = live[WPT_MAX_DEPTH];
    if (r->two_lanes_in_flight) {
        WPT_CUDA(cudaMemcpy(live, r->live_acc1, sizeof live, cudaMemcpyDeviceToHost));
        for (int j = 0; j < r->depth_in_flight; j++) { stats->live[j] += live[j]; stats->segments += live[j]; }
        stats->retraced += live[WPT_MAX_DEPTH];
    }
    stats->tail_from = r->tail_used;
    stats->passes = r->passes;
    stats->kernel_launches = r->launches;
    if (!r->captured) WPT_CUDA(cudaEventElapsedTime(&stats->ms_total, r->ev_begin, r->ev_end));
    if (r->profile) {
        for (size_t k = 0; k < r->prof_kind.size(); k++) {
            float ms = 0;
            WPT_CUDA(cudaEventElapsedTime(&ms, r->prof_events[2 * k], r->prof_events[2 * k + 1]));
            (r->prof_kind[k] == 0 ? stats->ms_raygen : r->prof_kind[k] == 1 ? stats->ms_trace : stats->ms_shade) += ms;
        }
    }
    return WPT_OK;
}

extern "C" int wpt_film_scale(float *film_r, float *film_g, float *film_b, uint32_t n, float scale, void *stream)
{
    if (!film_r || !film_g || !film_b) return fail(WPT_E_INVALID, "wpt_film_scale: null film");
    launch_film_scale(film_r, film_g, film_b, n, scale, (cudaStream_t)stream);
    WPT_CUDA(cudaGetLastError());
    return WPT_OK;
}

// Pinned host memory through the ABI: a film allocated here makes the three copies at the end of wpt_render_host
// asynchronous DMA transfers at PCIe/NVLink-C2C speed (a pageable destination is staged by the driver in chunks)
extern "C" int wpt_host_alloc(size_t bytes, void **out)
{
    if (!out || bytes == 0) return fail(WPT_E_INVALID, "wpt_host_alloc: bad argument");
    WPT_CUDA(cudaMallocHost(out, bytes));
    return WPT_OK;
}
extern "C" void wpt_host_free(void *p)
{
    if (p) cudaFreeHost(p);
}

// device buffers through the ABI, so that a C++ host (include/wpt_shim.hpp) needs no CUDA headers of its own — the
// reference's Enoki arrays allocate and copy for it (enoki/cuda.h:210-220)
extern "C" int wpt_device_alloc(int device, size_t bytes, void **out)
{
    if (!out || bytes == 0) return fail(WPT_E_INVALID, "wpt_device_alloc: bad argument");
    WPT_CUDA(cudaSetDevice(device));
    WPT_CUDA(cudaMalloc(out, bytes));
    return WPT_OK;
}
extern "C" void wpt_device_free(void *p)
{
    if (p) cudaFree(p);
}
extern "C" int wpt_device_upload(void *dst_device, const void *src_host, size_t bytes, void *stream)
{
    if (!dst_device || !src_host) return fail(WPT_E_INVALID, "wpt_device_upload: null argument");
    WPT_CUDA(cudaMemcpyAsync(dst_device, src_host, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return WPT_OK;
}
extern "C" int wpt_device_download(void *dst_host, const void *src_device, size_t bytes, void *stream)
{
    if (!dst_host || !src_device) return fail(WPT_E_INVALID, "wpt_device_download: null argument");
    WPT_CUDA(cudaMemcpyAsync(dst_host, src_device, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    WPT_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return WPT_OK;
}
extern "C" int wpt_device_sync(int device)
{
    WPT_CUDA(cudaSetDevice(device));
    WPT_CUDA(cudaDeviceSynchronize());
    return WPT_OK;
}
extern "C" int wpt_device_set(int device)
{
    WPT_CUDA(cudaSetDevice(device));
    return WPT_OK;
}
extern "C" int wpt_device_zero(void *dst_device, size_t bytes, void *stream)
{
    if (!dst_device) return fail(WPT_E_INVALID, "wpt_device_zero: null argument");
    WPT_CUDA(cudaMemsetAsync(dst_device, 0, bytes, (cudaStream_t)stream));
    return WPT_OK;
}
extern "C" int wpt_device_count(void)
{
    int n = 0;
    return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}

extern "C" int wpt_render_host(wpt_renderer *r, const wpt_camera *cam, const wpt_render_config *cfg,
                               float *film_r, float *film_g, float *film_b, wpt_stats *stats)
{
    if (!r || !cam || !cfg || !film_r || !film_g || !film_b) return fail(WPT_E_INVALID, "wpt_render_host: null argument");
    if (int rc = check_cfg(r, cfg)) return rc;
    NvtxRange nvtx_host("wpt_render_host");
    const size_t N = (size_t)r->W * r->H;
    WPT_CUDA(cudaSetDevice(r->scene->device));
    float *dr = r->film, *dg = r->film + N, *db = r->film + 2 * N;
    WPT_CUDA(cudaMemsetAsync(r->film, 0, sizeof(float) * 3 * N, nullptr));     // src/main.cpp:369
    if (int rc = wpt_render(r, cam, cfg, dr, dg, db, nullptr)) return rc;
    if (cfg->spp_count > 0)
        if (int rc = wpt_film_scale(dr, dg, db, (uint32_t)N, 1.0f / (float)cfg->spp_count, nullptr)) return rc;
    // src/main.cpp:437-439: three planar channels to the host, queued behind the frame; ONE synchronisation
    WPT_CUDA(cudaMemcpyAsync(film_r, dr, sizeof(float) * N, cudaMemcpyDeviceToHost, nullptr));
    WPT_CUDA(cudaMemcpyAsync(film_g, dg, sizeof(float) * N, cudaMemcpyDeviceToHost, nullptr));
    WPT_CUDA(cudaMemcpyAsync(film_b, db, sizeof(float) * N, cudaMemcpyDeviceToHost, nullptr));
    if (int rc = wpt_render_finish(r, stats, nullptr)) return rc;
    return WPT_OK;
}

// ------------------------------------------------------------------------------------------------
extern "C" int wpt_raygen(const wpt_camera *cam, int32_t width, int32_t height, int32_t spp_index, uint64_t seed,
                          float *ox, float *oy, float *oz, float *dx, float *dy, float *dz, void *stream)
{
    if (!cam || !ox || !oy || !oz || !dx || !dy || !dz) return fail(WPT_E_INVALID, "wpt_raygen: null argument");
    if (width <= 0 || height <= 0 || (uint64_t)width * height > (1u << 24)) return fail(WPT_E_LIMIT, "wpt_raygen: bad resolution");
    launch_raygen_streams(*cam, width, height, spp_index, seed, ox, oy, oz, dx, dy, dz, (cudaStream_t)stream);
    WPT_CUDA(cudaGetLastError());
    return WPT_OK;
}

extern "C" int wpt_shade_step(const wpt_scene *s, uint32_t n, const float *o_xyz[3], const float *d_xyz[3],
                              const int32_t *tri_id, const float *t, const float *u, const float *v,
                              const float *s0, const float *s1, float *next_o[3], float *next_d[3],
                              float *weight[3], void *stream)
{
    if (!s || !o_xyz || !d_xyz || !tri_id || !t || !u || !v || !s0 || !s1 || !next_o || !next_d || !weight)
        return fail(WPT_E_INVALID, "wpt_shade_step: null argument");
    if (!s->committed) return fail(WPT_E_STATE, "wpt_shade_step: scene not committed");
    WPT_CUDA(cudaSetDevice(s->device));
    launch_shade_streams(n, o_xyz, d_xyz, tri_id, t, u, v, s0, s1, reinterpret_cast<const float4 *>(s->d_attr),
                         s->d_albedo, next_o, next_d, weight, (cudaStream_t)stream, &s->tex);
    WPT_CUDA(cudaGetLastError());
    return WPT_OK;
}

extern "C" int wpt_philox(uint32_t n, uint32_t c1, uint32_t c2, uint64_t key, uint32_t *out, void *stream)
{
    if (!out) return fail(WPT_E_INVALID, "wpt_philox: null output");
    launch_philox(n, c1, c2, key, out, (cudaStream_t)stream);
    WPT_CUDA(cudaGetLastError());
    return WPT_OK;
}

// ------------------------------------------------------------------------------------------------
// multi-GPU: ONE reduce of the partial films.  NCCL is bound at run time (dlopen of libnccl.so.2:
// the copy already in the process — e.g. the one PyTorch loaded — or the system one), so libwpt.so itself has no
// link-time dependency on it and single-GPU hosts need no NCCL at all.
// ------------------------------------------------------------------------------------------------
extern "C" int wpt_nccl_unique_id(void *id128)
{
    if (!id128) return fail(WPT_E_INVALID, "wpt_nccl_unique_id: null argument");
    if (int rc = nccl_load()) return rc;
    static_assert(sizeof(ncclUniqueId) == 128, "wpt_nccl_unique_id hands out 128 bytes");
    WPT_NCCL(g_nccl.GetUniqueId(reinterpret_cast<ncclUniqueId *>(id128)));
    return WPT_OK;
}

extern "C" int wpt_nccl_comm_create(const void *id128, int n_ranks, int rank, int device, void **comm)
{
    if (!id128 || !comm || n_ranks <= 0 || rank < 0 || rank >= n_ranks) return fail(WPT_E_INVALID, "wpt_nccl_comm_create: bad argument");
    if (int rc = nccl_load()) return rc;
    WPT_CUDA(cudaSetDevice(device));
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof id);
    WPT_NCCL(g_nccl.CommInitRank(reinterpret_cast<ncclComm_t *>(comm), n_ranks, id, rank));
    return WPT_OK;
}

extern "C" int wpt_nccl_comm_create_all(int n_devices, const int *devices, void **comms)
{
    if (n_devices <= 0 || !comms) return fail(WPT_E_INVALID, "wpt_nccl_comm_create_all: bad argument");
    if (int rc = nccl_load()) return rc;
    WPT_NCCL(g_nccl.CommInitAll(reinterpret_cast<ncclComm_t *>(comms), n_devices, devices));
    return WPT_OK;
}

extern "C" void wpt_nccl_comm_destroy(void *comm)
{
    if (comm && g_nccl.lib) g_nccl.CommDestroy(reinterpret_cast<ncclComm_t>(comm));
}

extern "C" int wpt_nccl_group_start(void)
{
    if (int rc = nccl_load()) return rc;
    WPT_NCCL(g_nccl.GroupStart());
    return WPT_OK;
}
extern "C" int wpt_nccl_group_end(void)
{
    if (int rc = nccl_load()) return rc;
    WPT_NCCL(g_nccl.GroupEnd());
    return WPT_OK;
}

extern "C" int wpt_film_reduce(void *nccl_comm, float *film_r, float *film_g, float *film_b, uint32_t n, int root,
                               void *stream)
{
    if (!nccl_comm || !film_r || !film_g || !film_b) return fail(WPT_E_INVALID, "wpt_film_reduce: null argument");
    if (int rc = nccl_load()) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    ncclComm_t comm = reinterpret_cast<ncclComm_t>(nccl_comm);
    if (film_g == film_r + n && film_b == film_g + n) {
        // the usual planar film [3][W*H]: one collective of 3n floats, in place
        WPT_NCCL(g_nccl.Reduce(film_r, film_r, (size_t)3 * n, ncclFloat32, ncclSum, root, comm, st));
    } else {
        WPT_NCCL(g_nccl.GroupStart());
        WPT_NCCL(g_nccl.Reduce(film_r, film_r, n, ncclFloat32, ncclSum, root, comm, st));
        WPT_NCCL(g_nccl.Reduce(film_g, film_g, n, ncclFloat32, ncclSum, root, comm, st));
        WPT_NCCL(g_nccl.Reduce(film_b, film_b, n, ncclFloat32, ncclSum, root, comm, st));
        WPT_NCCL(g_nccl.GroupEnd());
    }
    return WPT_OK;
}

// The same exchange with the result left where it was computed: after ncclReduceScatter rank g holds the sum of floats
// [g*chunk, (g+1)*chunk) of the flattened film, in place.  No rank is special, each moves 1/n of what the root of
// wpt_film_reduce receives, and the 1/spp scale and the copy to the host that follow (wpt_film_shard_finish) touch
// 1/n of the film per rank instead of all of it on one.
extern "C" int wpt_film_shard_layout(uint32_t n, int rank, int n_ranks, uint64_t *padded_floats, uint64_t *first, uint64_t *count)
{
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(WPT_E_INVALID, "wpt_film_shard_layout: bad rank");
    const uint64_t total = 3ull * n, chunk = (total + (uint64_t)n_ranks - 1) / (uint64_t)n_ranks;
    const uint64_t lo = std::min(total, (uint64_t)rank * chunk), hi = std::min(total, lo + chunk);
    if (padded_floats) *padded_floats = chunk * (uint64_t)n_ranks;
    if (first) *first = lo;
    if (count) *count = hi - lo;
    return WPT_OK;
}

extern "C" int wpt_film_reduce_scatter(void *nccl_comm, float *film, uint32_t n, int rank, int n_ranks, void *stream)
{
    if (!nccl_comm || !film) return fail(WPT_E_INVALID, "wpt_film_reduce_scatter: null argument");
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(WPT_E_INVALID, "wpt_film_reduce_scatter: bad rank");
    if (int rc = nccl_load()) return rc;
    const uint64_t chunk = (3ull * n + (uint64_t)n_ranks - 1) / (uint64_t)n_ranks;
    // in place: NCCL takes recvbuff == sendbuff + rank * recvcount as the in-place form
    WPT_NCCL(g_nccl.ReduceScatter(film, film + (size_t)rank * chunk, chunk, ncclFloat32, ncclSum,
                                  reinterpret_cast<ncclComm_t>(nccl_comm), (cudaStream_t)stream));
    return WPT_OK;
}

extern "C" int wpt_film_shard_finish(float *film, uint32_t n, int rank, int n_ranks, float scale, float *host_film, void *stream)
{
    if (!film) return fail(WPT_E_INVALID, "wpt_film_shard_finish: null film");
    uint64_t lo = 0, cnt = 0;
    if (int rc = wpt_film_shard_layout(n, rank, n_ranks, nullptr, &lo, &cnt)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    launch_film_scale_flat(film + lo, (uint32_t)cnt, scale, st);
    WPT_CUDA(cudaGetLastError());
    if (host_film && cnt)
        WPT_CUDA(cudaMemcpyAsync(host_film + lo, film + lo, cnt * sizeof(float), cudaMemcpyDeviceToHost, st));
    return WPT_OK;
}

extern "C" int wpt_host_register(void *p, size_t bytes)
{
    if (!p || !bytes) return fail(WPT_E_INVALID, "wpt_host_register: null argument");
    WPT_CUDA(cudaHostRegister(p, bytes, cudaHostRegisterPortable));
    return WPT_OK;
}
extern "C" void wpt_host_unregister(void *p) { if (p) cudaHostUnregister(p); }

// ------------------------------------------------------------------------------------------------
// derived hit attributes (src/accel/optix_backend.h:463-486, src/main.cpp:325)
extern "C" int wpt_scene_set_texcoords(wpt_scene *s, const float *uvs, const int32_t *uv_triplets, uint32_t n_uvs)
{
    if (!s) return fail(WPT_E_INVALID, "wpt_scene_set_texcoords: null scene");
    if (s->committed) return fail(WPT_E_STATE, "wpt_scene_set_texcoords: call before wpt_scene_commit");
    if (!uvs || !uv_triplets || n_uvs == 0) { s->uvs.clear(); s->uvidx.clear(); return WPT_OK; }
    for (size_t k = 0; k < 3 * (size_t)s->n_tris; k++)
        if (uv_triplets[k] >= (int32_t)n_uvs) return fail(WPT_E_INVALID, "texcoord index out of range");
    s->uvs.assign(uvs, uvs + 2 * (size_t)n_uvs);
    s->uvidx.assign(uv_triplets, uv_triplets + 3 * (size_t)s->n_tris);
    return WPT_OK;
}

// ImageTexture of one material (src/main.cpp:35-80): the reference only ever builds 1x1 textures (ImageTexture::load is
// empty, :36-38), its eval is nevertheless the bilinear fetch — available here for real images
extern "C" int wpt_scene_set_texture(wpt_scene *s, uint32_t material, const float *rgb, uint32_t width, uint32_t height)
{
    if (!s || !rgb) return fail(WPT_E_INVALID, "wpt_scene_set_texture: null argument");
    if (s->committed) return fail(WPT_E_STATE, "wpt_scene_set_texture: call before wpt_scene_commit");
    const size_t n_mat = s->albedo.size() / 3;
    if (material >= n_mat) return fail(WPT_E_INVALID, "wpt_scene_set_texture: material out of range");
    if (width == 0 || height == 0 || (uint64_t)width * height > (1u << 26)) return fail(WPT_E_LIMIT, "wpt_scene_set_texture: bad size");
    for (size_t k = 0; k < 3 * (size_t)width * height; k++)
        if (!std::isfinite(rgb[k])) return fail(WPT_E_INVALID, "wpt_scene_set_texture: non-finite texel");
    if (s->tex_rgb.size() < n_mat) { s->tex_rgb.resize(n_mat); s->tex_w.resize(n_mat, 1u); s->tex_h.resize(n_mat, 1u); }
    s->tex_rgb[material].assign(rgb, rgb + 3 * (size_t)width * height);
    s->tex_w[material] = width; s->tex_h[material] = height;
    if (width * height == 1u) for (int k = 0; k < 3; k++) s->albedo[3 * (size_t)material + k] = rgb[k];   // ImageTexture(color)
    return WPT_OK;
}

extern "C" int wpt_hit_attributes(const wpt_scene *s, const wpt_ray_streams *rays, const wpt_hit_streams *hits,
                                  uint32_t n, const wpt_hit_attr_streams *out, void *stream)
{
    if (!s || !rays || !hits || !out) return fail(WPT_E_INVALID, "wpt_hit_attributes: null argument");
    if (!s->committed) return fail(WPT_E_STATE, "wpt_hit_attributes: scene not committed");
    if (n == 0) return WPT_OK;
    if (!hits->tri_id || !hits->u || !hits->v) return fail(WPT_E_INVALID, "wpt_hit_attributes: tri_id, u, v required");
    const bool want_p = out->px || out->py || out->pz, want_g = out->gnx || out->gny || out->gnz,
               want_s = out->snx || out->sny || out->snz, want_t = out->tu || out->tv;
    if ((want_p && !(out->px && out->py && out->pz)) || (want_g && !(out->gnx && out->gny && out->gnz)) ||
        (want_s && !(out->snx && out->sny && out->snz)) || (want_t && !(out->tu && out->tv)))
        return fail(WPT_E_INVALID, "wpt_hit_attributes: an output group must be given completely or not at all");
    if (want_p && !(rays->ox && rays->oy && rays->oz && rays->dx && rays->dy && rays->dz && hits->t))
        return fail(WPT_E_INVALID, "wpt_hit_attributes: position needs the ray streams and t");
    WPT_CUDA(cudaSetDevice(s->device));
    HitAttrArgs a;
    a.ox = rays->ox; a.oy = rays->oy; a.oz = rays->oz; a.dx = rays->dx; a.dy = rays->dy; a.dz = rays->dz;
    a.tri_id = hits->tri_id; a.t = hits->t; a.u = hits->u; a.v = hits->v;
    a.pos = s->d_pos; a.pidx = s->d_pidx; a.tri_attr = reinterpret_cast<const float4 *>(s->d_attr);
    a.uvs = s->d_uvs; a.uvidx = s->d_uvidx;
    a.p[0] = out->px; a.p[1] = out->py; a.p[2] = out->pz;
    a.ng[0] = out->gnx; a.ng[1] = out->gny; a.ng[2] = out->gnz;
    a.ns[0] = out->snx; a.ns[1] = out->sny; a.ns[2] = out->snz;
    a.uv[0] = out->tu; a.uv[1] = out->tv;
    a.mat = out->material_id;
    launch_hit_attributes(n, a, (cudaStream_t)stream);
    WPT_CUDA(cudaGetLastError());
    return WPT_OK;
}

// src/fimage.h:33-58
extern "C" int wpt_save_pfm(const char *path, const float *r, const float *g, const float *b, uint32_t width,
                            uint32_t height)
{
    if (!path || !r || !g || !b) return fail(WPT_E_INVALID, "wpt_save_pfm: null argument");
    FILE *f = std::fopen(path, "wb");
    if (!f) return fail(WPT_E_IO, std::string("wpt_save_pfm: cannot open ") + path);
    std::fprintf(f, "PF\n%u %u\n-1\n", width, height);
    std::vector<float> row(3 * (size_t)width);
    for (uint32_t y = 0; y < height; y++) {
        const size_t src = (size_t)(height - y - 1) * width;    // bottom-up (fimage.h:49)
        for (uint32_t x = 0; x < width; x++) {
            row[3 * x] = r[src + x]; row[3 * x + 1] = g[src + x]; row[3 * x + 2] = b[src + x];
        }
        if (std::fwrite(row.data(), sizeof(float), row.size(), f) != row.size()) {
            std::fclose(f);
            return fail(WPT_E_IO, "wpt_save_pfm: short write");
        }
    }
    std::fclose(f);
    return WPT_OK;
}

// ------------------------------------------------------------------------------------------------
// checkpoint of a render in progress: the UN-normalised film and how many samples it holds.  The reference
// keeps the film on the device until the end (src/main.cpp:434-442) and cannot resume; here a frame is additive over
// [spp_begin, spp_begin + spp_count) and the RNG is keyed on the global sample index, so "film so far + spp_done + seed"
// is all a later process needs to continue with spp_begin = spp_done.
namespace {
struct FilmFileHeader {
    uint64_t magic;
    uint32_t width, height;
    uint64_t spp_done, seed;
};
constexpr uint64_t kFilmMagic = 0x314d4c4946545057ull;   // "WPTFILM1"
int film_read_header(FILE *f, FilmFileHeader &h, const char *path)
{
    if (std::fread(&h, sizeof h, 1, f) != 1 || h.magic != kFilmMagic || h.width == 0 || h.height == 0)
        return fail(WPT_E_IO, std::string("wpt_film_load: not a film checkpoint: ") + path);
    return WPT_OK;
}
}  // namespace

extern "C" int wpt_film_save(const char *path, const float *r, const float *g, const float *b, uint32_t width, uint32_t height,
                             uint64_t spp_done, uint64_t seed)
{
    if (!path || !r || !g || !b || width == 0 || height == 0) return fail(WPT_E_INVALID, "wpt_film_save: null argument");
    const size_t n = (size_t)width * height;
    const std::string tmp = std::string(path) + ".tmp" + std::to_string((long)getpid());
    FILE *f = std::fopen(tmp.c_str(), "wb");
    if (!f) return fail(WPT_E_IO, std::string("wpt_film_save: cannot open ") + tmp);
    const FilmFileHeader h{kFilmMagic, width, height, spp_done, seed};
    uint64_t sum = fnv1a_words(&h, sizeof h, 0xcbf29ce484222325ull);
    sum = fnv1a_words(r, n * sizeof(float), sum);
    sum = fnv1a_words(g, n * sizeof(float), sum);
    sum = fnv1a_words(b, n * sizeof(float), sum);
    const bool ok = std::fwrite(&h, sizeof h, 1, f) == 1 && std::fwrite(r, sizeof(float), n, f) == n &&
                    std::fwrite(g, sizeof(float), n, f) == n && std::fwrite(b, sizeof(float), n, f) == n &&
                    std::fwrite(&sum, sizeof sum, 1, f) == 1;
    const bool closed = std::fclose(f) == 0;
    // a reader sees the previous checkpoint or this one, never half of it
    if (!ok || !closed || std::rename(tmp.c_str(), path) != 0) {
        std::remove(tmp.c_str());
        return fail(WPT_E_IO, std::string("wpt_film_save: cannot write ") + path);
    }
    return WPT_OK;
}

extern "C" int wpt_film_info(const char *path, uint32_t *width, uint32_t *height, uint64_t *spp_done, uint64_t *seed)
{
    if (!path) return fail(WPT_E_INVALID, "wpt_film_info: null path");
    FILE *f = std::fopen(path, "rb");
    if (!f) return fail(WPT_E_IO, std::string("wpt_film_info: cannot open ") + path);
    FilmFileHeader h;
    const int rc = film_read_header(f, h, path);
    std::fclose(f);
    if (rc) return rc;
    if (width) *width = h.width;
    if (height) *height = h.height;
    if (spp_done) *spp_done = h.spp_done;
    if (seed) *seed = h.seed;
    return WPT_OK;
}

extern "C" int wpt_film_load(const char *path, float *r, float *g, float *b, uint32_t width, uint32_t height, uint64_t *spp_done,
                             uint64_t *seed)
{
    if (!path || !r || !g || !b) return fail(WPT_E_INVALID, "wpt_film_load: null argument");
    FILE *f = std::fopen(path, "rb");
    if (!f) return fail(WPT_E_IO, std::string("wpt_film_load: cannot open ") + path);
    FilmFileHeader h;
    if (int rc = film_read_header(f, h, path)) { std::fclose(f); return rc; }
    if (h.width != width || h.height != height) {
        std::fclose(f);
        return fail(WPT_E_INVALID, "wpt_film_load: the checkpoint is " + std::to_string(h.width) + "x" + std::to_string(h.height) +
                                       ", the film given " + std::to_string(width) + "x" + std::to_string(height));
    }
    const size_t n = (size_t)width * height;
    uint64_t stored = 0;
    const bool ok = std::fread(r, sizeof(float), n, f) == n && std::fread(g, sizeof(float), n, f) == n &&
                    std::fread(b, sizeof(float), n, f) == n && std::fread(&stored, sizeof stored, 1, f) == 1;
    std::fclose(f);
    uint64_t sum = fnv1a_words(&h, sizeof h, 0xcbf29ce484222325ull);
    if (ok) { sum = fnv1a_words(r, n * sizeof(float), sum); sum = fnv1a_words(g, n * sizeof(float), sum); sum = fnv1a_words(b, n * sizeof(float), sum); }
    if (!ok || sum != stored) return fail(WPT_E_IO, std::string("wpt_film_load: truncated or damaged checkpoint: ") + path);
    if (spp_done) *spp_done = h.spp_done;
    if (seed) *seed = h.seed;
    return WPT_OK;
}
