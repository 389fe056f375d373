// lbvh_build.cu — device-side BVH build: Morton-order LBVH (Karras, HPG 2012) collapsed to the 7-wide 64-byte block
// layout of bvh.h, entirely on the GPU.
//
// SURVEY §8(f) rank 4.  The reference builds its acceleration structure on the device too (optixAccelBuild inside
// OptixBackend::set_triangles_soup, src/accel/optix_backend.h:283-364); the default builder of this library is the
// host SAH builder (bvh_build.cpp: better trees, 0.3 s per million triangles); this one is for scenes whose topology
// changes per frame (wpt_scene_commit_lbvh: 10 ms per million triangles on B200; traversal 1.2x slower than on the SAH
// tree, profiles/r2_lbvh.txt).
//
// Steps (all kernels below; the radix sort is CUB's, library code outside the render path):
//   1. per triangle: box, centroid -> 63-bit Morton code in the scene box          morton_kernel
//   2. sort (code, triangle)                                                         cub::DeviceRadixSort
//   3. binary radix tree over the sorted codes, ties broken by position             hierarchy_kernel
//   4. boxes bottom-up (second arrival at a node computes the union)                boxes_kernel
//   5. level by level from the root: a wide node takes a binary node's two children and keeps replacing the
//      child of largest area that still holds more than kMaxLeafTris triangles by its two children until it
//      has kWidth; children with <= kMaxLeafTris triangles become leaves (a contiguous run of the sorted order);
//      slots are assigned towards the box corners like the host builder does; blocks of the children are
//      allocated with one atomicAdd per node                                          emit_level_kernel
//   6. the geometry of the triangle blocks, the float boxes and the quantised planes are then written by the SAME
//      kernel that refits a tree after its vertices moved (refit.cu: launch_refit), deepest level first.
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <vector>

#include "kernels.cuh"

namespace wpt {

namespace {

constexpr uint32_t kLeafFlag = 0x80000000u;

struct Bin {                 // binary radix-tree node i (internal nodes 0 .. n-2)
    uint32_t left, right;    // child: index of an internal node, or kLeafFlag | sorted position of a triangle
    uint32_t first, last;    // range of sorted positions below this node
};

__device__ __forceinline__ uint64_t expand21(uint32_t v)   // 21 bits -> every third bit of 63
{
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// float min / max through the order-preserving integer image
__device__ __forceinline__ int fkey(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float funkey(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void scene_box_kernel(const float *__restrict__ pos, const int32_t *__restrict__ pidx, uint32_t n_tris,
                                 int *__restrict__ box /* [6] keys */)
{
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n_tris; t += gridDim.x * blockDim.x)
        for (int v = 0; v < 3; v++) {
            const float *p = pos + 3 * (size_t)pidx[3 * (size_t)t + v];
            for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], p[k]); hi[k] = fmaxf(hi[k], p[k]); }
        }
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
        if ((threadIdx.x & 31u) == 0) { atomicMin(box + k, fkey(lo[k])); atomicMax(box + 3 + k, fkey(hi[k])); }
    }
}

__global__ void morton_kernel(const float *__restrict__ pos, const int32_t *__restrict__ pidx, uint32_t n_tris,
                              const int *__restrict__ box, uint64_t *__restrict__ keys, uint32_t *__restrict__ vals)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tris) return;
    uint32_t q[3];
    for (int k = 0; k < 3; k++) {
        float lo = INFINITY, hi = -INFINITY;
        for (int v = 0; v < 3; v++) {
            const float x = pos[3 * (size_t)pidx[3 * (size_t)t + v] + k];
            lo = fminf(lo, x); hi = fmaxf(hi, x);
        }
        const float blo = funkey(box[k]), bhi = funkey(box[3 + k]);
        const float ext = fmaxf(bhi - blo, 1e-30f);
        const float c = (0.5f * (lo + hi) - blo) / ext;
        q[k] = (uint32_t)fminf(fmaxf(c * 2097152.0f, 0.0f), 2097151.0f);
    }
    keys[t] = expand21(q[0]) | (expand21(q[1]) << 1) | (expand21(q[2]) << 2);
    vals[t] = t;
}

// common prefix length of the keys at sorted positions i and j (-1 outside); equal codes continue with the positions
__device__ __forceinline__ int delta(const uint64_t *__restrict__ keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    return a == b ? 64 + __clz((uint32_t)i ^ (uint32_t)j) : __clzll((long long)(a ^ b));
}

__global__ void hierarchy_kernel(const uint64_t *__restrict__ keys, int n, Bin *__restrict__ bin,
                                 uint32_t *__restrict__ parent /* [2n-1]: internal i at i, leaf k at n-1+k */)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    Bin b;
    b.first = (uint32_t)lo; b.last = (uint32_t)hi;
    b.left = lo == gamma ? (kLeafFlag | (uint32_t)gamma) : (uint32_t)gamma;
    b.right = hi == gamma + 1 ? (kLeafFlag | (uint32_t)(gamma + 1)) : (uint32_t)(gamma + 1);
    bin[i] = b;
    parent[(b.left & kLeafFlag) ? (uint32_t)(n - 1) + (b.left & ~kLeafFlag) : b.left] = (uint32_t)i;
    parent[(b.right & kLeafFlag) ? (uint32_t)(n - 1) + (b.right & ~kLeafFlag) : b.right] = (uint32_t)i;
}

// boxes[6 * node]: internal node i at i, the triangle at sorted position k at n-1+k
__global__ void boxes_kernel(const float *__restrict__ pos, const int32_t *__restrict__ pidx,
                             const uint32_t *__restrict__ sorted_tri, int n, const Bin *__restrict__ bin,
                             const uint32_t *__restrict__ parent, uint32_t *__restrict__ arrived, float *__restrict__ boxes)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t tri = sorted_tri[k];
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int v = 0; v < 3; v++) {
        const float *p = pos + 3 * (size_t)pidx[3 * (size_t)tri + v];
        for (int c = 0; c < 3; c++) { lo[c] = fminf(lo[c], p[c]); hi[c] = fmaxf(hi[c], p[c]); }
    }
    float *b = boxes + 6 * (size_t)(n - 1 + k);
    for (int c = 0; c < 3; c++) { b[c] = lo[c]; b[3 + c] = hi[c]; }
    if (n == 1) return;
    uint32_t node = parent[n - 1 + k];
    for (;;) {
        __threadfence();
        if (atomicAdd(arrived + node, 1u) == 0u) return;          // the sibling subtree is not finished: it will go on
        const Bin bn = bin[node];
        const float *l = boxes + 6 * (size_t)((bn.left & kLeafFlag) ? (uint32_t)(n - 1) + (bn.left & ~kLeafFlag) : bn.left);
        const float *r = boxes + 6 * (size_t)((bn.right & kLeafFlag) ? (uint32_t)(n - 1) + (bn.right & ~kLeafFlag) : bn.right);
        float *o = boxes + 6 * (size_t)node;
        for (int c = 0; c < 3; c++) {
            o[c] = fminf(__ldcg(l + c), __ldcg(r + c));
            o[3 + c] = fmaxf(__ldcg(l + 3 + c), __ldcg(r + 3 + c));
        }
        if (node == 0) return;
        node = parent[node];
    }
}

struct Work { uint32_t bnode, block; };   // bnode: a binary child reference (kLeafFlag form allowed for the root only)

__device__ __forceinline__ uint32_t ref_count(uint32_t ref, const Bin *bin)
{
    return (ref & kLeafFlag) ? 1u : bin[ref].last - bin[ref].first + 1u;
}
__device__ __forceinline__ const float *ref_box(uint32_t ref, const float *boxes, int n)
{
    return boxes + 6 * (size_t)((ref & kLeafFlag) ? (uint32_t)(n - 1) + (ref & ~kLeafFlag) : ref);
}
__device__ __forceinline__ uint32_t ref_first(uint32_t ref, const Bin *bin)
{
    return (ref & kLeafFlag) ? (ref & ~kLeafFlag) : bin[ref].first;
}

__global__ void __launch_bounds__(128)
emit_level_kernel(const Work *__restrict__ in, uint32_t n_in, Work *__restrict__ out, uint32_t *__restrict__ n_out,
                  uint32_t *__restrict__ block_counter, uint32_t max_blocks, const Bin *__restrict__ bin,
                  const float *__restrict__ boxes, const uint32_t *__restrict__ sorted_tri, int n,
                  Block64 *__restrict__ blocks, uint32_t *__restrict__ level_nodes, int root_is_leaf)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_in) return;
    const Work w = in[i];
    level_nodes[i] = w.block;
    uint32_t kids[kWidth];
    int nk;
    if (root_is_leaf) {
        kids[0] = w.bnode; nk = 1;                       // a scene of <= kMaxLeafTris triangles: one leaf under the root
    } else {
        const Bin b = bin[w.bnode];
        kids[0] = b.left; kids[1] = b.right; nk = 2;
        while (nk < kWidth) {
            int best = -1;
            float best_a = -1.0f;
            for (int c = 0; c < nk; c++) {
                if (ref_count(kids[c], bin) <= (uint32_t)kMaxLeafTris) continue;
                const float *bx = ref_box(kids[c], boxes, n);
                const float dx = bx[3] - bx[0], dy = bx[4] - bx[1], dz = bx[5] - bx[2];
                const float a = dx * dy + dy * dz + dz * dx;
                if (a > best_a) { best_a = a; best = c; }
            }
            if (best < 0) break;
            const Bin cb = bin[kids[best]];
            kids[best] = cb.left;
            kids[nk++] = cb.right;
        }
    }
    // slot s should hold the child lying furthest towards corner s of the node box (traversal priority = slot XOR
    // ray octant): the host builder's greedy assignment (bvh_build.cpp)
    const float *nb = ref_box(w.bnode, boxes, n);
    float nc[3];
    for (int k = 0; k < 3; k++) nc[k] = 0.5f * (nb[k] + nb[3 + k]);
    float dc[kWidth][3];
    for (int c = 0; c < nk; c++) {
        const float *cb = ref_box(kids[c], boxes, n);
        for (int k = 0; k < 3; k++) dc[c][k] = 0.5f * (cb[k] + cb[3 + k]) - nc[k];
    }
    int child_in_slot[kWidth];
    for (int s = 0; s < kWidth; s++) child_in_slot[s] = -1;
    uint32_t cu = 0, su = 0;
    for (int it = 0; it < nk; it++) {
        float best = -INFINITY;
        int bc = -1, bs = -1;
        for (int c = 0; c < nk; c++) {
            if (cu >> c & 1u) continue;
            for (int s = 0; s < kWidth; s++) {
                if (su >> s & 1u) continue;
                const float cost = ((s & 1) ? dc[c][0] : -dc[c][0]) + ((s & 2) ? dc[c][1] : -dc[c][1]) + ((s & 4) ? dc[c][2] : -dc[c][2]);
                if (cost > best) { best = cost; bc = c; bs = s; }
            }
        }
        cu |= 1u << bc; su |= 1u << bs; child_in_slot[bs] = bc;
    }
    uint32_t imask = 0, cnt2 = 0, n_int = 0, n_leaf = 0;
    for (int s = 0; s < kWidth; s++) {
        const int c = child_in_slot[s];
        if (c < 0) continue;
        const uint32_t cnt = ref_count(kids[c], bin);
        if (cnt <= (uint32_t)kMaxLeafTris) { cnt2 |= cnt << (2 * s); n_leaf += cnt; }
        else { imask |= 1u << s; n_int++; }
    }
    const uint32_t base = atomicAdd(block_counter, n_int + n_leaf);
    if (base + n_int + n_leaf > max_blocks) return;      // cannot happen (a tree over n triangles has < 2n blocks)
    Node7 node;
    memset(&node, 0, sizeof node);
    node.base = base;
    node.yhi[7] = (uint8_t)imask;
    node.zlo[7] = (uint8_t)(cnt2 & 0xffu); node.zhi[7] = (uint8_t)(cnt2 >> 8);
    blocks[w.block].node = node;                          // planes, origin, exponents: launch_refit
    uint32_t next_node = base, next_tri = base + n_int;
    const uint32_t o0 = n_int ? atomicAdd(n_out, n_int) : 0u;
    uint32_t o = o0;
    for (int s = 0; s < kWidth; s++) {
        const int c = child_in_slot[s];
        if (c < 0) continue;
        const uint32_t ref = kids[c], cnt = ref_count(ref, bin);
        if (cnt <= (uint32_t)kMaxLeafTris) {
            const uint32_t first = ref_first(ref, bin);
            for (uint32_t t = 0; t < cnt; t++) blocks[next_tri++].tri.prim = (int32_t)sorted_tri[first + t];
        } else {
            out[o].bnode = ref; out[o].block = next_node++; o++;
        }
    }
}

}  // namespace

#define LBVH_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { *err = cudaGetErrorString(e_); goto fail; } } while (0)

// Builds the TOPOLOGY of a 7-wide tree over the triangles (device arrays pos / pidx) into a new device buffer of
// zeroed 64-byte blocks: node.base, internal mask and triangle counts, tri.prim.  The caller then runs launch_refit to
// fill the geometry (refit.cu).  Returns false + message on failure (CUDA error; tree deeper than the traversal stack).
bool lbvh_build_topology(const float *d_pos, const int32_t *d_pidx, uint32_t n_tris, Block64 **d_blocks_out,
                         uint32_t *n_blocks_out, uint32_t **d_level_nodes_out, std::vector<uint32_t> &level_begin,
                         uint32_t *n_nodes_out, uint32_t *max_depth_out, const char **err, cudaStream_t st)
{
    const int n = (int)n_tris;
    const uint32_t max_blocks = 2u * n_tris + 2u;
    uint64_t *keys = nullptr, *keys2 = nullptr;
    uint32_t *vals = nullptr, *vals2 = nullptr, *parent = nullptr, *arrived = nullptr, *counters = nullptr, *level_nodes = nullptr;
    int *box = nullptr;
    Bin *bin = nullptr;
    float *boxes = nullptr;
    void *tmp = nullptr;
    Work *q[2] = {nullptr, nullptr};
    Block64 *blocks = nullptr, *exact = nullptr;
    uint32_t *exact_levels = nullptr;
    size_t tmp_bytes = 0;
    level_begin.clear();
    {
        LBVH_CUDA(cudaMalloc(&keys, sizeof(uint64_t) * n_tris));
        LBVH_CUDA(cudaMalloc(&keys2, sizeof(uint64_t) * n_tris));
        LBVH_CUDA(cudaMalloc(&vals, sizeof(uint32_t) * n_tris));
        LBVH_CUDA(cudaMalloc(&vals2, sizeof(uint32_t) * n_tris));
        LBVH_CUDA(cudaMalloc(&box, sizeof(int) * 6));
        LBVH_CUDA(cudaMalloc(&bin, sizeof(Bin) * std::max(1u, n_tris - 1)));
        LBVH_CUDA(cudaMalloc(&parent, sizeof(uint32_t) * (2 * (size_t)n_tris)));
        LBVH_CUDA(cudaMalloc(&arrived, sizeof(uint32_t) * n_tris));
        LBVH_CUDA(cudaMalloc(&boxes, sizeof(float) * 6 * (2 * (size_t)n_tris)));
        LBVH_CUDA(cudaMalloc(&counters, sizeof(uint32_t) * 4));
        LBVH_CUDA(cudaMalloc(&q[0], sizeof(Work) * n_tris));
        LBVH_CUDA(cudaMalloc(&q[1], sizeof(Work) * n_tris));
        LBVH_CUDA(cudaMalloc(&level_nodes, sizeof(uint32_t) * n_tris));
        LBVH_CUDA(cudaMalloc(&blocks, sizeof(Block64) * (size_t)max_blocks));
        LBVH_CUDA(cudaMemsetAsync(blocks, 0, sizeof(Block64) * (size_t)max_blocks, st));
        LBVH_CUDA(cudaMemsetAsync(arrived, 0, sizeof(uint32_t) * n_tris, st));
        const int init_box[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
        LBVH_CUDA(cudaMemcpyAsync(box, init_box, sizeof init_box, cudaMemcpyHostToDevice, st));
        const uint32_t tb = 256, grid = (n_tris + tb - 1) / tb;
        scene_box_kernel<<<std::min(grid, 2048u), tb, 0, st>>>(d_pos, d_pidx, n_tris, box);
        morton_kernel<<<grid, tb, 0, st>>>(d_pos, d_pidx, n_tris, box, keys, vals);
        LBVH_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys2, vals, vals2, n, 0, 63, st));
        LBVH_CUDA(cudaMalloc(&tmp, tmp_bytes));
        LBVH_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, vals, vals2, n, 0, 63, st));
        if (n > 1) hierarchy_kernel<<<grid, tb, 0, st>>>(keys2, n, bin, parent);
        boxes_kernel<<<grid, tb, 0, st>>>(d_pos, d_pidx, vals2, n, bin, parent, arrived, boxes);
        // level-by-level emit
        const int root_is_leaf = n_tris <= (uint32_t)kMaxLeafTris;
        Work root;
        root.block = 0;
        root.bnode = 0;
        if (root_is_leaf && n == 1) root.bnode = kLeafFlag | 0u;
        LBVH_CUDA(cudaMemcpyAsync(q[0], &root, sizeof root, cudaMemcpyHostToDevice, st));
        uint32_t one = 1, zero = 0;
        LBVH_CUDA(cudaMemcpyAsync(counters, &one, sizeof one, cudaMemcpyHostToDevice, st));     // block counter: the root is block 0
        uint32_t n_in = 1, total_nodes = 0, depth = 0;
        int cur = 0;
        while (n_in) {
            if (depth >= (uint32_t)kStackSize - 1) { *err = "BVH deeper than the traversal stack"; goto fail; }
            LBVH_CUDA(cudaMemcpyAsync(counters + 1, &zero, sizeof zero, cudaMemcpyHostToDevice, st));
            level_begin.push_back(total_nodes);
            emit_level_kernel<<<(n_in + 127) / 128, 128, 0, st>>>(q[cur], n_in, q[cur ^ 1], counters + 1, counters, max_blocks, bin,
                                                                 boxes, vals2, n, blocks, level_nodes + total_nodes,
                                                                 root_is_leaf && depth == 0);
            total_nodes += n_in;
            depth++;
            LBVH_CUDA(cudaMemcpyAsync(&n_in, counters + 1, sizeof n_in, cudaMemcpyDeviceToHost, st));
            LBVH_CUDA(cudaStreamSynchronize(st));
            cur ^= 1;
        }
        level_begin.push_back(total_nodes);
        uint32_t n_blocks = 0;
        LBVH_CUDA(cudaMemcpyAsync(&n_blocks, counters, sizeof n_blocks, cudaMemcpyDeviceToHost, st));
        LBVH_CUDA(cudaStreamSynchronize(st));
        LBVH_CUDA(cudaGetLastError());
        if (n_blocks > max_blocks) { *err = "LBVH: block count overflow"; goto fail; }
        LBVH_CUDA(cudaMalloc(&exact, sizeof(Block64) * (size_t)n_blocks));
        LBVH_CUDA(cudaMemcpyAsync(exact, blocks, sizeof(Block64) * (size_t)n_blocks, cudaMemcpyDeviceToDevice, st));
        LBVH_CUDA(cudaMalloc(&exact_levels, sizeof(uint32_t) * total_nodes));
        LBVH_CUDA(cudaMemcpyAsync(exact_levels, level_nodes, sizeof(uint32_t) * total_nodes, cudaMemcpyDeviceToDevice, st));
        LBVH_CUDA(cudaStreamSynchronize(st));
        *d_blocks_out = exact; *n_blocks_out = n_blocks; *d_level_nodes_out = exact_levels;
        *n_nodes_out = total_nodes; *max_depth_out = depth;
    }
    cudaFree(keys); cudaFree(keys2); cudaFree(vals); cudaFree(vals2); cudaFree(box); cudaFree(bin); cudaFree(parent);
    cudaFree(arrived); cudaFree(boxes); cudaFree(counters); cudaFree(q[0]); cudaFree(q[1]); cudaFree(level_nodes);
    cudaFree(blocks); cudaFree(tmp);
    return true;
fail:
    cudaFree(keys); cudaFree(keys2); cudaFree(vals); cudaFree(vals2); cudaFree(box); cudaFree(bin); cudaFree(parent);
    cudaFree(arrived); cudaFree(boxes); cudaFree(counters); cudaFree(q[0]); cudaFree(q[1]); cudaFree(level_nodes);
    cudaFree(blocks); cudaFree(tmp); cudaFree(exact); cudaFree(exact_levels);
    return false;
}

}  // namespace wpt
