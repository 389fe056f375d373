// hosttest.cpp — TEST-ONLY library (libwpt_hosttest.so): runs the host builder and the
// __host__ __device__ traversal core on the CPU so the logic can be checked without a GPU.
// It is NOT linked into libwpt.so and no product entry point reaches it.
#include <cmath>
#include <cstdio>
#include <vector>
#include "bvh_build.h"
#include "trace_core.h"

using namespace wpt;

extern "C" {

// returns 0 ok; fills stats[0..7] = n_nodes, n_tris, max_depth, sah_cost, build_ms, n_problems, n_children, n_leaves
int wpt_hosttest_build(const float *pos, uint32_t n_pos, const int32_t *idx, uint32_t n_tris, double *stats)
{
    BuiltBvh b;
    const char *err = nullptr;
    if (!build_bvh7(pos, n_pos, idx, n_tris, b, &err)) { fprintf(stderr, "build failed: %s\n", err); return 1; }
    // structural validation: every triangle appears exactly once; every child's quantised box
    // contains the geometry below it; every block is referenced exactly once.
    std::vector<int> seen(n_tris, 0), block_refs(b.blocks.size(), 0);
    long problems = 0;
    std::vector<float> lo_of(b.blocks.size() * 3), hi_of(b.blocks.size() * 3);
    std::vector<uint32_t> st{0u}, order;
    block_refs[0] = 1;
    while (!st.empty()) {
        const uint32_t n = st.back(); st.pop_back();
        order.push_back(n);
        const Node7 &nd = b.blocks[n].node;
        int rel = 0;
        for (int s = 0; s < kWidth; s++)
            if (node_imask(nd) & (1u << s)) {
                const uint32_t c = nd.base + rel++;
                if (c >= b.blocks.size()) { problems++; continue; }
                block_refs[c]++;
                st.push_back(c);
            }
    }
    for (size_t oi = order.size(); oi-- > 0;) {
        const uint32_t n = order[oi];
        const Node7 &nd = b.blocks[n].node;
        float nlo[3] = {INFINITY, INFINITY, INFINITY}, nhi[3] = {-INFINITY, -INFINITY, -INFINITY};
        const uint32_t n_int = (uint32_t)pop_count(node_imask(nd));
        uint32_t rel = 0, tri_off = 0;
        for (int s = 0; s < kWidth; s++) {
            if (node_slot_empty(nd, s)) continue;
            float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
            if (node_imask(nd) & (1u << s)) {
                const uint32_t c = nd.base + rel++;
                for (int k = 0; k < 3; k++) { clo[k] = lo_of[3 * c + k]; chi[k] = hi_of[3 * c + k]; }
            } else {
                const uint32_t cnt = node_tri_count(nd, s);
                for (uint32_t i = 0; i < cnt; i++) {
                    const uint32_t blk = nd.base + n_int + tri_off + i;
                    if (blk >= b.blocks.size()) { problems++; continue; }
                    block_refs[blk]++;
                    const Tri64 &t = b.blocks[blk].tri;
                    const int32_t prim = t.prim;
                    if (prim < 0 || (uint32_t)prim >= n_tris) { problems++; continue; }
                    seen[prim]++;
                    for (int v = 0; v < 3; v++) {
                        const float *p = pos + 3 * (size_t)idx[3 * prim + v];
                        for (int k = 0; k < 3; k++) { clo[k] = fminf(clo[k], p[k]); chi[k] = fmaxf(chi[k], p[k]); }
                    }
                    const float *p0 = pos + 3 * (size_t)idx[3 * prim];
                    for (int k = 0; k < 3; k++) if (t.v0[k] != p0[k]) problems++;
                }
                tri_off += cnt;
            }
            float qlo[3], qhi[3];
            node_child_bounds(nd, s, qlo, qhi);
            for (int k = 0; k < 3; k++) {
                if (!(qlo[k] <= clo[k] && qhi[k] >= chi[k])) problems++;
                nlo[k] = fminf(nlo[k], clo[k]); nhi[k] = fmaxf(nhi[k], chi[k]);
            }
        }
        for (int k = 0; k < 3; k++) { lo_of[3 * n + k] = nlo[k]; hi_of[3 * n + k] = nhi[k]; }
    }
    double n_children = 0, n_leaves = 0;
    for (uint32_t n : order)
        for (int sl = 0; sl < kWidth; sl++) {
            if (node_slot_empty(b.blocks[n].node, sl)) continue;
            n_children++;
            if (!(node_imask(b.blocks[n].node) & (1u << sl))) n_leaves++;
        }
    stats[6] = n_children; stats[7] = n_leaves;
    for (uint32_t i = 0; i < n_tris; i++) if (seen[i] != 1) problems++;
    for (int r : block_refs) if (r != 1) problems++;
    if (order.size() != b.n_nodes || b.n_tris != n_tris) problems++;
    stats[0] = (double)b.n_nodes; stats[1] = (double)b.n_tris; stats[2] = b.max_depth;
    stats[3] = b.sah_cost; stats[4] = b.build_ms; stats[5] = (double)problems;
    return 0;
}

// LocalStack that remembers its high-water mark (the shared-memory-only trace kernels rely on
// "a traversal never holds more than max_depth - 1 entries")
struct CountingStack : LocalStack {
    int hi = 0;
    void push(uint2 v) { LocalStack::push(v); if (n > hi) hi = n; }
};

// counters (optional): [0] node visits, [1] triangle tests, summed over all rays; [2] deepest stack, [3] max_depth
int wpt_hosttest_trace_stats(const float *pos, uint32_t n_pos, const int32_t *idx, uint32_t n_tris,
                             const float *ox, const float *oy, const float *oz, const float *dx, const float *dy,
                             const float *dz, float tmin, float tmax, uint32_t n,
                             int32_t *id, float *t, float *u, float *v, double *counters)
{
    BuiltBvh b;
    const char *err = nullptr;
    if (!build_bvh7(pos, n_pos, idx, n_tris, b, &err)) { fprintf(stderr, "build failed: %s\n", err); return 1; }
    BvhView view;
    view.blocks = reinterpret_cast<const uint4 *>(b.blocks.data());
    view.n_blocks = (uint32_t)b.blocks.size();
    view.n_nodes = b.n_nodes;
    view.n_tris = b.n_tris;
    view.max_depth = b.max_depth;
    uint64_t nodes_visited = 0, tris_tested = 0;
    int stack_hi = 0;
#pragma omp parallel for schedule(dynamic, 1024) reduction(+ : nodes_visited, tris_tested) reduction(max : stack_hi)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        const RayCtx r = make_ray(ox[i], oy[i], oz[i], dx[i], dy[i], dz[i], tmin);
        CountingStack st;
        uint64_t c[2] = {0, 0};
        const Hit h = trace_ray(view, r, tmax, st, c);
        nodes_visited += c[0]; tris_tested += c[1];
        if (st.hi > stack_hi) stack_hi = st.hi;
        id[i] = h.prim;
        if (h.prim >= 0) { t[i] = h.t; u[i] = h.u; v[i] = h.v; }
    }
    if (counters) { counters[0] = (double)nodes_visited; counters[1] = (double)tris_tested; counters[2] = stack_hi; counters[3] = b.max_depth; }
    return 0;
}

// the block-count limit wpt_scene_commit enforces (bvh_build.h)
int wpt_hosttest_blocks_within_limit(uint64_t n_blocks) { return blocks_within_limit(n_blocks) ? 1 : 0; }

int wpt_hosttest_trace(const float *pos, uint32_t n_pos, const int32_t *idx, uint32_t n_tris,
                       const float *ox, const float *oy, const float *oz, const float *dx, const float *dy,
                       const float *dz, float tmin, float tmax, uint32_t n,
                       int32_t *id, float *t, float *u, float *v)
{
    return wpt_hosttest_trace_stats(pos, n_pos, idx, n_tris, ox, oy, oz, dx, dy, dz, tmin, tmax, n, id, t, u, v, nullptr);
}
}
