// bvh_build.h — host-side SAH BVH builder and wide-node flattener (C++, no CUDA).
// New work: the reference has no builder of its own (OptiX builds the GAS inside the driver,
// src/accel/optix_backend.h:309-358).
#pragma once
#include <vector>
#include "bvh.h"

namespace wpt {

struct BuiltBvh {
    std::vector<Block64> blocks;   // blocks[0] is the root node
    uint32_t n_nodes = 0, n_tris = 0;
    uint32_t max_depth = 0;        // wide-tree depth (root = 1)
    double   sah_cost = 0.0;
    double   build_ms = 0.0;
    // node block indices in breadth-first order and, per depth d (root = 0), where that level starts
    // (level_begin has max_depth + 1 entries): what a bottom-up refit sweeps, deepest level first
    std::vector<uint32_t> level_nodes, level_begin;
};

// positions f32[3*n_pos], triplets i32[3*n_tris].  Returns false + message on invalid input
// (index out of range, non-finite vertex, depth limit).
bool build_bvh7(const float *positions, uint32_t n_pos, const int32_t *triplets, uint32_t n_tris,
                BuiltBvh &out, const char **err, int threads = 0);

// The warp queue of the default trace kernel packs (owner lane << 27 | block index) into one 32-bit word
// (csrc/trace_kernels.cu), so a flattened tree may hold at most 2^27 blocks (8.6 GB; about 115 M triangles).
constexpr uint64_t kMaxBlocks = 1ull << 27;
inline bool blocks_within_limit(uint64_t n_blocks) { return n_blocks <= kMaxBlocks; }

// Decode helpers shared with the tests
uint32_t node_imask(const Node7 &n);
uint32_t node_tri_count(const Node7 &n, int slot);
bool     node_slot_empty(const Node7 &n, int slot);
// child box of slot i in world space (the bounds the traversal kernel effectively tests)
void node_child_bounds(const Node7 &n, int slot, float lo[3], float hi[3]);

}  // namespace wpt
