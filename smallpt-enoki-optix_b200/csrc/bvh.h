// bvh.h — the flattened acceleration structure shared by the host builder and the device kernels.
//
// Replaces the OptiX GAS the reference builds in OptixBackend::set_triangles_soup
// (src/accel/optix_backend.h:283-364).  ONE device buffer of 64-byte blocks; a block is either a
// 7-wide node or a triangle.  The layout follows two things measured on B200 (tools/mb/):
//  * lane-divergent traversal is bound by L1 wavefronts: one divergent load instruction costs the SM
//    ~11-33 cycles whatever its width, so a node visit must be as few load instructions as possible
//    -> 64-byte nodes fetched with two 256-bit loads (LDG.E.256), 64-byte triangles likewise;
//  * PRMT/LOP3/SHF/SEL/FSET share the half-rate ALU pipe, FFMA has twice that rate and IDP.4A runs
//    beside the ALU pipe -> child planes are 8-bit (compressed wide BVH, Ylitie/Karras/Laine HPG'17)
//    and one dp4a turns a plane byte into an FMA operand (trace_core.h).
// A node holds up to SEVEN children so that the eighth byte of each of its six plane arrays carries
// the header (three scale exponents, internal-child mask, leaf triangle counts) and the whole node is
// exactly half a cache line.  Triangles store v0, e1 = v0-v1, e2 = v2-v0, Ng = cross(e2,e1) — the
// operands of Embree's Moeller-Trumbore test (dependencies/embree/kernels/geometry/triangle.h:54,
// triangle_intersector_moeller.h:75-113) — and the original primitive index.
#pragma once
#include <stdint.h>
#include <vector_types.h>

namespace wpt {

constexpr int kWidth = 7;        // children per node
constexpr int kMaxLeafTris = 3;
constexpr int kStackSize = 32;   // traversal stack entries (one per tree level suffices)

// plane = p[k] + q * 2^(e[k]-127), q = 0..255.  Children of a node occupy consecutive blocks starting
// at `base`: first the internal children in slot order, then the triangles of the leaf children in
// slot order.  Slot s (0..6) holds the child lying furthest towards corner s of the node box.
struct alignas(64) Node7 {
    float    p[3];
    uint32_t base;
    uint8_t  xlo[8], xhi[8];   // [7]: e[0], e[1]
    uint8_t  ylo[8], yhi[8];   // [7]: e[2], imask
    uint8_t  zlo[8], zhi[8];   // [7]: triangle counts of slots 0..3 / 4..6, two bits each
};
static_assert(sizeof(Node7) == 64, "Node7 must be 64 bytes");

struct alignas(64) Tri64 {
    float   v0[3], e1[3], e2[3], ng[3];
    int32_t prim;              // original triangle index (tri_id of the reference, src/main.cpp:165-188)
    uint32_t pad[3];
};
static_assert(sizeof(Tri64) == 64, "Tri64 must be 64 bytes");

union alignas(64) Block64 {
    Node7 node;
    Tri64 tri;
    uint4 q[4];
};
static_assert(sizeof(Block64) == 64, "Block64 must be 64 bytes");

// per ORIGINAL triangle shading attributes, gathered once per hit by the shade kernel:
// three vertex normals (src/accel/optix_backend.h:412-420 reads them through two index levels)
// and the material id (src/main.cpp:325)
struct alignas(16) TriAttr48 {
    float   n0[3], n1[3], n2[3];
    int32_t mat;
    uint32_t pad[2];
};
static_assert(sizeof(TriAttr48) == 48, "TriAttr48 must be 48 bytes");

// device-side view (device memory)
struct BvhView {
    const uint4 *blocks;    // Block64 as 4 x uint4; block 0 is the root node
    uint32_t n_blocks, n_nodes, n_tris;
    uint32_t max_depth;     // node levels, root = 1: a traversal stack needs at most max_depth - 1 entries
};

}  // namespace wpt
