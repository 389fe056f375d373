// bvh.h — the flattened acceleration structure shared by the host builder and the device kernels.
//
// Replaces the OptiX GAS the reference builds in OptixBackend::set_triangles_soup
// (src/accel/optix_backend.h:283-364).  Layout: 8-wide nodes whose child boxes are quantised
// against a per-node origin and per-axis power-of-two scale (compressed wide BVH in the style of
// Ylitie, Karras, Laine, HPG 2017), but with the planes stored as bfloat16 PAIRS instead of bytes:
// one node is exactly one 128-byte cache line = eight 16-byte loads, and decoding a plane costs no
// ALU-pipe instruction on sm_100a (see trace_core.h: the odd child of a pair is used as a float as
// it lies in the register, the even child is one IMAD.SHL on the FMA pipe).  Triangles are 48-byte
// blocks (v0, e1 = v0-v1, e2 = v2-v0, Ng = cross(e2,e1): the operands of Embree's Moeller-Trumbore
// test, dependencies/embree/kernels/geometry/triangle.h:54, triangle_intersector_moeller.h:75-113).
#pragma once
#include <stdint.h>
#include <vector_types.h>

namespace wpt {

// planes[axis][side][j]: side 0 = lower planes, 1 = upper planes; word j holds child 2j in its low
// half and child 2j+1 in its high half, each a bfloat16 in quantisation units q: plane = p + q*2^(e-127).
struct alignas(128) Node8 {
    float    p[3];        // quantisation origin (slightly below the node's lower corner)
    uint8_t  e[3];        // biased exponents: scale_k = 2^(e[k]-127)
    uint8_t  imask;       // bit i set: slot i holds an internal child
    uint32_t child_base;  // index of the first internal child (children are contiguous, slot order)
    uint32_t tri_base;    // index of the first triangle block of this node's leaf children
    uint8_t  meta[8];     // 0 = empty; internal: 0xff; leaf: count<<5 | tri offset (count 1..3, offset 0..23)
    uint32_t planes[3][2][4];
};
static_assert(sizeof(Node8) == 128, "Node8 must be 128 bytes");

struct alignas(16) Tri48 {
    float v0[3], e1[3], e2[3], ng[3];
};
static_assert(sizeof(Tri48) == 48, "Tri48 must be 48 bytes");

// per ORIGINAL triangle shading attributes, gathered once per hit by the shade kernel:
// three vertex normals (src/accel/optix_backend.h:412-420 reads them through two index levels)
// and the material id (src/main.cpp:325)
struct alignas(16) TriAttr48 {
    float   n0[3], n1[3], n2[3];
    int32_t mat;
    uint32_t pad[2];
};
static_assert(sizeof(TriAttr48) == 48, "TriAttr48 must be 48 bytes");

constexpr int kMaxLeafTris = 3;
constexpr int kStackSize = 32;   // traversal stack entries (one per tree level suffices)

// device-side view (all pointers device memory)
struct BvhView {
    const uint4 *nodes;     // Node8 as 8 x uint4
    const float4 *tris;     // Tri48 as 3 x float4
    const int32_t *prim_id; // leaf slot -> original triangle index
    uint32_t n_nodes, n_tris;
};

}  // namespace wpt
