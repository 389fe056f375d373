// bvh_build.cpp — binned-SAH BVH2 (one triangle per leaf) -> SAH-optimal collapse to an 8-wide
// tree by dynamic programming -> octant-ordered slot assignment -> 8-bit quantisation.
// Host C++ only.  See bvh.h for the layout and bvh_build.h for the rationale.
#include "bvh_build.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>

namespace wpt {
namespace {

constexpr float kInf = std::numeric_limits<float>::infinity();
constexpr int   kBins = 16;
constexpr float kCostNode = 1.0f;  // c_n
constexpr float kCostTri = 0.3f;   // c_p

struct Box {
    float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
    void grow(const float p[3])
    {
        for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); }
    }
    void grow(const Box &b)
    {
        for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); }
    }
    float half_area() const
    {
        const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return dx * dy + dy * dz + dz * dx;
    }
    bool valid() const { return lo[0] <= hi[0]; }
};

struct PrimRef {
    Box   b;
    float c[3];
};

// BVH2 node, preorder layout: a subtree over n primitives owns exactly 2n-1 consecutive nodes,
// left child at self+1, right child at self + 2*n_left.  That makes allocation deterministic and
// lock-free under parallel recursion.
struct Node2 {
    Box      b;
    uint32_t first, count;  // primitive range in the permuted index array
    uint32_t right;         // 0 for leaves (count == 1)
};

struct Builder {
    const PrimRef *refs;
    uint32_t *idx;
    Node2 *nodes;
    std::atomic<int> spare_threads{0};

    void build(uint32_t self, uint32_t first, uint32_t count, int depth)
    {
        Node2 &nd = nodes[self];
        nd.first = first;
        nd.count = count;
        nd.right = 0;
        Box b, cb;
        for (uint32_t i = first; i < first + count; i++) {
            const PrimRef &r = refs[idx[i]];
            b.grow(r.b);
            cb.grow(r.c);
        }
        nd.b = b;
        if (count == 1) return;

        uint32_t mid = split_sah(first, count, cb);
        if (mid == first || mid == first + count) {
            // degenerate (coincident centroids or SAH found nothing): object median by index
            mid = first + count / 2;
        }
        const uint32_t nl = mid - first;
        const uint32_t left = self + 1, right = self + 2 * nl;
        nd.right = right;
        if (count > (1u << 14) && spare_threads.fetch_sub(1) > 0) {
            std::thread t([=] { build(left, first, nl, depth + 1); spare_threads.fetch_add(1); });
            build(right, mid, count - nl, depth + 1);
            t.join();
        } else {
            if (count > (1u << 14)) spare_threads.fetch_add(1);  // undo the failed reservation
            build(left, first, nl, depth + 1);
            build(right, mid, count - nl, depth + 1);
        }
    }

    // returns the partition point; idx[first..mid) goes left
    uint32_t split_sah(uint32_t first, uint32_t count, const Box &cb)
    {
        if (count == 2) {
            return first + 1;
        }
        float best_cost = kInf;
        int best_axis = -1, best_bin = -1;
        for (int axis = 0; axis < 3; axis++) {
            const float ext = cb.hi[axis] - cb.lo[axis];
            if (!(ext > 0.0f)) continue;
            const float scale = (float)kBins * (1.0f - 1e-6f) / ext;
            Box bb[kBins];
            uint32_t bc[kBins] = {0};
            for (uint32_t i = first; i < first + count; i++) {
                const PrimRef &r = refs[idx[i]];
                int bi = (int)((r.c[axis] - cb.lo[axis]) * scale);
                bi = bi < 0 ? 0 : (bi >= kBins ? kBins - 1 : bi);
                bb[bi].grow(r.b);
                bc[bi]++;
            }
            float right_area[kBins];
            uint32_t right_cnt[kBins];
            Box acc;
            uint32_t cnt = 0;
            for (int i = kBins - 1; i > 0; i--) {
                if (bc[i]) acc.grow(bb[i]);
                cnt += bc[i];
                right_area[i] = cnt ? acc.half_area() : 0.0f;
                right_cnt[i] = cnt;
            }
            acc = Box();
            cnt = 0;
            for (int i = 0; i < kBins - 1; i++) {
                if (bc[i]) acc.grow(bb[i]);
                cnt += bc[i];
                if (cnt == 0 || right_cnt[i + 1] == 0) continue;
                const float cost = acc.half_area() * (float)cnt + right_area[i + 1] * (float)right_cnt[i + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = i; }
            }
        }
        if (best_axis < 0) return first;
        const float ext = cb.hi[best_axis] - cb.lo[best_axis];
        const float scale = (float)kBins * (1.0f - 1e-6f) / ext;
        const float lo = cb.lo[best_axis];
        uint32_t *b = idx + first, *e = idx + first + count;
        uint32_t *m = std::partition(b, e, [&](uint32_t p) {
            int bi = (int)((refs[p].c[best_axis] - lo) * scale);
            bi = bi < 0 ? 0 : (bi >= kBins ? kBins - 1 : bi);
            return bi <= best_bin;
        });
        return first + (uint32_t)(m - b);
    }
};

inline void cross_fma(const float a[3], const float b[3], float o[3])
{
    // same op order as Embree's cross (common/math/vec3.h:221): msub(a.y,b.z, a.z*b.y) ...
    o[0] = std::fmaf(a[1], b[2], -(a[2] * b[1]));
    o[1] = std::fmaf(a[2], b[0], -(a[0] * b[2]));
    o[2] = std::fmaf(a[0], b[1], -(a[1] * b[0]));
}

}  // namespace

// ------------------------------------------------------------------------------------------
// Collapse: SAH-optimal BVH2 -> BVH8 by dynamic programming over "cost of representing the
// subtree of n with at most i roots" (Ylitie et al. 2017, §4.1).
//   C(n,1)   = min( A_n*c_p*P_n  [P_n <= 3],  A_n*c_n + Cd(n,8) )
//   C(n,i>1) = min( Cd(n,i), C(n,i-1) )
//   Cd(n,j)  = min_{0<k<j} C(left,k) + C(right,j-k)
namespace {

struct Dp {
    float   c[7];      // C(n,i), i = 1..7 at c[i-1]
    uint8_t k[7];      // for i>=2: split k of Cd(n,i) if chosen, 0 = "same as i-1"
    uint8_t k8;        // split of Cd(n,8)
    uint8_t leaf;      // C(n,1) picked the leaf
};

struct Collapser {
    const Node2 *n2;
    const Dp *dp;
    const uint32_t *idx;
    const float *pos;
    const int32_t *trip;
    BuiltBvh *out;

    // children of the wide node rooted at BVH2 node r
    int gather_root(uint32_t r, uint32_t *kids) const
    {
        int n = 0;
        const Node2 &nd = n2[r];
        if (nd.count == 1) { kids[n++] = r; return n; }
        const int k = dp[r].k8;
        gather(r + 1, k, kids, n);
        gather(nd.right, 8 - k, kids, n);
        return n;
    }
    void gather(uint32_t v, int i, uint32_t *kids, int &n) const
    {
        const Node2 &nd = n2[v];
        if (nd.count == 1) { kids[n++] = v; return; }
        while (i > 1 && dp[v].k[i - 1] == 0) i--;
        if (i == 1) { kids[n++] = v; return; }
        const int k = dp[v].k[i - 1];
        gather(v + 1, k, kids, n);
        gather(nd.right, i - k, kids, n);
    }
    bool is_leaf(uint32_t v) const { return n2[v].count == 1 || dp[v].leaf; }
};

inline uint8_t exp_for_extent(float extent, float maxabs)
{
    // smallest power of two s with 253*s >= extent (2 steps of head-room for the margins below) and
    // s >= 2^-17 * maxabs, so that even a flat node keeps a margin (s/32 ~ 2^-22 * |coordinate|) that
    // dominates the float rounding of (p - o) / d in the device's slab test.
    const double need = std::max((double)extent / 253.0, std::ldexp((double)maxabs, -17));
    if (!(need > 0.0)) return (uint8_t)(127 - 100);
    int e;
    const double m = std::frexp(need, &e);  // need = m * 2^e, m in [0.5,1)
    int ex = (m == 0.5) ? e - 1 : e;
    ex = std::max(-100, std::min(100, ex));
    return (uint8_t)(ex + 127);
}

}  // namespace

// ---- bfloat16 plane encoding (see bvh.h / trace_core.h) -------------------------------------------
namespace {
inline float bits_to_float(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
inline uint32_t float_to_bits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
// largest bfloat16 <= x, x >= 0
inline uint16_t bf16_floor_pos(double x)
{
    if (!(x > 0.0)) return 0;
    float f = (float)x;
    if ((double)f > x) f = std::nextafterf(f, -kInf);
    return (uint16_t)(float_to_bits(f) >> 16);   // truncation = floor for positive values
}
// what the device sees for a plane stored in the HIGH half of a word whose low half is `low`
inline double decode_high(uint16_t stored, uint16_t low) { return (double)bits_to_float(((uint32_t)stored << 16) | low); }
inline double decode_low(uint16_t stored) { return (double)bits_to_float((uint32_t)stored << 16); }

// lower plane: the largest encodable value whose DECODED value is <= x
inline uint16_t enc_lo(double x, bool high_half, uint16_t low)
{
    uint16_t s = bf16_floor_pos(x);
    if (!high_half) return s;
    while (s > 0 && decode_high(s, low) > x) s--;
    if (s == 0 && decode_high(0, low) > x) return 0x8000u;   // -0 | garbage = a tiny negative number
    return s;
}
// upper plane: the smallest encodable value whose decoded value is >= x
inline uint16_t enc_hi(double x, bool high_half, uint16_t low)
{
    uint16_t s = bf16_floor_pos(x);
    while ((high_half ? decode_high(s, low) : decode_low(s)) < x) s++;
    return s;
}
constexpr uint16_t kEmptyLo = 0x4680u;  // +16384: an inverted box, never hit
constexpr uint16_t kEmptyHi = 0xC680u;  // -16384
}  // namespace

// conservative world-space box the traversal kernel effectively tests for child `slot`
void node_child_bounds(const Node8 &n, int slot, float lo[3], float hi[3])
{
    const int j = slot >> 1;
    const bool high = slot & 1;
    for (int k = 0; k < 3; k++) {
        const double s = std::ldexp(1.0, (int)n.e[k] - 127);
        const uint32_t wl = n.planes[k][0][j], wh = n.planes[k][1][j];
        const double ql = high ? (double)bits_to_float(wl) : decode_low((uint16_t)(wl & 0xffffu));
        const double qh = high ? (double)bits_to_float(wh) : decode_low((uint16_t)(wh & 0xffffu));
        lo[k] = (float)((double)n.p[k] + s * ql);
        hi[k] = (float)((double)n.p[k] + s * qh);
        // the float conversions above round to nearest: push outwards so the check stays strict
        if ((double)lo[k] < (double)n.p[k] + s * ql) lo[k] = std::nextafterf(lo[k], kInf);
        if ((double)hi[k] > (double)n.p[k] + s * qh) hi[k] = std::nextafterf(hi[k], -kInf);
    }
}

bool build_bvh8(const float *positions, uint32_t n_pos, const int32_t *triplets, uint32_t n_tris,
                BuiltBvh &out, const char **err, int threads)
{
    static const char *e_empty = "scene has no triangles";
    static const char *e_index = "position index out of range";
    static const char *e_finite = "non-finite vertex position";
    static const char *e_depth = "BVH deeper than the traversal stack";
    const auto t0 = std::chrono::steady_clock::now();
    out = BuiltBvh();
    if (n_tris == 0) { *err = e_empty; return false; }

    std::vector<PrimRef> refs(n_tris);
    std::vector<uint32_t> idx(n_tris);
    for (uint32_t i = 0; i < n_tris; i++) {
        PrimRef &r = refs[i];
        for (int v = 0; v < 3; v++) {
            const int32_t pi = triplets[3 * i + v];
            if (pi < 0 || (uint32_t)pi >= n_pos) { *err = e_index; return false; }
            const float *p = positions + 3 * (size_t)pi;
            if (!(std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2]))) { *err = e_finite; return false; }
            r.b.grow(p);
        }
        for (int k = 0; k < 3; k++) r.c[k] = 0.5f * (r.b.lo[k] + r.b.hi[k]);
        idx[i] = i;
    }

    // ---- BVH2
    std::vector<Node2> n2(2 * (size_t)n_tris - 1);
    {
        Builder b;
        b.refs = refs.data();
        b.idx = idx.data();
        b.nodes = n2.data();
        int hw = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
        b.spare_threads = std::max(0, hw - 1);
        b.build(0, 0, n_tris, 0);
    }

    // ---- DP, children before parents (preorder layout => reverse index order)
    const size_t nn = n2.size();
    std::vector<Dp> dp(nn);
    const float root_area = std::max(n2[0].b.half_area(), 1e-30f);
    for (size_t v = nn; v-- > 0;) {
        const Node2 &nd = n2[v];
        Dp &d = dp[v];
        const float A = nd.b.half_area() / root_area;
        if (nd.count == 1) {
            for (int i = 0; i < 7; i++) { d.c[i] = A * kCostTri; d.k[i] = 0; }
            d.k8 = 0;
            d.leaf = 1;
            continue;
        }
        const Dp &L = dp[v + 1], &R = dp[nd.right];
        float cd[9];
        uint8_t kd[9];
        for (int j = 2; j <= 8; j++) {
            float best = kInf;
            uint8_t bk = 1;
            for (int k = 1; k < j; k++) {
                if (k > 7 || j - k > 7) continue;
                const float c = L.c[k - 1] + R.c[j - k - 1];
                if (c < best) { best = c; bk = (uint8_t)k; }
            }
            cd[j] = best;
            kd[j] = bk;
        }
        const float c_leaf = nd.count <= (uint32_t)kMaxLeafTris ? A * kCostTri * (float)nd.count : kInf;
        const float c_int = A * kCostNode + cd[8];
        d.k8 = kd[8];
        d.leaf = c_leaf <= c_int;
        d.c[0] = d.leaf ? c_leaf : c_int;
        d.k[0] = 0;
        for (int i = 2; i <= 7; i++) {
            if (cd[i] < d.c[i - 2]) { d.c[i - 1] = cd[i]; d.k[i - 1] = kd[i]; }
            else { d.c[i - 1] = d.c[i - 2]; d.k[i - 1] = 0; }
        }
    }

    // ---- emit, breadth first so siblings are contiguous
    Collapser col{n2.data(), dp.data(), idx.data(), positions, triplets, &out};
    struct Work { uint32_t bvh2, depth; };
    std::vector<Work> queue;
    queue.reserve(n_tris / 2 + 16);
    out.nodes.reserve(n_tris / 2 + 16);
    out.tris.reserve(n_tris);
    out.prim_id.reserve(n_tris);
    out.nodes.emplace_back();
    queue.push_back({0u, 1u});
    double sah = 0.0;
    for (size_t qi = 0; qi < queue.size(); qi++) {
        const Work w = queue[qi];
        out.max_depth = std::max(out.max_depth, w.depth);
        uint32_t kids[8];
        const int nk = col.gather_root(w.bvh2, kids);
        const Box nb = n2[w.bvh2].b;
        sah += (double)(nb.half_area() / root_area) * kCostNode;

        // slot assignment: slot s should hold the child lying furthest towards corner s
        // (bit k of s set = +axis k), so that traversal priority is slot XOR ray octant.
        int slot_of[8];
        {
            float cost[8][8];
            float nc[3];
            for (int k = 0; k < 3; k++) nc[k] = 0.5f * (nb.lo[k] + nb.hi[k]);
            for (int c = 0; c < nk; c++) {
                const Box &cb = n2[kids[c]].b;
                float dc[3];
                for (int k = 0; k < 3; k++) dc[k] = 0.5f * (cb.lo[k] + cb.hi[k]) - nc[k];
                for (int s = 0; s < 8; s++)
                    cost[c][s] = ((s & 1) ? dc[0] : -dc[0]) + ((s & 2) ? dc[1] : -dc[1]) + ((s & 4) ? dc[2] : -dc[2]);
            }
            bool cu[8] = {false}, su[8] = {false};
            for (int it = 0; it < nk; it++) {
                float best = -kInf;
                int bc = -1, bs = -1;
                for (int c = 0; c < nk; c++) {
                    if (cu[c]) continue;
                    for (int s = 0; s < 8; s++)
                        if (!su[s] && cost[c][s] > best) { best = cost[c][s]; bc = c; bs = s; }
                }
                cu[bc] = true; su[bs] = true; slot_of[bc] = bs;
            }
        }
        int child_in_slot[8];
        for (int s = 0; s < 8; s++) child_in_slot[s] = -1;
        for (int c = 0; c < nk; c++) child_in_slot[slot_of[c]] = c;

        Node8 node;
        std::memset(&node, 0, sizeof node);
        for (int k = 0; k < 3; k++) {
            const float maxabs = std::max(std::fabs(nb.lo[k]), std::fabs(nb.hi[k]));
            node.e[k] = exp_for_extent(nb.hi[k] - nb.lo[k], maxabs);
            // origin sits s/32 below the true lower corner: every quantised plane keeps a margin
            const double sc = std::ldexp(1.0, (int)node.e[k] - 127);
            float p = (float)((double)nb.lo[k] - sc / 32.0);
            while ((double)p > (double)nb.lo[k] - sc / 64.0) p = std::nextafterf(p, -kInf);
            node.p[k] = p;
        }
        node.child_base = (uint32_t)out.nodes.size();
        node.tri_base = (uint32_t)out.tris.size();
        uint32_t tri_off = 0;
        uint16_t qlo[3][8], qhi[3][8];
        for (int s = 0; s < 8; s++)
            for (int k = 0; k < 3; k++) { qlo[k][s] = kEmptyLo; qhi[k][s] = kEmptyHi; }
        // even slots first: the odd child of a pair is rounded against the bits its sibling leaves
        // in the low half of the shared word
        for (int pass = 0; pass < 2; pass++) {
            for (int s = pass; s < 8; s += 2) {
                const int c = child_in_slot[s];
                if (c < 0) continue;
                const Node2 &cn = n2[kids[c]];
                for (int k = 0; k < 3; k++) {
                    const double sc = std::ldexp(1.0, (int)node.e[k] - 127);
                    const double xl = ((double)cn.b.lo[k] - (double)node.p[k]) / sc - 1.0 / 64.0;
                    const double xh = ((double)cn.b.hi[k] - (double)node.p[k]) / sc + 1.0 / 64.0;
                    qlo[k][s] = enc_lo(xl, pass == 1, qlo[k][s & ~1]);
                    qhi[k][s] = enc_hi(xh, pass == 1, qhi[k][s & ~1]);
                }
            }
        }
        for (int k = 0; k < 3; k++)
            for (int j = 0; j < 4; j++) {
                node.planes[k][0][j] = (uint32_t)qlo[k][2 * j] | ((uint32_t)qlo[k][2 * j + 1] << 16);
                node.planes[k][1][j] = (uint32_t)qhi[k][2 * j] | ((uint32_t)qhi[k][2 * j + 1] << 16);
            }
        for (int s = 0; s < 8; s++) {
            const int c = child_in_slot[s];
            if (c < 0) continue;
            const uint32_t v = kids[c];
            const Node2 &cn = n2[v];
            if (col.is_leaf(v)) {
                const uint32_t cnt = cn.count;
                node.meta[s] = (uint8_t)((((1u << cnt) - 1u) << 5) | tri_off);
                for (uint32_t i = cn.first; i < cn.first + cnt; i++) {
                    const uint32_t prim = idx[i];
                    const float *a = positions + 3 * (size_t)triplets[3 * prim + 0];
                    const float *b = positions + 3 * (size_t)triplets[3 * prim + 1];
                    const float *c3 = positions + 3 * (size_t)triplets[3 * prim + 2];
                    Tri48 t;
                    for (int k = 0; k < 3; k++) { t.v0[k] = a[k]; t.e1[k] = a[k] - b[k]; t.e2[k] = c3[k] - a[k]; }
                    cross_fma(t.e2, t.e1, t.ng);
                    out.tris.push_back(t);
                    out.prim_id.push_back((int32_t)prim);
                }
                tri_off += cnt;
                sah += (double)(cn.b.half_area() / root_area) * kCostTri * cnt;
            } else {
                node.meta[s] = 0xffu;
                node.imask |= (uint8_t)(1u << s);
                out.nodes.emplace_back();
                queue.push_back({v, w.depth + 1});
            }
        }
        out.nodes[qi] = node;
    }
    out.sah_cost = sah;
    if (out.max_depth > (uint32_t)kStackSize - 1) { *err = e_depth; return false; }
    out.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return true;
}

}  // namespace wpt
