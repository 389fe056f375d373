// bvh_build.cpp — binned-SAH BVH2 (one triangle per leaf) -> SAH-optimal collapse to an 8-wide
// tree by dynamic programming -> octant-ordered slot assignment -> 8-bit quantisation.
// Host C++ only.  See bvh.h for the layout and bvh_build.h for the rationale.
#include "bvh_build.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>

namespace wpt {
namespace {

constexpr float kInf = std::numeric_limits<float>::infinity();
constexpr int   kBins = 32;
constexpr float kCostNode = 1.0f;  // c_n
// c_p: cost of one triangle relative to a node step; WPT_COST_TRI overrides it (tuning knob, tools/bvh_stats.py)
static float cost_tri_init() { const char *e = std::getenv("WPT_COST_TRI"); return e ? (float)std::atof(e) : 0.3f; }
static const float kCostTri = cost_tri_init();

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
// Collapse: SAH-optimal BVH2 -> BVH7 by dynamic programming over "cost of representing the
// subtree of n with at most i roots" (Ylitie et al. 2017, §4.1), W = kWidth:
//   C(n,1)   = min( A_n*c_p*P_n  [P_n <= 3],  A_n*c_n + Cd(n,W) )
//   C(n,i>1) = min( Cd(n,i), C(n,i-1) )
//   Cd(n,j)  = min_{0<k<j} C(left,k) + C(right,j-k)
namespace {

constexpr int W = kWidth;

struct Dp {
    float   c[W - 1];  // C(n,i), i = 1..W-1 at c[i-1]
    uint8_t k[W - 1];  // for i>=2: split k of Cd(n,i) if chosen, 0 = "same as i-1"
    uint8_t kw;        // split of Cd(n,W)
    uint8_t leaf;      // C(n,1) picked the leaf
};

struct Collapser {
    const Node2 *n2;
    const Dp *dp;

    // children of the wide node rooted at BVH2 node r
    int gather_root(uint32_t r, uint32_t *kids) const
    {
        int n = 0;
        const Node2 &nd = n2[r];
        if (nd.count == 1) { kids[n++] = r; return n; }
        const int k = dp[r].kw;
        gather(r + 1, k, kids, n);
        gather(nd.right, W - k, kids, n);
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

// biased exponent e of the quantisation step 2^(e-127): the smallest power of two s with 240*s >=
// extent (head-room for the margins below) and s >= 2^-17 * maxabs, so that even a flat node keeps a
// margin (s/8 ~ 2^-20 * |coordinate|) that dominates the float rounding of (p - o) / d in the
// device's slab test.
inline int exp_for_extent(float extent, float maxabs)
{
    const double need = std::max((double)extent / 240.0, std::ldexp((double)maxabs, -17));
    if (!(need > 0.0)) return 127 - 100;
    int e;
    const double m = std::frexp(need, &e);  // need = m * 2^e, m in [0.5,1)
    int ex = (m == 0.5) ? e - 1 : e;
    ex = std::max(-100, std::min(100, ex));
    return ex + 127;
}

inline double node_scale(const Node7 &n, int k)
{
    const uint8_t eb = k == 0 ? n.xlo[7] : k == 1 ? n.xhi[7] : n.ylo[7];   // stored as e + 8
    return std::ldexp(1.0, (int)eb - 8 - 127);
}
inline const uint8_t *plane_lo(const Node7 &n, int k) { return k == 0 ? n.xlo : k == 1 ? n.ylo : n.zlo; }
inline const uint8_t *plane_hi(const Node7 &n, int k) { return k == 0 ? n.xhi : k == 1 ? n.yhi : n.zhi; }

}  // namespace

uint32_t node_imask(const Node7 &n) { return n.yhi[7]; }
uint32_t node_tri_count(const Node7 &n, int slot)
{
    const uint32_t c = (uint32_t)n.zlo[7] | ((uint32_t)n.zhi[7] << 8);
    return (c >> (2 * slot)) & 3u;
}
bool node_slot_empty(const Node7 &n, int slot)
{
    return !((node_imask(n) >> slot) & 1u) && node_tri_count(n, slot) == 0;
}

// world-space box the traversal kernel tests for child `slot` (before float rounding of the slab test)
void node_child_bounds(const Node7 &n, int slot, float lo[3], float hi[3])
{
    for (int k = 0; k < 3; k++) {
        const double s = node_scale(n, k);
        const double l = (double)n.p[k] + s * plane_lo(n, k)[slot], h = (double)n.p[k] + s * plane_hi(n, k)[slot];
        lo[k] = (float)l; hi[k] = (float)h;
        // the float conversions round to nearest: push inwards so the containment check stays strict
        if ((double)lo[k] < l) lo[k] = std::nextafterf(lo[k], kInf);
        if ((double)hi[k] > h) hi[k] = std::nextafterf(hi[k], -kInf);
    }
}

bool build_bvh7(const float *positions, uint32_t n_pos, const int32_t *triplets, uint32_t n_tris,
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
            for (int i = 0; i < W - 1; i++) { d.c[i] = A * kCostTri; d.k[i] = 0; }
            d.kw = 0;
            d.leaf = 1;
            continue;
        }
        const Dp &L = dp[v + 1], &R = dp[nd.right];
        float cd[W + 1];
        uint8_t kd[W + 1];
        for (int j = 2; j <= W; j++) {
            float best = kInf;
            uint8_t bk = 1;
            for (int k = 1; k < j; k++) {
                if (k > W - 1 || j - k > W - 1) continue;
                const float c = L.c[k - 1] + R.c[j - k - 1];
                if (c < best) { best = c; bk = (uint8_t)k; }
            }
            cd[j] = best;
            kd[j] = bk;
        }
        const float c_leaf = nd.count <= (uint32_t)kMaxLeafTris ? A * kCostTri * (float)nd.count : kInf;
        const float c_int = A * kCostNode + cd[W];
        d.kw = kd[W];
        d.leaf = c_leaf <= c_int;
        d.c[0] = d.leaf ? c_leaf : c_int;
        d.k[0] = 0;
        for (int i = 2; i <= W - 1; i++) {
            if (cd[i] < d.c[i - 2]) { d.c[i - 1] = cd[i]; d.k[i - 1] = kd[i]; }
            else { d.c[i - 1] = d.c[i - 2]; d.k[i - 1] = 0; }
        }
    }

    // ---- emit, breadth first: the children of a node (internal nodes, then leaf triangles) are one
    // contiguous run of blocks
    Collapser col{n2.data(), dp.data()};
    struct Work { uint32_t bvh2, block, depth; };
    std::vector<Work> queue;
    queue.reserve(n_tris / 2 + 16);
    out.blocks.reserve((size_t)n_tris + n_tris / 3 + 16);
    out.blocks.emplace_back();
    queue.push_back({0u, 0u, 1u});
    double sah = 0.0;
    for (size_t qi = 0; qi < queue.size(); qi++) {
        const Work w = queue[qi];
        out.max_depth = std::max(out.max_depth, w.depth);
        out.n_nodes++;
        while (out.level_begin.size() < w.depth) out.level_begin.push_back((uint32_t)out.level_nodes.size());
        out.level_nodes.push_back(w.block);
        uint32_t kids[W];
        const int nk = col.gather_root(w.bvh2, kids);
        const Box nb = n2[w.bvh2].b;
        sah += (double)(nb.half_area() / root_area) * kCostNode;

        // slot assignment: slot s should hold the child lying furthest towards corner s
        // (bit k of s set = +axis k), so that traversal priority is slot XOR ray octant.
        int slot_of[W];
        {
            float cost[W][W];
            float nc[3];
            for (int k = 0; k < 3; k++) nc[k] = 0.5f * (nb.lo[k] + nb.hi[k]);
            for (int c = 0; c < nk; c++) {
                const Box &cb = n2[kids[c]].b;
                float dc[3];
                for (int k = 0; k < 3; k++) dc[k] = 0.5f * (cb.lo[k] + cb.hi[k]) - nc[k];
                for (int s = 0; s < W; s++)
                    cost[c][s] = ((s & 1) ? dc[0] : -dc[0]) + ((s & 2) ? dc[1] : -dc[1]) + ((s & 4) ? dc[2] : -dc[2]);
            }
            bool cu[W] = {false}, su[W] = {false};
            for (int it = 0; it < nk; it++) {
                float best = -kInf;
                int bc = -1, bs = -1;
                for (int c = 0; c < nk; c++) {
                    if (cu[c]) continue;
                    for (int s = 0; s < W; s++)
                        if (!su[s] && cost[c][s] > best) { best = cost[c][s]; bc = c; bs = s; }
                }
                cu[bc] = true; su[bs] = true; slot_of[bc] = bs;
            }
        }
        int child_in_slot[W];
        for (int s = 0; s < W; s++) child_in_slot[s] = -1;
        for (int c = 0; c < nk; c++) child_in_slot[slot_of[c]] = c;

        Node7 node;
        std::memset(&node, 0, sizeof node);
        int eb[3];
        for (int k = 0; k < 3; k++) {
            const float maxabs = std::max(std::fabs(nb.lo[k]), std::fabs(nb.hi[k]));
            eb[k] = exp_for_extent(nb.hi[k] - nb.lo[k], maxabs);
            // origin sits s/4 below the true lower corner: every quantised plane keeps a margin
            const double sc = std::ldexp(1.0, eb[k] - 127);
            float p = (float)((double)nb.lo[k] - sc / 4.0);
            while ((double)p > (double)nb.lo[k] - sc / 8.0) p = std::nextafterf(p, -kInf);
            node.p[k] = p;
        }
        uint8_t *plo[3] = {node.xlo, node.ylo, node.zlo}, *phi[3] = {node.xhi, node.yhi, node.zhi};
        uint32_t imask = 0, cnt2 = 0, n_int = 0, n_leaf_tris = 0;
        for (int s = 0; s < W; s++) {
            const int c = child_in_slot[s];
            if (c < 0) {
                for (int k = 0; k < 3; k++) { plo[k][s] = 255; phi[k][s] = 0; }   // inverted: never hit
                continue;
            }
            const Node2 &cn = n2[kids[c]];
            for (int k = 0; k < 3; k++) {
                const double sc = std::ldexp(1.0, eb[k] - 127);
                const double ql = std::floor(((double)cn.b.lo[k] - (double)node.p[k]) / sc - 1.0 / 8.0);
                const double qh = std::ceil(((double)cn.b.hi[k] - (double)node.p[k]) / sc + 1.0 / 8.0);
                plo[k][s] = (uint8_t)std::max(0.0, std::min(255.0, ql));
                phi[k][s] = (uint8_t)std::max(0.0, std::min(255.0, qh));
            }
            if (col.is_leaf(kids[c])) { cnt2 |= cn.count << (2 * s); n_leaf_tris += cn.count; }
            else { imask |= 1u << s; n_int++; }
        }
        node.xlo[7] = (uint8_t)(eb[0] + 8); node.xhi[7] = (uint8_t)(eb[1] + 8); node.ylo[7] = (uint8_t)(eb[2] + 8);
        node.yhi[7] = (uint8_t)imask;
        node.zlo[7] = (uint8_t)(cnt2 & 0xffu); node.zhi[7] = (uint8_t)(cnt2 >> 8);
        node.base = (uint32_t)out.blocks.size();
        out.blocks.resize(out.blocks.size() + n_int + n_leaf_tris);
        uint32_t next_node = node.base, next_tri = node.base + n_int;
        for (int s = 0; s < W; s++) {
            const int c = child_in_slot[s];
            if (c < 0) continue;
            const uint32_t v = kids[c];
            const Node2 &cn = n2[v];
            if (col.is_leaf(v)) {
                for (uint32_t i = cn.first; i < cn.first + cn.count; i++) {
                    const uint32_t prim = idx[i];
                    const float *a = positions + 3 * (size_t)triplets[3 * prim + 0];
                    const float *b = positions + 3 * (size_t)triplets[3 * prim + 1];
                    const float *c3 = positions + 3 * (size_t)triplets[3 * prim + 2];
                    Tri64 t;
                    std::memset(&t, 0, sizeof t);
                    for (int k = 0; k < 3; k++) { t.v0[k] = a[k]; t.e1[k] = a[k] - b[k]; t.e2[k] = c3[k] - a[k]; }
                    cross_fma(t.e2, t.e1, t.ng);
                    t.prim = (int32_t)prim;
                    out.blocks[next_tri++].tri = t;
                    out.n_tris++;
                }
                sah += (double)(cn.b.half_area() / root_area) * kCostTri * cn.count;
            } else {
                queue.push_back({v, next_node++, w.depth + 1});
            }
        }
        out.blocks[w.block].node = node;
    }
    out.level_begin.push_back((uint32_t)out.level_nodes.size());
    out.sah_cost = sah;
    if (out.max_depth > (uint32_t)kStackSize - 1) { *err = e_depth; return false; }
    out.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return true;
}

}  // namespace wpt
