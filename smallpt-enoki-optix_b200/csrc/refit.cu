// refit.cu — bottom-up refit of the BVH7-64 blocks on the device for moved vertices (same topology).
//
// SURVEY §8(f) rank 4 ("GPU-side refit for dynamic scenes").  The reference rebuilds its OptiX GAS from
// scratch in set_triangles_soup (src/accel/optix_backend.h:283-364) and has no update path; this is new
// work on the same boundary: wpt_scene_refit(scene, positions).
//
// One launch per tree level, deepest first (levels recorded by the host builder, bvh_build.h).  A thread
// owns one node: it rewrites the triangle blocks of its leaf children from the new vertices (same
// operands and operation order as the builder: v0, e1 = v0-v1, e2 = v2-v0, Ng = cross(e2,e1) with the
// fused multiply-subtract of Embree's cross), reads the float boxes its internal children left in
// `boxes`, and re-quantises its seven child boxes with the builder's rules (origin s/4 below the lower
// corner, floor/ceil with a 1/8 step margin) so that the slab test stays conservative.  Topology, slot
// assignment, `base`, the internal mask and the triangle counts do not change.
#include "kernels.cuh"

namespace wpt {

namespace {

__device__ __forceinline__ int exp_for_extent_dev(float extent, float maxabs)
{
    const double need = fmax((double)extent / 240.0, ldexp((double)maxabs, -17));
    if (!(need > 0.0)) return 127 - 100;
    int e;
    const double m = frexp(need, &e);
    int ex = (m == 0.5) ? e - 1 : e;
    ex = max(-100, min(100, ex));
    return ex + 127;
}

__global__ void __launch_bounds__(128)
refit_level_kernel(Block64 *__restrict__ blocks, const uint32_t *__restrict__ level_nodes, uint32_t count,
                   const float *__restrict__ pos, const int32_t *__restrict__ pidx, float *__restrict__ boxes,
                   TriAttr48 *__restrict__ attr, const uint8_t *__restrict__ geo_normal, uint32_t *__restrict__ bad_flag)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint32_t nb = level_nodes[i];
    Node7 node = blocks[nb].node;
    const uint32_t imask = node.yhi[7];
    const uint32_t cnt2 = (uint32_t)node.zlo[7] | ((uint32_t)node.zhi[7] << 8);
    const uint32_t n_int = (uint32_t)__popc(imask);
    float clo[kWidth][3], chi[kWidth][3];
    bool used[kWidth];
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    uint32_t tri_off = 0;
    for (int s = 0; s < kWidth; s++) {
        const uint32_t cnt = (cnt2 >> (2 * s)) & 3u;
        used[s] = true;
        if ((imask >> s) & 1u) {
            const uint32_t cb = node.base + (uint32_t)__popc(imask & ((1u << s) - 1u));
            for (int k = 0; k < 3; k++) { clo[s][k] = boxes[6 * (size_t)cb + k]; chi[s][k] = boxes[6 * (size_t)cb + 3 + k]; }
        } else if (cnt) {
            for (int k = 0; k < 3; k++) { clo[s][k] = INFINITY; chi[s][k] = -INFINITY; }
            for (uint32_t t = 0; t < cnt; t++) {
                Tri64 &tr = blocks[node.base + n_int + tri_off + t].tri;
                const int32_t prim = tr.prim;
                const float *a = pos + 3 * (size_t)pidx[3 * (size_t)prim], *b = pos + 3 * (size_t)pidx[3 * (size_t)prim + 1],
                            *c = pos + 3 * (size_t)pidx[3 * (size_t)prim + 2];
                float v0[3], e1[3], e2[3];
                for (int k = 0; k < 3; k++) {
                    v0[k] = a[k]; e1[k] = a[k] - b[k]; e2[k] = c[k] - a[k];
                    clo[s][k] = fminf(clo[s][k], fminf(a[k], fminf(b[k], c[k])));
                    chi[s][k] = fmaxf(chi[s][k], fmaxf(a[k], fmaxf(b[k], c[k])));
                }
                for (int k = 0; k < 3; k++) { tr.v0[k] = v0[k]; tr.e1[k] = e1[k]; tr.e2[k] = e2[k]; }
                tr.ng[0] = fmaf(e2[1], e1[2], -(e2[2] * e1[1]));
                tr.ng[1] = fmaf(e2[2], e1[0], -(e2[0] * e1[2]));
                tr.ng[2] = fmaf(e2[0], e1[1], -(e2[1] * e1[0]));
                if (geo_normal[prim]) {
                    // faces without vertex normals shade with the geometric normal (api.cu, same operations)
                    const float f1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]};
                    float n[3] = {f1[1] * e2[2] - f1[2] * e2[1], f1[2] * e2[0] - f1[0] * e2[2], f1[0] * e2[1] - f1[1] * e2[0]};
                    const float l = sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
                    if (l > 0) { n[0] /= l; n[1] /= l; n[2] /= l; } else { n[0] = 0; n[1] = 1; n[2] = 0; }
                    TriAttr48 &at = attr[prim];
                    for (int k = 0; k < 3; k++) at.n0[k] = at.n1[k] = at.n2[k] = n[k];
                }
            }
            tri_off += cnt;
        } else {
            used[s] = false;
        }
        if (used[s])
            for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], clo[s][k]); hi[k] = fmaxf(hi[k], chi[s][k]); }
    }
    for (int k = 0; k < 3; k++) { boxes[6 * (size_t)nb + k] = lo[k]; boxes[6 * (size_t)nb + 3 + k] = hi[k]; }
    // device-side positions cannot be validated by the host: a NaN / Inf vertex shows up here as a non-finite box
    // (fminf / fmaxf drop NaNs, so the triangle operands are checked as well) and raises the scene's flag word
    if (bad_flag) {
        bool ok = true;
        for (int k = 0; k < 3; k++) ok = ok && isfinite(lo[k]) && isfinite(hi[k]);
        uint32_t off = 0;
        for (int sl = 0; sl < kWidth && ok; sl++) {
            const uint32_t cnt = ((imask >> sl) & 1u) ? 0u : ((cnt2 >> (2 * sl)) & 3u);
            for (uint32_t t = 0; t < cnt && ok; t++) {
                const Tri64 &tr = blocks[node.base + n_int + off + t].tri;
                for (int k = 0; k < 3; k++) ok = ok && isfinite(tr.v0[k]) && isfinite(tr.e1[k]) && isfinite(tr.e2[k]);
            }
            off += cnt;
        }
        if (!ok) atomicOr(bad_flag, 1u);
    }

    uint8_t *plo[3] = {node.xlo, node.ylo, node.zlo}, *phi[3] = {node.xhi, node.yhi, node.zhi};
    int eb[3];
    for (int k = 0; k < 3; k++) {
        const float maxabs = fmaxf(fabsf(lo[k]), fabsf(hi[k]));
        eb[k] = exp_for_extent_dev(hi[k] - lo[k], maxabs);
        const double sc = ldexp(1.0, eb[k] - 127);
        float p = (float)((double)lo[k] - sc / 4.0);
        while ((double)p > (double)lo[k] - sc / 8.0) p = nextafterf(p, -INFINITY);
        node.p[k] = p;
        for (int s = 0; s < kWidth; s++) {
            if (!used[s]) { plo[k][s] = 255; phi[k][s] = 0; continue; }
            const double ql = floor(((double)clo[s][k] - (double)p) / sc - 1.0 / 8.0);
            const double qh = ceil(((double)chi[s][k] - (double)p) / sc + 1.0 / 8.0);
            plo[k][s] = (uint8_t)fmax(0.0, fmin(255.0, ql));
            phi[k][s] = (uint8_t)fmax(0.0, fmin(255.0, qh));
        }
    }
    node.xlo[7] = (uint8_t)(eb[0] + 8); node.xhi[7] = (uint8_t)(eb[1] + 8); node.ylo[7] = (uint8_t)(eb[2] + 8);
    node.yhi[7] = (uint8_t)imask;
    node.zlo[7] = (uint8_t)(cnt2 & 0xffu); node.zhi[7] = (uint8_t)(cnt2 >> 8);
    blocks[nb].node = node;
}

}  // namespace

void launch_refit(Block64 *blocks, const uint32_t *level_nodes, const uint32_t *level_begin_host, uint32_t n_levels,
                  const float *pos, const int32_t *pidx, float *boxes, TriAttr48 *attr, const uint8_t *geo_normal,
                  cudaStream_t st, uint32_t *bad_flag)
{
    for (uint32_t d = n_levels; d-- > 0;) {
        const uint32_t first = level_begin_host[d], count = level_begin_host[d + 1] - first;
        if (!count) continue;
        refit_level_kernel<<<(count + 127) / 128, 128, 0, st>>>(blocks, level_nodes + first, count, pos, pidx, boxes,
                                                                attr, geo_normal, bad_flag);
    }
}

}  // namespace wpt
