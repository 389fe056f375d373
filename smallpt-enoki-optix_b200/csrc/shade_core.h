// shade_core.h — camera, warps, frame, RNG: the per-lane arithmetic of the reference's sample/bounce
// loop (src/main.cpp:385-426) as __host__ __device__ functions.  Fused into the raygen and shade
// kernels on the device; the CPU unit tests compile it into the test-only library to check the logic
// against the oracle without a GPU.
#pragma once
#include <math.h>
#include <stdint.h>
#include "trace_core.h"   // WPT_HD, u2f
#include "../../include/wpt.h"

namespace wpt {

struct V3 { float x, y, z; };
WPT_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }

WPT_HD float rsqrt_fast(float x)
{
#if defined(__CUDA_ARCH__)
    return rsqrtf(x);              // the reference uses rsqrt.approx too (enoki array_router.h:663-665)
#else
    return 1.0f / sqrtf(x);
#endif
}
// enoki/array_static.h:924-935: x*x' then an fmadd chain
WPT_HD float dot3(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
WPT_HD V3 normalize3(V3 a)
{
    const float r = rsqrt_fast(dot3(a, a));
    return v3(a.x * r, a.y * r, a.z * r);
}

// ---------------------------------------------------------------- RNG
// Philox4x32-10 (Salmon et al., SC'11).  Counter-based: no per-pixel state in HBM (the reference
// keeps 16 B/pixel of PCG32 state, src/main.cpp:376).
WPT_HD void mulhilo(uint32_t a, uint32_t b, uint32_t &hi, uint32_t &lo)
{
#if defined(__CUDA_ARCH__)
    lo = a * b; hi = __umulhi(a, b);
#else
    const uint64_t p = (uint64_t)a * b; lo = (uint32_t)p; hi = (uint32_t)(p >> 32);
#endif
}
WPT_HD void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                          uint32_t out[4])
{
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t h0, l0, h1, l1;
        mulhilo(0xD2511F53u, c0, h0, l0);
        mulhilo(0xCD9E8D57u, c2, h1, l1);
        const uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// enoki random.h:116-118: (u >> 9 | 0x3f800000) - 1
WPT_HD float u32_to_unit(uint32_t u) { return u2f((u >> 9) | 0x3f800000u) - 1.0f; }

// ---------------------------------------------------------------- warps (src/mapping.h)
#define WPT_PI 3.14159265358979323846f
WPT_HD void sincos_(float x, float &s, float &c)
{
#if defined(__CUDA_ARCH__)
    sincosf(x, &s, &c);
#else
    s = sinf(x); c = cosf(x);
#endif
}
// mapping.h:5-11
WPT_HD V3 cosine_hemisphere(float sx, float sy)
{
    const float sin_phi = sqrtf(1.0f - sx);
    const float theta = WPT_PI * 2.0f * sy;
    float st, ct;
    sincos_(theta, st, ct);
    return v3(ct * sin_phi, sqrtf(sx), st * sin_phi);
}
// mapping.h:15-27.  The reference yields NaN at exactly (0.5,0.5) (0/0); we return the centre.
WPT_HD void disk_from_square(float sx, float sy, float &ox, float &oy)
{
    const float a = sx * 2.0f - 1.0f, b = sy * 2.0f - 1.0f;
    if (a == 0.0f && b == 0.0f) { ox = 0.0f; oy = 0.0f; return; }
    const bool sel = a * a > b * b;
    const float r = sel ? a : b;
    const float phi = sel ? (WPT_PI / 4.0f) * (b / a) : (WPT_PI / 2.0f) - (WPT_PI / 4.0f) * (a / b);
    float s, c;
    sincos_(phi, s, c);
    ox = r * c; oy = r * s;
}

// ---------------------------------------------------------------- frame (src/coordframe.h:17-30)
// world = bx*l.x + n*l.y + bz*l.z   (normal is the local Y axis; enoki matrix.h:177-184 op order)
WPT_HD V3 frame_to_world(V3 n, V3 l)
{
    const float sign = copysignf(1.0f, n.y);
    const float a = -1.0f / (sign + n.y);
    const float b = n.z * n.x * a;
    const V3 bx = v3(sign + n.x * n.x * a, -n.x, b);
    const V3 bz = v3(sign * b, -sign * n.z, 1.0f + sign * n.z * n.z * a);
    return v3(fmaf(bz.x, l.z, fmaf(n.x, l.y, bx.x * l.x)),
              fmaf(bz.y, l.z, fmaf(n.y, l.y, bx.y * l.x)),
              fmaf(bz.z, l.z, fmaf(n.z, l.y, bx.z * l.x)));
}

// ---------------------------------------------------------------- camera (src/pinhole.h:27-56)
// The two quantities of sample_dir that depend on the camera and the resolution only (pinhole.h:41-47): the film
// distance and the film width.  The kernels get them from the host once per launch (camera_frame_constants) instead of
// one tanf and one W/H division per ray; same expressions, same order.
struct CameraFrame { float dist, film_w; };
WPT_HD CameraFrame camera_frame_constants(const wpt_camera &c, int W, int H)
{
    CameraFrame f;
    f.dist = c.film_size_y * 0.5f / tanf(c.fov_y * 0.5f);
    const float ratio = (float)W / (float)H;
    f.film_w = ratio * c.film_size_y;
    return f;
}

WPT_HD void camera_ray(const wpt_camera &c, const CameraFrame &cf, int px, int py, int W, int H, const float xi[4], V3 &org, V3 &dir)
{
    // sample_pos: lens point in the local x-z plane (sic), pinhole.h:27-32
    float lx = 0.0f, lz = 0.0f;
    if (c.lens_radius != 0.0f) {
        disk_from_square(xi[0], xi[1], lx, lz);
        lx *= c.lens_radius; lz *= c.lens_radius;
    }
    // to_world(lens) + origin ; columns of the camera frame are x,y,z
    org = v3(fmaf(c.z[0], lz, fmaf(c.y[0], 0.0f, c.x[0] * lx)) + c.origin[0],
             fmaf(c.z[1], lz, fmaf(c.y[1], 0.0f, c.x[1] * lx)) + c.origin[1],
             fmaf(c.z[2], lz, fmaf(c.y[2], 0.0f, c.x[2] * lx)) + c.origin[2]);
    // sample_dir, pinhole.h:34-56
    const V3 rel = v3(org.x - c.origin[0], org.y - c.origin[1], org.z - c.origin[2]);
    const V3 lens_local = v3(fmaf(c.x[2], rel.z, fmaf(c.x[1], rel.y, c.x[0] * rel.x)),
                             fmaf(c.y[2], rel.z, fmaf(c.y[1], rel.y, c.y[0] * rel.x)),
                             fmaf(c.z[2], rel.z, fmaf(c.z[1], rel.y, c.z[0] * rel.x)));
    const float dist = cf.dist;
    const float ndc_x = ((float)px + xi[2]) / (float)W;
    const float ndc_y = ((float)py + xi[3]) / (float)H;
    const float fx = (0.5f - ndc_x) * cf.film_w;
    const float fy = (0.5f - ndc_y) * c.film_size_y;
    const V3 focal = v3(c.focal_dist * fx / dist, c.focal_dist * fy / dist, c.focal_dist * dist / dist);
    const V3 d = normalize3(v3(focal.x - lens_local.x, focal.y - lens_local.y, focal.z - lens_local.z));
    dir = v3(fmaf(c.z[0], d.z, fmaf(c.y[0], d.y, c.x[0] * d.x)),
             fmaf(c.z[1], d.z, fmaf(c.y[1], d.y, c.x[1] * d.x)),
             fmaf(c.z[2], d.z, fmaf(c.y[2], d.y, c.x[2] * d.x)));
}

// ---------------------------------------------------------------- ImageTexture::eval (src/main.cpp:62-76)
// Bilinear fetch with clamped coordinates (texel_fetch_wrap, :56-60).  The reference's texel_fetch computes
// pixel_index = y * m_image_size.y() + x (:52) — wrong for non-square images (it should be the ROW LENGTH, size.x);
// fixed here and in the oracle, noted in DESIGN.md.  Operation order as written there: scaled = uv * size - 0.5,
// pos = floor, d1 = scaled - pos, d2 = 1 - d1, sum of four texel * dx * dy products left to right, no FMA.
struct MatTex { uint32_t offset, w, h, pad; };     // texel pool window of one material; w * h == 1: constant albedo
#if defined(__CUDA_ARCH__)
#define WPT_MUL(a, b) __fmul_rn(a, b)
#define WPT_ADD(a, b) __fadd_rn(a, b)
#define WPT_SUB(a, b) __fsub_rn(a, b)
#else
#define WPT_MUL(a, b) ((a) * (b))
#define WPT_ADD(a, b) ((a) + (b))
#define WPT_SUB(a, b) ((a) - (b))
#endif
WPT_HD void texture_eval(const float4 *texels, const MatTex mt, float tu, float tv, float rgb[3])
{
    const float su = WPT_SUB(WPT_MUL(tu, (float)mt.w), 0.5f), sv = WPT_SUB(WPT_MUL(tv, (float)mt.h), 0.5f);
    const float fx = floorf(su), fy = floorf(sv);
    const int x = (int)fx, y = (int)fy;
    const float dx1 = WPT_SUB(su, fx), dy1 = WPT_SUB(sv, fy);
    const float dx2 = WPT_SUB(1.0f, dx1), dy2 = WPT_SUB(1.0f, dy1);
    const int W = (int)mt.w, H = (int)mt.h;
    const int x0 = x < 0 ? 0 : (x > W - 1 ? W - 1 : x), x1 = x + 1 < 0 ? 0 : (x + 1 > W - 1 ? W - 1 : x + 1);
    const int y0 = y < 0 ? 0 : (y > H - 1 ? H - 1 : y), y1 = y + 1 < 0 ? 0 : (y + 1 > H - 1 ? H - 1 : y + 1);
    const float4 *t = texels + mt.offset;
    const float4 t00 = t[y0 * W + x0], t01 = t[y1 * W + x0], t10 = t[y0 * W + x1], t11 = t[y1 * W + x1];
#define WPT_BILERP(C)                                                                                            \
    WPT_ADD(WPT_ADD(WPT_ADD(WPT_MUL(WPT_MUL(t00.C, dx2), dy2), WPT_MUL(WPT_MUL(t01.C, dx2), dy1)),                   \
                    WPT_MUL(WPT_MUL(t10.C, dx1), dy2)), WPT_MUL(WPT_MUL(t11.C, dx1), dy1))
    rgb[0] = WPT_BILERP(x); rgb[1] = WPT_BILERP(y); rgb[2] = WPT_BILERP(z);
#undef WPT_BILERP
}
// interpolated texture coordinate of a hit (src/accel/optix_backend.h:479-480, src/add_math.h:3-7)
WPT_HD void interp_uv(const float2 a, const float2 b, const float2 c, float u, float v, float &tu, float &tv)
{
    const float w = WPT_SUB(WPT_SUB(1.0f, u), v);
    tu = WPT_ADD(WPT_ADD(WPT_MUL(w, a.x), WPT_MUL(u, b.x)), WPT_MUL(v, c.x));
    tv = WPT_ADD(WPT_ADD(WPT_MUL(w, a.y), WPT_MUL(u, b.y)), WPT_MUL(v, c.y));
}

// ---------------------------------------------------------------- one shading step
// src/accel/optix_backend.h:469 (p = o + t d), :412-420 + src/add_math.h:3-7 (un-normalised
// interpolated shading normal), src/main.cpp:414-424.  No FMA where the reference has none.
WPT_HD void shade_step(V3 o, V3 d, float t, float u, float v, V3 n0, V3 n1, V3 n2, float s0, float s1,
                       V3 &next_o, V3 &next_d)
{
#if defined(__CUDA_ARCH__)
    next_o = v3(__fadd_rn(o.x, __fmul_rn(t, d.x)), __fadd_rn(o.y, __fmul_rn(t, d.y)), __fadd_rn(o.z, __fmul_rn(t, d.z)));
    const float w = __fsub_rn(__fsub_rn(1.0f, u), v);
    const V3 n = v3(__fadd_rn(__fadd_rn(__fmul_rn(w, n0.x), __fmul_rn(u, n1.x)), __fmul_rn(v, n2.x)),
                    __fadd_rn(__fadd_rn(__fmul_rn(w, n0.y), __fmul_rn(u, n1.y)), __fmul_rn(v, n2.y)),
                    __fadd_rn(__fadd_rn(__fmul_rn(w, n0.z), __fmul_rn(u, n1.z)), __fmul_rn(v, n2.z)));
#else
    next_o = v3(o.x + t * d.x, o.y + t * d.y, o.z + t * d.z);
    const float w = 1.0f - u - v;
    const V3 n = v3(w * n0.x + u * n1.x + v * n2.x, w * n0.y + u * n1.y + v * n2.y, w * n0.z + u * n1.z + v * n2.z);
#endif
    next_d = frame_to_world(n, cosine_hemisphere(s0, s1));
}

}  // namespace wpt
