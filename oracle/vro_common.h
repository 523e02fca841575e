/*
 * vro_common.h -- the fixed-function stages of the reference pipeline, restated once
 * for both oracle builds (TEST INFRASTRUCTURE ONLY, see vr_oracle.h).
 *
 * The reference has no CPU code for these stages: they are Vulkan pipeline state
 * executed by the driver/hardware.  Each function cites the state it restates and
 * the Vulkan 1.0 rule applied.  All per-pixel arithmetic is fp32, one rounding per
 * operation (compile with -ffp-contract=off), so the CUDA path can reproduce the
 * sample positions bit for bit.
 *
 *   S2  coverage + ray entry   offscreen_pass.cpp:55-90 (cube), :675-712 (cull/depth),
 *                              :190-205 (viewport), res/shaders/volume.vert:19-24
 *   S4  volume sampler         offscreen_pass.cpp:1014-1039 (LINEAR, CLAMP_TO_BORDER, transparent black)
 *   S6  LUT image + sampler    offscreen_pass.cpp:1049-1150 (R8G8B8A8_SRGB 1-D, LINEAR, CLAMP_TO_EDGE)
 *   S7  blend + clear + store  offscreen_pass.cpp:715-725, :170-173, :293
 */
#ifndef VRO_COMMON_H
#define VRO_COMMON_H

#include <math.h>
#include <stdint.h>
#include <string.h>

#include "vr_oracle.h"

/* ------------------------------------------------------------------ S2 -- */

/* Host-side, double precision: columns of inverse(proj * view), rounded to fp32. */
typedef struct vro_rayctx {
    float c0[4], c1[4], c2[4], c3[4]; /* Minv * (x, y, z, 1) = x*c0 + y*c1 + z*c2 + c3 */
    float sx, sy;                     /* 2/W, 2/H */
    float cam[3];
} vro_rayctx;

static inline void vro_mat4_mul_d(const double *a, const double *b, double *out) /* column-major */
{
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++) {
            double s = 0.0;
            for (int k = 0; k < 4; k++) s += a[k * 4 + r] * b[c * 4 + k];
            out[c * 4 + r] = s;
        }
}

/* Cofactor inverse of a column-major 4x4 (layout-agnostic: inverse(M^T) = inverse(M)^T). */
static inline int vro_mat4_inv_d(const double *m, double *inv)
{
    double t[16];
    t[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] +
           m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    t[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] -
           m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    t[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] +
           m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    t[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] -
            m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    t[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] -
           m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    t[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] +
           m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    t[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] -
           m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    t[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] +
            m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    t[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] +
           m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    t[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] -
           m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    t[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] +
            m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    t[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] -
            m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    t[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] -
           m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    t[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] +
           m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    t[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] -
            m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    t[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] +
            m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    double det = m[0] * t[0] + m[1] * t[4] + m[2] * t[8] + m[3] * t[12];
    if (det == 0.0) return -1;
    double r = 1.0 / det;
    for (int i = 0; i < 16; i++) inv[i] = t[i] * r;
    return 0;
}

static inline int vro_rayctx_init(vro_rayctx *rc, const vro_frame *f)
{
    double v[16], p[16], pv[16], inv[16];
    for (int i = 0; i < 16; i++) {
        v[i] = (double)f->view[i];
        p[i] = (double)f->proj[i];
    }
    vro_mat4_mul_d(p, v, pv); /* gl_Position = proj * view * vec4(pos, 1): volume.vert:23 */
    if (vro_mat4_inv_d(pv, inv)) return -1;
    for (int r = 0; r < 4; r++) {
        rc->c0[r] = (float)inv[0 + r];
        rc->c1[r] = (float)inv[4 + r];
        rc->c2[r] = (float)inv[8 + r];
        rc->c3[r] = (float)inv[12 + r];
    }
    rc->sx = 2.0f / (float)f->width;
    rc->sy = 2.0f / (float)f->height;
    for (int i = 0; i < 3; i++) rc->cam[i] = f->camera_position[i];
    return 0;
}

/* Does pixel (px,py) receive a fragment, and where does its ray enter the cube?
 *
 * Vulkan rule: sample point at the pixel centre (px+0.5, py+0.5), framebuffer y down,
 * viewport = full target, depth range [0,1] (offscreen_pass.cpp:190-205).  The cube
 * [-0.5,0.5]^3 (offscreen_pass.cpp:55-86) is convex and back faces are culled
 * (:680-681), so a covered pixel gets exactly one fragment: the front-face point of
 * the line through the pixel, provided it lies inside the clip volume 0 <= z_ndc <= 1
 * (no depth clamp, :676).  The line is pn + s*(pf-pn), pn/pf = unprojected points on the
 * near (z_ndc 0) and far (z_ndc 1) planes, so 0 <= s <= 1 <=> inside the clip volume.
 * frag = interpolated in_frag_position = entry point; the entry axis is snapped to the
 * face exactly (vertex table: tex = pos + 0.5).
 * Returns 1 if covered and fills frag[3] (model space) and tex[3]. */
static inline int vro_pixel_entry(const vro_rayctx *rc, uint32_t px, uint32_t py, float frag[3], float tex[3])
{
    float xn = ((float)px + 0.5f) * rc->sx - 1.0f;
    float yn = ((float)py + 0.5f) * rc->sy - 1.0f;
    float nh[4], fh[4];
    for (int i = 0; i < 4; i++) {
        nh[i] = (xn * rc->c0[i] + yn * rc->c1[i]) + rc->c3[i];
        fh[i] = nh[i] + rc->c2[i];
    }
    float pn[3], d[3];
    for (int i = 0; i < 3; i++) {
        pn[i] = nh[i] / nh[3];
        d[i] = fh[i] / fh[3] - pn[i];
    }
    float tnear = -INFINITY, tfar = INFINITY;
    int axis = 0;
    for (int i = 0; i < 3; i++) {
        float lo, hi;
        if (d[i] != 0.0f) {
            float inv = 1.0f / d[i];
            float ta = (-0.5f - pn[i]) * inv;
            float tb = (0.5f - pn[i]) * inv;
            lo = fminf(ta, tb);
            hi = fmaxf(ta, tb);
        } else {
            if (pn[i] < -0.5f || pn[i] > 0.5f) return 0;
            lo = -INFINITY;
            hi = INFINITY;
        }
        if (lo > tnear) {
            tnear = lo;
            axis = i;
        }
        if (hi < tfar) tfar = hi;
    }
    if (!(tnear < tfar) || !(tnear >= 0.0f) || !(tnear <= 1.0f)) return 0;
    for (int i = 0; i < 3; i++) frag[i] = pn[i] + tnear * d[i];
    frag[axis] = d[axis] > 0.0f ? -0.5f : 0.5f;
    for (int i = 0; i < 3; i++) tex[i] = frag[i] + 0.5f;
    return 1;
}

/* ------------------------------------------------------------------ S4 -- */

static inline float vro_texel(const vro_volume *v, int x, int y, int z)
{
    /* CLAMP_TO_BORDER + FLOAT_TRANSPARENT_BLACK: offscreen_pass.cpp:1020-1026 */
    if (x < 0 || y < 0 || z < 0 || (uint32_t)x >= v->nx || (uint32_t)y >= v->ny || (uint32_t)z >= v->nz)
        return 0.0f;
    size_t i = ((size_t)z * v->ny + (size_t)y) * v->nx + (size_t)x;
    switch (v->dtype) {
        case VRO_U8: return (float)((const uint8_t *)v->data)[i];
        case VRO_U16: return (float)((const uint16_t *)v->data)[i];
        default: return ((const float *)v->data)[i];
    }
}

static inline float vro_lerp(float p, float q, float w) { return p + w * (q - p); }

/* texture(sampler3D, pos).r with VK_FILTER_LINEAR, normalised coordinates, one mip:
 * texel i has its centre at (i+0.5)/N; u = s*N - 0.5; i0 = floor(u); weight = frac(u). */
static inline float vro_sample_volume(const vro_volume *v, float s, float t, float r)
{
    float u = s * (float)v->nx - 0.5f, vv = t * (float)v->ny - 0.5f, w = r * (float)v->nz - 0.5f;
    float fu = floorf(u), fv = floorf(vv), fw = floorf(w);
    int i = (int)fu, j = (int)fv, k = (int)fw;
    float a = u - fu, b = vv - fv, c = w - fw;
    float x00 = vro_lerp(vro_texel(v, i, j, k), vro_texel(v, i + 1, j, k), a);
    float x10 = vro_lerp(vro_texel(v, i, j + 1, k), vro_texel(v, i + 1, j + 1, k), a);
    float x01 = vro_lerp(vro_texel(v, i, j, k + 1), vro_texel(v, i + 1, j, k + 1), a);
    float x11 = vro_lerp(vro_texel(v, i, j + 1, k + 1), vro_texel(v, i + 1, j + 1, k + 1), a);
    return vro_lerp(vro_lerp(x00, x10, b), vro_lerp(x01, x11, b), c);
}

/* ------------------------------------------------------------------ S6 -- */

/* VK_FORMAT_R8G8B8A8_SRGB (offscreen_pass.cpp:1076): R,G,B are sRGB-decoded to linear
 * BEFORE filtering, A is a/255.  Packed R|G<<8|B<<16|A<<24 (imgui.h:2487-2490). */
static inline float vro_srgb_to_linear_u8(unsigned c8)
{
    double c = (double)c8 / 255.0;
    return (float)(c <= 0.04045 ? c / 12.92 : pow((c + 0.055) / 1.055, 2.4));
}

static inline void vro_decode_lut(const uint32_t *lut, uint32_t n, float *rgba /* n*4 */)
{
    for (uint32_t i = 0; i < n; i++) {
        uint32_t p = lut[i];
        rgba[4 * i + 0] = vro_srgb_to_linear_u8(p & 0xff);
        rgba[4 * i + 1] = vro_srgb_to_linear_u8((p >> 8) & 0xff);
        rgba[4 * i + 2] = vro_srgb_to_linear_u8((p >> 16) & 0xff);
        rgba[4 * i + 3] = (float)((p >> 24) & 0xff) / 255.0f;
    }
}

/* texture(sampler1D, t): LINEAR, CLAMP_TO_EDGE (offscreen_pass.cpp:1127-1143).
 * t = NaN (max_density == min_density, unguarded in volume.frag:42) is DEFINED here as
 * texel 0; +-inf clamp to the end texels. */
static inline void vro_sample_lut(const float *rgba, uint32_t n, float t, float out[4])
{
    float u = t * (float)n - 0.5f;
    u = (u > -1.0f) ? u : -1.0f; /* also maps NaN to -1 */
    u = (u < (float)n) ? u : (float)n;
    float fu = floorf(u);
    int i0 = (int)fu;
    float w = u - fu;
    int j0 = i0 < 0 ? 0 : (i0 > (int)n - 1 ? (int)n - 1 : i0);
    int j1 = i0 + 1 < 0 ? 0 : (i0 + 1 > (int)n - 1 ? (int)n - 1 : i0 + 1);
    for (int c = 0; c < 4; c++) out[c] = vro_lerp(rgba[4 * j0 + c], rgba[4 * j1 + c], w);
}

/* ------------------------------------------------------------------ S7 -- */

#define VRO_CLEAR_RGB 0.11f /* offscreen_pass.cpp:171 */

static inline uint8_t vro_unorm8(float x)
{
    x = x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x);
    if (!(x == x)) x = 0.0f;
    return (uint8_t)(x * 255.0f + 0.5f);
}

/* Blend the shader output over the cleared R8G8B8A8_UNORM target with
 * src=SRC_ALPHA, dst=ONE_MINUS_SRC_ALPHA for colour AND alpha (offscreen_pass.cpp:715-725). */
static inline void vro_blend_store(const float c[4], uint8_t out[4])
{
    float a = c[3], ia = 1.0f - c[3];
    out[0] = vro_unorm8(c[0] * a + VRO_CLEAR_RGB * ia);
    out[1] = vro_unorm8(c[1] * a + VRO_CLEAR_RGB * ia);
    out[2] = vro_unorm8(c[2] * a + VRO_CLEAR_RGB * ia);
    out[3] = vro_unorm8(a * a + 1.0f * ia);
}

static inline void vro_clear_store(uint8_t out[4])
{
    out[0] = out[1] = out[2] = vro_unorm8(VRO_CLEAR_RGB);
    out[3] = 255;
}

#endif /* VRO_COMMON_H */
