/*
 * march.cu -- K1, the ray-march kernel (sm_100a).
 *
 * One launch does what one draw of the reference's offscreen pass does
 * (src/rendering/offscreen_pass.cpp:163-230): per pixel, the coverage decision and ray
 * entry of the rasterised front faces of the unit cube (res/shaders/volume.vert:19-24,
 * offscreen_pass.cpp:55-90,:661-712), the fragment-shader loop (res/shaders/volume.frag:21-52),
 * and the fixed-function blend over the clear colour into RGBA8 (offscreen_pass.cpp:170-173,
 * :715-725,:293).
 *
 * Mapping: a CTA of 128 threads (tunable, at most VR_CTA_THREADS) owns a 16x8 pixel tile (tiles are walked in a spiral from the
 * centre of the cube's screen bounding box, the remaining CTAs clear the rest of the target); each warp an
 * 8x4 block of it, so the 32 rays of a warp stay within a few voxels of one another at every step and share
 * texture cache lines.  The transfer function is staged in shared memory (exact sampler: decoded sRGB ->
 * linear fp32, edge-padded so the two taps need no clamp; hardware sampler: fp16 midpoint/difference pairs,
 * see lut_eval).  The volume is read through texture objects: a 3-D CUDA array (hardware block-linear =
 * bricked layout, one trilinear fetch per sample) or one of the paired 2-D layered copies (one bilinear
 * fetch + one fp32 lerp per sample, see fetch_pair).  Empty-space skipping reads per-octant distance maps
 * from global memory (see skip_empty).
 *
 * This file is compiled with -fmad=false: every fp32 operation rounds once, in the order
 * written, so ray entry and the sample positions (volume.frag:47 accumulates them) are
 * bit-identical to the CPU oracle.  Fused multiply-adds are written explicitly (fmaf) where
 * the hardware-sampler path is allowed to differ in rounding.
 */
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "vr_ctx.h"

/* Blocks of 128 threads per SM the hardware-sampler build is compiled for: 5 x 256 threads = 10 blocks of 128
 * = 48 registers per thread (without the cap: 54 registers, 9 blocks). */
/* 1: the cell index of an empty-space look-up comes out of a magic-number addition instead of three float -> int and
 * three int -> float conversions (skip_empty) */
#ifndef VR_ESS_MAGIC
#define VR_ESS_MAGIC 1
#endif
#ifndef VR_MIN_BLOCKS
#define VR_MIN_BLOCKS 5
#endif

namespace
{

struct Entry {
    float frag[3];
    float tex[3];
};

/* S2: does pixel (px,py) get a fragment, and where does its ray enter the cube?
 * Pixel centre -> NDC -> line through the near/far plane points (inverse(proj*view)),
 * slab test against [-0.5,0.5]^3, front face inside the clip volume 0 <= z_ndc <= 1. */
__device__ __forceinline__ bool pixel_entry(const MarchParams &p, uint32_t px, uint32_t py, Entry &e)
{
    const float xn = ((float)px + 0.5f) * p.sx - 1.0f;
    const float yn = ((float)py + 0.5f) * p.sy - 1.0f;
    float nh[4], fh[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        nh[i] = (xn * p.c0[i] + yn * p.c1[i]) + p.c3[i];
        fh[i] = nh[i] + p.c2[i];
    }
    float pn[3], d[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        pn[i] = nh[i] / nh[3];
        d[i] = fh[i] / fh[3] - pn[i];
    }
    float tnear = -CUDART_INF_F, tfar = CUDART_INF_F;
    int axis = 0;
    bool miss = false;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        float lo, hi;
        if (d[i] != 0.0f) {
            const float inv = 1.0f / d[i];
            const float ta = (-0.5f - pn[i]) * inv;
            const float tb = (0.5f - pn[i]) * inv;
            lo = fminf(ta, tb);
            hi = fmaxf(ta, tb);
        } else {
            if (pn[i] < -0.5f || pn[i] > 0.5f) miss = true;
            lo = -CUDART_INF_F;
            hi = CUDART_INF_F;
        }
        if (lo > tnear) {
            tnear = lo;
            axis = i;
        }
        if (hi < tfar) tfar = hi;
    }
    if (miss || !(tnear < tfar) || !(tnear >= 0.0f) || !(tnear <= 1.0f)) return false;
#pragma unroll
    for (int i = 0; i < 3; i++) e.frag[i] = pn[i] + tnear * d[i];
    /* the entry coordinate is the face itself (vertex table: tex = pos + 0.5) */
#pragma unroll
    for (int i = 0; i < 3; i++)
        if (i == axis) e.frag[i] = d[i] > 0.0f ? -0.5f : 0.5f;
#pragma unroll
    for (int i = 0; i < 3; i++) e.tex[i] = e.frag[i] + 0.5f;
    return true;
}

__device__ __forceinline__ float lerp1(float a, float b, float w) { return a + w * (b - a); }

/* S4, verification sampler: VK_FILTER_LINEAR with CLAMP_TO_BORDER(0) restated as eight point
 * fetches (border mode returns 0 outside) and fp32 lerps in x, y, z order. */
__device__ __forceinline__ float sample_exact(const MarchParams &p, float s, float t, float r)
{
    const float u = s * p.dims[0] - 0.5f, v = t * p.dims[1] - 0.5f, w = r * p.dims[2] - 0.5f;
    const float fu = floorf(u), fv = floorf(v), fw = floorf(w);
    const float a = u - fu, b = v - fv, c = w - fw;
    const float x0 = (float)((int)fu - p.org[0]) + 0.5f, y0 = (float)((int)fv - p.org[1]) + 0.5f,
                z0 = (float)((int)fw - p.org[2]) + 0.5f;
    const float x1 = x0 + 1.0f, y1 = y0 + 1.0f, z1 = z0 + 1.0f;
    float v000, v100, v010, v110, v001, v101, v011, v111;
#define VR_FETCH8(T)                                                                                      \
    v000 = (float)tex3D<T>(p.tex_pt, x0, y0, z0), v100 = (float)tex3D<T>(p.tex_pt, x1, y0, z0),           \
    v010 = (float)tex3D<T>(p.tex_pt, x0, y1, z0), v110 = (float)tex3D<T>(p.tex_pt, x1, y1, z0),           \
    v001 = (float)tex3D<T>(p.tex_pt, x0, y0, z1), v101 = (float)tex3D<T>(p.tex_pt, x1, y0, z1),           \
    v011 = (float)tex3D<T>(p.tex_pt, x0, y1, z1), v111 = (float)tex3D<T>(p.tex_pt, x1, y1, z1)
    switch (p.dtype) {
        case VR_U8: VR_FETCH8(unsigned char); break;
        case VR_U16: VR_FETCH8(unsigned short); break;
        default: VR_FETCH8(float); break;
    }
#undef VR_FETCH8
    const float x00 = lerp1(v000, v100, a), x10 = lerp1(v010, v110, a);
    const float x01 = lerp1(v001, v101, a), x11 = lerp1(v011, v111, a);
    return lerp1(lerp1(x00, x10, b), lerp1(x01, x11, b), c);
}

/* Directional distance (in cells) of the cell holding this position, from the map of the ray's direction
 * octant (doff): 0 = the cell may contribute, sample here; d >= 1 = the cube of d cells that starts at this
 * cell and extends along the ray's octant is empty (volume.cu k_octant_distance).
 * A one-ulp slip across a cell face is harmless: every cell's range includes a 1-voxel apron. */
template <bool PART>
__device__ __forceinline__ int cell_distance(const MarchParams &p, uint32_t doff, float x, float y, float z)
{
    int cx = PART ? (int)(x * p.occ_scale[0] - p.occ_off[0]) : (int)(x * p.occ_scale[0]);
    int cy = PART ? (int)(y * p.occ_scale[1] - p.occ_off[1]) : (int)(y * p.occ_scale[1]);
    int cz = PART ? (int)(z * p.occ_scale[2] - p.occ_off[2]) : (int)(z * p.occ_scale[2]);
    cx = min(max(cx, 0), p.occ_n[0] - 1);
    cy = min(max(cy, 0), p.occ_n[1] - 1);
    cz = min(max(cz, 0), p.occ_n[2] - 1);
    /* 2 MiB per octant at most, read through L1 / L2 */
    return (int)__ldg(p.dist + doff + (cz * p.occ_stride[2] + cy * p.occ_stride[1] + cx * p.occ_stride[0]));
}

#define VR_MAGIC 12582912.0f /* 1.5 * 2^23: x + VR_MAGIC rounds x to an integer held in the low mantissa bits */
#define VR_MAGIC_BITS 0x4B400000u

/* Paired layout, one sample: the layer holding (v[fl], v[fl+1]) with fl = floor(w), w = z texel coordinate
 * - 0.5, is fl + 1 = round(w + 0.5), read from the mantissa of q_z * a + b (b carries 1.5*2^23; at a tie
 * either neighbour gives the same value because the z weight is then 0 or 1).  Returns the pair, and the
 * z weight w - fl evaluated with a single rounding. */
template <bool PART, bool SPLIT>
__device__ __forceinline__ float2 fetch_pair(const MarchParams &p, float qx, float qy, float qz, float &frac)
{
    const float tm = fmaf(qz, p.pair_a[2], p.pair_b[2]);
    const float fl1 = tm - VR_MAGIC;                        /* fl + 1 */
    const int layer = (int)(__float_as_uint(tm) & 0xffffu); /* 0 .. n_w (n_w <= 4095) */
    frac = fmaf(qz, p.pair_a[2], p.pair_zc - fl1);
    const float u = PART ? fmaf(qx, p.pair_a[0], p.pair_b[0]) : qx, v = PART ? fmaf(qy, p.pair_a[1], p.pair_b[1]) : qy;
    /* axes longer than 2047 voxels: the layers from pair_split on live in a second array */
    if (!SPLIT) return tex2DLayered<float2>(p.tex_pair, u, v, layer);
    return layer < p.pair_split ? tex2DLayered<float2>(p.tex_pair, u, v, layer)
                                : tex2DLayered<float2>(p.tex_pair_hi, u, v, layer - p.pair_split);
}

/* Hardware-sampler transfer function.  Table entry k (shared memory, 16 bytes = eight fp16) holds
 * M[k] = P[k] + D[k]/2 and D[k] = P[k+1] - P[k] of the edge-padded LUT, so with k = round(u - 1/2) and
 * wm = u - 1/2 - k in [-1/2, 1/2] the CLAMP_TO_EDGE / LINEAR result is M[k] + wm * D[k]: one LDS.128 and
 * two HFMA2, no float<->int conversion (k comes from the mantissa of u - 1/2 + 1.5*2^23). */
struct LutH {
    uint32_t base; /* shared-window address of entry 0, minus (VR_MAGIC_BITS << 4) */
    float A, Bh, hi;
};
__device__ __forceinline__ void lut_eval(const LutH &L, float dens, float &sr, float &sg, float &sb, float &sa)
{
    float uh = fmaf(dens, L.A, L.Bh); /* padded LUT coordinate - 1/2 */
    uh = fminf(fmaxf(uh, -0.5f), L.hi);
    const float tk = uh + VR_MAGIC;
    const float wm = uh - (tk - VR_MAGIC);
    uint4 e;
    asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(e.x), "=r"(e.y), "=r"(e.z), "=r"(e.w) : "r"((__float_as_uint(tk) << 4) + L.base));
    const __half2 w2 = __float2half2_rn(wm);
    const float2 rg = __half22float2(__hfma2(w2, *reinterpret_cast<const __half2 *>(&e.z), *reinterpret_cast<const __half2 *>(&e.x)));
    const float2 ba = __half22float2(__hfma2(w2, *reinterpret_cast<const __half2 *>(&e.w), *reinterpret_cast<const __half2 *>(&e.y)));
    sr = rg.x, sg = rg.y, sb = ba.x, sa = ba.y;
}

/* Per-thread state of one ray. */
struct Ray {
    float cr, cg, cb, T;   /* volume.frag:22 color (T = color.a = transmittance) */
    float px, py, pz;      /* ray_pos */
    float dx, dy, dz;      /* ray_dir * step_size */
    int i;                 /* loop counter */
    bool done;             /* no more positions to generate */
    bool stop;             /* early termination: composite nothing more */
    uint32_t nb, ns, nf;   /* counters (STATS) */
    float ifx, ify, ifz;   /* ESS: 1 / (d * occ_scale): steps per cell, per axis (1e30 where d = 0) */
    float sgx, sgy, sgz;   /* ESS: sign of the direction, per axis */
    uint32_t doff;         /* ESS: offset of the distance map of the direction's octant */
    bool in_empty;         /* ESS, hardware sampler: the lane sits the coming batch out (skip_empty) */
    /* sort-last, hardware sampler: positions are a pure function of the step index, pos_i = e + i * d (one FMA),
     * so that every rank computes the same position for step i wherever it starts walking (see outer_interval) */
    float ex, ey, ez, fi;
    int i_end;             /* last step that can still be sampled (sort-last: where the ray leaves the rank's box) */
    /* pre-integrated classification (PRE): table coordinate of the sample before this one, and whether there is one */
    float prev;
    bool prev_ok;
};

/* volume.frag:47, `ray_pos += ray_dir * step`.  FUNC (sort-last + hardware sampler): the same position evaluated as
 * e + i * d with a single rounding -- within steps * ulp / 2 of the accumulated one (the bound safe_interval
 * already allows for), identical on every rank. */
template <bool FUNC>
__device__ __forceinline__ void advance(Ray &r)
{
    if (FUNC) {
        r.fi += 1.0f;
        r.px = fmaf(r.fi, r.dx, r.ex), r.py = fmaf(r.fi, r.dy, r.ey), r.pz = fmaf(r.fi, r.dz, r.ez);
    } else {
        r.px += r.dx, r.py += r.dy, r.pz += r.dz;
    }
    r.i++;
}

/* The positions volume.frag:47 accumulates are monotone on every axis, and differ from
 * tex + i*d by at most steps * ulp(1)/2.  So the steps that pass the bounds test (:34-37), the
 * strict slice test (:39-40) and the sort-last ownership test form one interval of i, and the
 * analytic interval shrunk by that error bound and by one step at each end is inside it:
 * those steps can run without evaluating any of the tests, with identical results. */
__device__ __forceinline__ void safe_interval(const MarchParams &p, const Ray &r, float e0, float e1, float e2, int &fast_lo, int &fast_hi)
{
    const float eps = (float)(p.steps + 2) * 1.2e-7f;
    float tlo = 0.0f, thi = (float)(p.steps - 1);
    const float p0[3] = {e0, e1, e2}, d[3] = {r.dx, r.dy, r.dz}; /* the entry point: step 0 */
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float lo = p.box_lo[a] + eps, hi = p.box_hi[a] - eps;
        if (d[a] != 0.0f) {
            const float inv = __frcp_rn(d[a]);
            const float ta = (lo - p0[a]) * inv, tb = (hi - p0[a]) * inv;
            tlo = fmaxf(tlo, fminf(ta, tb));
            thi = fminf(thi, fmaxf(ta, tb));
        } else if (!(p0[a] > lo && p0[a] < hi)) {
            thi = -1.0f;
        }
    }
    if (tlo == tlo && thi == thi && thi - tlo >= 4.0f && tlo < 1.0e6f) {
        fast_lo = (int)ceilf(tlo) + 1;
        fast_hi = (int)floorf(thi) - 1;
    }
}

/* Sort-last: the steps that CAN fall into the rank's box (bounds, slices and ownership intersected) lie inside the
 * analytic interval of the box grown by the same error bound and by one step at each end.  A rank walks only
 * those: it starts at step lo with pos = e + lo * d instead of stepping through its neighbours' boxes. */
__device__ __forceinline__ void outer_interval(const MarchParams &p, const Ray &r, int &lo_i, int &hi_i)
{
    const float eps = (float)(p.steps + 2) * 1.2e-7f;
    float tlo = 0.0f, thi = (float)(p.steps - 1);
    const float p0[3] = {r.px, r.py, r.pz}, d[3] = {r.dx, r.dy, r.dz};
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float lo = p.box_lo[a] - eps, hi = p.box_hi[a] + eps;
        if (d[a] != 0.0f) {
            const float inv = __frcp_rn(d[a]);
            const float ta = (lo - p0[a]) * inv, tb = (hi - p0[a]) * inv;
            tlo = fmaxf(tlo, fminf(ta, tb));
            thi = fminf(thi, fmaxf(ta, tb));
        } else if (!(p0[a] >= lo && p0[a] <= hi)) {
            thi = -4.0f;
        }
    }
    lo_i = 0, hi_i = p.steps - 1;
    if (tlo == tlo && thi == thi && tlo < 1.0e6f) {
        lo_i = max((int)floorf(tlo) - 1, 0);
        hi_i = min((int)ceilf(thi) + 1, p.steps - 1);
    }
}

/* Empty-space skipping by directional distance map (hardware sampler only).  If the cell holding the
 * current position is empty with distance d, the cube of d cells that starts at this cell and extends along
 * the ray's octant is empty: the ray jumps k steps, to the first position that may lie outside the cube.
 * Every skipped position lies strictly inside the cube, i.e. is one the per-step test would have rejected
 * anyway.  The jump is one fused multiply-add per axis, pos += k * d, where volume.frag:47 makes k
 * additions: the positions after a jump differ from the reference's by the accumulated rounding of those
 * additions (< steps * ulp(1)/2 ~ 2e-5 texture units, far below the 1/256-texel weight resolution of the
 * filter) -- the exact sampler never jumps.  Jumps are taken only inside the ray's safe interval (see
 * safe_interval), where every position is inside the volume, the slices and the rank's own box: no
 * clipping of the cube, and every skipped step counts as visited for the statistics. */
template <bool STATS, bool PART>
__device__ __forceinline__ void skip_empty(const MarchParams &p, Ray &r, int fast_lo, int fast_hi)
{
    /* One look-up per BATCH, here, not one per step (hardware sampler): the batch that follows samples all its
     * steps when the cell of its first position may contribute, and none -- the lane sits the batch out, its position
     * unchanged -- when eight jumps did not reach such a cell.  A step sampled in an empty cell next to an occupied one
     * composites alpha 0: the result is unchanged, a blob costs at most U - 1 wasted fetches where the ray leaves it. */
    r.in_empty = false;
    if (r.done) return;
#pragma unroll 1
    for (int guard = 0; guard < p.ess_guard; guard++) {
        /* outside the safe interval (the first and last steps of the ray, at a slicing plane): sample every step */
        if (r.i < fast_lo || r.i > fast_hi) return;
        if (VR_ESS_MAGIC) {
            /* g = position in cells - 1/2; g + M rounds it to the cell index, held in the low mantissa bits (a position
             * exactly on a face may land in either cell: both hold it) -- no float <-> integer conversion for the index.
             * Inside the safe interval the position is strictly inside the grid: no clamp. */
            const float gx = fmaf(r.px, p.occ_scale[0], -p.occ_offh[0]), gy = fmaf(r.py, p.occ_scale[1], -p.occ_offh[1]),
                        gz = fmaf(r.pz, p.occ_scale[2], -p.occ_offh[2]);
            const float mx = gx + VR_MAGIC, my = gy + VR_MAGIC, mz = gz + VR_MAGIC;
            const int cx = __float_as_int(mx) - (int)VR_MAGIC_BITS, cy = __float_as_int(my) - (int)VR_MAGIC_BITS, cz = __float_as_int(mz) - (int)VR_MAGIC_BITS;
            const int d = (int)__ldg(p.dist + r.doff + (cz * p.occ_stride[2] + cy * p.occ_stride[1] + cx * p.occ_stride[0]));
            if (d == 0) return;
            /* as below: exit plane = cell centre +- (d - 1/2 - margin), and centre - position = c - g */
            const float R = (float)d - (0.5f + 2.0e-3f);
            const float tx = fmaf(r.sgx, R, (mx - VR_MAGIC) - gx) * r.ifx;
            const float ty = fmaf(r.sgy, R, (my - VR_MAGIC) - gy) * r.ify;
            const float tz = fmaf(r.sgz, R, (mz - VR_MAGIC) - gz) * r.ifz;
            int k = (int)fminf(fminf(fminf(tx, ty), tz), 65536.0f);
            k = min(max(k, 1), fast_hi - r.i + 1);
            r.i += k;
            if (STATS) r.nb += (uint32_t)k, r.ns += (uint32_t)k;
            const float kf = (float)k;
            if (PART) {
                r.fi += kf;
                r.px = fmaf(r.fi, r.dx, r.ex), r.py = fmaf(r.fi, r.dy, r.ey), r.pz = fmaf(r.fi, r.dz, r.ez);
            } else {
                r.px = fmaf(kf, r.dx, r.px), r.py = fmaf(kf, r.dy, r.py), r.pz = fmaf(kf, r.dz, r.pz);
            }
            continue;
        }
        /* position in cell units (> 0 and inside the grid: the safe interval is inside the box) */
        const float fx = PART ? fmaf(r.px, p.occ_scale[0], -p.occ_off[0]) : r.px * p.occ_scale[0];
        const float fy = PART ? fmaf(r.py, p.occ_scale[1], -p.occ_off[1]) : r.py * p.occ_scale[1];
        const float fz = PART ? fmaf(r.pz, p.occ_scale[2], -p.occ_off[2]) : r.pz * p.occ_scale[2];
        const int cx = min((int)fx, p.occ_n[0] - 1), cy = min((int)fy, p.occ_n[1] - 1), cz = min((int)fz, p.occ_n[2] - 1);
        const int d = (int)__ldg(p.dist + r.doff + (cz * p.occ_stride[2] + cy * p.occ_stride[1] + cx * p.occ_stride[0]));
        if (d == 0) return;
        /* steps until the ray leaves the cube: per axis the exit plane is the cell's entry face + d cells
         * along the direction, i.e. the cell centre +- (d - 1/2), pulled in by a margin that covers the
         * rounding of the positions (<= steps * ulp, far below 2e-3 cells); sg = direction sign,
         * if = steps per cell (1e30 on a zero component) */
        const float R = (float)d - (0.5f + 2.0e-3f);
        const float tx = (fmaf(r.sgx, R, (float)cx + 0.5f) - fx) * r.ifx;
        const float ty = (fmaf(r.sgy, R, (float)cy + 0.5f) - fy) * r.ify;
        const float tz = (fmaf(r.sgz, R, (float)cz + 0.5f) - fz) * r.ifz;
        /* positions 0..k-1 are inside the cube; position 0 is the current one, known to be empty */
        int k = (int)fminf(fminf(fminf(tx, ty), tz), 65536.0f);
        k = min(max(k, 1), fast_hi - r.i + 1);
        r.i += k;
        if (STATS) r.nb += (uint32_t)k, r.ns += (uint32_t)k;
        const float kf = (float)k;
        if (PART) {
            r.fi += kf;
            r.px = fmaf(r.fi, r.dx, r.ex), r.py = fmaf(r.fi, r.dy, r.ey), r.pz = fmaf(r.fi, r.dz, r.ez);
        } else {
            r.px = fmaf(kf, r.dx, r.px), r.py = fmaf(kf, r.dy, r.py), r.pz = fmaf(kf, r.dz, r.pz);
        }
    }
    r.in_empty = true; /* still in empty space: no step of the coming batch */
}

/* One batch of U consecutive iterations of volume.frag:33-48, split in two halves so that the
 * fetches of the next batch can be in flight while this one is composited. */
template <int U>
struct Batch {
    float val[U];
    float val2[U], frac[U]; /* paired layout: second slice of the pair and the z weight */
    bool ok[U];
    bool pred; /* warp-uniform: ok[] is meaningful (otherwise every lane samples every step) */
};

/* First half: positions (volume.frag:33-40,:47), then all U fetches back to back (:41).
 * CHECK = evaluate the per-step tests; otherwise every step of a live ray is sampled. */
template <bool EXACT, bool ESS, bool STATS, bool PART, int PAIR, int U, bool CHECK, bool PRED>
__device__ __forceinline__ void batch_issue(const MarchParams &p, Ray &r, Batch<U> &b)
{
    float qx[U], qy[U], qz[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
        bool s;
        const bool out = ESS && !EXACT && r.in_empty; /* sits this batch out (skip_empty) */
        if (CHECK) {
            /* volume.frag:33-37: loop bound and the strict bounds test */
            const bool inb = !r.done && !out && (PART ? r.i <= r.i_end : r.i < p.steps) &&
                             !(r.px > 1.0f || r.py > 1.0f || r.pz > 1.0f || r.px < 0.0f || r.py < 0.0f || r.pz < 0.0f);
            if (!inb && !out) r.done = true;
            /* volume.frag:39-40: strict on both sides */
            s = inb && r.px < p.max_slice[0] && r.py < p.max_slice[1] && r.pz < p.max_slice[2] &&
                r.px > p.min_slice[0] && r.py > p.min_slice[1] && r.pz > p.min_slice[2];
            if (PART)
                s = s && r.px >= p.own_lo[0] && r.py >= p.own_lo[1] && r.pz >= p.own_lo[2] && r.px < p.own_hi[0] &&
                    r.py < p.own_hi[1] && r.pz < p.own_hi[2];
            if (STATS) {
                r.nb += inb;
                r.ns += s;
            }
        } else {
            s = PRED ? (!r.done && !out) : true;
            if (STATS) {
                r.nb += !r.done && !out;
                r.ns += !r.done && !out;
            }
        }
        if (ESS && EXACT) {
            /* the exact sampler walks every step (bit-identical positions) and looks each one up */
            s = s && cell_distance<PART>(p, r.doff, r.px, r.py, r.pz) == 0;
        }
        if (CHECK || PRED) b.ok[u] = s;
        qx[u] = r.px, qy[u] = r.py, qz[u] = r.pz;
        if (ESS && !EXACT) {
            if (!r.in_empty) advance<PART>(r); /* a lane that sits the batch out stays where skip_empty left it */
        } else {
            advance<PART && !EXACT>(r);
        }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
        if (CHECK || PRED) b.val[u] = 0.0f;
        if (PAIR && (CHECK || PRED)) b.val2[u] = 0.0f, b.frac[u] = 0.0f;
        if ((CHECK || PRED) ? b.ok[u] : true) {
            if (EXACT) {
                b.val[u] = sample_exact(p, qx[u], qy[u], qz[u]);
            } else if (PAIR) {
                const float2 t = fetch_pair<PART, PAIR == 2>(p, qx[u], qy[u], qz[u], b.frac[u]);
                b.val[u] = t.x, b.val2[u] = t.y;
            } else if (PART) {
                b.val[u] = tex3D<float>(p.tex_lin, (qx[u] * p.dims[0] - p.forg[0]) * p.inv_ldims[0],
                                        (qy[u] * p.dims[1] - p.forg[1]) * p.inv_ldims[1],
                                        (qz[u] * p.dims[2] - p.forg[2]) * p.inv_ldims[2]);
            } else {
                b.val[u] = tex3D<float>(p.tex_lin, qx[u], qy[u], qz[u]);
            }
        }
    }
    b.pred = CHECK || PRED;
}

/* Second half: normalise, transfer function, composite (volume.frag:42-45). */
template <bool EXACT, bool STATS, int PAIR, int U, bool PRED, bool ERT = true, bool PROBE = false, bool PRE = false>
__device__ __forceinline__ void batch_consume(const MarchParams &p, Ray &r, Batch<U> &b, const float4 *sP,
                                              const LutH &L)
{
    const float lut_hi = (float)p.nlut;
#pragma unroll
    for (int u = 0; u < U; u++) {
        if (!PRED || (b.ok[u] && !r.stop)) {
            if (STATS) r.nf += PRED ? 1u : (r.done ? 0u : 1u);
            float sr, sg, sb, sa;
            if (EXACT) {
                /* volume.frag:42-43 and the LINEAR / CLAMP_TO_EDGE 1-D sampler */
                const float tt = (b.val[u] - p.min_density) / p.range;
                float uu = tt * lut_hi - 0.5f;
                uu = (uu > -1.0f) ? uu : -1.0f; /* NaN -> texel 0 */
                uu = (uu < lut_hi) ? uu : lut_hi;
                const float fu = floorf(uu);
                const int k = (int)fu + 1;
                const float w = uu - fu;
                const float4 a = sP[k], c = sP[k + 1];
                sr = lerp1(a.x, c.x, w), sg = lerp1(a.y, c.y, w), sb = lerp1(a.z, c.z, w), sa = lerp1(a.w, c.w, w);
                /* volume.frag:44-45 */
                r.cr += r.T * (sa * sr);
                r.cg += r.T * (sa * sg);
                r.cb += r.T * (sa * sb);
                r.T *= (1.0f - sa);
                /* early ray termination: what is left contributes < eps per channel */
                if (r.T < p.ert_eps) r.stop = true, r.done = true;
            } else {
                const float dens = PAIR ? fmaf(b.frac[u], b.val2[u] - b.val[u], b.val[u]) : b.val[u];
                if (PROBE) { /* measurement build: the fetched value is consumed, nothing else (vr_bench_fetch_only) */
                    r.cr += dens;
                    continue;
                }
                if (PRE) {
                    /* pre-integrated classification (no reference counterpart; defined in include/vr.h): the SEGMENT between the
                     * sample before and this one is looked up in the 2-D table -- x = front, y = back, unnormalised LUT
                     * coordinates t * n (= the padded-table coordinate of lut_eval), clamp to edge, NaN -> 0 */
                    const float tc = fminf(fmaxf(fmaf(dens, L.A, L.Bh), 0.0f), (float)p.nlut);
                    if (r.prev_ok) {
                        const float4 sg4 = tex2D<float4>(p.tex_pre, r.prev, tc);
                        r.cr = fmaf(r.T, sg4.x, r.cr);
                        r.cg = fmaf(r.T, sg4.y, r.cg);
                        r.cb = fmaf(r.T, sg4.z, r.cb);
                        r.T = fmaf(-r.T, sg4.w, r.T);
                    }
                    r.prev = tc, r.prev_ok = true;
                    continue;
                }
                lut_eval(L, dens, sr, sg, sb, sa);
                const float ta = r.T * sa;
                r.cr = fmaf(ta, sr, r.cr);
                r.cg = fmaf(ta, sg, r.cg);
                r.cb = fmaf(ta, sb, r.cb);
                r.T = fmaf(-r.T, sa, r.T);
            }
        } else if (PRE && PRED && !(b.ok[u])) {
            r.prev_ok = false; /* a step outside the slices (or past the end) breaks the chain of segments */
        }
    }
    /* the hardware path tests for early termination once per batch: the samples composited past
     * the threshold are real samples, the bound on what is dropped is unchanged */
    if (!EXACT && ERT && r.T < p.ert_eps) r.stop = true, r.done = true;
}

/* Which tile block b of the marching blocks works on; false = none (outside the rectangle / not this rank's). */
__device__ __forceinline__ bool tile_of_block(const MarchParams &p, uint32_t block, int &tx, int &ty)
{
    if (p.pt_enum) {
        /* sort-first: the marching blocks enumerate this rank's partition tiles only, row by row; partition tile
         * (x, y) belongs to rank (x + y * pt_m) % pt_nranks, so in row y the owned ones are every pt_nranks-th
         * from the first x >= pt_x0 with that residue */
        const uint32_t per_tile = p.pt_sub_x * p.pt_sub_y, per_row = p.pt_cnt * per_tile;
        const uint32_t row = block / per_row, rem = block - row * per_row;
        const uint32_t j = rem / per_tile, sub = rem - j * per_tile;
        /* rows from the middle of the box outwards (c, c+1, c-1, ...): the long rays start first, the short ones
         * at the top and bottom edge fill the tail */
        const uint32_t mid = (p.pt_rows - 1u) >> 1;
        const uint32_t pty = p.pt_y0 + ((row & 1u) ? mid + ((row + 1u) >> 1) : mid - (row >> 1));
        const uint32_t want = (p.pt_rank + p.pt_nranks - (pty * p.pt_m + p.pt_x0) % p.pt_nranks) % p.pt_nranks;
        const uint32_t ptx = p.pt_x0 + want + j * p.pt_nranks;
        if (ptx >= p.pt_x1) return false;
        const uint32_t sy = sub / p.pt_sub_x, sx = sub - sy * p.pt_sub_x;
        tx = (int)(ptx * p.pt_sub_x + sx), ty = (int)(pty * p.pt_sub_y + sy);
    } else {
        int ox = 0, oy = 0;
        const int b = (int)block;
        if (b > 0) {
            int r = (int)ceilf((sqrtf((float)(b + 1)) - 1.0f) * 0.5f);
            while ((2 * r - 1) * (2 * r - 1) > b) r--;
            while ((2 * r + 1) * (2 * r + 1) <= b) r++;
            /* ring r holds 8r tiles, four sides of 2r: no integer division */
            const int k = b - (2 * r - 1) * (2 * r - 1), r2 = 2 * r;
            const int seg = (k >= r2) + (k >= 2 * r2) + (k >= 3 * r2), off = k - seg * r2;
            if (seg == 0) ox = r, oy = -r + 1 + off;
            else if (seg == 1) ox = r - 1 - off, oy = r;
            else if (seg == 2) ox = -r, oy = r - 1 - off;
            else ox = -r + 1 + off, oy = -r;
        }
        tx = p.sp_cx + ox, ty = p.sp_cy + oy;
    }
    return !(tx < p.rect[0] || ty < p.rect[1] || tx >= p.rect[2] || ty >= p.rect[3]);
}

/* EXACT: verification sampler.  ESS: occupancy grid consulted.  STATS: sample counters.
 * PART: sort-last sub-volume.  PAIR: 0 = 3-D array, 1 = paired layout, 2 = paired layout whose copy spans
 * two layered arrays.  UB: steps per batch (HW sampler; two batches are in flight). */
template <bool EXACT, bool ESS, bool STATS, bool PART, int PAIR, int UB, bool PROBE = false, bool PRE = false>
__global__ void __launch_bounds__(VR_CTA_THREADS, (EXACT || ESS || STATS || PART || PRE || UB > 2) ? 1 : VR_MIN_BLOCKS) k_march(const __grid_constant__ MarchParams p)
{
    extern __shared__ float4 smem[];
    float4 *sP = smem;                              /* EXACT: nlut + 3 decoded texels */
    uint4 *sH = reinterpret_cast<uint4 *>(smem);    /* HW: nlut + 2 packed fp16 entries */

    /* Which tile.  The first sp_count blocks walk the tiles of the cube's screen bounding box in a
     * square spiral from its centre: the longest rays (through the middle of the cube) start
     * first and the short ones at the silhouette fill the tail.  The remaining blocks restore the
     * clear colour where the target does not hold it already (api.cu: dirty clear -- the whole target
     * on the first frame into it, afterwards only what the previous frame drew outside this frame's tiles). */
    const uint32_t nthreads = blockDim.x;
    int tx = 0, ty = 0;
    const bool in_bbox = blockIdx.x < p.sp_count;
    bool tile_empty = false;
    if (in_bbox) {
        /* one warp works the tile out (integer divisions, a square root), the others read it from shared memory */
        __shared__ int s_tile[3];
        if (threadIdx.x < 32) {
            int state = tile_of_block(p, blockIdx.x, tx, ty) ? 1 : 0;
            /* empty-space skipping in screen space: no ray of this tile meets an occupied cell (volume.cu k_tile_mask) --
             * every sample would have alpha 0, the pixels are the clear colour: no ray is set up at all */
            if (state && ESS && !STATS && p.tile_mask) {
                const uint32_t t = (uint32_t)ty * p.ntx + (uint32_t)tx;
                if (!((__ldg(p.tile_mask + (t >> 5)) >> (t & 31u)) & 1u) && __ldg(p.tile_mask + p.mask_words) == 0u) state = 2;
            }
            if (threadIdx.x == 0) s_tile[0] = tx, s_tile[1] = ty, s_tile[2] = state;
        }
        __syncthreads();
        tx = s_tile[0], ty = s_tile[1];
        if (s_tile[2] == 0) return;
        tile_empty = s_tile[2] == 2;
    } else {
        /* clear colour only (offscreen_pass.cpp:170-173): a block clears a 64x32 pixel region in groups of
         * four pixels (one 16-byte store), leaving out the tiles the spiral blocks own */
        const uint32_t j = blockIdx.x - p.sp_count;
        uint32_t rx, ry, rw = 64u, rh = 32u;
        if (p.pt_enum) {
            /* sort-first: one block per partition tile of this rank inside clr_t, row by row */
            const uint32_t row = j / p.pt_cnt_clear, jj = j - row * p.pt_cnt_clear;
            const uint32_t pty = p.clr_t[1] + row;
            const uint32_t want = (p.pt_rank + p.pt_nranks - (pty * p.pt_m + p.clr_t[0]) % p.pt_nranks) % p.pt_nranks;
            const uint32_t ptx = p.clr_t[0] + want + jj * p.pt_nranks;
            if (ptx >= p.clr_t[2]) return;
            rx = ptx * p.pt_px, ry = pty * p.pt_px, rw = rh = p.pt_px;
        } else {
            const uint32_t jy = j / p.nclr_x;
            ry = p.clr_org[1] + jy * 32u, rx = p.clr_org[0] + (j - jy * p.nclr_x) * 64u;
        }
        const uint32_t c8 = (uint32_t)(VR_CLEAR_RGB * 255.0f + 0.5f), cw = c8 | (c8 << 8) | (c8 << 16) | 0xff000000u;
        const bool wide = (p.W & 3u) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15u) == 0;
        const bool rgba = p.out != nullptr; /* sort-last: the float4 partial only */
        /* the spiral blocks' rectangle in pixels (a group of four pixels never straddles a 16-pixel tile column) */
        const int rpx0 = p.rect[0] * VR_TILE, rpx1 = p.rect[2] * VR_TILE, rpy0 = p.rect[1] * (int)p.tile_h, rpy1 = p.rect[3] * (int)p.tile_h;
        /* rows of rw / 4 groups of four pixels, 16 groups side by side per pass */
        for (uint32_t g = threadIdx.x; g < (rw >> 2) * rh; g += nthreads) {
            uint32_t col = g & 15u, rowi = g >> 4;
            if (rw != 64u) rowi = g / (rw >> 2), col = g - rowi * (rw >> 2);
            const uint32_t gx = rx + col * 4u, gy = ry + rowi;
            if (gx >= p.W || gy >= p.H) continue;
            if ((int)gx >= rpx0 && (int)gy >= rpy0 && (int)gx < rpx1 && (int)gy < rpy1) continue;
            if (!p.pt_enum && p.pt_nranks > 1 && ((gy / p.pt_px) * p.pt_m + gx / p.pt_px) % p.pt_nranks != p.pt_rank) continue;
            const size_t gp = (size_t)gy * p.W + gx;
            if (!rgba) {
            } else if (wide) {
                *reinterpret_cast<uint4 *>(p.out + gp) = make_uint4(cw, cw, cw, cw);
            } else {
                for (uint32_t k = 0; k < 4u && gx + k < p.W; k++) p.out[gp + k] = make_uchar4(c8, c8, c8, 255);
            }
            if (p.raw)
                for (uint32_t k = 0; k < 4u && gx + k < p.W; k++) p.raw[gp + k] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return;
    }
    const uint32_t x0 = (uint32_t)tx * VR_TILE, y0 = (uint32_t)ty * p.tile_h;
    if (x0 >= p.W || y0 >= p.H) return; /* uniform per CTA */
    /* sort-first: partition tiles of pt_px pixels dealt round-robin to ranks */
    if (p.pt_nranks > 1 && ((y0 / p.pt_px) * p.pt_m + x0 / p.pt_px) % p.pt_nranks != p.pt_rank) return;

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    /* warp footprint ww x (32/ww) pixels, warps tiled across the 16-pixel-wide CTA tile */
    /* ww is 4, 8 or 16 (api.cu): shifts and masks, no integer division */
    const uint32_t ws = 31u - (uint32_t)__clz((int)p.warp_w), ww = 1u << ws, wh = 32u >> ws, wprs = 4u - ws;
    const uint32_t px = x0 + ((warp & ((1u << wprs) - 1u)) << ws) + (lane & (ww - 1u));
    const uint32_t py = y0 + (warp >> wprs) * wh + (lane >> ws);
    const bool valid = px < p.W && py < p.H;
    const size_t pix = (size_t)py * p.W + px;

    Entry e;
    const bool covered = !tile_empty && valid && pixel_entry(p, px, py, e);

    if (tile_empty || !__syncthreads_or(covered)) {
        /* nothing to march in this tile: clear colour only (offscreen_pass.cpp:170-173) */
        if (valid) {
            const unsigned char c = (unsigned char)(VR_CLEAR_RGB * 255.0f + 0.5f);
            if (p.out) p.out[pix] = make_uchar4(c, c, c, 255);
            if (p.raw) p.raw[pix] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return;
    }

    if (EXACT)
        for (uint32_t i = threadIdx.x; i < p.nlut + 3; i += nthreads) sP[i] = p.lut_pad[i];
    else
        for (uint32_t i = threadIdx.x; i < p.nlut + 2; i += nthreads) sH[i] = p.lut_h[i];
    __syncthreads();

    /* volume.frag:22-24 */
    float cr = 0.0f, cg = 0.0f, cb = 0.0f, T = 1.0f;
    float posx = 0.f, posy = 0.f, posz = 0.f, dx = 0.f, dy = 0.f, dz = 0.f;
    if (covered) {
        /* (* ext, * inv_ext: non-cubic voxels -- the direction is normalised in physical space, the step is taken in the
         * cube's; both are (1, 1, 1) unless spacings are honoured, and a multiplication by 1 changes no bit) */
        const float vx = (e.frag[0] - p.cam[0]) * p.ext[0], vy = (e.frag[1] - p.cam[1]) * p.ext[1], vz = (e.frag[2] - p.cam[2]) * p.ext[2];
        const float d2 = (vx * vx + vy * vy) + vz * vz;
        const float inv = 1.0f / sqrtf(d2); /* glm::normalize = v * inversesqrt(dot(v,v)) */
        dx = ((vx * inv) * p.step_size) * p.inv_ext[0];
        dy = ((vy * inv) * p.step_size) * p.inv_ext[1];
        dz = ((vz * inv) * p.step_size) * p.inv_ext[2];
        posx = e.tex[0], posy = e.tex[1], posz = e.tex[2];
    }
    if (PAIR && p.pair_axis != 2) {
        /* sampling space (u, v, w) with w = the pair axis: (x, z, y) or (y, z, x).  Everything the march
         * tests or looks up per axis was stored in this order by the host. */
        float t = posy;
        posy = posz, posz = t;
        t = dy, dy = dz, dz = t;
        if (p.pair_axis == 0) {
            t = posx, posx = posz, posz = t; /* (x, z, y) -> (y, z, x) */
            t = dx, dx = dz, dz = t;
        }
    }

    Ray ray;
    ray.cr = cr, ray.cg = cg, ray.cb = cb, ray.T = T;
    ray.px = posx, ray.py = posy, ray.pz = posz, ray.dx = dx, ray.dy = dy, ray.dz = dz;
    ray.i = 0, ray.done = !covered, ray.stop = false, ray.ifx = ray.ify = ray.ifz = 0.0f, ray.sgx = ray.sgy = ray.sgz = 1.0f, ray.doff = 0u, ray.prev = 0.0f, ray.prev_ok = false, ray.in_empty = true, ray.nb = 0, ray.ns = 0, ray.nf = 0;

    ray.ex = posx, ray.ey = posy, ray.ez = posz, ray.fi = 0.0f, ray.i_end = p.steps - 1;
    if (PART && !EXACT && covered) {
        /* walk only the steps that can fall into this rank's box */
        int lo_i, hi_i;
        outer_interval(p, ray, lo_i, hi_i);
        if (hi_i < lo_i) {
            ray.done = true;
        } else {
            ray.i = lo_i, ray.fi = (float)lo_i, ray.i_end = hi_i;
            ray.px = fmaf(ray.fi, dx, posx), ray.py = fmaf(ray.fi, dy, posy), ray.pz = fmaf(ray.fi, dz, posz);
        }
    }

    bool parked = false; /* fast phase: this lane only keeps its warp company */
    float t_saved = 1.0f;
    /* Steps fast_lo..fast_hi need no bounds / slice / ownership test (see safe_interval). */
    int fast_lo = 0x7fffffff, fast_hi = -1;
    if (covered && !EXACT) safe_interval(p, ray, posx, posy, posz, fast_lo, fast_hi);

    constexpr int U = EXACT ? 1 : UB;
    LutH L;
    L.base = (uint32_t)__cvta_generic_to_shared(sH) - (VR_MAGIC_BITS << 4);
    asm volatile("" : "+r"(L.base)); /* keep it one register: the lookup address is a single shift-add */
    L.A = p.lutA, L.Bh = p.lutBh, L.hi = (float)p.nlut + 0.5f;
    if (ESS) {
        const float sx_ = ray.dx * p.occ_scale[0], sy_ = ray.dy * p.occ_scale[1], sz_ = ray.dz * p.occ_scale[2];
        ray.ifx = sx_ != 0.0f ? __frcp_rn(sx_) : 1.0e30f;
        ray.ify = sy_ != 0.0f ? __frcp_rn(sy_) : 1.0e30f;
        ray.ifz = sz_ != 0.0f ? __frcp_rn(sz_) : 1.0e30f;
        ray.sgx = ray.ifx < 0.0f ? -1.0f : 1.0f, ray.sgy = ray.ify < 0.0f ? -1.0f : 1.0f, ray.sgz = ray.ifz < 0.0f ? -1.0f : 1.0f;
        ray.doff = (ray.ifx < 0.0f ? p.occ_oct[0] : 0u) + (ray.ify < 0.0f ? p.occ_oct[1] : 0u) + (ray.ifz < 0.0f ? p.occ_oct[2] : 0u);
    }
    /* Warp-uniform choice per batch.  Checked: every test of volume.frag:34-40 is evaluated (always
     * correct).  Unchecked: every live lane is inside its safe interval; predicated if some lane is
     * finished or the occupancy grid is consulted, else no per-lane flag at all. */
#define VR_ISSUE(B_)                                                                                        \
    do {                                                                                                    \
        if (ESS && !EXACT) skip_empty<STATS, PART>(p, ray, fast_lo, fast_hi);                               \
        const bool in_safe = ray.i >= fast_lo && ray.i + U - 1 <= fast_hi;                                  \
        if (!EXACT && !ESS && __all_sync(0xffffffffu, !ray.done && in_safe))                                \
            batch_issue<EXACT, ESS, STATS, PART, PAIR, U, false, false>(p, ray, B_);                        \
        else if (!EXACT && __all_sync(0xffffffffu, ray.done || in_safe))                                    \
            batch_issue<EXACT, ESS, STATS, PART, PAIR, U, false, true>(p, ray, B_);                         \
        else                                                                                                \
            batch_issue<EXACT, ESS, STATS, PART, PAIR, U, true, true>(p, ray, B_);                          \
    } while (0)
#define VR_CONSUME(B_)                                                                                      \
    do {                                                                                                    \
        if (B_.pred) batch_consume<EXACT, STATS, PAIR, U, true, true, PROBE, PRE>(p, ray, B_, sP, L);       \
        else batch_consume<EXACT, STATS, PAIR, U, false, true, PROBE, PRE>(p, ray, B_, sP, L);              \
    } while (0)
    if (EXACT) {
        Batch<U> a;
        while (!ray.done) {
            VR_ISSUE(a);
            VR_CONSUME(a);
        }
    } else if (!ESS) {
        /* Fast phase.  While every lane of the warp is inside its safe interval (or has nothing left to
         * do), the warp runs `nb` double-batches with no test, no predicate and no vote except the
         * early-termination one.  Lanes that are finished (or never had a fragment) are parked: their
         * transmittance is saved and set to 0, so the unpredicated compositing adds exactly 0 to their
         * colour, and they step in place at the volume centre.  A lane that drops below the early-
         * termination threshold inside the phase keeps compositing real samples until the phase ends:
         * its result only gets closer to the unterminated one, the error bound is unchanged. */
        Batch<U> a, b;
        while (true) {
            const int room = ray.done ? 0x7fffffff : (ray.i >= fast_lo ? fast_hi - ray.i + 1 : 0);
            const int nmin = __reduce_min_sync(0xffffffffu, room);
            if (nmin == 0x7fffffff) break; /* every lane is done */
            const int nb = (nmin / U - 1) / 2;
            if (nb >= 1) {
                if (ray.done && !parked) {
                    parked = true, t_saved = ray.T, ray.T = 0.0f;
                    ray.px = ray.py = ray.pz = 0.5f, ray.dx = ray.dy = ray.dz = 0.0f;
                    ray.ex = ray.ey = ray.ez = 0.5f;
                }
                batch_issue<false, false, STATS, PART, PAIR, U, false, false>(p, ray, a);
#pragma unroll 1
                for (int it = 0; it < nb; it++) {
                    batch_issue<false, false, STATS, PART, PAIR, U, false, false>(p, ray, b);
                    batch_consume<false, STATS, PAIR, U, false, false, PROBE, PRE>(p, ray, a, sP, L);
                    batch_issue<false, false, STATS, PART, PAIR, U, false, false>(p, ray, a);
                    batch_consume<false, STATS, PAIR, U, false, false, PROBE, PRE>(p, ray, b, sP, L);
                    if (__all_sync(0xffffffffu, ray.T < p.ert_eps)) break; /* parked lanes: T = 0 (never true when ERT is off: nb bounds the loop) */
                }
                batch_consume<false, STATS, PAIR, U, false, false, PROBE, PRE>(p, ray, a, sP, L);
                if (!ray.done && ray.T < p.ert_eps) ray.stop = true, ray.done = true;
                continue;
            }
            /* head / tail: one double-batch with the tests the lanes need */
            VR_ISSUE(a);
            VR_ISSUE(b);
            VR_CONSUME(a);
            VR_CONSUME(b);
        }
    } else {
        /* two batches in rotation: the fetches of one are in flight while the other is composited */
        Batch<U> a, b;
        VR_ISSUE(a);
        while (true) {
            VR_ISSUE(b);
            VR_CONSUME(a);
            if (!__any_sync(0xffffffffu, !ray.done)) {
                VR_CONSUME(b);
                break;
            }
            VR_ISSUE(a);
            VR_CONSUME(b);
            if (!__any_sync(0xffffffffu, !ray.done)) {
                VR_CONSUME(a);
                break;
            }
        }
    }
#undef VR_ISSUE
#undef VR_CONSUME
    cr = ray.cr, cg = ray.cg, cb = ray.cb, T = parked ? t_saved : ray.T;
    uint32_t nb = ray.nb, ns = ray.ns, nf = ray.nf;

    if (valid) {
        uchar4 o;
        float4 rawv;
        if (covered) {
            /* volume.frag:50 then the blend: src*srcA + dst*(1-srcA) for colour and alpha */
            const float A = 1.0f - T, iA = 1.0f - A;
            rawv = make_float4(cr, cg, cb, A);
            float o0 = cr * A + VR_CLEAR_RGB * iA, o1 = cg * A + VR_CLEAR_RGB * iA, o2 = cb * A + VR_CLEAR_RGB * iA,
                  o3 = A * A + 1.0f * iA;
            float ov[4] = {o0, o1, o2, o3};
            unsigned char q[4];
#pragma unroll
            for (int c = 0; c < 4; c++) {
                float x = ov[c];
                x = x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x);
                if (!(x == x)) x = 0.0f;
                q[c] = (unsigned char)(x * 255.0f + 0.5f);
            }
            o = make_uchar4(q[0], q[1], q[2], q[3]);
        } else {
            const unsigned char c = (unsigned char)(VR_CLEAR_RGB * 255.0f + 0.5f);
            o = make_uchar4(c, c, c, 255);
            rawv = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (p.out) p.out[pix] = o;
        if (p.raw) p.raw[pix] = rawv;
    }

    if (STATS) {
        const uint32_t ncov = __reduce_add_sync(0xffffffffu, covered ? 1u : 0u);
        nb = __reduce_add_sync(0xffffffffu, nb);
        ns = __reduce_add_sync(0xffffffffu, ns);
        nf = __reduce_add_sync(0xffffffffu, nf);
        if (lane == 0 && ncov) {
            atomicAdd(&p.counters[0], (unsigned long long)ncov);
            atomicAdd(&p.counters[1], (unsigned long long)nb);
            atomicAdd(&p.counters[2], (unsigned long long)ns);
            atomicAdd(&p.counters[3], (unsigned long long)nf);
        }
    }
}

template <bool EXACT, bool ESS, bool STATS, bool PART, int PAIR, int UB>
cudaError_t launch(const MarchParams &p, size_t smem, dim3 grid, int threads, cudaStream_t s)
{
    auto kern = k_march<EXACT, ESS, STATS, PART, PAIR, UB>;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    if (p.carveout >= 0) cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, p.carveout);
    kern<<<grid, threads, smem, s>>>(p);
    return cudaGetLastError();
}

template <bool EXACT, bool ESS, int PAIR, int UB>
cudaError_t launch2(const MarchParams &p, bool stats, bool part, size_t smem, dim3 grid, int threads, cudaStream_t s)
{
    if (stats)
        return part ? launch<EXACT, ESS, true, true, PAIR, UB>(p, smem, grid, threads, s)
                    : launch<EXACT, ESS, true, false, PAIR, UB>(p, smem, grid, threads, s);
    return part ? launch<EXACT, ESS, false, true, PAIR, UB>(p, smem, grid, threads, s)
                : launch<EXACT, ESS, false, false, PAIR, UB>(p, smem, grid, threads, s);
}

template <bool ESS, int PAIR>
cudaError_t launch_hw(const MarchParams &p, bool stats, bool part, int ub, size_t smem, dim3 grid, int threads,
                      cudaStream_t s)
{
    switch (ub) {
        case 4: return launch2<false, ESS, PAIR, 4>(p, stats, part, smem, grid, threads, s);
        default: return launch2<false, ESS, PAIR, 2>(p, stats, part, smem, grid, threads, s);
    }
}

}  // namespace

size_t vr_march_smem_bytes(const MarchParams &p, bool exact)
{
    /* EXACT: nlut + 3 float4 texels; HW: nlut + 2 packed fp16 entries.  The distance map is read from global memory. */
    return exact ? sizeof(float4) * (size_t)(p.nlut + 3) : sizeof(uint4) * (size_t)(p.nlut + 2);
}

cudaError_t vr_launch_march_probe(const MarchParams &p, bool pair, int ub, int threads, cudaStream_t s)
{
    const uint64_t ctas = (uint64_t)p.sp_count + (uint64_t)p.nclr_x * p.nclr_y;
    if (ctas == 0) return cudaSuccess;
    if (ctas > 0x7fffffffull || (pair && p.pair_split != 0x10000)) return cudaErrorInvalidConfiguration;
    const size_t smem = vr_march_smem_bytes(p, false);
    /* same steps per batch as the product launch */
    if (ub == 4) {
        if (pair) k_march<false, false, false, false, 1, 4, true><<<(unsigned)ctas, threads, smem, s>>>(p);
        else k_march<false, false, false, false, 0, 4, true><<<(unsigned)ctas, threads, smem, s>>>(p);
    } else {
        if (pair) k_march<false, false, false, false, 1, 2, true><<<(unsigned)ctas, threads, smem, s>>>(p);
        else k_march<false, false, false, false, 0, 2, true><<<(unsigned)ctas, threads, smem, s>>>(p);
    }
    return cudaGetLastError();
}

/* pre-integrated classification: hardware sampler, no skipping, no statistics, no sub-volume; 2 steps per batch */
cudaError_t vr_launch_march_preint(const MarchParams &p, bool pair, int threads, cudaStream_t s)
{
    const uint64_t ctas = (uint64_t)p.sp_count + (uint64_t)p.nclr_x * p.nclr_y;
    if (ctas == 0) return cudaSuccess;
    if (ctas > 0x7fffffffull) return cudaErrorInvalidConfiguration;
    const size_t smem = vr_march_smem_bytes(p, false);
    if (pair && p.pair_split != 0x10000) k_march<false, false, false, false, 2, 2, false, true><<<(unsigned)ctas, threads, smem, s>>>(p);
    else if (pair) k_march<false, false, false, false, 1, 2, false, true><<<(unsigned)ctas, threads, smem, s>>>(p);
    else k_march<false, false, false, false, 0, 2, false, true><<<(unsigned)ctas, threads, smem, s>>>(p);
    return cudaGetLastError();
}

cudaError_t vr_launch_march(const MarchParams &p, bool exact, bool ess, bool stats, bool part, bool pair, int ub,
                            int threads, cudaStream_t s)
{
    const uint64_t ctas = (uint64_t)p.sp_count + (uint64_t)p.nclr_x * p.nclr_y;
    if (ctas == 0) return cudaSuccess;
    if (ctas > 0x7fffffffull) return cudaErrorInvalidConfiguration;
    const size_t smem = vr_march_smem_bytes(p, exact);
    const dim3 grid((unsigned)ctas);
    if (exact)
        return ess ? launch2<true, true, 0, 1>(p, stats, part, smem, grid, threads, s)
                   : launch2<true, false, 0, 1>(p, stats, part, smem, grid, threads, s);
    if (pair && p.pair_split != 0x10000)
        return ess ? launch_hw<true, 2>(p, stats, part, ub, smem, grid, threads, s)
                   : launch_hw<false, 2>(p, stats, part, ub, smem, grid, threads, s);
    if (pair)
        return ess ? launch_hw<true, 1>(p, stats, part, ub, smem, grid, threads, s)
                   : launch_hw<false, 1>(p, stats, part, ub, smem, grid, threads, s);
    return ess ? launch_hw<true, 0>(p, stats, part, ub, smem, grid, threads, s)
               : launch_hw<false, 0>(p, stats, part, ub, smem, grid, threads, s);
}
