/*
 * preint.c -- CPU restatement of the PRE-INTEGRATED transfer-function mode (optional quality mode).
 *
 * TEST INFRASTRUCTURE ONLY (see vr_oracle.h).  Port build only: the mode has NO counterpart in the reference -- its
 * shader classifies each sample on its own (res/shaders/volume.frag:41-45), which misses thin features of the transfer
 * function that fall between two samples.  Pre-integration (Engel, Kraus, Ertl 2001) classifies the SEGMENT between two
 * consecutive samples instead: a 2-D table indexed by (front scalar, back scalar) holds the colour and opacity of a
 * segment along which the scalar varies linearly.  Everything else is the reference's loop: same positions, same
 * bounds and slice tests (volume.frag:33-40, :47), same blend and store.
 *
 * Table, n x n entries for an n-texel LUT.  Entry (back j, front i), texel centres i + 1/2 and j + 1/2 in unnormalised
 * LUT coordinates: M sub-samples at s_m = (i + 1/2) + (m + 1/2) / M * (j - i), each classified by the reference's 1-D
 * sampler (vro_sample_lut: LINEAR, CLAMP_TO_EDGE, sRGB-decoded texels) and composited front to back with the
 * opacity correction a' = 1 - (1 - a)^(1/M) -- so that a segment of constant scalar gives exactly the per-sample
 * result of the reference: (rgb * a, a).  Entry = (premultiplied rgb, opacity).
 *
 * March: c_k = t_k * n is the table coordinate of sample k (t_k of volume.frag:42; NaN -> 0).  Consecutive in-slice
 * samples k-1, k contribute the bilinearly filtered entry at (x = c_{k-1}, y = c_k), unnormalised coordinates, clamp to
 * edge; a sample outside the slices breaks the chain.  The first sample of a chain contributes nothing by itself.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "vr_oracle.h"
#include "vro_common.h"

int vro_preint_table(const uint32_t *lut, uint32_t n, uint32_t substeps, float *table /* n * n * 4 */)
{
    if (!lut || !table || n == 0 || substeps == 0) return -1;
    float *lin = (float *)malloc(sizeof(float) * 4 * n);
    if (!lin) return -3;
    vro_decode_lut(lut, n, lin);
    const double inv_m = 1.0 / (double)substeps;
#pragma omp parallel for schedule(static)
    for (long j = 0; j < (long)n; j++)
        for (uint32_t i = 0; i < n; i++) {
            double c[3] = {0.0, 0.0, 0.0}, T = 1.0;
            for (uint32_t m = 0; m < substeps; m++) {
                const double s = ((double)i + 0.5) + ((double)m + 0.5) * inv_m * ((double)j - (double)i);
                float smp[4];
                vro_sample_lut(lin, n, (float)(s / (double)n), smp);
                const double a = 1.0 - pow(1.0 - (double)smp[3], inv_m);
                for (int k = 0; k < 3; k++) c[k] += T * a * (double)smp[k];
                T *= 1.0 - a;
            }
            float *e = table + 4 * ((size_t)j * n + i);
            e[0] = (float)c[0], e[1] = (float)c[1], e[2] = (float)c[2], e[3] = (float)(1.0 - T);
        }
    free(lin);
    return 0;
}

/* bilinear, unnormalised coordinates, clamp to edge (the filter of a CUDA / Vulkan LINEAR sampler, fp32 weights) */
static void table_sample(const float *tab, uint32_t n, float x, float y, float out[4])
{
    const float u = x - 0.5f, v = y - 0.5f;
    const float fu = floorf(u), fv = floorf(v);
    const float a = u - fu, b = v - fv;
    const long i0 = (long)fu, j0 = (long)fv;
    long ix[2] = {i0, i0 + 1}, jy[2] = {j0, j0 + 1};
    for (int k = 0; k < 2; k++) {
        ix[k] = ix[k] < 0 ? 0 : (ix[k] > (long)n - 1 ? (long)n - 1 : ix[k]);
        jy[k] = jy[k] < 0 ? 0 : (jy[k] > (long)n - 1 ? (long)n - 1 : jy[k]);
    }
    for (int c = 0; c < 4; c++) {
        const float r0 = vro_lerp(tab[4 * ((size_t)jy[0] * n + ix[0]) + c], tab[4 * ((size_t)jy[0] * n + ix[1]) + c], a);
        const float r1 = vro_lerp(tab[4 * ((size_t)jy[1] * n + ix[0]) + c], tab[4 * ((size_t)jy[1] * n + ix[1]) + c], a);
        out[c] = vro_lerp(r0, r1, b);
    }
}

static void march_pixel_preint(const vro_frame *f, const vro_volume *vol, const float *tab, uint32_t n, const float frag[3],
                               const float tex[3], float out[4])
{
    float color[4] = {0.0f, 0.0f, 0.0f, 1.0f};
    /* ray set-up and stepping exactly as vr_oracle.c march_pixel (volume.frag:23-24, :29-37, :47) */
    float v[3] = {frag[0] - f->camera_position[0], frag[1] - f->camera_position[1], frag[2] - f->camera_position[2]};
    float d2 = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
    float inv = 1.0f / sqrtf(d2);
    float dir[3] = {v[0] * inv, v[1] * inv, v[2] * inv};
    float pos[3] = {tex[0], tex[1], tex[2]};
    int steps = (int)(f->ray_dist / f->step_size);
    float range = f->max_density - f->min_density;
    int have_prev = 0;
    float prev = 0.0f;
    for (int i = 0; i < steps; i++) {
        if (pos[0] > 1.0f || pos[1] > 1.0f || pos[2] > 1.0f || pos[0] < 0.0f || pos[1] < 0.0f || pos[2] < 0.0f) break;
        if (pos[0] < f->max_slice[0] && pos[1] < f->max_slice[1] && pos[2] < f->max_slice[2] && pos[0] > f->min_slice[0] &&
            pos[1] > f->min_slice[1] && pos[2] > f->min_slice[2]) {
            float density = vro_sample_volume(vol, pos[0], pos[1], pos[2]);
            float t = (density - f->min_density) / range;
            float cur = t * (float)n;
            cur = (cur > 0.0f) ? cur : 0.0f; /* also NaN -> 0 */
            cur = (cur < (float)n) ? cur : (float)n;
            if (have_prev) {
                float s[4];
                table_sample(tab, n, prev, cur, s);
                color[0] += color[3] * s[0];
                color[1] += color[3] * s[1];
                color[2] += color[3] * s[2];
                color[3] *= (1.0f - s[3]);
            }
            prev = cur, have_prev = 1;
        } else {
            have_prev = 0;
        }
        pos[0] += dir[0] * f->step_size;
        pos[1] += dir[1] * f->step_size;
        pos[2] += dir[2] * f->step_size;
    }
    out[0] = color[0], out[1] = color[1], out[2] = color[2], out[3] = 1.0f - color[3];
}

int vro_render_preint(const vro_frame *f, const vro_volume *vol, const uint32_t *lut, uint32_t nlut, uint32_t substeps, uint8_t *rgba8,
                      float *raw, int nthreads)
{
    if (!f || !vol || !lut || nlut == 0 || substeps == 0 || !rgba8 || f->width == 0 || f->height == 0) return -1;
    vro_rayctx rc;
    if (vro_rayctx_init(&rc, f)) return -2;
    float *tab = (float *)malloc(sizeof(float) * 4 * (size_t)nlut * nlut);
    if (!tab) return -3;
    if (vro_preint_table(lut, nlut, substeps, tab)) {
        free(tab);
        return -3;
    }
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_num_procs();
#else
    nthreads = 1;
#endif
    const long W = f->width, H = f->height;
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads)
    for (long py = 0; py < H; py++)
        for (long px = 0; px < W; px++) {
            size_t p = (size_t)py * W + px;
            float frag[3], tex[3];
            if (!vro_pixel_entry(&rc, (uint32_t)px, (uint32_t)py, frag, tex)) {
                vro_clear_store(rgba8 + 4 * p);
                if (raw) raw[4 * p] = raw[4 * p + 1] = raw[4 * p + 2] = raw[4 * p + 3] = 0.0f;
                continue;
            }
            float c[4];
            march_pixel_preint(f, vol, tab, nlut, frag, tex, c);
            vro_blend_store(c, rgba8 + 4 * p);
            if (raw) memcpy(raw + 4 * p, c, sizeof(c));
        }
    free(tab);
    return 0;
}
