/*
 * spacing.c -- CPU restatement of rendering a volume whose voxels are not cubes (NRRD `spacings` / `space directions`).
 *
 * TEST INFRASTRUCTURE ONLY (see vr_oracle.h).  Port build only: the reference ignores the spacings of a file
 * (src/data/nrrd_file_parser.cpp:32-33) and always draws the unit cube; honouring them is SURVEY.md 8(f).1's extension and
 * is off unless the caller asks (vr_set_volume_spacing).  With extents e = n_i s_i / max_j(n_j s_j) <= 1 the volume
 * fills the box [-e/2, e/2]; this is the reference's loop on that box, written in the unit cube's coordinates so that
 * e = (1, 1, 1) is the reference bit for bit:
 *   - the cube is drawn with view * diag(e) and seen from camera_position / e (fp32, one rounding each): coverage, entry
 *     point and texture coordinates (volume.vert, rasteriser) follow as before;
 *   - ray_dir (volume.frag:23) is the normalised PHYSICAL direction, (frag - camera) * e, and a step of step_size along
 *     it advances the texture position by dir * step_size / e (volume.frag:47), the same physical distance whichever way
 *     the ray points.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "vr_oracle.h"
#include "vro_common.h"

static void march_pixel_ext(const vro_frame *f, const float ext[3], const float inv_ext[3], const vro_volume *vol, const float *lut_lin,
                            uint32_t nlut, const float frag[3], const float tex[3], float out[4], uint32_t *n_bounds, uint32_t *n_slice)
{
    float color[4] = {0.0f, 0.0f, 0.0f, 1.0f};
    float v[3] = {(frag[0] - f->camera_position[0]) * ext[0], (frag[1] - f->camera_position[1]) * ext[1],
                  (frag[2] - f->camera_position[2]) * ext[2]};
    float d2 = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
    float inv = 1.0f / sqrtf(d2);
    float dir[3] = {v[0] * inv, v[1] * inv, v[2] * inv};
    float pos[3] = {tex[0], tex[1], tex[2]};
    int steps = (int)(f->ray_dist / f->step_size);
    float range = f->max_density - f->min_density;
    uint32_t nb = 0, ns = 0;
    for (int i = 0; i < steps; i++) {
        if (pos[0] > 1.0f || pos[1] > 1.0f || pos[2] > 1.0f || pos[0] < 0.0f || pos[1] < 0.0f || pos[2] < 0.0f) break;
        nb++;
        if (pos[0] < f->max_slice[0] && pos[1] < f->max_slice[1] && pos[2] < f->max_slice[2] && pos[0] > f->min_slice[0] &&
            pos[1] > f->min_slice[1] && pos[2] > f->min_slice[2]) {
            float density = vro_sample_volume(vol, pos[0], pos[1], pos[2]);
            float t = (density - f->min_density) / range;
            float s[4];
            vro_sample_lut(lut_lin, nlut, t, s);
            color[0] += color[3] * (s[3] * s[0]);
            color[1] += color[3] * (s[3] * s[1]);
            color[2] += color[3] * (s[3] * s[2]);
            color[3] *= (1.0f - s[3]);
            ns++;
        }
        pos[0] += (dir[0] * f->step_size) * inv_ext[0];
        pos[1] += (dir[1] * f->step_size) * inv_ext[1];
        pos[2] += (dir[2] * f->step_size) * inv_ext[2];
    }
    out[0] = color[0], out[1] = color[1], out[2] = color[2], out[3] = 1.0f - color[3];
    *n_bounds = nb, *n_slice = ns;
}

int vro_render_spacing(const vro_frame *frame, const float extent[3], const vro_volume *vol, const uint32_t *lut, uint32_t nlut,
                       uint8_t *rgba8, float *raw, vro_stats *stats, int nthreads)
{
    if (!frame || !extent || !vol || !lut || nlut == 0 || !rgba8 || frame->width == 0 || frame->height == 0) return -1;
    for (int i = 0; i < 3; i++)
        if (!(extent[i] > 0.0f) || extent[i] > 1.0f) return -1;
    /* the frame as the unit cube sees it */
    vro_frame g = *frame;
    float inv_ext[3];
    for (int c = 0; c < 3; c++) {
        for (int r = 0; r < 4; r++) g.view[c * 4 + r] = frame->view[c * 4 + r] * extent[c];
        g.camera_position[c] = frame->camera_position[c] / extent[c];
        inv_ext[c] = 1.0f / extent[c];
    }
    vro_rayctx rc;
    if (vro_rayctx_init(&rc, &g)) return -2;
    float *lut_lin = (float *)malloc(sizeof(float) * 4 * nlut);
    if (!lut_lin) return -3;
    vro_decode_lut(lut, nlut, lut_lin);
    uint64_t covered = 0, n_ref = 0, n_in = 0;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_num_procs();
#else
    nthreads = 1;
#endif
    const long W = g.width, H = g.height;
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads) reduction(+ : covered, n_ref, n_in)
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
            uint32_t nb, ns;
            march_pixel_ext(&g, extent, inv_ext, vol, lut_lin, nlut, frag, tex, c, &nb, &ns);
            vro_blend_store(c, rgba8 + 4 * p);
            if (raw) memcpy(raw + 4 * p, c, sizeof(c));
            covered++, n_ref += nb, n_in += ns;
        }
    free(lut_lin);
    if (stats) stats->covered_pixels = covered, stats->n_ref = n_ref, stats->n_inslice = n_in;
    return 0;
}
