/*
 * raster.c -- an INDEPENDENT check of stage S2 (coverage + ray entry) of the oracle.
 *
 * TEST INFRASTRUCTURE ONLY (see vr_oracle.h).
 *
 * The oracle and the CUDA kernel decide coverage and the entry point analytically (vro_pixel_entry: unproject the
 * pixel centre, slab test).  The reference does something else: it DRAWS the cube -- 24 vertices / 36 indices
 * (src/rendering/offscreen_pass.cpp:55-90) through res/shaders/volume.vert:19-24 with the fixed-function state at
 * offscreen_pass.cpp:661-712 (triangle list, back-face cull, counter-clockwise front face, depth test LESS, no
 * depth clamp, full-target viewport :190-205).  This file restates THAT: a software rasteriser of the twelve
 * triangles under Vulkan's rules --
 *   - clip volume 0 <= z_c <= w_c (VK depth range 0..1), polygon clipping against near and far;
 *   - viewport transform x_f = (x_ndc + 1) W/2, y_f = (y_ndc + 1) H/2, framebuffer y down;
 *   - facing by the sign of a = -1/2 sum (x_i y_{i+1} - x_{i+1} y_i) in framebuffer coordinates, positive =
 *     front for VK_FRONT_FACE_COUNTER_CLOCKWISE; back faces culled;
 *   - sample point at the pixel centre, top-left fill rule;
 *   - depth test LESS on the screen-linear z_ndc;
 *   - perspective-correct interpolation of the vertex attribute in_tex_coords (= position + 0.5).
 * tests/test_oracle.py compares its covered set and interpolated texture coordinates with vro_pixel_entry's.
 * Arithmetic: vertex transform in fp32 like the shader, everything after it in double.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "vr_oracle.h"
#include "vro_common.h"

typedef struct {
    double c[4]; /* clip-space position */
    double a[3]; /* attribute: in_tex_coords */
} rv;

/* The cube as the reference's vertex table lays it out: six faces, each a quad o, o+e1, o+e1+e2, o+e2 drawn as the
 * triangles (0,1,2), (0,2,3); e1 x e2 is the outward normal on every face (offscreen_pass.cpp:55-90). */
static const struct {
    float o[3], e1[3], e2[3];
} k_faces[6] = {
    {{-0.5f, -0.5f, 0.5f}, {1, 0, 0}, {0, 1, 0}},   /* top    +z */
    {{-0.5f, -0.5f, -0.5f}, {1, 0, 0}, {0, 0, 1}},  /* front  -y */
    {{0.5f, -0.5f, -0.5f}, {0, 1, 0}, {0, 0, 1}},   /* right  +x */
    {{0.5f, 0.5f, -0.5f}, {-1, 0, 0}, {0, 0, 1}},   /* back   +y */
    {{-0.5f, 0.5f, -0.5f}, {0, -1, 0}, {0, 0, 1}},  /* left   -x */
    {{-0.5f, 0.5f, -0.5f}, {1, 0, 0}, {0, -1, 0}},  /* bottom -z */
};

static void mat4_mul_f(const float *a, const float *b, float *out)
{
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++) {
            float s = 0.0f;
            for (int k = 0; k < 4; k++) s += a[k * 4 + r] * b[c * 4 + k];
            out[c * 4 + r] = s;
        }
}

/* Sutherland-Hodgman against one plane: keep dist >= 0 */
static int clip_plane(const rv *in, int n, rv *out, int which)
{
    int m = 0;
    for (int i = 0; i < n; i++) {
        const rv *p = &in[i], *q = &in[(i + 1) % n];
        const double dp = which == 0 ? p->c[2] : p->c[3] - p->c[2];
        const double dq = which == 0 ? q->c[2] : q->c[3] - q->c[2];
        if (dp >= 0.0) out[m++] = *p;
        if ((dp >= 0.0) != (dq >= 0.0)) {
            const double t = dp / (dp - dq);
            rv r;
            for (int k = 0; k < 4; k++) r.c[k] = p->c[k] + t * (q->c[k] - p->c[k]);
            for (int k = 0; k < 3; k++) r.a[k] = p->a[k] + t * (q->a[k] - p->a[k]);
            out[m++] = r;
        }
    }
    return m;
}

static int top_left(double ax, double ay, double bx, double by, double cx, double cy)
{
    /* edge a->b of a triangle whose third vertex is c, framebuffer y down */
    if (ay == by) return cy > ay; /* horizontal: a top edge has the interior below it */
    const double t = (cy - ay) / (by - ay);
    return ax + t * (bx - ax) < cx; /* a left edge has the interior to its right */
}

int vro_raster_cube(const vro_frame *f, uint8_t *mask, float *tex, float *depth_out)
{
    const uint32_t W = f->width, H = f->height;
    const size_t npix = (size_t)W * H;
    memset(mask, 0, npix);
    double *depth = (double *)malloc(npix * sizeof(double));
    if (!depth) return -1;
    for (size_t i = 0; i < npix; i++) depth[i] = 1.0e300;
    float pv[16];
    mat4_mul_f(f->proj, f->view, pv); /* volume.vert:23, fp32 */
    int covered = 0;
    for (int face = 0; face < 6; face++) {
        rv quad[4];
        for (int v = 0; v < 4; v++) {
            const float s1 = (v == 1 || v == 2) ? 1.0f : 0.0f, s2 = (v >= 2) ? 1.0f : 0.0f;
            float pos[3];
            for (int k = 0; k < 3; k++) pos[k] = k_faces[face].o[k] + s1 * k_faces[face].e1[k] + s2 * k_faces[face].e2[k];
            for (int r = 0; r < 4; r++)
                quad[v].c[r] = (double)(((pv[r] * pos[0] + pv[4 + r] * pos[1]) + pv[8 + r] * pos[2]) + pv[12 + r]);
            for (int k = 0; k < 3; k++) quad[v].a[k] = (double)(pos[k] + 0.5f);
        }
        static const int tri_idx[2][3] = {{0, 1, 2}, {0, 2, 3}};
        for (int t = 0; t < 2; t++) {
            rv poly[8], tmp[8];
            for (int k = 0; k < 3; k++) poly[k] = quad[tri_idx[t][k]];
            int n = clip_plane(poly, 3, tmp, 0);
            n = clip_plane(tmp, n, poly, 1);
            if (n < 3) continue;
            /* framebuffer coordinates */
            double fx[8], fy[8], fz[8], iw[8];
            for (int k = 0; k < n; k++) {
                iw[k] = 1.0 / poly[k].c[3];
                fx[k] = (poly[k].c[0] * iw[k] + 1.0) * 0.5 * (double)W;
                fy[k] = (poly[k].c[1] * iw[k] + 1.0) * 0.5 * (double)H;
                fz[k] = poly[k].c[2] * iw[k];
            }
            double area = 0.0;
            for (int k = 0; k < n; k++) area += fx[k] * fy[(k + 1) % n] - fx[(k + 1) % n] * fy[k];
            area *= -0.5;
            if (!(area > 0.0)) continue; /* back-facing (or degenerate): culled */
            for (int k = 1; k + 1 < n; k++) {
                const int id[3] = {0, k, k + 1};
                const double x0 = fx[id[0]], y0 = fy[id[0]], x1 = fx[id[1]], y1 = fy[id[1]], x2 = fx[id[2]], y2 = fy[id[2]];
                const double den = (x1 - x0) * (y2 - y0) - (x2 - x0) * (y1 - y0);
                if (den == 0.0) continue;
                double xmin = fmin(x0, fmin(x1, x2)), xmax = fmax(x0, fmax(x1, x2));
                double ymin = fmin(y0, fmin(y1, y2)), ymax = fmax(y0, fmax(y1, y2));
                long px0 = (long)floor(xmin - 0.5), px1 = (long)ceil(xmax - 0.5), py0 = (long)floor(ymin - 0.5), py1 = (long)ceil(ymax - 0.5);
                if (px0 < 0) px0 = 0;
                if (py0 < 0) py0 = 0;
                if (px1 > (long)W - 1) px1 = (long)W - 1;
                if (py1 > (long)H - 1) py1 = (long)H - 1;
                const int tl0 = top_left(x1, y1, x2, y2, x0, y0), tl1 = top_left(x2, y2, x0, y0, x1, y1), tl2 = top_left(x0, y0, x1, y1, x2, y2);
                for (long py = py0; py <= py1; py++)
                    for (long px = px0; px <= px1; px++) {
                        const double sx = (double)px + 0.5, sy = (double)py + 0.5;
                        /* barycentrics; sign-normalised by den so that inside is >= 0 */
                        const double b0 = ((x1 - sx) * (y2 - sy) - (x2 - sx) * (y1 - sy)) / den;
                        const double b1 = ((x2 - sx) * (y0 - sy) - (x0 - sx) * (y2 - sy)) / den;
                        const double b2 = 1.0 - b0 - b1;
                        if (b0 < 0.0 || b1 < 0.0 || b2 < 0.0) continue;
                        if ((b0 == 0.0 && !tl0) || (b1 == 0.0 && !tl1) || (b2 == 0.0 && !tl2)) continue;
                        const double z = b0 * fz[id[0]] + b1 * fz[id[1]] + b2 * fz[id[2]];
                        const size_t pix = (size_t)py * W + (size_t)px;
                        if (!(z < depth[pix])) continue; /* VK_COMPARE_OP_LESS */
                        depth[pix] = z;
                        if (!mask[pix]) covered++;
                        mask[pix] = (uint8_t)(1 + face);
                        const double w0 = b0 * iw[id[0]], w1 = b1 * iw[id[1]], w2 = b2 * iw[id[2]], ws = w0 + w1 + w2;
                        for (int c = 0; c < 3; c++)
                            tex[pix * 3 + c] = (float)((w0 * poly[id[0]].a[c] + w1 * poly[id[1]].a[c] + w2 * poly[id[2]].a[c]) / ws);
                    }
            }
        }
    }
    if (depth_out)
        for (size_t i = 0; i < npix; i++) depth_out[i] = mask[i] ? (float)depth[i] : 1.0f;
    free(depth);
    return covered;
}

/* The analytic decision the oracle and the kernel share, pixel by pixel (for the comparison). */
int vro_entry_map(const vro_frame *f, uint8_t *mask, float *tex)
{
    vro_rayctx rc;
    const uint32_t W = f->width, H = f->height;
    memset(mask, 0, (size_t)W * H);
    if (vro_rayctx_init(&rc, f)) return -1;
    int covered = 0;
    for (uint32_t py = 0; py < H; py++)
        for (uint32_t px = 0; px < W; px++) {
            float frag[3], t[3];
            if (vro_pixel_entry(&rc, px, py, frag, t)) {
                const size_t pix = (size_t)py * W + px;
                mask[pix] = 1;
                covered++;
                for (int c = 0; c < 3; c++) tex[pix * 3 + c] = t[c];
            }
        }
    return covered;
}
