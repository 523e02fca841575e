/*
 * vr_oracle.c -- plain-C restatement ("port") of the reference's ray-march path.
 *
 * TEST INFRASTRUCTURE ONLY (see vr_oracle.h).  Never linked into the product.
 * Parity status: unpinned by reference fixtures (the reference has none); pinned
 * against oracle/_ref (reference sources compiled here) by tests/test_oracle_ref.py
 * and against tests/golden/ vectors generated from _ref.
 *
 * Each function cites the reference lines it follows (paths relative to /root/reference).
 * Build: gcc -O2 -ffp-contract=off -fopenmp (oracle/Makefile).  fp32 throughout, one
 * rounding per operation, operation order as glm 0.9.9.9 evaluates the GLSL.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "vr_oracle.h"
#include "vro_common.h"

const char *vro_kind(void) { return "port"; }

/* ---------------------------------------------------------------- S3 ---- */

/* res/shaders/volume.frag:21-52.  color[3] is TRANSMITTANCE until line 50. */
static void march_pixel(const vro_frame *f, const vro_volume *vol, const float *lut_lin, uint32_t nlut,
                        const float frag[3], const float tex[3], float out[4], uint32_t *n_bounds,
                        uint32_t *n_slice)
{
    float color[4] = {0.0f, 0.0f, 0.0f, 1.0f}; /* :22 */
    /* :23 ray_dir = normalize(in_frag_position - camera_position);
     * glm: v * inversesqrt(dot(v,v)), dot = (x*x + y*y) + z*z, inversesqrt = 1/sqrt */
    float v[3] = {frag[0] - f->camera_position[0], frag[1] - f->camera_position[1],
                  frag[2] - f->camera_position[2]};
    float d2 = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
    float inv = 1.0f / sqrtf(d2);
    float dir[3] = {v[0] * inv, v[1] * inv, v[2] * inv};
    float pos[3] = {tex[0], tex[1], tex[2]}; /* :24 */

    int steps = (int)(f->ray_dist / f->step_size); /* :29-31 */
    float range = f->max_density - f->min_density;
    uint32_t nb = 0, ns = 0;
    for (int i = 0; i < steps; i++) { /* :33 */
        /* :34-37 strict comparisons against (0,0,0)/(1,1,1) */
        if (pos[0] > 1.0f || pos[1] > 1.0f || pos[2] > 1.0f || pos[0] < 0.0f || pos[1] < 0.0f || pos[2] < 0.0f)
            break;
        nb++;
        /* :39-40 strict on both sides */
        if (pos[0] < f->max_slice[0] && pos[1] < f->max_slice[1] && pos[2] < f->max_slice[2] &&
            pos[0] > f->min_slice[0] && pos[1] > f->min_slice[1] && pos[2] > f->min_slice[2]) {
            float density = vro_sample_volume(vol, pos[0], pos[1], pos[2]); /* :41 */
            float t = (density - f->min_density) / range;                  /* :42 */
            float s[4];
            vro_sample_lut(lut_lin, nlut, t, s); /* :43 */
            /* :44 color.rgb += color.a * (sample.a * sample.rgb) */
            color[0] += color[3] * (s[3] * s[0]);
            color[1] += color[3] * (s[3] * s[1]);
            color[2] += color[3] * (s[3] * s[2]);
            color[3] *= (1.0f - s[3]); /* :45 */
            ns++;
        }
        /* :47 ray_pos += ray_dir * step_size (accumulated) */
        pos[0] += dir[0] * f->step_size;
        pos[1] += dir[1] * f->step_size;
        pos[2] += dir[2] * f->step_size;
    }
    out[0] = color[0];
    out[1] = color[1];
    out[2] = color[2];
    out[3] = 1.0f - color[3]; /* :50 */
    *n_bounds = nb;
    *n_slice = ns;
}

int vro_render(const vro_frame *f, const vro_volume *vol, const uint32_t *lut, uint32_t nlut, uint8_t *rgba8,
               float *raw, uint16_t *iters, vro_stats *stats, int nthreads)
{
    return vro_render_rows(f, vol, lut, nlut, rgba8, raw, iters, stats, nthreads, 0, 1);
}

int vro_render_rows(const vro_frame *f, const vro_volume *vol, const uint32_t *lut, uint32_t nlut, uint8_t *rgba8,
                    float *raw, uint16_t *iters, vro_stats *stats, int nthreads, uint32_t row_start,
                    uint32_t row_stride)
{
    if (!f || !vol || !lut || nlut == 0 || !rgba8 || f->width == 0 || f->height == 0 || row_stride == 0) return -1;
    vro_rayctx rc;
    if (vro_rayctx_init(&rc, f)) return -2;
    float *lut_lin = (float *)malloc(sizeof(float) * 4 * nlut);
    if (!lut_lin) return -3;
    vro_decode_lut(lut, nlut, lut_lin);
    uint64_t covered = 0, n_ref = 0, n_in = 0;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_num_procs();
#else
    nthreads = 1;
#endif
    const long W = f->width, H = f->height;
    const long nrows = ((long)H - (long)row_start + (long)row_stride - 1) / (long)row_stride;
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads) reduction(+ : covered, n_ref, n_in)
    for (long row = 0; row < nrows; row++) {
        const long py = (long)row_start + row * (long)row_stride;
        for (long px = 0; px < W; px++) {
            size_t p = (size_t)py * W + px;
            float frag[3], tex[3];
            if (!vro_pixel_entry(&rc, (uint32_t)px, (uint32_t)py, frag, tex)) {
                vro_clear_store(rgba8 + 4 * p);
                if (raw) raw[4 * p] = raw[4 * p + 1] = raw[4 * p + 2] = raw[4 * p + 3] = 0.0f;
                if (iters) iters[p] = 0;
                continue;
            }
            float c[4];
            uint32_t nb, ns;
            march_pixel(f, vol, lut_lin, nlut, frag, tex, c, &nb, &ns);
            vro_blend_store(c, rgba8 + 4 * p);
            if (raw) memcpy(raw + 4 * p, c, sizeof(c));
            if (iters) iters[p] = (uint16_t)nb;
            covered++;
            n_ref += nb;
            n_in += ns;
        }
    }
    free(lut_lin);
    if (stats) {
        stats->covered_pixels = covered;
        stats->n_ref = n_ref;
        stats->n_inslice = n_in;
    }
    return 0;
}

/* ---------------------------------------------------------------- S1 ---- */

struct vro_camera {
    float center[3];
    float q[4]; /* w, x, y, z */
    float radius;
};

#define VRO_DEG2RAD 0.01745329251994329576923690768489f /* glm::radians */

/* glm::angleAxis (glm/ext/quaternion_trigonometric.inl) */
static void quat_angle_axis(float angle, const float axis[3], float q[4])
{
    float s = sinf(angle * 0.5f);
    q[0] = cosf(angle * 0.5f);
    q[1] = axis[0] * s;
    q[2] = axis[1] * s;
    q[3] = axis[2] * s;
}

/* glm quat * quat (glm/detail/type_quat.inl operator*=) */
static void quat_mul(const float p[4], const float q[4], float r[4])
{
    r[0] = p[0] * q[0] - p[1] * q[1] - p[2] * q[2] - p[3] * q[3];
    r[1] = p[0] * q[1] + p[1] * q[0] + p[2] * q[3] - p[3] * q[2];
    r[2] = p[0] * q[2] + p[2] * q[0] + p[3] * q[1] - p[1] * q[3];
    r[3] = p[0] * q[3] + p[3] * q[0] + p[1] * q[2] - p[2] * q[1];
}

static void cross3(const float a[3], const float b[3], float r[3])
{
    r[0] = a[1] * b[2] - b[1] * a[2];
    r[1] = a[2] * b[0] - b[2] * a[0];
    r[2] = a[0] * b[1] - b[0] * a[1];
}

/* glm quat * vec3: v + ((uv * q.w) + uuv) * 2 */
static void quat_rotate(const float q[4], const float v[3], float r[3])
{
    float uv[3], uuv[3];
    cross3(q + 1, v, uv);
    cross3(q + 1, uv, uuv);
    for (int i = 0; i < 3; i++) r[i] = v[i] + ((uv[i] * q[0]) + uuv[i]) * 2.0f;
}

vro_camera *vro_camera_new(void)
{
    /* src/scene/camera.cpp:7-13 */
    vro_camera *c = (vro_camera *)calloc(1, sizeof(*c));
    const float z[3] = {0.0f, 0.0f, 1.0f};
    quat_angle_axis(VRO_DEG2RAD * 180.0f, z, c->q);
    c->radius = 3.0f;
    return c;
}

void vro_camera_free(vro_camera *c) { free(c); }

void vro_camera_rotate(vro_camera *c, float dx, float dy)
{
    /* src/scene/camera.cpp:15-29 */
    float ax = dx * 0.25f, ay = dy * 0.25f;
    const float z[3] = {0.0f, 0.0f, 1.0f}, x[3] = {1.0f, 0.0f, 0.0f};
    float yaw[4], t[4], right[3], pitch[4];
    quat_angle_axis(VRO_DEG2RAD * (-ax), z, yaw);
    quat_mul(yaw, c->q, t);
    memcpy(c->q, t, sizeof(t));
    quat_rotate(c->q, x, right);
    quat_angle_axis(VRO_DEG2RAD * ay, right, pitch);
    quat_mul(pitch, c->q, t);
    memcpy(c->q, t, sizeof(t));
}

void vro_camera_zoom(vro_camera *c, float delta)
{
    /* src/scene/camera.cpp:31-34 */
    float r = c->radius - delta;
    c->radius = r < 0.1f ? 0.1f : (r > 10.0f ? 10.0f : r);
}

void vro_camera_get_position(const vro_camera *c, float out[3])
{
    /* src/scene/camera.cpp:36-40: center + radius * -forward */
    const float f0[3] = {0.0f, -1.0f, 0.0f};
    float fwd[3];
    quat_rotate(c->q, f0, fwd);
    for (int i = 0; i < 3; i++) out[i] = c->center[i] + c->radius * -fwd[i];
}

/* glm mat4 * mat4, column-major: col_j = ((A0*b0j + A1*b1j) + A2*b2j) + A3*b3j */
static void mat4_mul(const float *a, const float *b, float *out)
{
    float r[16];
    for (int j = 0; j < 4; j++)
        for (int i = 0; i < 4; i++)
            r[4 * j + i] = ((a[i] * b[4 * j] + a[4 + i] * b[4 * j + 1]) + a[8 + i] * b[4 * j + 2]) +
                           a[12 + i] * b[4 * j + 3];
    memcpy(out, r, sizeof(r));
}

void vro_camera_get_view(const vro_camera *c, float out[16])
{
    /* src/scene/camera.cpp:42-48: transpose(mat4_cast(q)) * translate(I, -position) */
    float pos[3];
    vro_camera_get_position(c, pos);
    float tr[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    tr[12] = -pos[0];
    tr[13] = -pos[1];
    tr[14] = -pos[2];
    const float w = c->q[0], x = c->q[1], y = c->q[2], z = c->q[3];
    float qxx = x * x, qyy = y * y, qzz = z * z, qxz = x * z, qxy = x * y, qyz = y * z, qwx = w * x,
          qwy = w * y, qwz = w * z;
    /* glm::mat3_cast, m[col][row] */
    float m[3][3];
    m[0][0] = 1.0f - 2.0f * (qyy + qzz);
    m[0][1] = 2.0f * (qxy + qwz);
    m[0][2] = 2.0f * (qxz - qwy);
    m[1][0] = 2.0f * (qxy - qwz);
    m[1][1] = 1.0f - 2.0f * (qxx + qzz);
    m[1][2] = 2.0f * (qyz + qwx);
    m[2][0] = 2.0f * (qxz + qwy);
    m[2][1] = 2.0f * (qyz - qwx);
    m[2][2] = 1.0f - 2.0f * (qxx + qyy);
    float rot[16] = {0};
    for (int col = 0; col < 3; col++)
        for (int row = 0; row < 3; row++) rot[4 * col + row] = m[row][col]; /* transpose */
    rot[15] = 1.0f;
    mat4_mul(rot, tr, out);
}

void vro_projection(uint32_t width, uint32_t height, float out[16])
{
    /* src/rendering/offscreen_pass.cpp:1154-1168 */
    float aspect = (float)width / (float)height;
    /* glm::rotate(I, radians(90), (1,0,0)) (glm/ext/matrix_transform.inl) */
    float a = VRO_DEG2RAD * 90.0f, c = cosf(a), s = sinf(a);
    float axis[3] = {1.0f, 0.0f, 0.0f}; /* normalize((1,0,0)) */
    float temp[3] = {(1.0f - c) * axis[0], (1.0f - c) * axis[1], (1.0f - c) * axis[2]};
    float R[3][3];
    R[0][0] = c + temp[0] * axis[0];
    R[0][1] = temp[0] * axis[1] + s * axis[2];
    R[0][2] = temp[0] * axis[2] - s * axis[1];
    R[1][0] = temp[1] * axis[0] - s * axis[2];
    R[1][1] = c + temp[1] * axis[1];
    R[1][2] = temp[1] * axis[2] + s * axis[0];
    R[2][0] = temp[2] * axis[0] + s * axis[1];
    R[2][1] = temp[2] * axis[1] - s * axis[0];
    R[2][2] = c + temp[2] * axis[2];
    float rot[16] = {0};
    for (int col = 0; col < 3; col++)
        for (int row = 0; row < 3; row++) rot[4 * col + row] = R[col][row]; /* I * Rotate */
    rot[15] = 1.0f;
    float sc[16] = {-1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}; /* glm::scale(I, (-1,1,1)) */
    float conv[16];
    mat4_mul(rot, sc, conv);
    /* glm::perspectiveRH_ZO (glm/ext/matrix_clip_space.inl:233-246), GLM_FORCE_DEPTH_ZERO_TO_ONE
     * defined at offscreen_pass.cpp:3 */
    float fovy = VRO_DEG2RAD * 40.0f, zn = 0.1f, zf = 10.0f;
    float th = tanf(fovy / 2.0f);
    float P[16] = {0};
    P[0] = 1.0f / (aspect * th);
    P[5] = 1.0f / th;
    P[10] = zf / (zn - zf);
    P[11] = -1.0f;
    P[14] = -(zf * zn) / (zf - zn);
    mat4_mul(P, conv, out);
}

void vro_frame_from_camera(vro_frame *f, const vro_camera *cam)
{
    /* OffscreenPass::update_uniform_buffer, offscreen_pass.cpp:1152-1171 */
    vro_camera_get_view(cam, f->view);
    vro_projection(f->width, f->height, f->proj);
    vro_camera_get_position(cam, f->camera_position);
}

/* ------------------------------------------------- transfer function ---- */

/* sample_markers, src/ui/components/gradient.cpp:471-485; lerp :13-17 = a*(1-t) + b*t.
 * markers: n * stride floats, location first. */
static void sample_markers(const float *m, uint32_t n, uint32_t stride, float location, float *out)
{
    location = location < 0.0f ? 0.0f : (location > 1.0f ? 1.0f : location);
    uint32_t it = 0; /* std::lower_bound with marker.location < location (gradient.h:56-59) */
    while (it < n && m[it * stride] < location) it++;
    if (it == 0) {
        for (uint32_t c = 1; c < stride; c++) out[c - 1] = m[c];
        return;
    }
    if (it == n) {
        for (uint32_t c = 1; c < stride; c++) out[c - 1] = m[(n - 1) * stride + c];
        return;
    }
    float curr = m[it * stride], prev = m[(it - 1) * stride];
    float t = (location - prev) / (curr - prev);
    for (uint32_t c = 1; c < stride; c++)
        out[c - 1] = m[(it - 1) * stride + c] * (1.0f - t) + m[it * stride + c] * t;
}

static uint32_t f32_to_int8_sat(float v) /* imgui_internal.h:268, :472 */
{
    v = (v < 0.0f) ? 0.0f : (v > 1.0f) ? 1.0f : v;
    return (uint32_t)(int)(v * 255.0f + 0.5f);
}

void vro_gradient_discretize(const float *color_markers, uint32_t nc, const float *alpha_markers, uint32_t na,
                             uint32_t count, uint32_t *out)
{
    /* Gradient::discretize, src/ui/components/gradient.cpp:90-108 */
    float stride = 1.0f / count;
    float offset = stride / 2.0f;
    float location = offset;
    for (uint32_t i = 0; i < count; i++) {
        float rgb[3], a;
        sample_markers(color_markers, nc, 4, location, rgb);
        sample_markers(alpha_markers, na, 2, location, &a);
        /* ImGui::ColorConvertFloat4ToU32, extern/imgui/imgui.cpp:2321-2329; shifts imgui.h:2487-2490 */
        out[i] = f32_to_int8_sat(rgb[0]) | (f32_to_int8_sat(rgb[1]) << 8) | (f32_to_int8_sat(rgb[2]) << 16) |
                 (f32_to_int8_sat(a) << 24);
        location += stride;
    }
}

/* offscreen_pass.cpp:361-375: magFilter = minFilter = LINEAR, CLAMP_TO_EDGE; image format :293 */
int vro_present_bilinear(const uint8_t *src, uint32_t sw, uint32_t sh, uint8_t *dst, uint32_t dw, uint32_t dh)
{
    if (!src || !dst || !sw || !sh) return -1;
    const float kx = (float)sw / (float)dw, ky = (float)sh / (float)dh;
    for (uint32_t y = 0; y < dh; y++) {
        for (uint32_t x = 0; x < dw; x++) {
            const float u = ((float)x + 0.5f) * kx - 0.5f, v = ((float)y + 0.5f) * ky - 0.5f;
            const float fu = floorf(u), fv = floorf(v);
            const float a = u - fu, b = v - fv;
            int i0 = (int)fu, i1 = i0 + 1, j0 = (int)fv, j1 = j0 + 1;
            i0 = i0 < 0 ? 0 : (i0 > (int)sw - 1 ? (int)sw - 1 : i0);
            i1 = i1 < 0 ? 0 : (i1 > (int)sw - 1 ? (int)sw - 1 : i1);
            j0 = j0 < 0 ? 0 : (j0 > (int)sh - 1 ? (int)sh - 1 : j0);
            j1 = j1 < 0 ? 0 : (j1 > (int)sh - 1 ? (int)sh - 1 : j1);
            const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
            for (int c = 0; c < 4; c++) {
                float t = w00 * ((float)src[((size_t)j0 * sw + i0) * 4 + c] / 255.0f);
                t = t + w10 * ((float)src[((size_t)j0 * sw + i1) * 4 + c] / 255.0f);
                t = t + w01 * ((float)src[((size_t)j1 * sw + i0) * 4 + c] / 255.0f);
                t = t + w11 * ((float)src[((size_t)j1 * sw + i1) * 4 + c] / 255.0f);
                t = t < 0.0f ? 0.0f : (t > 1.0f ? 1.0f : t);
                dst[((size_t)y * dw + x) * 4 + c] = (uint8_t)(t * 255.0f + 0.5f);
            }
        }
    }
    return 0;
}

int vro_load_nrrd(const char *path, uint32_t dims[3], float *min, float *max, float **data)
{
    (void)path, (void)dims, (void)min, (void)max, (void)data;
    return -1; /* reference build only */
}

int vro_load_csv(const char *const *paths, uint32_t npaths, uint32_t dims[3], float *min, float *max, float **data,
                 char *msg, size_t msg_len)
{
    (void)paths, (void)npaths, (void)dims, (void)min, (void)max, (void)data, (void)msg, (void)msg_len;
    return -1; /* reference build only */
}

void vro_free(void *p) { free(p); }
