/*
 * hostside.cpp -- the app's input surfaces that feed the ray-march pass, as plain host code
 * behind the C ABI: the orbit camera, the projection of update_uniform_buffer, and the
 * transfer-function discretisation.  No GPU involved; these produce the 192-byte uniform
 * block and the 256 LUT texels the kernels consume, so they must give the reference's bits.
 *
 *   Vol::Scene::Camera                      src/scene/camera.cpp:7-48
 *   OffscreenPass::update_uniform_buffer    src/rendering/offscreen_pass.cpp:1152-1171
 *   Gradient::discretize / sample_markers   src/ui/components/gradient.cpp:90-108, :471-485
 *
 * The reference evaluates these with glm 0.9.9.9 in fp32; the small vector types below keep
 * glm's evaluation order (each helper names the glm routine whose order it keeps).  Built
 * with -ffp-contract=off.
 */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <new>

#include "../../include/vr.h"

namespace
{
struct V3 {
    float x, y, z;
};
struct Quat {
    float w, x, y, z;
};
/* column-major 4x4: m[col*4 + row] */
struct M4 {
    float m[16];
};

const float kDegToRad = 0.01745329251994329576923690768489f; /* glm::radians */

V3 cross(const V3 &a, const V3 &b) /* glm::cross */
{
    return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}

Quat angle_axis(float angle, const V3 &axis) /* glm::angleAxis */
{
    const float s = std::sin(angle * 0.5f);
    return {std::cos(angle * 0.5f), axis.x * s, axis.y * s, axis.z * s};
}

Quat mul(const Quat &p, const Quat &q) /* glm::qua::operator*= */
{
    return {p.w * q.w - p.x * q.x - p.y * q.y - p.z * q.z, p.w * q.x + p.x * q.w + p.y * q.z - p.z * q.y,
            p.w * q.y + p.y * q.w + p.z * q.x - p.x * q.z, p.w * q.z + p.z * q.w + p.x * q.y - p.y * q.x};
}

V3 rotate(const Quat &q, const V3 &v) /* glm: quat * vec3 = v + ((uv * w) + uuv) * 2 */
{
    const V3 qv{q.x, q.y, q.z};
    const V3 uv = cross(qv, v);
    const V3 uuv = cross(qv, uv);
    return {v.x + ((uv.x * q.w) + uuv.x) * 2.0f, v.y + ((uv.y * q.w) + uuv.y) * 2.0f,
            v.z + ((uv.z * q.w) + uuv.z) * 2.0f};
}

M4 identity()
{
    M4 r;
    std::memset(r.m, 0, sizeof(r.m));
    r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1.0f;
    return r;
}

M4 mul(const M4 &a, const M4 &b) /* glm mat4*mat4: ((A0*b0 + A1*b1) + A2*b2) + A3*b3 per column */
{
    M4 r;
    for (int c = 0; c < 4; c++)
        for (int i = 0; i < 4; i++)
            r.m[4 * c + i] = ((a.m[i] * b.m[4 * c] + a.m[4 + i] * b.m[4 * c + 1]) + a.m[8 + i] * b.m[4 * c + 2]) +
                             a.m[12 + i] * b.m[4 * c + 3];
    return r;
}

/* transpose(mat4_cast(q)) in one go (glm::mat3_cast element formulas) */
M4 rotation_transposed(const Quat &q)
{
    const float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z, qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z,
                qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
    /* R[col][row] of mat3_cast */
    const float R00 = 1.0f - 2.0f * (qyy + qzz), R01 = 2.0f * (qxy + qwz), R02 = 2.0f * (qxz - qwy);
    const float R10 = 2.0f * (qxy - qwz), R11 = 1.0f - 2.0f * (qxx + qzz), R12 = 2.0f * (qyz + qwx);
    const float R20 = 2.0f * (qxz + qwy), R21 = 2.0f * (qyz - qwx), R22 = 1.0f - 2.0f * (qxx + qyy);
    M4 t = identity();
    /* transposed: t[col][row] = R[row][col] */
    t.m[0] = R00, t.m[1] = R10, t.m[2] = R20;
    t.m[4] = R01, t.m[5] = R11, t.m[6] = R21;
    t.m[8] = R02, t.m[9] = R12, t.m[10] = R22;
    return t;
}
}  // namespace

struct vr_camera {
    V3 center;
    Quat orientation;
    float radius;
};

extern "C" vr_camera *vr_camera_new(void)
{
    /* camera.cpp:7-13 */
    vr_camera *c = new (std::nothrow) vr_camera;
    if (!c) return nullptr;
    c->center = {0.0f, 0.0f, 0.0f};
    c->orientation = angle_axis(kDegToRad * 180.0f, {0.0f, 0.0f, 1.0f});
    c->radius = 3.0f;
    return c;
}

extern "C" void vr_camera_free(vr_camera *c) { delete c; }

extern "C" void vr_camera_rotate(vr_camera *c, float dx, float dy)
{
    /* camera.cpp:15-29: sensitivity 0.25 deg per unit; yaw about world +z, then pitch about
     * the camera's right vector */
    const float ax = dx * 0.25f, ay = dy * 0.25f;
    const Quat yaw = angle_axis(kDegToRad * (-ax), {0.0f, 0.0f, 1.0f});
    c->orientation = mul(yaw, c->orientation);
    const V3 right = rotate(c->orientation, {1.0f, 0.0f, 0.0f});
    const Quat pitch = angle_axis(kDegToRad * ay, right);
    c->orientation = mul(pitch, c->orientation);
}

extern "C" void vr_camera_zoom(vr_camera *c, float delta)
{
    /* camera.cpp:31-34 */
    const float r = c->radius - delta;
    c->radius = r < 0.1f ? 0.1f : (r > 10.0f ? 10.0f : r);
}

extern "C" void vr_camera_get_position(const vr_camera *c, float out[3])
{
    /* camera.cpp:36-40 */
    const V3 fwd = rotate(c->orientation, {0.0f, -1.0f, 0.0f});
    out[0] = c->center.x + c->radius * -fwd.x;
    out[1] = c->center.y + c->radius * -fwd.y;
    out[2] = c->center.z + c->radius * -fwd.z;
}

extern "C" void vr_camera_get_view(const vr_camera *c, float out[16])
{
    /* camera.cpp:42-48: transpose(mat4_cast(q)) * translate(I, -position) */
    float pos[3];
    vr_camera_get_position(c, pos);
    M4 tr = identity();
    tr.m[12] = -pos[0], tr.m[13] = -pos[1], tr.m[14] = -pos[2];
    const M4 v = mul(rotation_transposed(c->orientation), tr);
    std::memcpy(out, v.m, sizeof(v.m));
}

extern "C" void vr_projection(uint32_t width, uint32_t height, float out[16])
{
    /* offscreen_pass.cpp:1154-1168:
     * perspectiveRH_ZO(radians(40), w/h, 0.1, 10) * (rotate(I, radians(90), +x) * scale(I, (-1,1,1))) */
    const float aspect = static_cast<float>(width) / static_cast<float>(height);
    /* glm::rotate(I, a, (1,0,0)): axis already unit; temp = (1-c)*axis */
    const float a = kDegToRad * 90.0f, c = std::cos(a), s = std::sin(a);
    const float ax = 1.0f, ay = 0.0f, az = 0.0f;
    const float tx = (1.0f - c) * ax, ty = (1.0f - c) * ay, tz = (1.0f - c) * az;
    M4 rot = identity();
    rot.m[0] = c + tx * ax, rot.m[1] = tx * ay + s * az, rot.m[2] = tx * az - s * ay;
    rot.m[4] = ty * ax - s * az, rot.m[5] = c + ty * ay, rot.m[6] = ty * az + s * ax;
    rot.m[8] = tz * ax + s * ay, rot.m[9] = tz * ay - s * ax, rot.m[10] = c + tz * az;
    M4 sc = identity();
    sc.m[0] = -1.0f;
    const M4 conv = mul(rot, sc);
    /* glm::perspectiveRH_ZO (GLM_FORCE_DEPTH_ZERO_TO_ONE, offscreen_pass.cpp:3) */
    const float fovy = kDegToRad * 40.0f, zn = 0.1f, zf = 10.0f;
    const float th = std::tan(fovy / 2.0f);
    M4 P;
    std::memset(P.m, 0, sizeof(P.m));
    P.m[0] = 1.0f / (aspect * th);
    P.m[5] = 1.0f / th;
    P.m[10] = zf / (zn - zf);
    P.m[11] = -1.0f;
    P.m[14] = -(zf * zn) / (zf - zn);
    const M4 r = mul(P, conv);
    std::memcpy(out, r.m, sizeof(r.m));
}

/* ------------------------------------------------------------ transfer function -- */

namespace
{
/* sample_markers (gradient.cpp:471-485): clamp, std::lower_bound on location, lerp
 * a*(1-t) + b*t (gradient.cpp:13-17).  Markers are rows of `stride` floats, location first. */
void sample_markers(const float *m, uint32_t n, uint32_t stride, float location, float *out)
{
    location = location < 0.0f ? 0.0f : (location > 1.0f ? 1.0f : location);
    uint32_t lo = 0, hi = n; /* first marker with location >= query */
    while (lo < hi) {
        const uint32_t mid = lo + (hi - lo) / 2;
        if (m[mid * stride] < location) lo = mid + 1;
        else hi = mid;
    }
    const uint32_t it = lo;
    if (it == 0 || it == n) {
        const float *row = m + (it == 0 ? 0 : (n - 1)) * stride;
        for (uint32_t c = 1; c < stride; c++) out[c - 1] = row[c];
        return;
    }
    const float *prev = m + (it - 1) * stride, *curr = m + it * stride;
    const float t = (location - prev[0]) / (curr[0] - prev[0]);
    for (uint32_t c = 1; c < stride; c++) out[c - 1] = prev[c] * (1.0f - t) + curr[c] * t;
}

uint32_t unorm8_sat(float v) /* IM_F32_TO_INT8_SAT: extern/imgui/imgui_internal.h:268,:472 */
{
    v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
    return static_cast<uint32_t>(static_cast<int>(v * 255.0f + 0.5f));
}
}  // namespace

extern "C" int vr_gradient_discretize(const float *color_markers, uint32_t nc, const float *alpha_markers, uint32_t na,
                                      uint32_t count, uint32_t *out)
{
    if (!color_markers || !alpha_markers || !out || nc == 0 || na == 0 || count == 0) return VR_ERR_INVALID;
    /* gradient.cpp:90-108: texel centres, location accumulated by repeated addition */
    const float stride = 1.0f / count;
    float location = stride / 2.0f;
    for (uint32_t i = 0; i < count; i++) {
        float rgb[3], a;
        sample_markers(color_markers, nc, 4, location, rgb);
        sample_markers(alpha_markers, na, 2, location, &a);
        /* ImGui::ColorConvertFloat4ToU32 (extern/imgui/imgui.cpp:2321-2329), R in the low byte */
        out[i] = unorm8_sat(rgb[0]) | (unorm8_sat(rgb[1]) << 8) | (unorm8_sat(rgb[2]) << 16) | (unorm8_sat(a) << 24);
        location += stride;
    }
    return VR_OK;
}
