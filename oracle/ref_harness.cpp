/*
 * ref_harness.cpp -- builds oracle/_ref/libvr_oracle_ref.so: the reference's OWN sources
 * for the ray-march path compiled as CPU code, behind the vr_oracle.h API.
 *
 * TEST INFRASTRUCTURE ONLY (see vr_oracle.h).  Compiled by oracle/Makefile from the
 * sources where they lie under /root/reference (nothing is copied into the repo):
 *   res/shaders/volume.frag        body of main() (lines 21-52) extracted by sed at build
 *                                  time into _ref/volume_frag_body.inc and compiled against
 *                                  extern/glm as C++  -> S3, the loop
 *   src/scene/camera.cpp           Vol::Scene::Camera, unmodified            -> S1
 *   src/ui/components/gradient.cpp Gradient::discretize, unmodified (+ imgui core)
 *   src/data/nrrd_file_parser.cpp  NrrdFileParser, unmodified (+ extern/NrrdIO/*.c)
 * What the reference leaves to the Vulkan driver/hardware (coverage, samplers, blend)
 * comes from vro_common.h, shared with the port.
 *
 * The shader itself lives in ref_shader.cpp (its glm configuration differs).
 */
#define GLM_FORCE_DEPTH_ZERO_TO_ONE /* as src/rendering/offscreen_pass.cpp:3 */
#include <glm/glm.hpp>
#include <glm/gtc/matrix_transform.hpp>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <stdexcept>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "data/csv_file_parser.h"
#include "data/nrrd_file_parser.h"
#include "scene/camera.h"
#include "ui/components/gradient.h"

#include "ref_shader.h"
#include "vr_oracle.h"
#include "vro_common.h"

extern "C" const char *vro_kind(void) { return "reference"; }

/* Iterations that pass the bounds test are not observable from outside the shader;
 * the harness re-walks the positions the way volume.frag:23-24,:34-37,:47 does. */
static uint32_t count_bounds_iterations(const glm::vec3 &frag, const glm::vec3 &tex, const glm::vec3 &cam)
{
    glm::vec3 dir = glm::normalize(frag - cam);
    glm::vec3 pos = tex;
    float ray_dist = 1.8f, step_size = 0.005f;
    int steps = int(ray_dist / step_size);
    uint32_t n = 0;
    for (int i = 0; i < steps; i++) {
        if (glm::any(glm::greaterThan(pos, glm::vec3(1.0f))) || glm::any(glm::lessThan(pos, glm::vec3(0.0f)))) break;
        n++;
        pos += dir * step_size;
    }
    return n;
}

extern "C" int vro_render(const vro_frame *f, const vro_volume *vol, const uint32_t *lut, uint32_t nlut,
                          uint8_t *rgba8, float *raw, uint16_t *iters, vro_stats *stats, int nthreads)
{
    return vro_render_rows(f, vol, lut, nlut, rgba8, raw, iters, stats, nthreads, 0, 1);
}

extern "C" int vro_render_rows(const vro_frame *f, const vro_volume *vol, const uint32_t *lut, uint32_t nlut,
                               uint8_t *rgba8, float *raw, uint16_t *iters, vro_stats *stats, int nthreads,
                               uint32_t row_start, uint32_t row_stride)
{
    if (!f || !vol || !lut || nlut == 0 || !rgba8 || f->width == 0 || f->height == 0 || row_stride == 0) return -1;
    /* the shader's march constants are literals (volume.frag:29-30) */
    if (f->ray_dist != 1.8f || f->step_size != 0.005f) return -4;
    vro_rayctx rc;
    if (vro_rayctx_init(&rc, f)) return -2;
    std::vector<float> lut_lin(4 * (size_t)nlut);
    vro_decode_lut(lut, nlut, lut_lin.data());

    ref_shader_bind(f, vol, lut_lin.data(), nlut);

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
            uint32_t fetches = ref_shader_run(tex, frag, c);
            vro_blend_store(c, rgba8 + 4 * p);
            if (raw) std::memcpy(raw + 4 * p, c, sizeof(c));
            uint32_t nb = count_bounds_iterations(glm::vec3(frag[0], frag[1], frag[2]), glm::vec3(tex[0], tex[1], tex[2]),
                                                  glm::vec3(f->camera_position[0], f->camera_position[1], f->camera_position[2]));
            if (iters) iters[p] = (uint16_t)nb;
            covered++;
            n_ref += nb;
            n_in += fetches;
        }
    }
    if (stats) {
        stats->covered_pixels = covered;
        stats->n_ref = n_ref;
        stats->n_inslice = n_in;
    }
    return 0;
}

/* ---------------------------------------------------------------- camera -- */

struct vro_camera {
    Vol::Scene::Camera cam;
};

extern "C" vro_camera *vro_camera_new(void) { return new vro_camera(); }
extern "C" void vro_camera_free(vro_camera *c) { delete c; }
extern "C" void vro_camera_rotate(vro_camera *c, float dx, float dy) { c->cam.rotate(glm::vec2(dx, dy)); }
extern "C" void vro_camera_zoom(vro_camera *c, float delta) { c->cam.zoom(delta); }
extern "C" void vro_camera_get_position(const vro_camera *c, float out[3])
{
    glm::vec3 p = c->cam.get_position();
    out[0] = p.x, out[1] = p.y, out[2] = p.z;
}
extern "C" void vro_camera_get_view(const vro_camera *c, float out[16])
{
    glm::mat4 v = c->cam.get_view();
    std::memcpy(out, &v[0][0], sizeof(float) * 16);
}

extern "C" void vro_projection(uint32_t width, uint32_t height, float out[16])
{
    /* the statements of OffscreenPass::update_uniform_buffer, src/rendering/offscreen_pass.cpp:1154-1168,
     * evaluated by the reference's own glm */
    float aspect = static_cast<float>(width) / static_cast<float>(height);
    glm::mat4 coordinate_conversion =
        glm::rotate(glm::mat4(1.0f), glm::radians(90.0f), glm::vec3(1.0f, 0.0f, 0.0f)) *
        glm::scale(glm::mat4(1.0f), glm::vec3(-1.0f, 1.0f, 1.0f));
    glm::mat4 proj = glm::perspectiveRH(glm::radians(40.0f), aspect, 0.1f, 10.0f) * coordinate_conversion;
    std::memcpy(out, &proj[0][0], sizeof(float) * 16);
}

extern "C" void vro_frame_from_camera(vro_frame *f, const vro_camera *cam)
{
    vro_camera_get_view(cam, f->view);
    vro_projection(f->width, f->height, f->proj);
    vro_camera_get_position(cam, f->camera_position);
}

/* -------------------------------------------------------------- gradient -- */

extern "C" void vro_gradient_discretize(const float *color_markers, uint32_t nc, const float *alpha_markers,
                                        uint32_t na, uint32_t count, uint32_t *out)
{
    Vol::UI::Components::Gradient g;
    g.color_markers.clear();
    g.alpha_markers.clear();
    for (uint32_t i = 0; i < nc; i++)
        g.color_markers.push_back({color_markers[4 * i],
                                   glm::vec3(color_markers[4 * i + 1], color_markers[4 * i + 2], color_markers[4 * i + 3])});
    for (uint32_t i = 0; i < na; i++) g.alpha_markers.push_back({alpha_markers[2 * i], alpha_markers[2 * i + 1]});
    std::vector<uint32_t> d = g.discretize(count);
    std::memcpy(out, d.data(), sizeof(uint32_t) * count);
}

/* ------------------------------------------------------------------ nrrd -- */

extern "C" int vro_load_nrrd(const char *path, uint32_t dims[3], float *min, float *max, float **data)
{
    try {
        Vol::Data::NrrdFileParser parser{std::filesystem::path(path)};
        Vol::Data::Dataset ds = parser.parse();
        dims[0] = ds.dimensions.x, dims[1] = ds.dimensions.y, dims[2] = ds.dimensions.z;
        *min = ds.min;
        *max = ds.max;
        *data = static_cast<float *>(std::malloc(sizeof(float) * ds.data.size()));
        if (!*data) return -3;
        std::memcpy(*data, ds.data.data(), sizeof(float) * ds.data.size());
        return 0;
    } catch (const std::exception &) {
        return -1; /* importer.cpp:47-49 shows a popup; here: error code */
    }
}

/* ------------------------------------------------------------------- csv -- */

/* The reference's own CsvFileParser (src/data/csv_file_parser.cpp).  msg receives e.what() on failure. */
extern "C" int vro_load_csv(const char *const *paths, uint32_t npaths, uint32_t dims[3], float *min, float *max,
                            float **data, char *msg, size_t msg_len)
{
    try {
        std::vector<std::filesystem::path> v;
        for (uint32_t i = 0; i < npaths; i++) v.emplace_back(paths[i]);
        Vol::Data::CsvFileParser parser{v};
        Vol::Data::Dataset ds = parser.parse();
        dims[0] = ds.dimensions.x, dims[1] = ds.dimensions.y, dims[2] = ds.dimensions.z;
        *min = ds.min;
        *max = ds.max;
        *data = static_cast<float *>(std::malloc(sizeof(float) * (ds.data.size() + 1)));
        if (!*data) return -3;
        if (!ds.data.empty()) std::memcpy(*data, ds.data.data(), sizeof(float) * ds.data.size());
        return 0;
    } catch (const std::exception &e) {
        if (msg && msg_len) std::snprintf(msg, msg_len, "%s", e.what());
        return -1; /* importer.cpp:47-49 shows a popup with this text */
    }
}

extern "C" void vro_free(void *p) { std::free(p); }

extern "C" int vro_synth(int, uint32_t, int, uint32_t, uint32_t, uint32_t, void *, int) { return -1; /* port build only */ }
