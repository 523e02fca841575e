/*
 * ref_shader.cpp -- the reference fragment shader compiled as C++ (one translation unit).
 *
 * TEST INFRASTRUCTURE ONLY (see vr_oracle.h).  The body of main() of
 * /root/reference/res/shaders/volume.frag (lines 21-52) is extracted by oracle/Makefile at
 * build time and #included below, unmodified, against the reference's vendored glm.
 * This file supplies what the GLSL front end would: the interface block (volume.frag:8-16),
 * the two samplers with their texture() overloads (S4/S6 from vro_common.h), and the
 * per-fragment inputs/outputs (volume.frag:3-6).
 *
 * Compiled with -fsingle-precision-constant so the shader's unsuffixed literals are fp32
 * as in GLSL (otherwise C++ would evaluate `1.0 - sample_color.a` in double).
 * The glm configuration macros below change glm's vector types, so this TU shares no
 * glm-typed interface with the rest of the library: plain floats only.
 */
#define GLM_FORCE_SWIZZLE
#define GLM_FORCE_INTRINSICS /* only so glm enables .rgb swizzle operators on GCC (glm/detail/setup.hpp) */
#include <glm/glm.hpp>

#include <cstring>

#include "ref_shader.h"
#include "vro_common.h"

namespace shader
{
using namespace glm;

struct sampler3D {
    const vro_volume *vol;
};
struct sampler1D {
    const float *rgba;
    uint32_t n;
};

struct UniformBufferObject { /* volume.frag:8-16 */
    mat4 view;
    mat4 proj;
    vec3 camera_position;
    float min_density;
    float max_density;
    vec3 min_slice;
    vec3 max_slice;
};

static UniformBufferObject u_ubo;
static sampler3D u_volume;
static sampler1D u_transfer_func;
static thread_local vec3 in_tex_coords;
static thread_local vec3 in_frag_position;
static thread_local vec4 out_color;
static thread_local uint32_t tl_fetches; /* texture(u_volume, ..) calls = in-slice iterations */

static inline vec4 texture(const sampler3D &s, const vec3 &p)
{
    tl_fetches++;
    return vec4(vro_sample_volume(s.vol, p.x, p.y, p.z), 0.0f, 0.0f, 1.0f);
}

static inline vec4 texture(const sampler1D &s, float t)
{
    float o[4];
    vro_sample_lut(s.rgba, s.n, t, o);
    return vec4(o[0], o[1], o[2], o[3]);
}

#define main shader_main
#include "_ref/volume_frag_body.inc"
#undef main
}  // namespace shader

void ref_shader_bind(const vro_frame *f, const vro_volume *vol, const float *lut_rgba, uint32_t nlut)
{
    std::memcpy(&shader::u_ubo.view, f->view, sizeof(float) * 16);
    std::memcpy(&shader::u_ubo.proj, f->proj, sizeof(float) * 16);
    shader::u_ubo.camera_position = glm::vec3(f->camera_position[0], f->camera_position[1], f->camera_position[2]);
    shader::u_ubo.min_density = f->min_density;
    shader::u_ubo.max_density = f->max_density;
    shader::u_ubo.min_slice = glm::vec3(f->min_slice[0], f->min_slice[1], f->min_slice[2]);
    shader::u_ubo.max_slice = glm::vec3(f->max_slice[0], f->max_slice[1], f->max_slice[2]);
    shader::u_volume.vol = vol;
    shader::u_transfer_func.rgba = lut_rgba;
    shader::u_transfer_func.n = nlut;
}

uint32_t ref_shader_run(const float tex[3], const float frag[3], float out[4])
{
    /* volume.vert:20-21: out_tex_coords = in_tex_coords, out_frag_position = in_position */
    shader::in_tex_coords = glm::vec3(tex[0], tex[1], tex[2]);
    shader::in_frag_position = glm::vec3(frag[0], frag[1], frag[2]);
    shader::tl_fetches = 0;
    shader::shader_main();
    out[0] = shader::out_color.x, out[1] = shader::out_color.y, out[2] = shader::out_color.z, out[3] = shader::out_color.w;
    return shader::tl_fetches;
}
