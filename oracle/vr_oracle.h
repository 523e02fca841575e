/*
 * vr_oracle.h -- C API shared by the two CPU oracles of the ray-march path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (volumetric-renderer_b200/,
 * include/vr.h) may include, link or call anything under oracle/.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * Two builds export exactly this API:
 *   oracle/libvr_oracle_port.so      plain-C restatement (oracle/vr_oracle.c, oracle/synth.c)
 *   oracle/_ref/libvr_oracle_ref.so  the reference's own sources compiled as CPU code
 *                                    (oracle/ref_harness.cpp + /root/reference/...; see oracle/Makefile)
 *
 * Parity status: the reference holds no golden vectors or tests for this path
 * (SURVEY.md section 4), so parity is UNPINNED by reference fixtures; the pin is
 * the _ref build (reference shader body + camera + gradient + NRRD parser compiled
 * from /root/reference), against which the port is checked bit-for-bit, and the
 * golden vectors in tests/golden/ generated from _ref by tests/golden/make_golden.py.
 *
 * Semantics: SURVEY.md section 3A stages S1..S7.
 */
#ifndef VR_ORACLE_H
#define VR_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { VRO_U8 = 0, VRO_U16 = 1, VRO_F32 = 2 };

/* One frame's uniform state: the reference's UniformBufferObject
 * (src/rendering/offscreen_pass.h:29-37) plus the target size and the two
 * shader literals (res/shaders/volume.frag:29-30) made parameters. */
typedef struct vro_frame {
    uint32_t width, height;
    float view[16]; /* column-major, glm layout */
    float proj[16];
    float camera_position[3];
    float min_density, max_density;
    float min_slice[3], max_slice[3];
    float ray_dist, step_size; /* reference: 1.8f, 0.005f */
} vro_frame;

typedef struct vro_volume {
    const void *data; /* x fastest, tightly packed (nrrd_file_parser.cpp:33) */
    int dtype;        /* VRO_U8 / VRO_U16 / VRO_F32 */
    uint32_t nx, ny, nz;
} vro_volume;

typedef struct vro_stats {
    uint64_t covered_pixels; /* pixels that received a fragment (S2) */
    uint64_t n_ref;          /* loop iterations that passed the bounds test (volume.frag:34-37) */
    uint64_t n_inslice;      /* iterations that passed the slice test = fetch pairs (volume.frag:39-40) */
} vro_stats;

/* Which build is this?  "port" or "reference". */
const char *vro_kind(void);

/* Render one frame.  rgba8: width*height*4 bytes, row 0 at the top (display
 * equivalent, S7).  raw: optional width*height*4 floats = the shader's out_color
 * (premultiplied rgb, alpha = 1 - transmittance) with (0,0,0,0) where no fragment.
 * iters: optional width*height uint16 = iterations that passed the bounds test.
 * nthreads <= 0: all cores.  Returns 0 on success. */
int vro_render(const vro_frame *frame, const vro_volume *vol, const uint32_t *lut, uint32_t nlut,
               uint8_t *rgba8, float *raw, uint16_t *iters, vro_stats *stats, int nthreads);

/* Same, but only rows row_start, row_start + row_stride, ... are rendered (other rows of the
 * outputs are left untouched; stats count the rendered rows only).  A bounded sample of a frame
 * for the timed CPU baseline of bench.py. */
int vro_render_rows(const vro_frame *frame, const vro_volume *vol, const uint32_t *lut, uint32_t nlut,
                    uint8_t *rgba8, float *raw, uint16_t *iters, vro_stats *stats, int nthreads,
                    uint32_t row_start, uint32_t row_stride);

/* Pre-integrated classification (oracle/preint.c; port build only -- the mode has no counterpart in the reference).
 * table: n * n entries of (premultiplied rgb, opacity), entry (back j, front i) at table[4 * (j * n + i)]. */
int vro_preint_table(const uint32_t *lut, uint32_t nlut, uint32_t substeps, float *table);
int vro_render_preint(const vro_frame *frame, const vro_volume *vol, const uint32_t *lut, uint32_t nlut, uint32_t substeps,
                      uint8_t *rgba8, float *raw, int nthreads);

/* Non-cubic voxels (oracle/spacing.c; port build only -- the reference ignores a file's spacings): the volume fills the box
 * [-extent/2, extent/2], extent in (0, 1].  extent = (1, 1, 1) is vro_render bit for bit. */
int vro_render_spacing(const vro_frame *frame, const float extent[3], const vro_volume *vol, const uint32_t *lut, uint32_t nlut,
                       uint8_t *rgba8, float *raw, vro_stats *stats, int nthreads);

/* Orbit camera (src/scene/camera.cpp:7-48). */
typedef struct vro_camera vro_camera;
vro_camera *vro_camera_new(void);
void vro_camera_free(vro_camera *);
void vro_camera_rotate(vro_camera *, float dx, float dy);
void vro_camera_zoom(vro_camera *, float delta);
void vro_camera_get_position(const vro_camera *, float out[3]);
void vro_camera_get_view(const vro_camera *, float out[16]);

/* proj = perspectiveRH_ZO(40 deg, w/h, 0.1, 10) * rotate(90 deg, +x) * scale(-1,1,1)
 * (src/rendering/offscreen_pass.cpp:1154-1168). */
void vro_projection(uint32_t width, uint32_t height, float out[16]);

/* Fill view/proj/camera_position of a frame from a camera, like
 * OffscreenPass::update_uniform_buffer (offscreen_pass.cpp:1152-1171). */
void vro_frame_from_camera(vro_frame *frame, const vro_camera *cam);

/* Gradient::discretize (src/ui/components/gradient.cpp:90-108, :471-485).
 * color_markers: nc * {location, r, g, b}; alpha_markers: na * {location, a};
 * both sorted by location.  out: count packed R|G<<8|B<<16|A<<24. */
void vro_gradient_discretize(const float *color_markers, uint32_t nc, const float *alpha_markers,
                             uint32_t na, uint32_t count, uint32_t *out);

/* NRRD load through the reference loader (reference build only; the port returns -1):
 * NrrdFileParser::parse (src/data/nrrd_file_parser.cpp:21-47).  On success *data is a
 * malloc'd float array (caller frees with vro_free). */
/* Display side: LINEAR / CLAMP_TO_EDGE resample of an R8G8B8A8_UNORM image (the viewport sampler,
 * src/rendering/offscreen_pass.cpp:361-375; filter equation: Vulkan 1.0 section 15.8.3), fp32, single roundings. */
int vro_present_bilinear(const uint8_t *src_rgba8, uint32_t sw, uint32_t sh, uint8_t *dst_rgba8, uint32_t dw, uint32_t dh);
int vro_load_nrrd(const char *path, uint32_t dims[3], float *min, float *max, float **data);
/* reference build only (port: -1): CsvFileParser::parse, src/data/csv_file_parser.cpp:14-50 */
int vro_load_csv(const char *const *paths, uint32_t npaths, uint32_t dims[3], float *min, float *max, float **data,
                 char *msg, size_t msg_len);
void vro_free(void *);

/* Independent check of S2 (port build only, oracle/raster.c): the cube's twelve triangles
 * (src/rendering/offscreen_pass.cpp:55-90) rasterised under Vulkan's rules with the state at :661-712.
 * mask[H*W]: 0 = no fragment, else 1 + face index; tex[H*W*3]: interpolated in_tex_coords; depth[H*W] or NULL.
 * Returns the number of covered pixels (< 0 on failure). */
int vro_raster_cube(const vro_frame *frame, uint8_t *mask, float *tex, float *depth);
/* The analytic decision of vro_pixel_entry for every pixel, in the same form. */
int vro_entry_map(const vro_frame *frame, uint8_t *mask, float *tex);

/* Synthetic volumes of SURVEY.md section 8(d) (port build only; reference returns -1).
 * kind: 0 = Marschner-Lobb, 1 = fBm value noise, 2 = sparse Gaussian blobs. */
int vro_synth(int kind, uint32_t seed, int dtype, uint32_t nx, uint32_t ny, uint32_t nz, void *out,
              int nthreads);

#ifdef __cplusplus
}
#endif
#endif /* VR_ORACLE_H */
