/*
 * vr.h -- C ABI of the B200-native ray-march backend (libvr_b200.so).
 *
 * This is the drop-in boundary for the ONE hot path of TomBeacock/volumetric-renderer:
 * the offscreen ray-march pass.  The reference has no FFI layer; the seam is the public
 * C++ surface of Vol::Rendering::OffscreenPass (src/rendering/offscreen_pass.h:39-54)
 * plus the camera it pulls through the application singleton
 * (src/rendering/offscreen_pass.cpp:1152-1171).  Every entry point below names the
 * reference member it replaces.  Paths are relative to the reference checkout.
 *
 * Conventions
 *   - plain C types only; no C++/torch/CUDA types in any signature (a CUDA stream is
 *     passed as void*, NULL = the context's own stream);
 *   - every function returns a vr_status (0 = ok); no exception crosses the ABI.  The
 *     reference throws std::runtime_error on any failure (e.g. offscreen_pass.cpp:309-313);
 *     here the message is kept per context and read with vr_last_error();
 *   - a context is bound to one CUDA device and is not thread-safe (the reference is
 *     single-threaded, src/application.cpp:59-95);
 *   - matrices are column-major float[16] in glm's memory layout;
 *   - there is NO CPU fallback: creating a context without a usable CUDA device fails.
 */
#ifndef VR_H
#define VR_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VR_ABI_VERSION 2

typedef enum vr_status {
    VR_OK = 0,
    VR_ERR_INVALID = 1,  /* bad argument */
    VR_ERR_CUDA = 2,     /* a CUDA runtime call failed (message in vr_last_error) */
    VR_ERR_IO = 3,       /* file could not be read */
    VR_ERR_FORMAT = 4,   /* file parsed but is not a supported 3-D scalar NRRD */
    VR_ERR_NOMEM = 5,
    VR_ERR_STATE = 6     /* call not valid in the current state */
} vr_status;

/* Device storage type of the volume.  The reference widens every NRRD type to float on
 * the host (src/data/nrrd_file_parser.cpp:49-77) and uploads VK_FORMAT_R32_SFLOAT
 * (offscreen_pass.cpp:943,:968); this backend keeps u8/u16 native and converts on fetch
 * (the values seen by the march are the same floats). */
typedef enum vr_dtype { VR_U8 = 0, VR_U16 = 1, VR_F32 = 2 } vr_dtype;

/* How the density is fetched (res/shaders/volume.frag:41 `texture(u_volume, ray_pos)`). */
typedef enum vr_sampler {
    VR_SAMPLER_HW = 0,    /* one hardware-filtered trilinear fetch (9-bit fixed-point weights,
                             as a GPU running the reference shader would do) -- default */
    VR_SAMPLER_EXACT = 1  /* eight point fetches + fp32 lerp, one rounding per operation:
                             bit-exact with the CPU oracle; the verification mode */
} vr_sampler;

/* How the volume is laid out in HBM for the hardware sampler.  The reference has one layout
 * (an optimal-tiling VK_IMAGE_TYPE_3D image, offscreen_pass.cpp:943-989); here the loader may add a
 * second one.  Results of VR_SAMPLER_EXACT do not depend on it. */
typedef enum vr_layout {
    VR_LAYOUT_AUTO = 0,    /* both layouts when the volume qualifies (u8/u16, nz <= 2047, within the memory
                              budget); each frame samples the one that is cheaper for it (vr_render_frame) */
    VR_LAYOUT_ARRAY3D = 1, /* 3-D block-linear array only, hardware trilinear fetch (half texture rate) */
    VR_LAYOUT_PAIRED = 2   /* + per axis w a 2-D layered array whose layer l holds the texel pairs (v[w=l-1], v[w=l]):
                              one full-rate bilinear fetch returns both slices, the lerp along w is one fp32
                              FMA; the hardware sampler always uses them (7x the volume's bytes in HBM) */
} vr_layout;

typedef struct vr_ctx vr_ctx;

typedef struct vr_options {
    int sampler;          /* vr_sampler */
    int empty_skip;       /* 1: skip samples in cells whose transfer-function alpha is exactly 0 (result-identical) */
    int early_term;       /* 1: stop a ray once its transmittance < early_term_eps */
    float early_term_eps; /* default 1/1024: bounds the error of every channel by eps */
    int collect_stats;    /* 1: count samples per frame (vr_get_stats) */
    int write_raw;        /* 1: also keep the shader's float4 output (vr_get_raw) */
} vr_options;

typedef struct vr_stats {
    uint64_t covered_pixels; /* pixels that received a fragment */
    uint64_t n_ref;          /* loop iterations passing the bounds test of volume.frag:34-37 that this
                                frame accounts for (exact only when early_term = 0) */
    uint64_t n_inslice;      /* iterations that passed the slice test (volume.frag:39-40) and were visited */
    uint64_t n_fetch;        /* volume fetches actually executed (after empty-space skipping) */
} vr_stats;

/* What the last vr_render_frame launched (every cost decision of the context, for harnesses and logs). */
typedef struct vr_frame_info {
    int32_t layout;            /* VR_LAYOUT_ARRAY3D or VR_LAYOUT_PAIRED: the copy the frame sampled */
    int32_t pair_axis;         /* model axis 0..2 the pairs ran along; -1 on the 3-D array */
    int32_t empty_skip_active; /* 1: the occupancy grid was consulted (options.empty_skip AND enough empty cells) */
    int32_t exact_sampler;     /* 1: the verification sampler ran */
    uint32_t bbox[4];          /* x0, y0, x1, y1 (exclusive): pixels the cube can cover */
    uint32_t march_blocks;     /* CTAs launched over tiles of the bounding box */
    uint32_t clear_blocks;     /* CTAs launched only to restore the clear colour (0 when the target already holds it) */
    uint64_t target_bytes;     /* upper bound of the bytes the launch stores into the RGBA8 target (peer memory under sort-first) */
} vr_frame_info;

/* ---- lifetime ---------------------------------------------------------------------- */

/* OffscreenPass::OffscreenPass(context, width, height) (offscreen_pass.cpp:112-134).
 * Starts, like the reference, with a 1x1x1 zero volume (:118,:130) and a one-texel
 * 0xFFFFFFFF transfer function (:119,:131), default camera, slices [0,1]^3. */
int vr_create(int device, uint32_t width, uint32_t height, vr_ctx **out);
/* OffscreenPass::~OffscreenPass (offscreen_pass.cpp:136-161). */
int vr_destroy(vr_ctx *ctx);
/* Message of the last failed call on this context ("" if none).  ctx may be NULL for
 * the message of a failed vr_create. */
const char *vr_last_error(const vr_ctx *ctx);
int vr_abi_version(void);
/* Version tag of the kernels in this build (what profiles/ and bench lines are stamped with). */
const char *vr_kernel_version(void);

/* ---- per-frame ----------------------------------------------------------------------- */

/* OffscreenPass::record(command_buffer, frame_index) (offscreen_pass.cpp:163-230):
 * clear to (0.11,0.11,0.11,1), draw the cube's front faces with volume.frag, blend,
 * store RGBA8.  Asynchronous on `cuda_stream` like command recording; results are
 * complete after the stream is synchronised (vr_sync / vr_download do that). */
int vr_render(vr_ctx *ctx, void *cuda_stream);
/* The same with the reference's frame_index argument: the reference keeps MAX_FRAMES_IN_FLIGHT = 2
 * frames in flight (src/rendering/vulkan_context.h:17, main_pass.cpp:55-57,:115-117).  Each
 * frame_index < VR_MAX_FRAMES_IN_FLIGHT has its own RGBA8 target, so frame k+1 can be recorded on
 * another stream while frame k is still rendering or being downloaded.  vr_render = frame 0. */
#define VR_MAX_FRAMES_IN_FLIGHT 4
int vr_render_frame(vr_ctx *ctx, uint32_t frame_index, void *cuda_stream);
/* The decisions vr_render_frame took for the frame it enqueued last. */
int vr_get_last_frame_info(const vr_ctx *ctx, vr_frame_info *out);
/* Empty-space skipping in screen space (no reference counterpart; the frame is unchanged).  When skipping is active
 * and no statistics are collected, a small kernel ahead of the march projects every occupied cell of the volume and
 * marks the 16 x tile_h pixel tiles its screen box touches; the march sets no rays up in the other tiles and stores
 * the clear colour there.  This returns the mask of the frame recorded last into `frame_index` (synchronises):
 * bit (ty * tiles_x + tx) of `words`, plus one trailing word that is non-zero when some cell could not be
 * projected and the march ignored the mask.  `words` may be NULL to query the shape; capacity is in words and must
 * be >= (tiles_x * tiles_y + 31) / 32 + 1.  VR_ERR_STATE when that frame used no mask. */
int vr_get_tile_mask(vr_ctx *ctx, uint32_t frame_index, uint32_t *words, uint32_t capacity, uint32_t *tiles_x, uint32_t *tiles_y,
                     uint32_t *tile_h);
int vr_get_image_frame(vr_ctx *ctx, uint32_t frame_index, void **dev_rgba8, size_t *pitch_bytes);
int vr_download_frame_async(vr_ctx *ctx, uint32_t frame_index, void *host_rgba8, void *cuda_stream);
/* vr_download_frame_async that moves only what can differ from what `host_rgba8` already holds.  Outside the
 * screen bounding box of the cube every pixel of a frame is the clear colour (offscreen_pass.cpp:170-173), so
 * after one full delivery into a host buffer, later frames need only the union of their box and the box of the
 * frame delivered there before (one strided copy).  The host image is complete and identical to a full download
 * as long as the caller leaves the buffer alone between calls; the context remembers the last 8 host buffers
 * (address + frame size).  Frames of a partitioned context are always delivered in full.  bytes_copied may be NULL. */
int vr_download_frame_dirty_async(vr_ctx *ctx, uint32_t frame_index, void *host_rgba8, void *cuda_stream, size_t *bytes_copied);
/* The dirty-rectangle path remembers host buffers by ADDRESS (and frame size).  A buffer that is freed and another
 * one allocated at the same address, or written by anyone else, must be forgotten first (NULL: forget all); full
 * downloads into a known buffer update what is remembered about it. */
int vr_forget_host_buffer(vr_ctx *ctx, const void *host_rgba8);
int vr_sync(vr_ctx *ctx);

/* The uniform block the reference fills by pulling the camera inside record()
 * (OffscreenPass::update_uniform_buffer, offscreen_pass.cpp:1152-1171): pull -> push. */
int vr_set_camera(vr_ctx *ctx, const float view[16], const float proj[16], const float camera_position[3]);

/* ---- mutators ------------------------------------------------------------------------ */

/* OffscreenPass::framebuffer_size_changed(width, height) (offscreen_pass.cpp:232-255):
 * a zero size is ignored (:237-239); otherwise synchronous re-creation of the target.
 * The projection is NOT touched: the caller pushes a new one, as record() would. */
int vr_resize(vr_ctx *ctx, uint32_t width, uint32_t height);

/* OffscreenPass::volume_dataset_changed(dataset) (offscreen_pass.cpp:257-269).
 * data: host pointer, x fastest, tightly packed (nrrd_file_parser.cpp:33); copied.
 * min/max: the Dataset's min/max (offscreen_pass.cpp:265-266); pass NULL to have them
 * computed on the device (replaces nrrd_file_parser.cpp:39-40).  Synchronous. */
int vr_set_volume(vr_ctx *ctx, const void *data, int dtype, uint32_t nx, uint32_t ny, uint32_t nz,
                  const float *min_density, const float *max_density);
/* Same, `data` is a device pointer on the context's device (x fastest, tightly packed). */
int vr_set_volume_device(vr_ctx *ctx, const void *dev_data, int dtype, uint32_t nx, uint32_t ny, uint32_t nz,
                         const float *min_density, const float *max_density);
/* Importer::import -> NrrdFileParser::parse -> volume_dataset_changed
 * (src/data/importer.cpp:36-46, src/data/nrrd_file_parser.cpp:21-47) in one call with a
 * native reader: 3-D scalar NRRD/NHDR, raw/ascii/hex encoding, any of the ten scalar
 * types (u8/u16 stay native on the device, the rest are widened to float like :49-77). */
int vr_load_nrrd(vr_ctx *ctx, const char *path);
/* Non-cubic voxels (SURVEY.md 8(f).1's extension; off by default).  The reference ignores the `spacings` /
 * `space directions` of a file (src/data/nrrd_file_parser.cpp:32-33) and always draws the unit cube, and so do an
 * upload and vr_load_nrrd here.  vr_set_volume_spacing(ctx, s) makes the volume fill the box [-e/2, e/2] with
 * e_i = n_i s_i / max_j (n_j s_j) instead (n = the volume's dimensions, also of later uploads; NULL = back to the
 * unit cube): the cube is drawn scaled, rays are normalised and stepped in physical space -- step_size is the same
 * physical distance whichever way a ray points -- and everything else is the reference's loop (with s such that
 * e = (1, 1, 1) the frame is the reference's bit for bit).  vr_get_volume_spacing returns what the NRRD loaded last
 * states per axis (NaN where it states nothing) and the extents in use.  The partition helpers (vr_partition_order,
 * vr_partition_balance) take the camera in the unit cube's coordinates: camera_position / e. */
int vr_set_volume_spacing(vr_ctx *ctx, const float spacing[3]);
int vr_get_volume_spacing(const vr_ctx *ctx, float file_spacing[3], float extent[3]);
/* The reader alone, no context and no GPU: dtype/dims/data as vr_load_nrrd would upload them.
 * *data is malloc'd (release with vr_free_host); errbuf receives the message on failure. */
int vr_nrrd_read_host(const char *path, int *dtype, uint32_t dims[3], void **data, char *errbuf, size_t errbuf_len);
/* The header's per-axis sample distances alone: `spacings`, or the length of the axis's `space directions` vector; NaN
 * where the header states none (what vr_get_volume_spacing reports after vr_load_nrrd). */
int vr_nrrd_read_spacing(const char *path, double spacing[3], char *errbuf, size_t errbuf_len);
void vr_free_host(void *p);
/* CsvFileParser::parse + volume_dataset_changed (src/data/importer.cpp:30-46, src/data/csv_file_parser.cpp:14-50):
 * one CSV file per z slice, one line per y row, comma-separated floats along x.  min/max are the
 * reference's (they start at 0, so min <= 0 <= max).  Errors: VR_ERR_FORMAT with the reference's message
 * ("Inconsistant dimensions", or std::stof's for a field that is not a number). */
int vr_load_csv(vr_ctx *ctx, const char *const *paths, uint32_t npaths);
/* The parser alone, no context and no GPU.  *data is malloc'd (release with vr_free_host). */
int vr_csv_read_host(const char *const *paths, uint32_t npaths, uint32_t dims[3], float **data, float *min_density,
                     float *max_density, char *errbuf, size_t errbuf_len);
int vr_get_volume_info(const vr_ctx *ctx, int *dtype, uint32_t dims[3], float *min_density, float *max_density);
/* Overwrite ubo.min_density / max_density (offscreen_pass.cpp:265-266), e.g. with the global
 * range after each sort-last rank measured its own sub-volume. */
int vr_set_density_range(vr_ctx *ctx, float min_density, float max_density);

/* OffscreenPass::slicing_changed(min, max) (offscreen_pass.cpp:271-277). */
int vr_set_slicing(vr_ctx *ctx, const float min_slice[3], const float max_slice[3]);

/* OffscreenPass::transfer_function_changed(data) (offscreen_pass.cpp:279-288): n >= 1
 * texels packed R | G<<8 | B<<16 | A<<24, interpreted as VK_FORMAT_R8G8B8A8_SRGB (:1076). */
int vr_set_transfer_function(vr_ctx *ctx, const uint32_t *rgba8_srgb, uint32_t n);
/* The same with the texels already in device memory (no host round trip).  Either way the tables the march reads
 * -- sRGB decode, edge padding, the fp16 table of the hardware sampler, the running alpha count the occupancy
 * classification uses -- are built on the device (csrc/lut.cu). */
int vr_set_transfer_function_device(vr_ctx *ctx, const void *dev_rgba8_srgb, uint32_t n);
/* Gradient::discretize (src/ui/components/gradient.cpp:90-108) + transfer_function_changed in one call, the
 * discretisation done on the device: the texels are bit-identical to vr_gradient_discretize's.  Markers as there:
 * color_markers nc x {location, r, g, b}, alpha_markers na x {location, a}, sorted by location. */
int vr_set_gradient(vr_ctx *ctx, const float *color_markers, uint32_t nc, const float *alpha_markers, uint32_t na, uint32_t count);
/* The packed texels in use (rgba8_srgb may be NULL to query n). */
int vr_get_transfer_function(vr_ctx *ctx, uint32_t *rgba8_srgb, uint32_t capacity, uint32_t *n);

/* Pre-integrated classification: an OPTIONAL quality mode with no counterpart in the reference; it
 * CHANGES the image and is off by default.  The reference classifies every sample on its own
 * (res/shaders/volume.frag:41-45), which misses features of the transfer function thinner than the scalar difference
 * between two samples.  With substeps >= 1 the march classifies the SEGMENT between two consecutive samples from an
 * n x n table (front scalar, back scalar) built on the device from the current transfer function: `substeps`
 * sub-samples of a linearly varying scalar, opacity-corrected so that a segment of constant scalar gives the
 * reference's per-sample result exactly.  Entry (back j, front i): sub-samples at LUT coordinates
 * s_m = (i + 1/2) + (m + 1/2) / substeps * (j - i), each classified by the 1-D sampler (LINEAR, CLAMP_TO_EDGE,
 * sRGB-decoded), opacity a' = 1 - (1 - a)^(1 / substeps), composited front to back; the march looks the table up
 * bilinearly at (x = t_front * n, y = t_back * n), clamp to edge, for consecutive in-slice samples (a sample outside
 * the slices breaks the chain; the first sample of a chain contributes nothing by itself).  Positions, bounds and
 * slice tests, blend and store are the reference's.  Needs the hardware sampler, a transfer function of at most 1024 texels, the whole
 * volume (no sort-last partition) and no statistics -- vr_render fails with VR_ERR_STATE otherwise; empty-space
 * skipping is not used in this mode.  0 switches it off.  vr_get_preintegration_table returns the n * n entries
 * of (premultiplied rgb, opacity), entry (back j, front i) at table[4 * (j * n + i)]; table may be NULL to query n. */
int vr_set_preintegration(vr_ctx *ctx, uint32_t substeps);
int vr_get_preintegration_table(vr_ctx *ctx, float *table, uint32_t capacity_texels, uint32_t *n);

/* The two shader literals ray_dist = 1.8, step_size = 0.005 (res/shaders/volume.frag:29-30)
 * as parameters; steps = int(ray_dist / step_size) in fp32 (:31). */
int vr_set_march(vr_ctx *ctx, float ray_dist, float step_size);

/* Layout used by the NEXT volume upload (vr_set_volume*, vr_load_nrrd, vr_synth_*); default AUTO.
 * VR_LAYOUT_PAIRED on a volume that does not qualify makes that upload fail with VR_ERR_INVALID. */
int vr_set_volume_layout(vr_ctx *ctx, int layout);
/* Layouts resident for the current volume: VR_LAYOUT_ARRAY3D (only that) or VR_LAYOUT_PAIRED (both).
 * last_frame_paired (may be NULL): 0 if the last vr_render* sampled the 3-D array, else 1 + the model
 * axis (0 x, 1 y, 2 z) its pairs ran along -- the axis that keeps a warp's fetch in the fewest layers
 * (the axis the camera looks down, or the third axis for a view between two axes). */
int vr_get_volume_layout(const vr_ctx *ctx, int *layout, int *last_frame_paired);

int vr_get_options(const vr_ctx *ctx, vr_options *out);
int vr_set_options(vr_ctx *ctx, const vr_options *opt);

/* ---- outputs ------------------------------------------------------------------------- */

/* OffscreenPass::get_image_view() / get_sampler() (offscreen_pass.h:53-54): the colour
 * attachment, here a headless device buffer of height rows of width RGBA8 pixels, row 0
 * at the top (Vulkan framebuffer orientation).  Re-created by vr_resize. */
int vr_get_image(vr_ctx *ctx, void **dev_rgba8, size_t *pitch_bytes);
/* Synchronise and copy the image to host memory (width*height*4 bytes). */
int vr_download(vr_ctx *ctx, void *host_rgba8);
/* Asynchronous copy on the stream used by the last vr_render (host should be pinned). */
int vr_download_async(vr_ctx *ctx, void *host_rgba8, void *cuda_stream);
/* The fragment shader's own output before the blend: float4 per pixel, premultiplied rgb
 * and alpha = 1 - transmittance, (0,0,0,0) where no fragment.  Needs options.write_raw. */
int vr_get_raw(vr_ctx *ctx, void **dev_rgba_f32);
/* The float4 output of any frame in flight (options.write_raw; allocated on first use).  Under a sort-last partition a
 * frame's output IS this partial: vr_render_frame then stores no RGBA8 image (vr_composite* produces it). */
int vr_get_raw_frame(vr_ctx *ctx, uint32_t frame_index, void **dev_rgba_f32);
int vr_download_raw(vr_ctx *ctx, float *host_rgba_f32);
/* The display image as a file: binary PPM (P6, the alpha channel is dropped) or, if `path` ends in
 * ".pam", PAM with all four channels.  Stands where the reference hands the image to ImGui for display
 * (src/ui/main_window.cpp:159-165); there is no window on a headless box.  Synchronises. */
/* The step after the path: the viewport samples the offscreen image with LINEAR filters and CLAMP_TO_EDGE
 * addressing (offscreen_pass.cpp:361-375, main_window.cpp:157-165).  Resamples frame `frame_index` into a
 * device RGBA8 buffer of dst_width x dst_height (dst_pitch_bytes = 0: tightly packed) with that filter, on
 * cuda_stream (NULL: the stream of the last render).  Same size = copy.  Zero size = no-op. */
int vr_present_frame(vr_ctx *ctx, uint32_t frame_index, void *dst_dev_rgba8, uint32_t dst_width, uint32_t dst_height,
                     size_t dst_pitch_bytes, void *cuda_stream);
int vr_write_image(vr_ctx *ctx, const char *path);
int vr_get_stats(vr_ctx *ctx, vr_stats *out);
int vr_get_size(const vr_ctx *ctx, uint32_t *width, uint32_t *height);
/* Empty-space-skipping grid: cells per axis, voxels per cell edge, cells whose samples are all
 * skipped under the current transfer function. */
int vr_get_occupancy_info(vr_ctx *ctx, uint32_t cells[3], uint32_t *cell_voxels, uint64_t *empty_cells);

/* ---- input surfaces kept from the app (host-side, no GPU needed) ---------------------- */

/* Vol::Scene::Camera (src/scene/camera.h:9-23, camera.cpp:7-48). */
typedef struct vr_camera vr_camera;
vr_camera *vr_camera_new(void);
void vr_camera_free(vr_camera *cam);
void vr_camera_rotate(vr_camera *cam, float dx, float dy); /* Camera::rotate(vec2) */
void vr_camera_zoom(vr_camera *cam, float delta);          /* Camera::zoom(float) */
void vr_camera_get_position(const vr_camera *cam, float out[3]);
void vr_camera_get_view(const vr_camera *cam, float out[16]);
/* proj of update_uniform_buffer (offscreen_pass.cpp:1154-1168). */
void vr_projection(uint32_t width, uint32_t height, float out[16]);
/* update_uniform_buffer in one call: view/proj/position from `cam` at the context's size. */
int vr_set_camera_orbit(vr_ctx *ctx, const vr_camera *cam);

/* Gradient::discretize (src/ui/components/gradient.cpp:90-108,:471-485).
 * color_markers: nc x {location, r, g, b}; alpha_markers: na x {location, a}; sorted. */
int vr_gradient_discretize(const float *color_markers, uint32_t nc, const float *alpha_markers, uint32_t na,
                           uint32_t count, uint32_t *out);

/* ---- synthetic volumes (BASELINE.json configs), generated on the device ---------------- */

typedef enum vr_synth_kind {
    VR_SYNTH_MARSCHNER_LOBB = 0,
    VR_SYNTH_FBM = 1,
    VR_SYNTH_BLOBS = 2,
    VR_SYNTH_BLOBS_FBM = 3 /* config 5: the blobs as an envelope of the fractal noise (background exactly 0) */
} vr_synth_kind;
/* Generate the volume on the device and make it the context's volume (min/max computed on
 * the device). */
int vr_synth_volume(vr_ctx *ctx, int kind, uint32_t seed, int dtype, uint32_t nx, uint32_t ny, uint32_t nz);
/* Copy the current volume back to the host (tests feed it to the CPU oracle). */
int vr_download_volume(vr_ctx *ctx, void *host_out);

/* ---- multi-GPU partitioning (new; the reference is single-GPU) -------------------------- */

/* Sort-first: this context renders only the screen tiles t with t % nranks == rank
 * (tiles of tile_px x tile_px in row-major order); other pixels are left untouched.
 * nranks = 1 restores the whole frame. */
int vr_set_partition_sort_first(vr_ctx *ctx, uint32_t rank, uint32_t nranks, uint32_t tile_px);
/* The rank that renders partition tile (tile_x, tile_y) (pixels / tile_px): tiles are dealt along diagonals,
 * (tile_x + tile_y * m) % nranks with m the smallest number >= 3 coprime to nranks, so that every rank gets an even
 * share of any screen rectangle whatever the number of tiles per row. */
uint32_t vr_sort_first_owner(uint32_t tile_x, uint32_t tile_y, uint32_t nranks);

/* Sort-last: this context holds voxels [org, org+n) (+ ghost layers) of a
 * global grid `full` and accumulates only samples whose position lies in its half-open
 * sub-box.  The volume passed to vr_set_volume* / vr_synth_volume afterwards is the
 * sub-volume INCLUDING ghosts. */
int vr_set_partition_sort_last(vr_ctx *ctx, const uint32_t full_dims[3], const uint32_t own_org[3],
                               const uint32_t own_dims[3]);
/* Generate this rank's sub-volume (+ ghosts) of the full synthetic grid on the device.  The
 * density range is the LOCAL one: reduce it across ranks and call vr_set_density_range. */
int vr_synth_subvolume(vr_ctx *ctx, int kind, uint32_t seed, int dtype);
int vr_clear_partition(vr_ctx *ctx);

/* K2: ordered `over` of nparts partial images (float4: premultiplied rgb, alpha = 1 - T)
 * for pixels [pixel_begin, pixel_end) of the frame, front to back in the order given,
 * fused with the display blend + RGBA8 store into this context's image.  parts[i] are
 * device pointers readable from this context's device (local or NVLink peer memory),
 * each pointing at that partial's pixel 0. */
int vr_composite(vr_ctx *ctx, const void *const *parts, uint32_t nparts, uint64_t pixel_begin, uint64_t pixel_end,
                 void *cuda_stream);
/* vr_composite with an explicit destination (device pointer to the RGBA8 frame, NULL = this
 * context's image): lets every rank merge its own pixel range straight into the display rank's
 * image over NVLink. */
int vr_composite_to(vr_ctx *ctx, const void *const *parts, uint32_t nparts, uint64_t pixel_begin, uint64_t pixel_end,
                    void *out_rgba8, void *cuda_stream);

/* Sort-last exchanges bounding rectangles, not full frames.  A context under vr_set_partition_sort_last marches and
 * stores only the pixels inside the screen bounding box of ITS sub-box; everywhere else its float4 output stays
 * (0,0,0,0), the identity of the merge.  vr_partition_screen_rect returns that box (x0, y0, x1, y1, exclusive) for
 * any sub-box of the grid under this context's camera and target size -- every rank computes every rank's box --
 * and vr_composite_rects is vr_composite_to that neither reads nor composites partial k outside rects[4k..4k+3]
 * (rects = NULL: every partial covers the frame). */
int vr_partition_screen_rect(const vr_ctx *ctx, const uint32_t full_dims[3], const uint32_t org[3], const uint32_t dims[3],
                             uint32_t rect[4]);
int vr_composite_rects(vr_ctx *ctx, const void *const *parts, const uint32_t *rects, uint32_t nparts, uint64_t pixel_begin,
                       uint64_t pixel_end, void *out_rgba8, void *cuda_stream);

/* Peer memory for one-process-per-GPU runs (CUDA IPC; all GPUs of one NVLink/NVSwitch box).
 * vr_ipc_export: 64-byte handle of this context's RGBA8 image (which = 0) or float4 output
 * (which = 1).  vr_ipc_open maps a peer's buffer into this process (peer access is enabled on
 * first use); the pointer can be passed to vr_set_image_target / vr_composite(_to).
 * vr_set_image_target redirects frame 0 of vr_render into such a pointer (NULL restores): with
 * sort-first partitioning every rank writes its tiles directly into the display rank's image --
 * the tile gather is the kernel's own store. */
int vr_ipc_export(vr_ctx *ctx, int which, unsigned char handle[64]);
int vr_ipc_open(vr_ctx *ctx, const unsigned char handle[64], void **dev_ptr);
int vr_ipc_close(vr_ctx *ctx, void *dev_ptr);
int vr_set_image_target(vr_ctx *ctx, void *dev_rgba8);

/* Regular split of a grid into grid[0]*grid[1]*grid[2] sub-boxes: box `index` (x fastest)
 * owns voxels [org, org+dims) with faces at floor(k * full / grid). */
int vr_partition_box(const uint32_t grid[3], const uint32_t full_dims[3], uint32_t index, uint32_t org[3],
                     uint32_t dims[3]);
/* Front-to-back visiting order of those sub-boxes for a camera position (model space), valid
 * for every pixel of the frame: out[k] = box index visited k-th. */
int vr_partition_order(const uint32_t grid[3], const uint32_t full_dims[3], const float camera_position[3],
                       uint32_t *out);

/* ---- sort-last load balance: split planes that follow the camera -------------------------------------------
 * Under perspective the sub-boxes nearest the eye hold up to 1.4x the samples of the farthest, and a frame waits for
 * its slowest rank.  A rank may therefore HOLD more than it owns: vr_partition_stored_box is the regular box grown by
 * `margin` voxels on its interior faces (pass it to vr_set_partition_sort_last, then upload / synthesise), and per
 * frame every rank computes the same planes from the camera (vr_partition_balance: per axis the interior planes at
 * the quantiles of a 1 / r^2 sample density, at most `margin` voxels from the regular split; layout: the grid[0]-1
 * planes of x, then y, then z, in voxels), takes its box from them (vr_partition_box_planes; planes NULL = the
 * regular split) and declares it with vr_set_owned_box.  Every sample still has exactly one owner: the planes are
 * integers, the same on all ranks.  vr_partition_order_planes is vr_partition_order for those planes. */
int vr_set_owned_box(vr_ctx *ctx, const uint32_t org[3], const uint32_t dims[3]);
int vr_partition_stored_box(const uint32_t grid[3], const uint32_t full_dims[3], uint32_t index, const uint32_t margin[3],
                            uint32_t org[3], uint32_t dims[3]);
int vr_partition_balance(const uint32_t grid[3], const uint32_t full_dims[3], const float camera_position[3], const uint32_t margin[3],
                         uint32_t *planes);
int vr_partition_box_planes(const uint32_t grid[3], const uint32_t full_dims[3], const uint32_t *planes, uint32_t index,
                            uint32_t org[3], uint32_t dims[3]);
int vr_partition_order_planes(const uint32_t grid[3], const uint32_t full_dims[3], const uint32_t *planes,
                              const float camera_position[3], uint32_t *out);

/* vr_composite_rects over a pixel REGION (x0, y0, x1, y1) dealt in bands of 8 rows: band b of the region is merged by
 * the caller with band_index == b % band_count.  vr_sort_last_region gives the region a sort-last frame has to merge
 * into an image slot -- the bounding rectangle of this frame's sub-box rectangles and of those of the frame merged
 * into that slot before (slot_region: the caller's four words per slot, updated; first_into_slot: the whole frame):
 * outside it every partial is the identity and the slot already holds the clear colour. */
int vr_composite_region(vr_ctx *ctx, const void *const *parts, const uint32_t *rects, uint32_t nparts, const uint32_t region[4],
                        uint32_t band_index, uint32_t band_count, void *out_rgba8, void *cuda_stream);
void vr_sort_last_region(const uint32_t *rects, uint32_t nparts, uint32_t width, uint32_t height, int first_into_slot, uint32_t slot_region[4],
                         uint32_t region[4]);

/* ---- one sort-last frame of one rank in ONE call (one process per GPU) --------------------------------------
 * What multigpu.SortLast.render enqueues per frame -- split planes, owned box, view order, screen rectangles, the
 * waits / march / signals on the march stream, the waits / merge / signals on the merge stream, the display rank's
 * present hand-off -- as one host call: at eight ranks a frame is 0.2 ms, and a dozen calls through a scripting
 * host cost as much.  Counters (indices into each rank's flag block, the layout multigpu.py documents):
 * ready 256 + slot * 32 + r, read-done 384 + slot * 32 + r, range-done 512 + slot * 32 + r, consumed 640 + slot. */
typedef struct vr_sort_last_wiring {
    uint32_t rank, world, display, slots;
    uint32_t grid[3], full[3], margin[3]; /* margin 0: fixed regular split (vr_partition_balance otherwise) */
    void *const *flags;                    /* [world]: every rank's flag block, local or peer-mapped */
    void *const *raw;                      /* [slots * world]: float4 partial of rank r in slot s at raw[s * world + r] */
    void *const *image;                    /* [slots]: the display rank's RGBA8 target of slot s, local or peer-mapped */
    uint64_t pixel_begin, pixel_end;       /* (not used: the merge deals 8-row bands of the region that can differ from the clear colour to the ranks) */
} vr_sort_last_wiring;
int vr_sort_last_frame(vr_ctx *ctx, const vr_sort_last_wiring *wiring, uint64_t frame, const float camera_position[3],
                       void *march_stream, void *merge_stream, void *present_stream);

/* ---- one process driving several GPUs (the C++ multi-GPU host; csrc/group.cu) ------------------------------
 * The reference constructs one pass (src/rendering/vulkan_context.cpp:51) and records it once per frame
 * (src/rendering/main_pass.cpp:91); a group stands in the same place with one context per device.  Mutators apply
 * to every rank; vr_group_render enqueues one frame on all devices and returns (frames are handed between the
 * devices by CUDA events: no host synchronisation per frame, `slots` frames in flight); the finished frame is in the
 * display rank's (rank 0) image.  devs may name a device more than once (logical ranks on one GPU: how the
 * single-GPU tests run it).  vr_group_context lends a rank's context for per-rank calls (replicas: its camera). */
typedef struct vr_group vr_group;
typedef enum vr_partition {
    VR_PARTITION_REPLICAS = 0,   /* independent frames: every rank renders its own camera */
    VR_PARTITION_SORT_FIRST = 1, /* image tiles dealt to the ranks, volume replicated (BASELINE.json config 4) */
    VR_PARTITION_SORT_LAST = 2   /* one sub-box per rank, view-ordered merge of the partials (config 5) */
} vr_partition;
int vr_create_group(uint32_t ndev, const int *devs, int partition, uint32_t width, uint32_t height, uint32_t slots, vr_group **out);
int vr_group_destroy(vr_group *group);
const char *vr_group_last_error(const vr_group *group); /* NULL: of a failed vr_create_group */
uint32_t vr_group_size(const vr_group *group);
vr_ctx *vr_group_context(vr_group *group, uint32_t rank);
int vr_group_set_camera(vr_group *group, const float view[16], const float proj[16], const float camera_position[3]);
int vr_group_set_transfer_function(vr_group *group, const uint32_t *rgba8_srgb, uint32_t n);
int vr_group_set_slicing(vr_group *group, const float min_slice[3], const float max_slice[3]);
int vr_group_set_options(vr_group *group, const vr_options *options);
/* volume_dataset_changed for the group: replicated (replicas, sort-first) or cut into one sub-box per rank plus a
 * ghost voxel towards each neighbour (sort-last; min/max NULL: measured per rank on the device and reduced). */
int vr_group_set_volume(vr_group *group, const void *data, int dtype, uint32_t nx, uint32_t ny, uint32_t nz,
                        const float *min_density, const float *max_density);
int vr_group_synth_volume(vr_group *group, int kind, uint32_t seed, int dtype, uint32_t nx, uint32_t ny, uint32_t nz);
/* Sort-last load balance: how far (fraction of the volume's edge; default 0 = fixed regular split) the split
 * planes may follow the camera; the ranks hold that margin around their sub-boxes.  Applies to the next volume. */
int vr_group_set_balance(vr_group *group, float fraction);
int vr_group_render(vr_group *group);
int vr_group_sync(vr_group *group);
/* The frame recorded last, complete (waits for it).  rank: replicas only (else the display rank's image). */
int vr_group_download(vr_group *group, uint32_t rank, void *host_rgba8);

/* ---- no-allocation entry point --------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * min / max of `count` values of type `dtype` (VR_U8 / VR_U16 / VR_F32) in device memory -- what vr_set_volume* does
 * with NULL min/max (std::min_element / max_element of src/data/nrrd_file_parser.cpp:39-40) -- written as two
 * floats {min, max} to DEVICE memory at dev_min_max.  Two kernel launches on `cuda_stream` (NULL = the default
 * stream) and nothing else: no cudaMalloc / cudaFree, no host allocation, no copy, no synchronisation; the result is
 * there when the stream has run.  The per-block partials live in a file-scope device array of 16 KiB that every GPU
 * holding the library keeps until exit and that ALL calls on a device share: calls that overlap on the GPU
 * (different streams without ordering) race -- serialise them.  Limits: 1 <= count <= vr_minmax_static_max_count()
 * = 2^36 (a 4096^3 volume), dev_values 16-byte aligned; anything else is refused with VR_ERR_INVALID and a message
 * in errbuf (may be NULL). */
#define VR_MINMAX_STATIC_MAX_COUNT (1ull << 36)
uint64_t vr_minmax_static_max_count(void);
int vr_minmax_static(int device, const void *dev_values, int dtype, uint64_t count, float *dev_min_max, void *cuda_stream,
                     char *errbuf, size_t errbuf_len);

/* ---- measurement ----------------------------------------------------------------------- */

/* ---- stream-ordered frame hand-off between ranks (one process per GPU) --------------------
 * The reference orders its two frames in flight with per-frame fences (src/rendering/main_pass.cpp:55-57,
 * :115-117); across GPUs the same ordering is kept by 32-bit counters in device memory, written and
 * awaited in stream order by one-warp kernels: no host synchronisation and no collective per frame. */
#define VR_FLAG_COUNT 1024
/* The context's block of VR_FLAG_COUNT zero-initialised uint32 counters (export with vr_ipc_export(ctx, 2, ..)). */
int vr_get_flags(vr_ctx *ctx, void **dev_u32);
/* After everything enqueued so far on the stream: store `value` into the n <= 32 counters `flags[i]`
 * (device pointers, local or peer-mapped), with system-scope release semantics. */
int vr_stream_signal(vr_ctx *ctx, void *const *flags, uint32_t n, uint32_t value, void *cuda_stream);
/* Hold the stream until the n <= 256 consecutive LOCAL counters at `flags` have all reached `value`
 * (counters only grow).  After timeout_ms the wait gives up and the next vr_sync fails with VR_ERR_STATE. */
int vr_stream_wait(vr_ctx *ctx, const void *flags, uint32_t n, uint32_t value, uint32_t timeout_ms, void *cuda_stream);
/* 1 when the two calls above use the driver's stream memory operations (cuStreamWaitValue32 / cuStreamWriteValue32:
 * environment VR_SYNC_MEMOPS=1, an experiment -- measured equal to the one-warp kernels, and without their time-out)
 * instead of the kernels. */
int vr_sync_uses_memops(void);
/* vr_ipc_export also accepts which = 32 + frame_index: the float4 output of that frame slot (vr_get_raw_frame).
 * vr_set_image_target for any frame in flight; vr_ipc_export accepts which = 16 + frame_index for those
 * targets (allocated on first use). */
int vr_set_image_target_frame(vr_ctx *ctx, uint32_t frame_index, void *dev_rgba8);

/* Texture-fetch roofline micro-benchmark: filtered fetches of an L1-resident volume of `dtype` from all
 * SMs; returns fetches per second (device-timed).  hw_filter: 0 = point tex3D, 1 = trilinear tex3D
 * (VR_LAYOUT_ARRAY3D's fetch), 2 = bilinear fetch of the paired layered array (VR_LAYOUT_PAIRED's fetch).
 * DRAM-fed ceilings of the same path (u8/u16 only): 3 = a 1 GiB paired layered array (8x the L2) streamed exactly
 * once with ideal quad locality, one bilinear fetch per texel (4 new bytes per fetch for u16: K1's ratio on
 * config 2); 4 = the same array with one fetch per 2x2 texels (16 new bytes per fetch: the byte ceiling). */
int vr_bench_tex_peak(int device, int dtype, int hw_filter, double *fetches_per_s);
/* Measurement aid (bench.py, roofline_tex.fetch_only): while on, vr_render / vr_render_frame run a build of
 * the march kernel that does everything up to and including the volume fetches of the frame (ray setup,
 * stepping, the same fetches in the same order) and then only sums the fetched values: no transfer function,
 * no compositing, no early termination.  Its frame time is what the memory system alone allows for this
 * frame with this layout.  The image holds no picture while on.  Requires the hardware sampler with
 * empty-space skipping and statistics off, and a whole (not sort-last) volume; otherwise vr_render fails
 * with VR_ERR_STATE.  Not part of the reference interface. */
int vr_bench_fetch_only(vr_ctx *ctx, int on);
/* Test hook: the 16 floats (4 columns) of inverse(proj * view) the kernel uses. */
int vr_debug_ray_context(const vr_ctx *ctx, float out[16]);

#ifdef __cplusplus
}
#endif
#endif /* VR_H */
