/*
 * api.cu -- the C ABI of include/vr.h: the render context that stands where
 * Vol::Rendering::OffscreenPass stands in the reference (src/rendering/offscreen_pass.{h,cpp}).
 *
 * The reference's pass owns Vulkan objects (colour/depth attachments, pipeline, descriptor
 * sets, a 3-D image + sampler, a 1-D sRGB image + sampler, two uniform buffers); this context
 * owns their CUDA equivalents: an RGBA8 device buffer, a 3-D CUDA array with two texture
 * objects, the decoded LUT, the occupancy grid, and one stream.  Every mutator keeps the
 * reference's semantics (synchronous, copy-in), every failure becomes a status + message.
 */
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <new>
#include <vector>

#include "nrrd.h"
#include "vr_ctx.h"

namespace
{
thread_local std::string g_create_error;

int fail(vr_ctx *c, int status, const std::string &msg)
{
    if (c) c->err = msg;
    else g_create_error = msg;
    return status;
}

int cuda_fail(vr_ctx *c, cudaError_t e, const char *what)
{
    return fail(c, VR_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

#define VR_CUDA(c, call)                                     \
    do {                                                     \
        cudaError_t e__ = (call);                            \
        if (e__ != cudaSuccess) return cuda_fail(c, e__, #call); \
    } while (0)

size_t dtype_size(int dtype) { return dtype == VR_U8 ? 1 : (dtype == VR_U16 ? 2 : 4); }

/* ---- S2 host part: inverse(proj * view) in double, rounded to fp32 -------------------- */

void mat4_mul_d(const double *a, const double *b, double *out)
{
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++) {
            double s = 0.0;
            for (int k = 0; k < 4; k++) s += a[k * 4 + r] * b[c * 4 + k];
            out[c * 4 + r] = s;
        }
}

/* 4x4 inverse through 2x2 sub-determinants (Laplace expansion by complementary minors) */
bool mat4_inverse_d(const double *m, double *inv)
{
    const double a00 = m[0], a01 = m[1], a02 = m[2], a03 = m[3], a10 = m[4], a11 = m[5], a12 = m[6], a13 = m[7],
                 a20 = m[8], a21 = m[9], a22 = m[10], a23 = m[11], a30 = m[12], a31 = m[13], a32 = m[14], a33 = m[15];
    const double b00 = a00 * a11 - a01 * a10, b01 = a00 * a12 - a02 * a10, b02 = a00 * a13 - a03 * a10,
                 b03 = a01 * a12 - a02 * a11, b04 = a01 * a13 - a03 * a11, b05 = a02 * a13 - a03 * a12,
                 b06 = a20 * a31 - a21 * a30, b07 = a20 * a32 - a22 * a30, b08 = a20 * a33 - a23 * a30,
                 b09 = a21 * a32 - a22 * a31, b10 = a21 * a33 - a23 * a31, b11 = a22 * a33 - a23 * a32;
    const double det = b00 * b11 - b01 * b10 + b02 * b09 + b03 * b08 - b04 * b07 + b05 * b06;
    if (det == 0.0 || !(det == det)) return false;
    const double r = 1.0 / det;
    inv[0] = (a11 * b11 - a12 * b10 + a13 * b09) * r;
    inv[1] = (a02 * b10 - a01 * b11 - a03 * b09) * r;
    inv[2] = (a31 * b05 - a32 * b04 + a33 * b03) * r;
    inv[3] = (a22 * b04 - a21 * b05 - a23 * b03) * r;
    inv[4] = (a12 * b08 - a10 * b11 - a13 * b07) * r;
    inv[5] = (a00 * b11 - a02 * b08 + a03 * b07) * r;
    inv[6] = (a32 * b02 - a30 * b05 - a33 * b01) * r;
    inv[7] = (a20 * b05 - a22 * b02 + a23 * b01) * r;
    inv[8] = (a10 * b10 - a11 * b08 + a13 * b06) * r;
    inv[9] = (a01 * b08 - a00 * b10 - a03 * b06) * r;
    inv[10] = (a30 * b04 - a31 * b02 + a33 * b00) * r;
    inv[11] = (a21 * b02 - a20 * b04 - a23 * b00) * r;
    inv[12] = (a11 * b07 - a10 * b09 - a12 * b06) * r;
    inv[13] = (a00 * b09 - a01 * b07 + a02 * b06) * r;
    inv[14] = (a31 * b01 - a30 * b03 - a32 * b00) * r;
    inv[15] = (a20 * b03 - a21 * b01 + a22 * b00) * r;
    return true;
}

bool update_ray_context(vr_ctx *c);

/* The box the volume fills and the camera as the unit cube sees it (vr_set_volume_spacing).  Called when the camera, the
 * spacing or the volume's dimensions change. */
bool apply_extent(vr_ctx *c)
{
    for (int i = 0; i < 3; i++) c->ext[i] = 1.0f;
    if (c->honour_spacing) {
        const uint32_t d[3] = {c->sort_last ? c->full[0] : c->nx, c->sort_last ? c->full[1] : c->ny, c->sort_last ? c->full[2] : c->nz};
        double phys[3], m = 0.0;
        for (int i = 0; i < 3; i++) phys[i] = (double)d[i] * (double)c->spacing[i], m = std::max(m, phys[i]);
        if (m > 0.0)
            for (int i = 0; i < 3; i++) c->ext[i] = phys[i] > 0.0 ? (float)(phys[i] / m) : 1.0f;
    }
    for (int col = 0; col < 4; col++)
        for (int r = 0; r < 4; r++) c->view[col * 4 + r] = col < 3 ? c->view_in[col * 4 + r] * c->ext[col] : c->view_in[col * 4 + r];
    for (int i = 0; i < 3; i++) c->cam_pos[i] = c->cam_in[i] / c->ext[i];
    return update_ray_context(c);
}

bool update_ray_context(vr_ctx *c)
{
    double v[16], p[16], pv[16], inv[16];
    for (int i = 0; i < 16; i++) v[i] = (double)c->view[i], p[i] = (double)c->proj[i];
    mat4_mul_d(p, v, pv); /* gl_Position = proj * view * vec4(pos, 1): volume.vert:23 */
    c->ray_valid = mat4_inverse_d(pv, inv);
    if (!c->ray_valid) return false;
    for (int col = 0; col < 4; col++)
        for (int r = 0; r < 4; r++) c->rc[col][r] = (float)inv[col * 4 + r];
    return true;
}

/* ---- S6 host part: VK_FORMAT_R8G8B8A8_SRGB decode, edge padding ------------------------ */

float srgb8_to_linear(unsigned c8)
{
    const double c = (double)c8 / 255.0;
    return (float)(c <= 0.04045 ? c / 12.92 : std::pow((c + 0.055) / 1.055, 2.4));
}

/* (Re)build everything the march reads of the transfer function from the packed texels in d_lut_rgba8 (lut.cu). */
int upload_lut(vr_ctx *c)
{
    const uint32_t n = c->nlut;
    if (c->lut_tables_cap < n) {
        if (c->d_lut_pad) cudaFree(c->d_lut_pad), c->d_lut_pad = nullptr;
        if (c->d_lut_dif) cudaFree(c->d_lut_dif), c->d_lut_dif = nullptr;
        if (c->d_lut_h) cudaFree(c->d_lut_h), c->d_lut_h = nullptr;
        if (c->d_lut_nz) cudaFree(c->d_lut_nz), c->d_lut_nz = nullptr;
        c->lut_tables_cap = 0;
        VR_CUDA(c, cudaMalloc(&c->d_lut_h, sizeof(uint4) * (n + 2)));
        VR_CUDA(c, cudaMalloc(&c->d_lut_pad, sizeof(float4) * (n + 3)));
        VR_CUDA(c, cudaMalloc(&c->d_lut_dif, sizeof(float4) * (n + 2)));
        VR_CUDA(c, cudaMalloc(&c->d_lut_nz, sizeof(uint32_t) * (n + 4)));
        c->lut_tables_cap = n;
    }
    VR_CUDA(c, vr_lut_tables(c->d_lut_rgba8, n, c->d_lut_pad, c->d_lut_dif, c->d_lut_h, c->d_lut_nz, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    c->occ_dirty = true;
    c->preint_dirty = true;
    return VR_OK;
}

/* room for n packed texels (+ the texel centres of a device-side discretize) */
int reserve_lut(vr_ctx *c, uint32_t n)
{
    if (c->lut_texels_cap >= n) return VR_OK;
    if (c->d_lut_rgba8) cudaFree(c->d_lut_rgba8), c->d_lut_rgba8 = nullptr;
    if (c->d_lut_loc) cudaFree(c->d_lut_loc), c->d_lut_loc = nullptr;
    c->lut_texels_cap = 0;
    VR_CUDA(c, cudaMalloc(&c->d_lut_rgba8, sizeof(uint32_t) * n));
    VR_CUDA(c, cudaMalloc(&c->d_lut_loc, sizeof(float) * n));
    c->lut_texels_cap = n;
    return VR_OK;
}

void release_volume(vr_ctx *c)
{
    if (c->tex_lin) cudaDestroyTextureObject(c->tex_lin), c->tex_lin = 0;
    if (c->tex_pt) cudaDestroyTextureObject(c->tex_pt), c->tex_pt = 0;
    for (int a = 0; a < 3; a++) {
        if (c->tex_pair[a]) cudaDestroyTextureObject(c->tex_pair[a]), c->tex_pair[a] = 0;
        if (c->tex_pair_hi[a]) cudaDestroyTextureObject(c->tex_pair_hi[a]), c->tex_pair_hi[a] = 0;
        if (c->pair_array[a]) cudaFreeArray(c->pair_array[a]), c->pair_array[a] = nullptr;
        if (c->pair_array_hi[a]) cudaFreeArray(c->pair_array_hi[a]), c->pair_array_hi[a] = nullptr;
        c->pair_split[a] = 0x10000;
    }
    if (c->array) cudaFreeArray(c->array), c->array = nullptr;
    if (c->d_cell_range) cudaFree(c->d_cell_range), c->d_cell_range = nullptr;
    if (c->d_occ) cudaFree(c->d_occ), c->d_occ = nullptr;
    if (c->d_dist) cudaFree(c->d_dist), c->d_dist = nullptr;
    if (c->d_sat) cudaFree(c->d_sat), c->d_sat = nullptr;
    if (c->d_occ_list) cudaFree(c->d_occ_list), c->d_occ_list = nullptr;
    if (c->d_occ_count) cudaFree(c->d_occ_count), c->d_occ_count = nullptr;
}

int alloc_target(vr_ctx *c, uint32_t w, uint32_t h)
{
    if (c->d_image) cudaFree(c->d_image), c->d_image = nullptr;
    if (c->d_raw) cudaFree(c->d_raw), c->d_raw = nullptr;
    for (int k = 1; k < VR_MAX_FRAMES_IN_FLIGHT; k++) {
        if (c->d_image_slot[k]) cudaFree(c->d_image_slot[k]), c->d_image_slot[k] = nullptr;
        if (c->d_raw_slot[k]) cudaFree(c->d_raw_slot[k]), c->d_raw_slot[k] = nullptr;
    }
    for (int k = 0; k < VR_MAX_FRAMES_IN_FLIGHT; k++)
        if (c->d_tile_mask[k]) cudaFree(c->d_tile_mask[k]), c->d_tile_mask[k] = nullptr;
    c->W = w, c->H = h;
    VR_CUDA(c, cudaMalloc(&c->d_image, (size_t)w * h * sizeof(uchar4)));
    VR_CUDA(c, cudaMemset(c->d_image, 0, (size_t)w * h * sizeof(uchar4)));
    if (c->opt.write_raw) {
        VR_CUDA(c, cudaMalloc(&c->d_raw, (size_t)w * h * sizeof(float4)));
        VR_CUDA(c, cudaMemset(c->d_raw, 0, (size_t)w * h * sizeof(float4)));
    }
    return VR_OK;
}

/* Can this volume be held in the paired layout at all?  (u8/u16, every axis within two layered arrays) */
bool volume_can_pair(const vr_ctx *c, int dtype, uint32_t nx, uint32_t ny, uint32_t nz)
{
    const uint32_t nd[3] = {nx, ny, nz};
    const int split = std::min(std::max(c->tune_pair_split, 2), 2048);
    bool can_pair = dtype == VR_U8 || dtype == VR_U16;
    for (int a = 0; a < 3; a++) can_pair = can_pair && nd[a] + 1 <= 2u * (uint32_t)split && nd[a] <= 32768;
    return can_pair;
}

int install_volume_unchecked(vr_ctx *c, const void *dev, int dtype, uint32_t nx, uint32_t ny, uint32_t nz, const float *mn,
                             const float *mx);

/* Make `dev` (linear, x fastest, local extent) the context's volume.  A request that cannot be honoured is refused
 * before anything is touched (the old volume stays); a failure while the new one is being built (out of memory, a
 * CUDA error) leaves the context without a volume -- vr_render then reports "no volume" -- never half-installed. */
int install_volume(vr_ctx *c, const void *dev, int dtype, uint32_t nx, uint32_t ny, uint32_t nz, const float *mn,
                   const float *mx)
{
    if (c->layout_req == VR_LAYOUT_PAIRED && !volume_can_pair(c, dtype, nx, ny, nz))
        return fail(c, VR_ERR_INVALID, "VR_LAYOUT_PAIRED needs a u8/u16 volume of at most 4095 voxels per axis");
    const int rc = install_volume_unchecked(c, dev, dtype, nx, ny, nz, mn, mx);
    if (rc != VR_OK) {
        const std::string why = c->err;
        cudaStreamSynchronize(c->stream);
        release_volume(c);
        c->nx = c->ny = c->nz = 0;
        c->occ_n[0] = c->occ_n[1] = c->occ_n[2] = 0, c->occ_words = 0, c->occ_empty_cells = 0;
        c->occ_dirty = true;
        c->err = why;
    }
    return rc;
}

int install_volume_unchecked(vr_ctx *c, const void *dev, int dtype, uint32_t nx, uint32_t ny, uint32_t nz, const float *mn,
                             const float *mx)
{
    release_volume(c);
    c->dtype = dtype, c->nx = nx, c->ny = ny, c->nz = nz;
    if (c->honour_spacing) apply_extent(c); /* the box follows the dimensions */
    c->layout_used = c->layout_req;
    for (int i = 0; i < VR_MAX_FRAMES_IN_FLIGHT; i++) c->slot_clear[i].valid = false; /* (cheap) a new volume starts from a full clear */
    cudaChannelFormatDesc cd = dtype == VR_U8    ? cudaCreateChannelDesc<unsigned char>()
                               : dtype == VR_U16 ? cudaCreateChannelDesc<unsigned short>()
                                                 : cudaCreateChannelDesc<float>();
    VR_CUDA(c, cudaMalloc3DArray(&c->array, &cd, make_cudaExtent(nx, ny, nz)));
    cudaMemcpy3DParms cp = {};
    cp.srcPtr = make_cudaPitchedPtr(const_cast<void *>(dev), (size_t)nx * dtype_size(dtype), nx, ny);
    cp.dstArray = c->array;
    cp.extent = make_cudaExtent(nx, ny, nz);
    cp.kind = cudaMemcpyDeviceToDevice;
    VR_CUDA(c, cudaMemcpy3DAsync(&cp, c->stream));

    cudaResourceDesc rd = {};
    rd.resType = cudaResourceTypeArray;
    rd.res.array.array = c->array;
    /* offscreen_pass.cpp:1016-1032: LINEAR, CLAMP_TO_BORDER, FLOAT_TRANSPARENT_BLACK, normalised */
    cudaTextureDesc td = {};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeBorder;
    td.filterMode = cudaFilterModeLinear;
    td.readMode = dtype == VR_F32 ? cudaReadModeElementType : cudaReadModeNormalizedFloat;
    td.normalizedCoords = 1;
    VR_CUDA(c, cudaCreateTextureObject(&c->tex_lin, &rd, &td, nullptr));
    cudaTextureDesc tp = {};
    tp.addressMode[0] = tp.addressMode[1] = tp.addressMode[2] = cudaAddressModeBorder;
    tp.filterMode = cudaFilterModePoint;
    tp.readMode = cudaReadModeElementType;
    tp.normalizedCoords = 0;
    VR_CUDA(c, cudaCreateTextureObject(&c->tex_pt, &rd, &tp, nullptr));

    /* second layout for the hardware sampler (vr_layout): pairs along each axis in 2-D layered arrays;
     * a layered array holds at most 2048 layers, so an axis longer than 2047 voxels is split over two */
    const uint32_t nd[3] = {nx, ny, nz};
    const int split = std::min(std::max(c->tune_pair_split, 2), 2048);
    const bool can_pair = volume_can_pair(c, dtype, nx, ny, nz);
    size_t pair_bytes = 0;
    for (int a = 0; a < 3; a++) pair_bytes += (size_t)nx * ny * nz / nd[a] * ((size_t)nd[a] + 1) * 2 * dtype_size(dtype);
    bool want_pairs = c->layout_req == VR_LAYOUT_PAIRED;
    if (c->layout_req == VR_LAYOUT_AUTO && can_pair && pair_bytes <= c->tune_pair_max_bytes) {
        size_t free_b = 0, total_b = 0;
        VR_CUDA(c, cudaMemGetInfo(&free_b, &total_b));
        want_pairs = pair_bytes + ((size_t)6 << 30) <= free_b; /* leave room for targets, partials and the caller */
    }
    if (want_pairs) {
        cudaChannelFormatDesc cd2 = dtype == VR_U8 ? cudaCreateChannelDesc<uchar2>() : cudaCreateChannelDesc<ushort2>();
        for (int a = 0; a < 3; a++) {
            /* in-plane extent (u, v), layers along a: a = 2 -> (x, y), a = 1 -> (x, z), a = 0 -> (y, z) */
            const uint32_t eu = a == 0 ? ny : nx, ev = a == 2 ? ny : nz;
            const uint32_t layers = nd[a] + 1, lo = std::min(layers, (uint32_t)split), hi = layers - lo;
            c->pair_split[a] = hi ? (int)lo : 0x10000;
            cudaArray_t *arrs[2] = {&c->pair_array[a], &c->pair_array_hi[a]};
            cudaTextureObject_t *texs[2] = {&c->tex_pair[a], &c->tex_pair_hi[a]};
            const uint32_t cnt[2] = {lo, hi};
            for (int k = 0; k < 2; k++)
                if (cnt[k])
                    VR_CUDA(c, cudaMalloc3DArray(arrs[k], &cd2, make_cudaExtent(eu, ev, cnt[k]), cudaArrayLayered | cudaArraySurfaceLoadStore));
            VR_CUDA(c, vr_k0_pack_pairs(dev, dtype, nx, ny, nz, a, c->pair_array[a], c->pair_array_hi[a], c->pair_split[a], c->stream));
            for (int k = 0; k < 2; k++) {
                if (!cnt[k]) continue;
                cudaResourceDesc rp = {};
                rp.resType = cudaResourceTypeArray;
                rp.res.array.array = *arrs[k];
                cudaTextureDesc tq = {};
                tq.addressMode[0] = tq.addressMode[1] = tq.addressMode[2] = cudaAddressModeBorder;
                tq.filterMode = cudaFilterModeLinear;
                tq.readMode = cudaReadModeNormalizedFloat;
                tq.normalizedCoords = 1;
                VR_CUDA(c, cudaCreateTextureObject(texs[k], &rp, &tq, nullptr));
            }
        }
    }

    /* ubo.min_density / max_density = dataset.min / max (offscreen_pass.cpp:265-266) */
    if (mn && mx) {
        c->min_density = *mn, c->max_density = *mx;
    } else {
        float a = 0.f, b = 0.f;
        VR_CUDA(c, vr_k0_global_minmax(dev, dtype, (size_t)nx * ny * nz, &a, &b, c->stream));
        c->min_density = mn ? *mn : a;
        c->max_density = mx ? *mx : b;
    }

    /* occupancy grid: power-of-two cells, at most tune_occ_max_axis (128) per axis */
    uint32_t maxdim = nx > ny ? nx : ny;
    maxdim = maxdim > nz ? maxdim : nz;
    int cell = 4;
    while ((maxdim + cell - 1) / cell > (uint32_t)c->tune_occ_max_axis) cell *= 2;
    c->occ_cell = cell;
    c->occ_n[0] = (int)((nx + cell - 1) / cell), c->occ_n[1] = (int)((ny + cell - 1) / cell),
    c->occ_n[2] = (int)((nz + cell - 1) / cell);
    const int ncells = c->occ_n[0] * c->occ_n[1] * c->occ_n[2];
    c->occ_words = (ncells + 31) / 32;
    VR_CUDA(c, cudaMalloc(&c->d_cell_range, sizeof(float2) * ncells));
    VR_CUDA(c, cudaMalloc(&c->d_occ, sizeof(uint32_t) * c->occ_words));
    VR_CUDA(c, cudaMalloc(&c->d_dist, (size_t)ncells * 8));
    VR_CUDA(c, cudaMalloc(&c->d_sat, sizeof(uint32_t) * (size_t)(c->occ_n[0] + 1) * (c->occ_n[1] + 1) * (c->occ_n[2] + 1)));
    VR_CUDA(c, cudaMalloc(&c->d_occ_list, sizeof(uint32_t) * (size_t)ncells));
    VR_CUDA(c, cudaMalloc(&c->d_occ_count, sizeof(uint32_t)));
    VR_CUDA(c, vr_k0_cell_ranges(dev, dtype, nx, ny, nz, cell, c->occ_n, c->d_cell_range, c->stream));
    c->occ_dirty = true;
    VR_CUDA(c, cudaStreamSynchronize(c->stream)); /* the reference's upload is synchronous too */
    return VR_OK;
}

int default_state(vr_ctx *c)
{
    /* offscreen_pass.cpp:118-119,:130-131: 1x1x1 zero volume, one white opaque LUT texel */
    float *d0 = nullptr;
    VR_CUDA(c, cudaMalloc(&d0, sizeof(float)));
    VR_CUDA(c, cudaMemset(d0, 0, sizeof(float)));
    const float zero = 0.0f;
    const int want = c->layout_req; /* the start-up volume is f32: never paired */
    c->layout_req = VR_LAYOUT_AUTO;
    int rc = install_volume(c, d0, VR_F32, 1, 1, 1, &zero, &zero);
    c->layout_req = want;
    cudaFree(d0);
    if (rc != VR_OK) return rc;
    c->nlut = 1;
    {   /* the sRGB decode of the LUT image (offscreen_pass.cpp:1076) as a table, made with the oracle's arithmetic */
        float table[256];
        for (unsigned i = 0; i < 256; i++) table[i] = srgb8_to_linear(i);
        VR_CUDA(c, vr_lut_upload_srgb_table(table));
    }
    rc = reserve_lut(c, 1);
    if (rc != VR_OK) return rc;
    const uint32_t white = 0xFFFFFFFFu;
    VR_CUDA(c, cudaMemcpy(c->d_lut_rgba8, &white, sizeof(white), cudaMemcpyHostToDevice));
    rc = upload_lut(c);
    if (rc != VR_OK) return rc;
    vr_camera *cam = vr_camera_new();
    if (!cam) return fail(c, VR_ERR_NOMEM, "out of host memory");
    vr_camera_get_view(cam, c->view_in);
    vr_camera_get_position(cam, c->cam_in);
    vr_camera_free(cam);
    vr_projection(c->W, c->H, c->proj);
    apply_extent(c);
    return VR_OK;
}
}  // namespace

/* =============================================================== lifetime == */

extern "C" int vr_abi_version(void) { return VR_ABI_VERSION; }
extern "C" const char *vr_kernel_version(void) { return VR_KERNEL_VERSION; }

extern "C" const char *vr_last_error(const vr_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

extern "C" int vr_create(int device, uint32_t width, uint32_t height, vr_ctx **out)
{
    if (!out || width == 0 || height == 0) return fail(nullptr, VR_ERR_INVALID, "vr_create: bad arguments");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, VR_ERR_CUDA,
                    std::string("vr_create: no CUDA device (this backend has no CPU fallback): ") +
                        (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
    if (device < 0 || device >= ndev) return fail(nullptr, VR_ERR_INVALID, "vr_create: device index out of range");
    vr_ctx *c = new (std::nothrow) vr_ctx;
    if (!c) return fail(nullptr, VR_ERR_NOMEM, "out of host memory");
    c->device = device;
    c->opt.sampler = VR_SAMPLER_HW;
    c->opt.empty_skip = 1;
    c->opt.early_term = 0;
    c->opt.early_term_eps = 1.0f / 1024.0f;
    c->opt.collect_stats = 0;
    c->opt.write_raw = 0;
    if (const char *e = std::getenv("VR_BATCH")) c->tune_batch = c->tune_batch_ess = std::atoi(e);
    if (const char *e = std::getenv("VR_BATCH_ESS")) c->tune_batch_ess = std::atoi(e);
    if (const char *e = std::getenv("VR_ESS_GUARD")) c->tune_ess_guard = std::max(1, std::atoi(e));
    if (const char *e = std::getenv("VR_ESS_MIN_EMPTY")) c->tune_ess_min_empty = (float)std::atof(e);
    if (const char *e = std::getenv("VR_CTA_THREADS")) c->tune_threads = std::atoi(e) == 256 ? 256 : 128;
    if (const char *e = std::getenv("VR_BBOX")) c->tune_bbox = std::atoi(e);
    if (const char *e = std::getenv("VR_DIRTY_CLEAR")) c->tune_dirty_clear = std::atoi(e);
    if (const char *e = std::getenv("VR_PAIR_LOCAL")) c->tune_pair_local = std::atoi(e);
    if (const char *e = std::getenv("VR_SL_REGION")) c->tune_sl_region = std::atoi(e);
    if (const char *e = std::getenv("VR_TILE_MASK")) c->tune_tile_mask = std::atoi(e);
    if (const char *e = std::getenv("VR_WARP_W")) {
        const int w = std::atoi(e);
        if (w == 4 || w == 8 || w == 16) c->tune_warp_w = w;
    }
    if (const char *e = std::getenv("VR_CARVEOUT")) c->tune_carveout = std::atoi(e);
    if (const char *e = std::getenv("VR_LAYOUT")) c->layout_req = std::atoi(e) == 1 ? VR_LAYOUT_ARRAY3D : (std::atoi(e) == 2 ? VR_LAYOUT_PAIRED : VR_LAYOUT_AUTO);
    if (const char *e = std::getenv("VR_PAIR_POLICY")) c->tune_pair_policy = std::atoi(e);
    if (const char *e = std::getenv("VR_PAIR_SPLIT")) c->tune_pair_split = std::atoi(e);
    if (const char *e = std::getenv("VR_OCC_MAX_AXIS")) c->tune_occ_max_axis = std::min(std::max(std::atoi(e), 8), 256);
    if (const char *e = std::getenv("VR_PAIR_AXIS")) c->tune_pair_axis = std::atoi(e) >= 0 && std::atoi(e) <= 2 ? std::atoi(e) : -1;
    if (const char *e = std::getenv("VR_PAIR_MAX_MB")) c->tune_pair_max_bytes = (size_t)std::atoll(e) << 20;
    if (const char *e = std::getenv("VR_L2_GRAN")) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)std::atoi(e));
    int rc = VR_OK;
    do {
        if ((e = cudaSetDevice(device)) != cudaSuccess) { rc = cuda_fail(c, e, "cudaSetDevice"); break; }
        if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) { rc = cuda_fail(c, e, "cudaStreamCreate"); break; }
        if ((e = cudaMalloc(&c->d_counters, 8 * sizeof(unsigned long long))) != cudaSuccess) { rc = cuda_fail(c, e, "cudaMalloc"); break; }
        if ((e = cudaMallocHost(&c->h_counters, 8 * sizeof(unsigned long long))) != cudaSuccess) { rc = cuda_fail(c, e, "cudaMallocHost"); break; }
        std::memset(c->h_counters, 0, 8 * sizeof(unsigned long long));
        if ((e = cudaMalloc(&c->d_flags, VR_FLAG_COUNT * sizeof(uint32_t))) != cudaSuccess) { rc = cuda_fail(c, e, "cudaMalloc"); break; }
        if ((e = cudaMemset(c->d_flags, 0, VR_FLAG_COUNT * sizeof(uint32_t))) != cudaSuccess) { rc = cuda_fail(c, e, "cudaMemset"); break; }
        if ((rc = alloc_target(c, width, height)) != VR_OK) break;
        if ((rc = default_state(c)) != VR_OK) break;
    } while (0);
    if (rc != VR_OK) {
        g_create_error = c->err;
        vr_destroy(c);
        return rc;
    }
    c->last_stream = c->stream;
    *out = c;
    return VR_OK;
}

extern "C" int vr_destroy(vr_ctx *c)
{
    if (!c) return VR_OK;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream); /* vkDeviceWaitIdle, offscreen_pass.cpp:138 */
    release_volume(c);
    if (c->d_lut_pad) cudaFree(c->d_lut_pad);
    if (c->d_lut_dif) cudaFree(c->d_lut_dif);
    if (c->d_lut_h) cudaFree(c->d_lut_h);
    if (c->d_lut_nz) cudaFree(c->d_lut_nz);
    if (c->preint_tex) cudaDestroyTextureObject(c->preint_tex);
    if (c->preint_array) cudaFreeArray(c->preint_array);
    if (c->d_preint) cudaFree(c->d_preint);
    if (c->d_preint_h) cudaFree(c->d_preint_h);
    if (c->d_lut_rgba8) cudaFree(c->d_lut_rgba8);
    if (c->d_lut_loc) cudaFree(c->d_lut_loc);
    if (c->d_markers) cudaFree(c->d_markers);
    if (c->d_image) cudaFree(c->d_image);
    if (c->d_raw) cudaFree(c->d_raw);
    for (int k = 1; k < VR_MAX_FRAMES_IN_FLIGHT; k++) {
        if (c->d_image_slot[k]) cudaFree(c->d_image_slot[k]);
        if (c->d_raw_slot[k]) cudaFree(c->d_raw_slot[k]);
    }
    for (int k = 0; k < VR_MAX_FRAMES_IN_FLIGHT; k++)
        if (c->d_tile_mask[k]) cudaFree(c->d_tile_mask[k]);
    if (c->d_counters) cudaFree(c->d_counters);
    if (c->d_flags) cudaFree(c->d_flags);
    if (c->h_counters) cudaFreeHost(c->h_counters);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return VR_OK;
}

/* =============================================================== mutators == */

extern "C" int vr_set_camera(vr_ctx *c, const float view[16], const float proj[16], const float camera_position[3])
{
    if (!c || !view || !proj || !camera_position) return fail(c, VR_ERR_INVALID, "vr_set_camera: null argument");
    std::memcpy(c->view_in, view, sizeof(c->view_in));
    std::memcpy(c->proj, proj, sizeof(c->proj));
    std::memcpy(c->cam_in, camera_position, sizeof(c->cam_in));
    if (!apply_extent(c)) return fail(c, VR_ERR_INVALID, "vr_set_camera: proj * view is singular");
    return VR_OK;
}

extern "C" int vr_set_volume_spacing(vr_ctx *c, const float spacing[3])
{
    if (!c) return VR_ERR_INVALID;
    if (spacing)
        for (int i = 0; i < 3; i++)
            if (!(spacing[i] > 0.0f) || !(spacing[i] <= 3.0e38f)) return fail(c, VR_ERR_INVALID, "vr_set_volume_spacing: spacings must be positive and finite");
    c->honour_spacing = spacing != nullptr;
    for (int i = 0; i < 3; i++) c->spacing[i] = spacing ? spacing[i] : 1.0f;
    for (int i = 0; i < VR_MAX_FRAMES_IN_FLIGHT; i++) c->slot_clear[i].valid = false;
    if (!apply_extent(c)) return fail(c, VR_ERR_INVALID, "vr_set_volume_spacing: proj * view is singular");
    return VR_OK;
}

extern "C" int vr_get_volume_spacing(const vr_ctx *c, float file_spacing[3], float extent[3])
{
    if (!c) return VR_ERR_INVALID;
    for (int i = 0; i < 3; i++) {
        if (file_spacing) file_spacing[i] = c->file_spacing[i];
        if (extent) extent[i] = c->ext[i];
    }
    return VR_OK;
}

extern "C" int vr_set_camera_orbit(vr_ctx *c, const vr_camera *cam)
{
    if (!c || !cam) return fail(c, VR_ERR_INVALID, "vr_set_camera_orbit: null argument");
    float view[16], proj[16], pos[3];
    vr_camera_get_view(cam, view);
    vr_projection(c->W, c->H, proj);
    vr_camera_get_position(cam, pos);
    return vr_set_camera(c, view, proj, pos);
}

extern "C" int vr_resize(vr_ctx *c, uint32_t width, uint32_t height)
{
    if (!c) return VR_ERR_INVALID;
    if (width == 0 || height == 0) return VR_OK; /* offscreen_pass.cpp:237-239 */
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize()); /* vkDeviceWaitIdle, :241 */
    for (int i = 0; i < VR_MAX_FRAMES_IN_FLIGHT; i++) c->slot_bbox_valid[i] = false, c->slot_clear[i].valid = false; /* new targets: nothing rendered yet */
    for (auto &t : c->host_targets) t.ptr = nullptr;
    return alloc_target(c, width, height);
}

extern "C" int vr_set_volume_device(vr_ctx *c, const void *dev, int dtype, uint32_t nx, uint32_t ny, uint32_t nz,
                                    const float *mn, const float *mx)
{
    if (!c || !dev || dtype < VR_U8 || dtype > VR_F32 || nx == 0 || ny == 0 || nz == 0)
        return fail(c, VR_ERR_INVALID, "vr_set_volume: bad arguments");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize());
    return install_volume(c, dev, dtype, nx, ny, nz, mn, mx);
}

extern "C" int vr_set_volume(vr_ctx *c, const void *data, int dtype, uint32_t nx, uint32_t ny, uint32_t nz,
                             const float *mn, const float *mx)
{
    if (!c || !data || dtype < VR_U8 || dtype > VR_F32 || nx == 0 || ny == 0 || nz == 0)
        return fail(c, VR_ERR_INVALID, "vr_set_volume: bad arguments");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize());
    const size_t bytes = (size_t)nx * ny * nz * dtype_size(dtype);
    void *stage = nullptr;
    VR_CUDA(c, cudaMalloc(&stage, bytes));
    cudaError_t e = cudaMemcpy(stage, data, bytes, cudaMemcpyHostToDevice); /* staging copy, offscreen_pass.cpp:955-959 */
    if (e != cudaSuccess) {
        cudaFree(stage);
        return cuda_fail(c, e, "cudaMemcpy(volume)");
    }
    const int rc = install_volume(c, stage, dtype, nx, ny, nz, mn, mx);
    cudaFree(stage);
    return rc;
}

/* Loader fast path: a raw payload in a type the device keeps natively is mapped (page cache) and handed to
 * the driver's host-to-device copy as it lies in the file: one pass over the bytes.  The three serial O(N)
 * host passes of the reference (widen to float, min, max: nrrd_file_parser.cpp:35-41,:67-77) do not exist
 * here at all -- min/max, cell ranges and the layouts are built on the device by K0.  (Measured against
 * pread() into pinned double buffers overlapped with cudaMemcpyAsync: that is 3x slower, the kernel's
 * copy out of the page cache runs at 3 GB/s, the driver's staging copy at 9 GB/s.) */
static int stream_payload_to_device(vr_ctx *c, const VrNrrdPayload &pl, void *stage)
{
    const int fd = ::open(pl.path.c_str(), O_RDONLY);
    if (fd < 0) return fail(c, VR_ERR_IO, "Failed to read file: " + pl.path);
    const uint64_t page = (uint64_t)sysconf(_SC_PAGESIZE), base = pl.offset / page * page, lead = pl.offset - base;
    void *map = ::mmap(nullptr, (size_t)(pl.bytes + lead), PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, (off_t)base);
    ::close(fd);
    if (map == MAP_FAILED) return fail(c, VR_ERR_IO, "Failed to read file: cannot map " + pl.path);
    ::madvise(map, (size_t)(pl.bytes + lead), MADV_SEQUENTIAL);
    int rc = VR_OK;
    const size_t chunk = (size_t)256 << 20;
    for (size_t done = 0; done < pl.bytes && rc == VR_OK; done += chunk) {
        const size_t n = std::min(chunk, (size_t)pl.bytes - done);
        const cudaError_t e = cudaMemcpyAsync((char *)stage + done, (const char *)map + lead + done, n, cudaMemcpyHostToDevice, c->stream);
        if (e != cudaSuccess) rc = cuda_fail(c, e, "cudaMemcpy(volume)");
    }
    const cudaError_t es = cudaStreamSynchronize(c->stream);
    if (rc == VR_OK && es != cudaSuccess) rc = cuda_fail(c, es, "cudaStreamSynchronize");
    ::munmap(map, (size_t)(pl.bytes + lead));
    return rc;
}

extern "C" int vr_load_nrrd(vr_ctx *c, const char *path)
{
    if (!c || !path) return fail(c, VR_ERR_INVALID, "vr_load_nrrd: null argument");
    VrNrrd n;
    VrNrrdPayload pl;
    std::string err;
    const int rc = vr_nrrd_read(path, n, err, &pl);
    if (rc != VR_OK) return fail(c, rc, err);
    for (int i = 0; i < 3; i++) c->file_spacing[i] = (float)n.spacing[i]; /* ignored like the reference does, unless vr_set_volume_spacing */
    if (!pl.direct) return vr_set_volume(c, n.bytes.data(), n.dtype, n.dims[0], n.dims[1], n.dims[2], nullptr, nullptr);
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize());
    void *stage = nullptr;
    VR_CUDA(c, cudaMalloc(&stage, pl.bytes));
    int rc2 = stream_payload_to_device(c, pl, stage);
    if (rc2 == VR_OK) rc2 = install_volume(c, stage, n.dtype, n.dims[0], n.dims[1], n.dims[2], nullptr, nullptr);
    cudaFree(stage);
    return rc2;
}

extern "C" int vr_get_volume_info(const vr_ctx *c, int *dtype, uint32_t dims[3], float *mn, float *mx)
{
    if (!c) return VR_ERR_INVALID;
    if (dtype) *dtype = c->dtype;
    if (dims) dims[0] = c->nx, dims[1] = c->ny, dims[2] = c->nz;
    if (mn) *mn = c->min_density;
    if (mx) *mx = c->max_density;
    return VR_OK;
}

extern "C" int vr_set_density_range(vr_ctx *c, float mn, float mx)
{
    if (!c) return VR_ERR_INVALID;
    c->min_density = mn, c->max_density = mx;
    c->occ_dirty = true;
    return VR_OK;
}

extern "C" int vr_set_slicing(vr_ctx *c, const float mn[3], const float mx[3])
{
    if (!c || !mn || !mx) return fail(c, VR_ERR_INVALID, "vr_set_slicing: null argument");
    std::memcpy(c->min_slice, mn, sizeof(c->min_slice)); /* offscreen_pass.cpp:271-277 */
    std::memcpy(c->max_slice, mx, sizeof(c->max_slice));
    return VR_OK;
}

extern "C" int vr_set_transfer_function(vr_ctx *c, const uint32_t *rgba8, uint32_t n)
{
    if (!c || !rgba8 || n == 0) return fail(c, VR_ERR_INVALID, "vr_set_transfer_function: need n >= 1 texels");
    if (n > 4096) return fail(c, VR_ERR_INVALID, "vr_set_transfer_function: at most 4096 texels (shared-memory staged)");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize()); /* wait_till_idle, offscreen_pass.cpp:282 */
    const int rc = reserve_lut(c, n);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaMemcpy(c->d_lut_rgba8, rgba8, sizeof(uint32_t) * n, cudaMemcpyHostToDevice));
    c->nlut = n;
    return upload_lut(c);
}

extern "C" int vr_set_transfer_function_device(vr_ctx *c, const void *dev_rgba8, uint32_t n)
{
    if (!c || !dev_rgba8 || n == 0) return fail(c, VR_ERR_INVALID, "vr_set_transfer_function_device: need n >= 1 texels");
    if (n > 4096) return fail(c, VR_ERR_INVALID, "vr_set_transfer_function_device: at most 4096 texels (shared-memory staged)");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize());
    const int rc = reserve_lut(c, n);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaMemcpy(c->d_lut_rgba8, dev_rgba8, sizeof(uint32_t) * n, cudaMemcpyDeviceToDevice));
    c->nlut = n;
    return upload_lut(c);
}

extern "C" int vr_set_gradient(vr_ctx *c, const float *color_markers, uint32_t nc, const float *alpha_markers, uint32_t na, uint32_t count)
{
    if (!c || !color_markers || !alpha_markers || nc == 0 || na == 0 || count == 0)
        return fail(c, VR_ERR_INVALID, "vr_set_gradient: need markers and count >= 1");
    if (count > 4096) return fail(c, VR_ERR_INVALID, "vr_set_gradient: at most 4096 texels (shared-memory staged)");
    if (nc > 4096 || na > 4096) return fail(c, VR_ERR_INVALID, "vr_set_gradient: at most 4096 markers of a kind");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize());
    int rc = reserve_lut(c, count);
    if (rc != VR_OK) return rc;
    const size_t need = (size_t)4 * nc + (size_t)2 * na;
    if (c->markers_cap < need) {
        if (c->d_markers) cudaFree(c->d_markers), c->d_markers = nullptr;
        c->markers_cap = 0;
        VR_CUDA(c, cudaMalloc(&c->d_markers, sizeof(float) * need));
        c->markers_cap = need;
    }
    VR_CUDA(c, cudaMemcpy(c->d_markers, color_markers, sizeof(float) * 4 * nc, cudaMemcpyHostToDevice));
    VR_CUDA(c, cudaMemcpy(c->d_markers + 4 * nc, alpha_markers, sizeof(float) * 2 * na, cudaMemcpyHostToDevice));
    VR_CUDA(c, vr_lut_gradient(c->d_markers, nc, na, count, c->d_lut_loc, c->d_lut_rgba8, c->stream));
    c->nlut = count;
    return upload_lut(c);
}

extern "C" int vr_get_transfer_function(vr_ctx *c, uint32_t *rgba8, uint32_t capacity, uint32_t *n)
{
    if (!c) return VR_ERR_INVALID;
    if (n) *n = c->nlut;
    if (!rgba8) return VR_OK;
    if (capacity < c->nlut) return fail(c, VR_ERR_INVALID, "vr_get_transfer_function: capacity below the texel count");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaMemcpy(rgba8, c->d_lut_rgba8, sizeof(uint32_t) * c->nlut, cudaMemcpyDeviceToHost));
    return VR_OK;
}

namespace
{
int prepare_preint(vr_ctx *c, cudaStream_t s);
}

extern "C" int vr_set_preintegration(vr_ctx *c, uint32_t substeps)
{
    if (!c || substeps > 64) return fail(c, VR_ERR_INVALID, "vr_set_preintegration: 0 (off) or 1..64 sub-samples per table entry");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize());
    if (substeps != c->preint_substeps) c->preint_dirty = true;
    c->preint_substeps = substeps;
    return VR_OK;
}

extern "C" int vr_get_preintegration_table(vr_ctx *c, float *table, uint32_t capacity_texels, uint32_t *n)
{
    if (!c) return VR_ERR_INVALID;
    if (!c->preint_substeps) return fail(c, VR_ERR_STATE, "vr_get_preintegration_table: pre-integration is off");
    if (c->nlut > VR_PREINT_MAX_TEXELS) return fail(c, VR_ERR_STATE, "vr_get_preintegration_table: at most 1024 texels");
    if (n) *n = c->nlut;
    if (!table) return VR_OK;
    if ((uint64_t)capacity_texels < (uint64_t)c->nlut * c->nlut) return fail(c, VR_ERR_INVALID, "vr_get_preintegration_table: capacity below n * n entries");
    VR_CUDA(c, cudaSetDevice(c->device));
    const int rc = prepare_preint(c, c->stream);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaMemcpy(table, c->d_preint, sizeof(float4) * (size_t)c->nlut * c->nlut, cudaMemcpyDeviceToHost));
    return VR_OK;
}

extern "C" int vr_set_march(vr_ctx *c, float ray_dist, float step_size)
{
    if (!c || !(step_size > 0.0f) || !(ray_dist >= 0.0f)) return fail(c, VR_ERR_INVALID, "vr_set_march: bad arguments");
    c->ray_dist = ray_dist, c->step_size = step_size;
    return VR_OK;
}

extern "C" int vr_set_volume_layout(vr_ctx *c, int layout)
{
    if (!c || layout < VR_LAYOUT_AUTO || layout > VR_LAYOUT_PAIRED) return fail(c, VR_ERR_INVALID, "vr_set_volume_layout: bad layout");
    c->layout_req = layout;
    return VR_OK;
}

extern "C" int vr_get_volume_layout(const vr_ctx *c, int *layout, int *last_frame_paired)
{
    if (!c || !layout) return VR_ERR_INVALID;
    *layout = c->tex_pair[2] ? VR_LAYOUT_PAIRED : VR_LAYOUT_ARRAY3D;
    if (last_frame_paired) *last_frame_paired = c->last_pair ? 1 + c->last_pair_axis : 0;
    return VR_OK;
}

extern "C" int vr_get_options(const vr_ctx *c, vr_options *out)
{
    if (!c || !out) return VR_ERR_INVALID;
    *out = c->opt;
    return VR_OK;
}

extern "C" int vr_set_options(vr_ctx *c, const vr_options *o)
{
    if (!c || !o || (o->sampler != VR_SAMPLER_HW && o->sampler != VR_SAMPLER_EXACT) || !(o->early_term_eps >= 0.0f))
        return fail(c, VR_ERR_INVALID, "vr_set_options: bad arguments");
    VR_CUDA(c, cudaSetDevice(c->device));
    const bool need_raw = o->write_raw && !c->d_raw;
    c->opt = *o;
    if (need_raw) {
        VR_CUDA(c, cudaStreamSynchronize(c->last_stream));
        VR_CUDA(c, cudaMalloc(&c->d_raw, (size_t)c->W * c->H * sizeof(float4)));
        VR_CUDA(c, cudaMemset(c->d_raw, 0, (size_t)c->W * c->H * sizeof(float4)));
    }
    return VR_OK;
}

/* ================================================================ render == */

namespace
{
int prepare_occupancy(vr_ctx *c, cudaStream_t s)
{
    if (!c->occ_dirty) return VR_OK;
    VR_CUDA(c, vr_k0_classify(c->d_cell_range, c->occ_n, c->d_lut_nz, c->nlut, c->min_density, c->max_density, c->d_occ,
                              c->occ_words, c->d_counters + 4, s));
    VR_CUDA(c, vr_k0_distance(c->d_occ, c->occ_n, c->d_dist, c->d_sat, s));
    VR_CUDA(c, vr_k0_list_occupied(c->d_occ, c->occ_n, c->d_occ_list, c->d_occ_count, s));
    VR_CUDA(c, cudaMemcpyAsync(&c->h_counters[4], c->d_counters + 4, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    VR_CUDA(c, cudaStreamSynchronize(s));
    c->occ_empty_cells = c->h_counters[4];
    c->occ_dirty = false;
    return VR_OK;
}

/* (re)build the pre-integration table when the transfer function or the sub-sample count changed */
int prepare_preint(vr_ctx *c, cudaStream_t s)
{
    if (!c->preint_dirty && c->preint_tex && c->preint_n == c->nlut) return VR_OK;
    const uint32_t n = c->nlut;
    if (c->preint_n != n || !c->preint_tex) {
        VR_CUDA(c, cudaDeviceSynchronize());
        if (c->preint_tex) cudaDestroyTextureObject(c->preint_tex), c->preint_tex = 0;
        if (c->preint_array) cudaFreeArray(c->preint_array), c->preint_array = nullptr;
        if (c->d_preint) cudaFree(c->d_preint), c->d_preint = nullptr;
        if (c->d_preint_h) cudaFree(c->d_preint_h), c->d_preint_h = nullptr;
        c->preint_n = 0;
        VR_CUDA(c, cudaMalloc(&c->d_preint, sizeof(float4) * (size_t)n * n));
        VR_CUDA(c, cudaMalloc(&c->d_preint_h, sizeof(uint2) * (size_t)n * n));
        const cudaChannelFormatDesc fmt = cudaCreateChannelDescHalf4();
        VR_CUDA(c, cudaMallocArray(&c->preint_array, &fmt, n, n));
        cudaResourceDesc rd{};
        rd.resType = cudaResourceTypeArray, rd.res.array.array = c->preint_array;
        cudaTextureDesc td{};
        td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModeLinear, td.readMode = cudaReadModeElementType, td.normalizedCoords = 0;
        VR_CUDA(c, cudaCreateTextureObject(&c->preint_tex, &rd, &td, nullptr));
        c->preint_n = n;
    }
    VR_CUDA(c, vr_lut_preint_table(c->d_lut_pad, n, c->preint_substeps, c->d_preint, c->d_preint_h, s));
    VR_CUDA(c, cudaMemcpy2DToArrayAsync(c->preint_array, 0, 0, c->d_preint_h, sizeof(uint2) * n, sizeof(uint2) * n, n, cudaMemcpyDeviceToDevice, s));
    /* frames in flight on other streams may still sample the old table: like a transfer-function change, this waits */
    VR_CUDA(c, cudaStreamSynchronize(s));
    c->preint_dirty = false;
    return VR_OK;
}

int ensure_slot(vr_ctx *c, uint32_t frame_index)
{
    if (frame_index == 0 || c->d_image_slot[frame_index]) return VR_OK;
    VR_CUDA(c, cudaMalloc(&c->d_image_slot[frame_index], (size_t)c->W * c->H * sizeof(uchar4)));
    VR_CUDA(c, cudaMemset(c->d_image_slot[frame_index], 0, (size_t)c->W * c->H * sizeof(uchar4)));
    return VR_OK;
}

int ensure_raw_slot(vr_ctx *c, uint32_t frame_index)
{
    if (frame_index == 0 || c->d_raw_slot[frame_index]) return VR_OK;
    VR_CUDA(c, cudaMalloc(&c->d_raw_slot[frame_index], (size_t)c->W * c->H * sizeof(float4)));
    VR_CUDA(c, cudaMemset(c->d_raw_slot[frame_index], 0, (size_t)c->W * c->H * sizeof(float4)));
    return VR_OK;
}

/* Pixels the cube can cover: the projections of its 8 corners bound its silhouette when all of
 * them are in front of the eye; otherwise (camera inside or near the cube) the whole target. */
void screen_bbox(const vr_ctx *c, uint32_t bbox[4], const double *box_lo = nullptr, const double *box_hi = nullptr)
{
    bbox[0] = 0, bbox[1] = 0, bbox[2] = c->W, bbox[3] = c->H;
    if (!c->tune_bbox) return;
    const double unit_lo[3] = {-0.5, -0.5, -0.5}, unit_hi[3] = {0.5, 0.5, 0.5};
    if (!box_lo) box_lo = unit_lo;
    if (!box_hi) box_hi = unit_hi;
    double v[16], p[16], pv[16];
    for (int i = 0; i < 16; i++) v[i] = (double)c->view[i], p[i] = (double)c->proj[i];
    mat4_mul_d(p, v, pv);
    double x0 = 1e30, y0 = 1e30, x1 = -1e30, y1 = -1e30;
    for (int k = 0; k < 8; k++) {
        const double q[3] = {(k & 1) ? box_hi[0] : box_lo[0], (k & 2) ? box_hi[1] : box_lo[1], (k & 4) ? box_hi[2] : box_lo[2]};
        double h[4];
        for (int r = 0; r < 4; r++) h[r] = pv[r] * q[0] + pv[4 + r] * q[1] + pv[8 + r] * q[2] + pv[12 + r];
        if (!(h[3] > 1e-4)) return;
        const double px = (h[0] / h[3] + 1.0) * 0.5 * c->W, py = (h[1] / h[3] + 1.0) * 0.5 * c->H;
        x0 = std::min(x0, px), x1 = std::max(x1, px), y0 = std::min(y0, py), y1 = std::max(y1, py);
    }
    if (!(x0 == x0) || !(x1 == x1) || !(y0 == y0) || !(y1 == y1)) return;
    const double m = 2.0; /* pixels of slack for the fp32 coverage test */
    const double lx = std::floor(x0 - m), ly = std::floor(y0 - m), hx = std::ceil(x1 + m), hy = std::ceil(y1 + m);
    bbox[0] = (uint32_t)std::min(std::max(lx, 0.0), (double)c->W);
    bbox[1] = (uint32_t)std::min(std::max(ly, 0.0), (double)c->H);
    bbox[2] = (uint32_t)std::min(std::max(hx, 0.0), (double)c->W);
    bbox[3] = (uint32_t)std::min(std::max(hy, 0.0), (double)c->H);
}

/* Pixels the rays through a sub-box of the grid can cover (sort-last: what a rank marches and stores; outside it
 * the rank's partial is the identity).  The box is the half-open ownership region widened by half a voxel: a sample
 * owned by the rank lies inside it. */
void subbox_screen_rect(const vr_ctx *c, const uint32_t full[3], const uint32_t org[3], const uint32_t dims[3], uint32_t rect[4])
{
    double lo[3], hi[3];
    for (int i = 0; i < 3; i++) {
        lo[i] = std::max(((double)org[i] - 0.5) / (double)full[i], 0.0) - 0.5;
        hi[i] = std::min(((double)org[i] + (double)dims[i] + 0.5) / (double)full[i], 1.0) - 0.5;
    }
    screen_bbox(c, rect, lo, hi);
}

template <typename T>
void permute3(T *v, const int perm[3])
{
    const T t[3] = {v[0], v[1], v[2]};
    for (int j = 0; j < 3; j++) v[j] = t[perm[j]];
}

/* pair_axis < 0: the frame samples the 3-D array; else the model axis the pairs run along.
 * clear: pixel rectangle (x0, y0, x1, y1) of the target that may hold something else than the clear colour and must
 * be restored outside this frame's tiles; nullptr = the whole target. */
void fill_params(const vr_ctx *c, MarchParams &p, int pair_axis, const uint32_t *clear = nullptr)
{
    std::memset(&p, 0, sizeof(p));
    for (int r = 0; r < 4; r++) p.c0[r] = c->rc[0][r], p.c1[r] = c->rc[1][r], p.c2[r] = c->rc[2][r], p.c3[r] = c->rc[3][r];
    p.sx = 2.0f / (float)c->W, p.sy = 2.0f / (float)c->H;
    for (int i = 0; i < 3; i++) p.cam[i] = c->cam_pos[i], p.min_slice[i] = c->min_slice[i], p.max_slice[i] = c->max_slice[i];
    for (int i = 0; i < 3; i++) p.ext[i] = c->ext[i], p.inv_ext[i] = 1.0f / c->ext[i];
    p.W = c->W, p.H = c->H;
    p.min_density = c->min_density, p.max_density = c->max_density;
    p.range = c->max_density - c->min_density;
    p.steps = (int)(c->ray_dist / c->step_size); /* volume.frag:31, fp32 divide */
    p.step_size = c->step_size;
    p.tex_lin = c->tex_lin, p.tex_pt = c->tex_pt;
    const uint32_t ldims[3] = {c->nx, c->ny, c->nz};
    const float inf = std::numeric_limits<float>::infinity();
    for (int i = 0; i < 3; i++) {
        const uint32_t full = c->sort_last ? c->full[i] : ldims[i];
        p.dims[i] = (float)full;
        p.org[i] = c->sort_last ? c->tex_org[i] : 0;
        p.forg[i] = (float)p.org[i];
        p.inv_ldims[i] = 1.0f / (float)ldims[i];
        p.own_lo[i] = -inf, p.own_hi[i] = inf;
        if (c->sort_last) {
            /* both neighbours evaluate the shared face with the same integers -> same float */
            if (c->own_org[i] > 0) p.own_lo[i] = (float)c->own_org[i] / (float)full;
            if (c->own_org[i] + c->own_dims[i] < full) p.own_hi[i] = (float)(c->own_org[i] + c->own_dims[i]) / (float)full;
        }
        p.box_lo[i] = std::max(0.0f, std::max(c->min_slice[i], p.own_lo[i]));
        p.box_hi[i] = std::min(1.0f, std::min(c->max_slice[i], p.own_hi[i]));
        if (!(c->min_slice[i] == c->min_slice[i]) || !(c->max_slice[i] == c->max_slice[i]))
            p.box_lo[i] = inf, p.box_hi[i] = -inf; /* NaN slice bound: every strict test fails */
        p.occ_n[i] = c->occ_n[i];
        p.occ_f_hi[i] = (float)ldims[i] / (float)c->occ_cell;
        p.occ_scale[i] = (float)full / (float)c->occ_cell;
        p.occ_off[i] = (float)p.org[i] / (float)c->occ_cell;
        p.occ_offh[i] = p.occ_off[i] + 0.5f;
    }
    p.dtype = c->dtype;
    p.dscale = c->dtype == VR_U8 ? 255.0f : (c->dtype == VR_U16 ? 65535.0f : 1.0f);
    p.lut_pad = c->d_lut_pad, p.lut_dif = c->d_lut_dif, p.lut_h = c->d_lut_h, p.nlut = c->nlut;
    p.occ_stride[0] = 1, p.occ_stride[1] = c->occ_n[0], p.occ_stride[2] = c->occ_n[0] * c->occ_n[1];
    for (int i = 0; i < 3; i++) p.occ_oct[i] = (uint32_t)(c->occ_n[0] * c->occ_n[1] * c->occ_n[2]) << i;
    p.pair_axis = 2;
    p.tex_pair = 0, p.tex_pair_hi = 0, p.pair_split = 0x10000;
    if (pair_axis >= 0) {
        /* sampling space (u, v, w) = model axes perm[0..2], w = the pair axis: every per-axis field is
         * stored in that order; the kernel permutes the ray once (march.cu) */
        const int perm[3] = {pair_axis == 0 ? 1 : 0, pair_axis == 2 ? 1 : 2, pair_axis};
        p.pair_axis = pair_axis;
        p.tex_pair = c->tex_pair[pair_axis];
        p.tex_pair_hi = c->tex_pair_hi[pair_axis] ? c->tex_pair_hi[pair_axis] : c->tex_pair[pair_axis];
        p.pair_split = c->pair_split[pair_axis];
        permute3(p.min_slice, perm), permute3(p.max_slice, perm);
        permute3(p.dims, perm), permute3(p.org, perm), permute3(p.forg, perm), permute3(p.inv_ldims, perm);
        permute3(p.own_lo, perm), permute3(p.own_hi, perm), permute3(p.box_lo, perm), permute3(p.box_hi, perm);
        permute3(p.occ_n, perm), permute3(p.occ_stride, perm), permute3(p.occ_scale, perm), permute3(p.occ_off, perm);
        permute3(p.occ_f_hi, perm), permute3(p.occ_oct, perm), permute3(p.occ_offh, perm);
    }
    for (int i = 0; i < 2; i++) {
        p.pair_a[i] = p.dims[i] * p.inv_ldims[i];
        p.pair_b[i] = -p.forg[i] * p.inv_ldims[i];
    }
    p.pair_a[2] = p.dims[2];
    p.pair_b[2] = 12582912.0f - p.forg[2]; /* 1.5 * 2^23: the FMA rounds the local z texel coordinate to an integer = floor(w) + 1 */
    p.pair_zc = 0.5f - p.forg[2];          /* w - floor(w) = q_z * dims_z + (pair_zc - (floor(w) + 1)) */
    const double range = (double)p.range, n = (double)c->nlut;
    p.lutA = (float)((double)p.dscale * n / range);
    p.lutB = (float)(0.5 - (double)c->min_density * n / range);
    p.lutBh = (float)(0.0 - (double)c->min_density * n / range);
    p.occ = c->d_occ, p.occ_words = c->occ_words;
    p.dist = c->d_dist;
    p.out = c->d_image;
    p.raw = c->opt.write_raw ? c->d_raw : nullptr;
    p.counters = c->d_counters;
    p.ert_eps = c->opt.early_term ? c->opt.early_term_eps : -1.0f;
    p.pt_px = c->pt_px, p.pt_rank = c->pt_rank, p.pt_nranks = c->pt_nranks;
    p.pt_nx = (c->W + c->pt_px - 1) / c->pt_px, p.pt_ny = (c->H + c->pt_px - 1) / c->pt_px;
    p.pt_m = vr_sort_first_stride(c->pt_nranks);
    p.warp_w = (uint32_t)c->tune_warp_w;
    p.carveout = c->tune_carveout;
    p.tile_h = (uint32_t)c->tune_threads / VR_TILE;
    if (c->sort_last) subbox_screen_rect(c, c->full, c->own_org, c->own_dims, p.bbox);
    else screen_bbox(c, p.bbox);
    p.ntx = (c->W + VR_TILE - 1) / VR_TILE, p.nty = (c->H + p.tile_h - 1) / p.tile_h;
    uint32_t clr[4] = {0, 0, c->W, c->H};
    if (clear)
        for (int i = 0; i < 4; i++) clr[i] = std::min(clear[i], i & 1 ? c->H : c->W);
    const bool clr_any = clr[2] > clr[0] && clr[3] > clr[1];
    p.clr_org[0] = clr[0] & ~63u, p.clr_org[1] = clr[1] & ~31u;
    p.nclr_x = clr_any ? (clr[2] - p.clr_org[0] + 63) / 64 : 0, p.nclr_y = clr_any ? (clr[3] - p.clr_org[1] + 31) / 32 : 0;
    p.rect[0] = (int)(p.bbox[0] / VR_TILE), p.rect[1] = (int)(p.bbox[1] / p.tile_h);
    p.rect[2] = (int)((p.bbox[2] + VR_TILE - 1) / VR_TILE), p.rect[3] = (int)((p.bbox[3] + p.tile_h - 1) / p.tile_h);
    if (p.bbox[2] <= p.bbox[0] || p.bbox[3] <= p.bbox[1]) {
        p.rect[0] = p.rect[1] = p.rect[2] = p.rect[3] = 0;
        p.sp_count = 0, p.sp_cx = p.sp_cy = 0;
    } else {
        const int w = p.rect[2] - p.rect[0], h = p.rect[3] - p.rect[1];
        const int R = (std::max(w, h) + 1) / 2; /* the (2R+1)^2 spiral square covers the rectangle */
        p.sp_cx = p.rect[0] + w / 2, p.sp_cy = p.rect[1] + h / 2;
        p.sp_count = (uint32_t)((2 * R + 1) * (2 * R + 1));
        /* sort-first: launch this rank's partition tiles only (7/8 of a spiral over the whole box would be blocks
         * that find out they own nothing: 12 % of the frame time at 8 ranks, profiles/r1_exp_c4_balance.txt) */
        if (p.pt_nranks > 1 && p.pt_px % VR_TILE == 0 && p.pt_px % p.tile_h == 0) {
            p.pt_enum = 1;
            p.pt_sub_x = p.pt_px / VR_TILE, p.pt_sub_y = p.pt_px / p.tile_h;
            p.pt_x0 = (uint32_t)p.rect[0] / p.pt_sub_x, p.pt_x1 = ((uint32_t)p.rect[2] + p.pt_sub_x - 1) / p.pt_sub_x;
            p.pt_y0 = (uint32_t)p.rect[1] / p.pt_sub_y;
            const uint32_t rows = ((uint32_t)p.rect[3] + p.pt_sub_y - 1) / p.pt_sub_y - p.pt_y0;
            p.pt_cnt = (p.pt_x1 - p.pt_x0 + p.pt_nranks - 1) / p.pt_nranks;
            p.pt_rows = rows;
            p.sp_count = rows * p.pt_cnt * p.pt_sub_x * p.pt_sub_y;
        }
    }
    if (p.pt_nranks > 1 && p.pt_px % VR_TILE == 0 && p.pt_px % p.tile_h == 0) {
        /* ... and one clear block per partition tile of this rank */
        p.pt_enum = 1;
        p.clr_t[0] = clr[0] / p.pt_px, p.clr_t[1] = clr[1] / p.pt_px;
        p.clr_t[2] = std::min((clr[2] + p.pt_px - 1) / p.pt_px, p.pt_nx), p.clr_t[3] = std::min((clr[3] + p.pt_px - 1) / p.pt_px, p.pt_ny);
        const bool any = clr_any && p.clr_t[2] > p.clr_t[0] && p.clr_t[3] > p.clr_t[1];
        p.pt_cnt_clear = any ? (p.clr_t[2] - p.clr_t[0] + p.pt_nranks - 1) / p.pt_nranks : 0;
        p.nclr_x = any ? p.pt_cnt_clear * (p.clr_t[3] - p.clr_t[1]) : 0, p.nclr_y = any ? 1 : 0;
    }
}

/* the pixel rectangle whose tiles the march blocks of this launch write completely */
void drawn_rect(const vr_ctx *c, const MarchParams &p, uint32_t out[4])
{
    out[0] = std::min((uint32_t)p.rect[0] * VR_TILE, c->W), out[1] = std::min((uint32_t)p.rect[1] * p.tile_h, c->H);
    out[2] = std::min((uint32_t)p.rect[2] * VR_TILE, c->W), out[3] = std::min((uint32_t)p.rect[3] * p.tile_h, c->H);
}
}  // namespace

extern "C" int vr_render(vr_ctx *c, void *cuda_stream) { return vr_render_frame(c, 0, cuda_stream); }

extern "C" int vr_render_frame(vr_ctx *c, uint32_t frame_index, void *cuda_stream)
{
    if (!c) return VR_ERR_INVALID;
    if (frame_index >= VR_MAX_FRAMES_IN_FLIGHT) return fail(c, VR_ERR_INVALID, "vr_render_frame: frame_index out of range");
    if (!c->ray_valid) return fail(c, VR_ERR_STATE, "vr_render: no valid camera");
    if (!c->array) return fail(c, VR_ERR_STATE, "vr_render: no volume");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
    c->last_stream = s;
    const float range = c->max_density - c->min_density;
    const bool nondegenerate = range > 0.0f;
    /* max_density == min_density divides by zero in volume.frag:42; only the exact path
     * reproduces the oracle's definition of that case */
    const bool exact = c->opt.sampler == VR_SAMPLER_EXACT || !nondegenerate;
    bool ess = c->opt.empty_skip && nondegenerate;
    const bool preint = c->preint_substeps != 0;
    if (preint) {
        /* pre-integrated classification (vr_set_preintegration): segments between consecutive samples, so every step is
         * walked (no skipping); hardware sampler, whole volume, no statistics */
        if (exact || c->sort_last || c->opt.collect_stats || c->bench_fetch_only)
            return fail(c, VR_ERR_STATE, "vr_render: pre-integration needs the hardware sampler, a density range > 0, the whole volume (no sort-last "
                                         "partition) and no statistics");
        if (c->nlut > VR_PREINT_MAX_TEXELS) return fail(c, VR_ERR_STATE, "vr_render: pre-integration takes transfer functions of at most 1024 texels");
        ess = false;
        const int rc = prepare_preint(c, s);
        if (rc != VR_OK) return rc;
    }
    if (ess) {
        const int rc = prepare_occupancy(c, s);
        if (rc != VR_OK) return rc;
        /* skipping is result-identical, so whether to consult the grid is purely a cost question:
         * with few empty cells the per-sample lookup costs more than the fetches it saves */
        const double ncells = (double)c->occ_n[0] * c->occ_n[1] * c->occ_n[2];
        ess = (double)c->occ_empty_cells >= (double)c->tune_ess_min_empty * ncells;
    }
    uchar4 *target = c->ext_image ? c->ext_image : c->d_image;
    if (frame_index > 0) {
        if (c->ext_image_slot[frame_index]) {
            target = c->ext_image_slot[frame_index];
        } else {
            const int rc = ensure_slot(c, frame_index);
            if (rc != VR_OK) return rc;
            target = c->d_image_slot[frame_index];
        }
    }
    /* Both layouts are resident; which one this frame samples is a cost question (the images agree within
     * the hardware-sampler tolerance either way).  Small frames (the cube's screen bounding box holds less
     * than one wave of CTAs) are latency-bound and sparse volumes are bound by the skipping: both measured
     * faster on the 3-D array.  Otherwise a two-term estimate per candidate, calibrated on
     * profiles/r1_exp_axis.txt / r1_exp_stream_layouts.txt (the DRAM term is only partly hidden):
     *   t = max(tex, dram) + 0.3 min(tex, dram)
     *   3-D array:      tex = S / 3.6e11,              dram = V / 5e12
     *   pairs along w:  tex = S / 5.7e11 * pen(L_w),   dram = 2 V f_w / 5e12
     * S = samples of the frame (0.37 * bounding-box pixels * steps), V = bytes of the volume (this rank's share),
     * L_w = layers one warp fetch spans = 1 + pitch (ww |r_w| + wh |s_w|) with r, s the model-space directions
     * of screen x, y (columns 0, 1 of inverse(proj * view)) and pitch the voxels per pixel,
     * pen(L) = 1 + 0.2 max(0, L - 4.5), and f_w = min(1, 1 / (step in voxels along w)) the share of the layers
     * a ray touches: looking down w, consecutive samples fall into different layers and only every
     * step-th layer is read from HBM. */
    bool pair = !exact && c->tex_pair[2] != 0;
    int pair_axis = 2;
    const int policy = c->layout_used == VR_LAYOUT_PAIRED ? 1 : c->tune_pair_policy; /* asked for by name: always */
    if (pair && policy == 0) pair = false;
    MarchParams p;
    fill_params(c, p, -1);
    if (pair) {
        const int64_t bw = std::max<int64_t>((int64_t)p.bbox[2] - (int64_t)p.bbox[0], 0), bh = std::max<int64_t>((int64_t)p.bbox[3] - (int64_t)p.bbox[1], 0);
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
        const uint32_t nd[3] = {c->nx, c->ny, c->nz};
        double share = 1.0 / (double)std::max(1u, c->pt_nranks); /* this rank's part of the frame's work */
        if (c->sort_last) share *= ((double)c->own_dims[0] * c->own_dims[1] * c->own_dims[2]) / ((double)c->full[0] * c->full[1] * c->full[2]);
        const double S = 0.37 * (double)(bw * bh) * (double)p.steps * share;
        const double V = (double)nd[0] * nd[1] * nd[2] * (double)dtype_size(c->dtype) / (double)std::max(1u, c->pt_nranks);
        auto combine = [](double a, double b) { return std::max(a, b) + 0.3 * std::min(a, b); };
        const double t3d = combine(S / 3.6e11, V / 5.0e12);
        const float *r = c->rc[0], *u = c->rc[1];
        const double rn = std::sqrt((double)r[0] * r[0] + (double)r[1] * r[1] + (double)r[2] * r[2]);
        const double un = std::sqrt((double)u[0] * u[0] + (double)u[1] * u[1] + (double)u[2] * u[2]);
        const double dn = std::sqrt((double)c->cam_pos[0] * c->cam_pos[0] + (double)c->cam_pos[1] * c->cam_pos[1] + (double)c->cam_pos[2] * c->cam_pos[2]);
        const double ww = (double)c->tune_warp_w, wh = 32.0 / ww;
        double best = 1e30, cost[3] = {1e30, 1e30, 1e30};
        for (int a = 2; a >= 0; a--) { /* ties go to z, then y */
            const double full_a = c->sort_last ? (double)c->full[a] : (double)nd[a];
            const double pitch = 1.2 * full_a / (double)std::max<int64_t>(std::max(bw, bh), 1); /* voxels per pixel along a */
            const double L = 1.0 + pitch * ((rn > 0 ? ww * std::fabs(r[a]) / rn : 0.0) + (un > 0 ? wh * std::fabs(u[a]) / un : 0.0));
            const double step_vox = (double)c->step_size * full_a * (dn > 0 ? std::fabs(c->cam_pos[a]) / dn : 0.0);
            const double f = step_vox > 1.0 ? 1.0 / step_vox : 1.0;
            const double t = combine(S / 5.7e11 * (1.0 + 0.2 * std::max(0.0, L - 4.5)), 2.0 * V * f / 5.0e12) *
                             (1.0 + 1e-6 * L); /* equal cost: the axis with fewer layers per fetch */
            cost[a] = t;
            if (t < best - 1e-12) best = t, pair_axis = a;
        }
        /* Sort-last: what a warp fetches together are the positions entry(pixel) + i d(pixel) of neighbouring pixels at the
         * same step index i.  Rays that entered through the cube face perpendicular to w share their w coordinate at the
         * face exactly and drift apart only by the divergence of their directions along the path inside the cube: behind
         * that face the pairs along w cost what they cost when looking straight down w, whatever the view angle, while
         * the ranks behind the other front face want the other axis.  On the frames between two axes this is a factor of
         * two between the ranks of one frame (profiles/r2_exp_sl_stats.txt, r2_exp_sl_axis_table_n8.txt), and the frame
         * waits for the slowest.  So where the estimate above rates two axes close, a rank takes the axis of the face the
         * central ray of its screen rectangle enters the cube through. */
        if (c->sort_last && c->ray_valid && c->tune_pair_local && bw > 0 && bh > 0) {
            const double px = 0.5 * ((double)p.bbox[0] + (double)p.bbox[2]), py = 0.5 * ((double)p.bbox[1] + (double)p.bbox[3]);
            const double xn = (px + 0.5) * (2.0 / (double)c->W) - 1.0, yn = (py + 0.5) * (2.0 / (double)c->H) - 1.0;
            double nh[4], fh[4], tn = -1e300;
            int face = -1;
            for (int i = 0; i < 4; i++) nh[i] = xn * (double)c->rc[0][i] + yn * (double)c->rc[1][i] + (double)c->rc[3][i], fh[i] = nh[i] + (double)c->rc[2][i];
            for (int i = 0; i < 3; i++) { /* slab test of march.cu pixel_entry: the axis of the entry face */
                const double o = nh[i] / nh[3], d = fh[i] / fh[3] - o;
                if (d != 0.0) {
                    const double lo = std::min((-0.5 - o) / d, (0.5 - o) / d);
                    if (lo > tn) tn = lo, face = i;
                }
            }
            if (face >= 0 && face != pair_axis && cost[face] <= 1.35 * best) pair_axis = face;
        }
        if (c->tune_pair_axis >= 0) pair_axis = c->tune_pair_axis;
        if (policy == 2) pair = !ess && bw * bh >= (int64_t)sms * 10 * 128 && best < t3d;
    }
    /* Dirty clear.  The reference clears the whole attachment every frame (offscreen_pass.cpp:170-173); here a
     * target keeps the clear colour outside the tiles the last frame into it marched, so a frame restores only
     * those of them that lie outside its own tiles -- on an orbit a thin frame of pixels instead of the 83 % of the
     * target outside the box (under sort-first: bytes that would cross NVLink into the display rank). */
    float4 *raw_target = nullptr;
    if (c->opt.write_raw) {
        const int rc = ensure_raw_slot(c, frame_index);
        if (rc != VR_OK) return rc;
        raw_target = frame_index == 0 ? c->d_raw : c->d_raw_slot[frame_index];
    }
    /* a sort-last rank's frame is its float4 partial; the RGBA8 image comes out of the merge (vr_composite*) */
    if (c->sort_last && raw_target) target = nullptr;
    vr_ctx::SlotClear &sc = c->slot_clear[frame_index];
    const bool keep = c->tune_dirty_clear && sc.valid && sc.target == (const void *)target && sc.raw == (const void *)raw_target &&
                      sc.W == c->W && sc.H == c->H && sc.pt_rank == c->pt_rank && sc.pt_nranks == c->pt_nranks && sc.pt_px == c->pt_px;
    fill_params(c, p, pair ? pair_axis : -1, keep ? sc.drawn : nullptr);
    {
        uint32_t now[4];
        drawn_rect(c, p, now);
        /* nothing to restore when the previous frame's tiles all lie inside this frame's */
        if (keep && (!(sc.drawn[2] > sc.drawn[0] && sc.drawn[3] > sc.drawn[1]) ||
                     (now[0] <= sc.drawn[0] && now[1] <= sc.drawn[1] && now[2] >= sc.drawn[2] && now[3] >= sc.drawn[3])))
            p.nclr_x = p.nclr_y = 0;
        sc.valid = true, sc.target = target, sc.raw = raw_target, sc.W = c->W, sc.H = c->H;
        sc.pt_rank = c->pt_rank, sc.pt_nranks = c->pt_nranks, sc.pt_px = c->pt_px;
        std::memcpy(sc.drawn, now, sizeof now);
    }
    c->last_pair = pair, c->last_pair_axis = pair_axis;
    {
        vr_frame_info &fi = c->last_info;
        fi.layout = pair ? VR_LAYOUT_PAIRED : VR_LAYOUT_ARRAY3D, fi.pair_axis = pair ? pair_axis : -1;
        fi.empty_skip_active = ess ? 1 : 0, fi.exact_sampler = exact ? 1 : 0;
        for (int i = 0; i < 4; i++) fi.bbox[i] = p.bbox[i];
        fi.march_blocks = p.sp_count, fi.clear_blocks = p.nclr_x * p.nclr_y;
        const uint64_t tile_px = (uint64_t)VR_TILE * p.tile_h, clear_px = p.pt_enum ? (uint64_t)p.pt_px * p.pt_px : 64ull * 32ull;
        const uint64_t rect_tiles = (uint64_t)std::max(p.rect[2] - p.rect[0], 0) * (uint64_t)std::max(p.rect[3] - p.rect[1], 0);
        fi.target_bytes = 4ull * (std::min<uint64_t>(p.sp_count, (rect_tiles + c->pt_nranks - 1) / c->pt_nranks + p.pt_cnt) * tile_px + (uint64_t)fi.clear_blocks * clear_px);
    }
    /* outside p.bbox every pixel of this frame is the clear colour (screen_bbox is conservative); with a
     * partition the frame is only partly this context's */
    for (int i = 0; i < 4; i++) c->slot_bbox[frame_index][i] = p.bbox[i];
    c->slot_bbox_valid[frame_index] = p.pt_nranks <= 1 && !c->sort_last &&
                                      target == (frame_index == 0 ? c->d_image : c->d_image_slot[frame_index]); /* not a redirected target */
    p.out = target;
    p.raw = raw_target;
    if (c->opt.collect_stats) VR_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 4 * sizeof(unsigned long long), s));
    p.tile_mask = nullptr, p.mask_words = 0;
    c->slot_mask_valid[frame_index] = false;
    if (ess && !exact && !c->opt.collect_stats && c->tune_tile_mask && p.sp_count) {
        /* screen-space mask of the tiles that can meet an occupied cell (one small launch ahead of K1, same stream) */
        const uint32_t words = (p.ntx * p.nty + 31u) / 32u;
        if (c->d_tile_mask[frame_index] && c->tile_mask_words != words)
            for (int k = 0; k < VR_MAX_FRAMES_IN_FLIGHT; k++)
                if (c->d_tile_mask[k]) {
                    VR_CUDA(c, cudaDeviceSynchronize());
                    cudaFree(c->d_tile_mask[k]), c->d_tile_mask[k] = nullptr;
                }
        if (!c->d_tile_mask[frame_index]) {
            VR_CUDA(c, cudaMalloc(&c->d_tile_mask[frame_index], sizeof(uint32_t) * 2 * (words + 1)));
            VR_CUDA(c, cudaMemsetAsync(c->d_tile_mask[frame_index], 0, sizeof(uint32_t) * 2 * (words + 1), s));
            c->tile_mask_cur[frame_index] = 0;
        }
        c->tile_mask_words = words;
        float pv[16], lo[3], cell[3];
        {
            double v[16], pr[16], m[16];
            for (int i = 0; i < 16; i++) v[i] = (double)c->view[i], pr[i] = (double)c->proj[i];
            mat4_mul_d(pr, v, m);
            for (int i = 0; i < 16; i++) pv[i] = (float)m[i];
        }
        const uint32_t ld[3] = {c->nx, c->ny, c->nz};
        for (int i = 0; i < 3; i++) {
            const double full = c->sort_last ? (double)c->full[i] : (double)ld[i], org = c->sort_last ? (double)c->tex_org[i] : 0.0;
            lo[i] = (float)(org / full - 0.5), cell[i] = (float)((double)c->occ_cell / full);
        }
        const uint32_t cur = c->tile_mask_cur[frame_index];
        uint32_t *mask = c->d_tile_mask[frame_index] + (size_t)cur * (words + 1), *other = c->d_tile_mask[frame_index] + (size_t)(cur ^ 1u) * (words + 1);
        const uint64_t ncells = (uint64_t)c->occ_n[0] * c->occ_n[1] * c->occ_n[2];
        VR_CUDA(c, vr_k0_tile_mask(pv, lo, cell, c->occ_n, c->d_occ_list, c->d_occ_count, (uint32_t)(ncells - c->occ_empty_cells), c->W, c->H,
                                   p.ntx, p.tile_h, mask, other, words, c->tune_tile_mask != 2, s));
        c->tile_mask_cur[frame_index] = cur ^ 1u;
        p.tile_mask = mask, p.mask_words = words;
        c->slot_mask_valid[frame_index] = true;
        c->slot_mask_shape[frame_index][0] = p.ntx, c->slot_mask_shape[frame_index][1] = p.nty, c->slot_mask_shape[frame_index][2] = p.tile_h;
    }
    if (preint) {
        p.tex_pre = c->preint_tex;
        VR_CUDA(c, vr_launch_march_preint(p, pair, c->tune_threads, s));
        return VR_OK;
    }
    if (c->bench_fetch_only) {
        if (exact || ess || c->opt.collect_stats || c->sort_last || (pair && p.pair_split != 0x10000))
            return fail(c, VR_ERR_STATE, "vr_bench_fetch_only: hardware sampler, no skipping, no statistics, whole volume, pair copy in one array");
        VR_CUDA(c, vr_launch_march_probe(p, pair, c->tune_batch, c->tune_threads, s));
        return VR_OK;
    }
    p.ess_guard = c->tune_ess_guard;
    VR_CUDA(c, vr_launch_march(p, exact, ess, c->opt.collect_stats != 0, c->sort_last, pair, ess ? c->tune_batch_ess : c->tune_batch, c->tune_threads, s));
    if (c->opt.collect_stats)
        VR_CUDA(c, cudaMemcpyAsync(c->h_counters, c->d_counters, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    return VR_OK;
}

extern "C" int vr_bench_fetch_only(vr_ctx *c, int on)
{
    if (!c) return VR_ERR_INVALID;
    c->bench_fetch_only = on != 0;
    return VR_OK;
}

extern "C" int vr_get_last_frame_info(const vr_ctx *c, vr_frame_info *out)
{
    if (!c || !out) return VR_ERR_INVALID;
    *out = c->last_info;
    return VR_OK;
}

extern "C" int vr_get_tile_mask(vr_ctx *c, uint32_t frame_index, uint32_t *words, uint32_t capacity, uint32_t *tiles_x, uint32_t *tiles_y,
                                uint32_t *tile_h)
{
    if (!c || frame_index >= VR_MAX_FRAMES_IN_FLIGHT) return VR_ERR_INVALID;
    if (!c->d_tile_mask[frame_index] || !c->slot_mask_valid[frame_index])
        return fail(c, VR_ERR_STATE, "vr_get_tile_mask: the frame recorded last into this slot used no tile mask");
    const uint32_t n = c->tile_mask_words + 1;
    if (tiles_x) *tiles_x = c->slot_mask_shape[frame_index][0];
    if (tiles_y) *tiles_y = c->slot_mask_shape[frame_index][1];
    if (tile_h) *tile_h = c->slot_mask_shape[frame_index][2];
    if (!words) return VR_OK;
    if (capacity < n) return fail(c, VR_ERR_INVALID, "vr_get_tile_mask: capacity " + std::to_string(capacity) + " < " + std::to_string(n) + " words");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize());
    /* the buffer the frame recorded last used is the one that is not next in turn */
    VR_CUDA(c, cudaMemcpy(words, c->d_tile_mask[frame_index] + (size_t)(c->tile_mask_cur[frame_index] ^ 1u) * n, sizeof(uint32_t) * n,
                          cudaMemcpyDeviceToHost));
    return VR_OK;
}

extern "C" int vr_sync(vr_ctx *c)
{
    if (!c) return VR_ERR_INVALID;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize());
    if (c->h_counters[7]) { /* raised by a vr_stream_wait that gave up (sync.cu) */
        const unsigned long long which = c->h_counters[7] - 1;
        c->h_counters[7] = 0;
        return fail(c, VR_ERR_STATE, "vr_stream_wait timed out waiting for counter " + std::to_string(which) + " of its range");
    }
    return VR_OK;
}

/* =============================================================== outputs == */

extern "C" int vr_get_size(const vr_ctx *c, uint32_t *w, uint32_t *h)
{
    if (!c) return VR_ERR_INVALID;
    if (w) *w = c->W;
    if (h) *h = c->H;
    return VR_OK;
}

extern "C" int vr_get_image(vr_ctx *c, void **dev, size_t *pitch)
{
    if (!c || !dev) return VR_ERR_INVALID;
    *dev = c->d_image;
    if (pitch) *pitch = (size_t)c->W * sizeof(uchar4);
    return VR_OK;
}

extern "C" int vr_get_image_frame(vr_ctx *c, uint32_t frame_index, void **dev, size_t *pitch)
{
    if (!c || !dev || frame_index >= VR_MAX_FRAMES_IN_FLIGHT) return VR_ERR_INVALID;
    *dev = frame_index == 0 ? c->d_image : c->d_image_slot[frame_index];
    if (!*dev) return fail(c, VR_ERR_STATE, "vr_get_image_frame: that frame was never rendered");
    if (pitch) *pitch = (size_t)c->W * sizeof(uchar4);
    return VR_OK;
}

/* the viewport's LINEAR / CLAMP_TO_EDGE sampling of the offscreen image (offscreen_pass.cpp:361-375) */
extern "C" int vr_present_frame(vr_ctx *c, uint32_t frame_index, void *dst_dev_rgba8, uint32_t dst_width, uint32_t dst_height,
                                size_t dst_pitch_bytes, void *cuda_stream)
{
    if (!c || frame_index >= VR_MAX_FRAMES_IN_FLIGHT) return fail(c, VR_ERR_INVALID, "vr_present_frame: bad arguments");
    if (dst_width == 0 || dst_height == 0) return VR_OK; /* like framebuffer_size_changed(0, 0): nothing to show */
    if (!dst_dev_rgba8 || (dst_pitch_bytes && (dst_pitch_bytes % 4 || dst_pitch_bytes / 4 < dst_width)))
        return fail(c, VR_ERR_INVALID, "vr_present_frame: bad destination");
    const uchar4 *src = frame_index == 0 ? (c->ext_image ? c->ext_image : c->d_image)
                                         : (c->ext_image_slot[frame_index] ? c->ext_image_slot[frame_index] : c->d_image_slot[frame_index]);
    if (!src) return fail(c, VR_ERR_STATE, "vr_present_frame: no frame has been rendered into this slot");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : c->last_stream; /* after the frame's render */
    VR_CUDA(c, vr_launch_present(src, c->W, c->H, reinterpret_cast<uchar4 *>(dst_dev_rgba8), dst_width, dst_height,
                                 dst_pitch_bytes ? dst_pitch_bytes / 4 : dst_width, s));
    return VR_OK;
}

/* A full delivery into a host buffer the dirty-rectangle path knows: the buffer now holds exactly that frame. */
static void note_full_delivery(vr_ctx *c, uint32_t frame_index, void *host)
{
    for (auto &t : c->host_targets)
        if (t.ptr == host) {
            if (t.W == c->W && t.H == c->H && c->slot_bbox_valid[frame_index]) {
                for (int i = 0; i < 4; i++) t.rect[i] = c->slot_bbox[frame_index][i];
                t.stamp = ++c->host_target_clock;
            } else {
                t.ptr = nullptr;
            }
        }
}

extern "C" int vr_forget_host_buffer(vr_ctx *c, const void *host)
{
    if (!c) return VR_ERR_INVALID;
    for (auto &t : c->host_targets)
        if (!host || t.ptr == host) t.ptr = nullptr;
    return VR_OK;
}

extern "C" int vr_download_frame_async(vr_ctx *c, uint32_t frame_index, void *host, void *cuda_stream)
{
    if (!c || !host || frame_index >= VR_MAX_FRAMES_IN_FLIGHT) return fail(c, VR_ERR_INVALID, "vr_download_frame_async: bad arguments");
    const uchar4 *src = frame_index == 0 ? c->d_image : c->d_image_slot[frame_index];
    if (!src) return fail(c, VR_ERR_STATE, "vr_download_frame_async: that frame was never rendered");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->last_stream;
    VR_CUDA(c, cudaMemcpyAsync(host, src, (size_t)c->W * c->H * sizeof(uchar4), cudaMemcpyDeviceToHost, s));
    note_full_delivery(c, frame_index, host);
    return VR_OK;
}

/* Copy only what can have changed.  Invariant per known host buffer: outside `rect` it holds the clear colour. */
extern "C" int vr_download_frame_dirty_async(vr_ctx *c, uint32_t frame_index, void *host, void *cuda_stream, size_t *bytes_copied)
{
    if (!c || !host || frame_index >= VR_MAX_FRAMES_IN_FLIGHT) return fail(c, VR_ERR_INVALID, "vr_download_frame_dirty_async: bad arguments");
    const uchar4 *src = frame_index == 0 ? c->d_image : c->d_image_slot[frame_index];
    if (!src) return fail(c, VR_ERR_STATE, "vr_download_frame_dirty_async: that frame was never rendered");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->last_stream;
    vr_ctx::HostTarget *e = nullptr, *lru = &c->host_targets[0];
    for (auto &t : c->host_targets) {
        if (t.ptr == host && t.W == c->W && t.H == c->H) e = &t;
        if (t.stamp < lru->stamp) lru = &t;
    }
    const uint32_t *now = c->slot_bbox[frame_index];
    const size_t all = (size_t)c->W * c->H * sizeof(uchar4);
    size_t bytes = 0;
    if (!c->slot_bbox_valid[frame_index]) {
        /* a frame this context does not own alone (partitioned), or not rendered since a resize: everything */
        VR_CUDA(c, cudaMemcpyAsync(host, src, all, cudaMemcpyDeviceToHost, s));
        if (e) e->ptr = nullptr; /* what the buffer holds is no longer known */
        if (bytes_copied) *bytes_copied = all;
        return VR_OK;
    }
    if (!e) {
        /* first delivery into this buffer: everything */
        bytes = all;
        VR_CUDA(c, cudaMemcpyAsync(host, src, all, cudaMemcpyDeviceToHost, s));
        e = lru;
        e->ptr = host, e->W = c->W, e->H = c->H;
    } else {
        const bool a = e->rect[2] > e->rect[0] && e->rect[3] > e->rect[1], b = now[2] > now[0] && now[3] > now[1];
        uint32_t r[4] = {0, 0, 0, 0};
        if (a && b) r[0] = std::min(e->rect[0], now[0]), r[1] = std::min(e->rect[1], now[1]), r[2] = std::max(e->rect[2], now[2]), r[3] = std::max(e->rect[3], now[3]);
        else if (a) std::memcpy(r, e->rect, sizeof r);
        else if (b) std::memcpy(r, now, sizeof r);
        if (r[2] > r[0] && r[3] > r[1]) {
            const size_t off = (size_t)r[1] * c->W + r[0], pitch = (size_t)c->W * sizeof(uchar4);
            bytes = (size_t)(r[2] - r[0]) * (r[3] - r[1]) * sizeof(uchar4);
            VR_CUDA(c, cudaMemcpy2DAsync(static_cast<uchar4 *>(host) + off, pitch, src + off, pitch, (size_t)(r[2] - r[0]) * sizeof(uchar4),
                                         r[3] - r[1], cudaMemcpyDeviceToHost, s));
        }
    }
    for (int i = 0; i < 4; i++) e->rect[i] = now[i];
    e->stamp = ++c->host_target_clock;
    if (bytes_copied) *bytes_copied = bytes;
    return VR_OK;
}

extern "C" int vr_download_async(vr_ctx *c, void *host, void *cuda_stream)
{
    if (!c || !host) return fail(c, VR_ERR_INVALID, "vr_download: null argument");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->last_stream;
    VR_CUDA(c, cudaMemcpyAsync(host, c->d_image, (size_t)c->W * c->H * sizeof(uchar4), cudaMemcpyDeviceToHost, s));
    note_full_delivery(c, 0, host);
    return VR_OK;
}

extern "C" int vr_download(vr_ctx *c, void *host)
{
    const int rc = vr_download_async(c, host, nullptr);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaStreamSynchronize(c->last_stream));
    return VR_OK;
}

extern "C" int vr_write_image(vr_ctx *c, const char *path)
{
    if (!c || !path) return fail(c, VR_ERR_INVALID, "vr_write_image: null argument");
    std::vector<unsigned char> img((size_t)c->W * c->H * 4);
    const int rc = vr_download(c, img.data());
    if (rc != VR_OK) return rc;
    const size_t len = std::strlen(path);
    const bool pam = len >= 4 && std::strcmp(path + len - 4, ".pam") == 0;
    std::FILE *f = std::fopen(path, "wb");
    if (!f) return fail(c, VR_ERR_IO, std::string("vr_write_image: cannot open ") + path);
    bool ok;
    if (pam) {
        ok = std::fprintf(f, "P7\nWIDTH %u\nHEIGHT %u\nDEPTH 4\nMAXVAL 255\nTUPLTYPE RGB_ALPHA\nENDHDR\n", c->W, c->H) > 0 &&
             std::fwrite(img.data(), 1, img.size(), f) == img.size();
    } else {
        std::vector<unsigned char> rgb((size_t)c->W * c->H * 3);
        for (size_t i = 0; i < (size_t)c->W * c->H; i++) rgb[3 * i] = img[4 * i], rgb[3 * i + 1] = img[4 * i + 1], rgb[3 * i + 2] = img[4 * i + 2];
        ok = std::fprintf(f, "P6\n%u %u\n255\n", c->W, c->H) > 0 && std::fwrite(rgb.data(), 1, rgb.size(), f) == rgb.size();
    }
    ok = (std::fclose(f) == 0) && ok;
    return ok ? VR_OK : fail(c, VR_ERR_IO, std::string("vr_write_image: write failed: ") + path);
}

extern "C" int vr_get_raw(vr_ctx *c, void **dev)
{
    if (!c || !dev) return VR_ERR_INVALID;
    if (!c->d_raw) return fail(c, VR_ERR_STATE, "vr_get_raw: options.write_raw is off");
    *dev = c->d_raw;
    return VR_OK;
}

extern "C" int vr_get_raw_frame(vr_ctx *c, uint32_t frame_index, void **dev)
{
    if (!c || !dev || frame_index >= VR_MAX_FRAMES_IN_FLIGHT) return VR_ERR_INVALID;
    if (!c->opt.write_raw || !c->d_raw) return fail(c, VR_ERR_STATE, "vr_get_raw_frame: options.write_raw is off");
    VR_CUDA(c, cudaSetDevice(c->device));
    const int rc = ensure_raw_slot(c, frame_index);
    if (rc != VR_OK) return rc;
    *dev = frame_index == 0 ? c->d_raw : c->d_raw_slot[frame_index];
    return VR_OK;
}

extern "C" int vr_download_raw(vr_ctx *c, float *host)
{
    if (!c || !host) return VR_ERR_INVALID;
    if (!c->d_raw) return fail(c, VR_ERR_STATE, "vr_download_raw: options.write_raw is off");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaMemcpyAsync(host, c->d_raw, (size_t)c->W * c->H * sizeof(float4), cudaMemcpyDeviceToHost, c->last_stream));
    VR_CUDA(c, cudaStreamSynchronize(c->last_stream));
    return VR_OK;
}

extern "C" int vr_get_stats(vr_ctx *c, vr_stats *out)
{
    if (!c || !out) return VR_ERR_INVALID;
    if (!c->opt.collect_stats) return fail(c, VR_ERR_STATE, "vr_get_stats: options.collect_stats is off");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->last_stream));
    out->covered_pixels = c->h_counters[0], out->n_ref = c->h_counters[1];
    out->n_inslice = c->h_counters[2], out->n_fetch = c->h_counters[3];
    return VR_OK;
}

extern "C" int vr_get_occupancy_info(vr_ctx *c, uint32_t cells[3], uint32_t *cell_voxels, uint64_t *empty_cells)
{
    if (!c) return VR_ERR_INVALID;
    VR_CUDA(c, cudaSetDevice(c->device));
    if (c->max_density - c->min_density > 0.0f) {
        const int rc = prepare_occupancy(c, c->last_stream);
        if (rc != VR_OK) return rc;
    }
    if (cells) cells[0] = c->occ_n[0], cells[1] = c->occ_n[1], cells[2] = c->occ_n[2];
    if (cell_voxels) *cell_voxels = (uint32_t)c->occ_cell;
    if (empty_cells) *empty_cells = c->occ_dirty ? 0 : c->occ_empty_cells;
    return VR_OK;
}

extern "C" int vr_download_volume(vr_ctx *c, void *host)
{
    if (!c || !host) return VR_ERR_INVALID;
    if (!c->array) return fail(c, VR_ERR_STATE, "no volume");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaMemcpy3DParms cp = {};
    cp.srcArray = c->array;
    cp.dstPtr = make_cudaPitchedPtr(host, (size_t)c->nx * dtype_size(c->dtype), c->nx, c->ny);
    cp.extent = make_cudaExtent(c->nx, c->ny, c->nz);
    cp.kind = cudaMemcpyDeviceToHost;
    VR_CUDA(c, cudaMemcpy3D(&cp));
    return VR_OK;
}

/* ============================================================ synthetic == */

extern "C" int vr_synth_volume(vr_ctx *c, int kind, uint32_t seed, int dtype, uint32_t nx, uint32_t ny, uint32_t nz)
{
    if (!c || kind < 0 || kind > VR_SYNTH_BLOBS_FBM || dtype < VR_U8 || dtype > VR_F32 || nx == 0 || ny == 0 || nz == 0)
        return fail(c, VR_ERR_INVALID, "vr_synth_volume: bad arguments");
    if (c->sort_last) return fail(c, VR_ERR_STATE, "vr_synth_volume: context is sort-last partitioned (use vr_synth_subvolume)");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize());
    void *lin = nullptr;
    VR_CUDA(c, cudaMalloc(&lin, (size_t)nx * ny * nz * dtype_size(dtype)));
    const int org[3] = {0, 0, 0};
    const uint32_t full[3] = {nx, ny, nz};
    cudaError_t e = vr_k0_synth(lin, kind, seed, dtype, nx, ny, nz, org, full, c->stream);
    if (e != cudaSuccess) {
        cudaFree(lin);
        return cuda_fail(c, e, "k_synth");
    }
    const int rc = install_volume(c, lin, dtype, nx, ny, nz, nullptr, nullptr);
    cudaFree(lin);
    return rc;
}

extern "C" int vr_synth_subvolume(vr_ctx *c, int kind, uint32_t seed, int dtype)
{
    if (!c || kind < 0 || kind > VR_SYNTH_BLOBS_FBM || dtype < VR_U8 || dtype > VR_F32) return fail(c, VR_ERR_INVALID, "vr_synth_subvolume: bad arguments");
    if (!c->sort_last) return fail(c, VR_ERR_STATE, "vr_synth_subvolume: call vr_set_partition_sort_last first");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaDeviceSynchronize());
    uint32_t l[3];
    for (int i = 0; i < 3; i++) { /* the stored box plus one ghost voxel towards each neighbour */
        const uint32_t end = c->st_org[i] + c->st_dims[i];
        l[i] = c->st_dims[i] + (c->st_org[i] > 0 ? 1 : 0) + (end < c->full[i] ? 1 : 0);
    }
    void *lin = nullptr;
    VR_CUDA(c, cudaMalloc(&lin, (size_t)l[0] * l[1] * l[2] * dtype_size(dtype)));
    cudaError_t e = vr_k0_synth(lin, kind, seed, dtype, l[0], l[1], l[2], c->tex_org, c->full, c->stream);
    if (e != cudaSuccess) {
        cudaFree(lin);
        return cuda_fail(c, e, "k_synth");
    }
    const int rc = install_volume(c, lin, dtype, l[0], l[1], l[2], nullptr, nullptr);
    cudaFree(lin);
    return rc;
}

/* =========================================================== partitioning == */

extern "C" int vr_set_partition_sort_first(vr_ctx *c, uint32_t rank, uint32_t nranks, uint32_t tile_px)
{
    if (!c || nranks == 0 || rank >= nranks || tile_px == 0 || tile_px % VR_TILE != 0)
        return fail(c, VR_ERR_INVALID, "vr_set_partition_sort_first: need rank < nranks and tile_px a multiple of 16");
    c->pt_rank = rank, c->pt_nranks = nranks, c->pt_px = tile_px;
    return VR_OK;
}

extern "C" uint32_t vr_sort_first_owner(uint32_t tile_x, uint32_t tile_y, uint32_t nranks)
{
    return nranks ? (tile_x + tile_y * vr_sort_first_stride(nranks)) % nranks : 0;
}

extern "C" int vr_set_partition_sort_last(vr_ctx *c, const uint32_t full_dims[3], const uint32_t own_org[3],
                                          const uint32_t own_dims[3])
{
    if (!c || !full_dims || !own_org || !own_dims) return fail(c, VR_ERR_INVALID, "vr_set_partition_sort_last: null argument");
    for (int i = 0; i < 3; i++)
        if (own_dims[i] == 0 || own_org[i] + own_dims[i] > full_dims[i])
            return fail(c, VR_ERR_INVALID, "vr_set_partition_sort_last: sub-box outside the grid");
    c->sort_last = true;
    for (int i = 0; i < 3; i++) {
        c->full[i] = full_dims[i], c->own_org[i] = c->st_org[i] = own_org[i], c->own_dims[i] = c->st_dims[i] = own_dims[i];
        c->tex_org[i] = (int)own_org[i] - (own_org[i] > 0 ? 1 : 0); /* one ghost voxel towards each neighbour */
    }
    return VR_OK;
}

extern "C" int vr_set_owned_box(vr_ctx *c, const uint32_t org[3], const uint32_t dims[3])
{
    if (!c || !org || !dims) return fail(c, VR_ERR_INVALID, "vr_set_owned_box: null argument");
    if (!c->sort_last) return fail(c, VR_ERR_STATE, "vr_set_owned_box: call vr_set_partition_sort_last first");
    for (int i = 0; i < 3; i++)
        if (dims[i] == 0 || org[i] < c->st_org[i] || org[i] + dims[i] > c->st_org[i] + c->st_dims[i])
            return fail(c, VR_ERR_INVALID, "vr_set_owned_box: the box leaves the sub-volume this rank holds");
    for (int i = 0; i < 3; i++) c->own_org[i] = org[i], c->own_dims[i] = dims[i];
    return VR_OK;
}

extern "C" int vr_clear_partition(vr_ctx *c)
{
    if (!c) return VR_ERR_INVALID;
    c->sort_last = false;
    c->pt_rank = 0, c->pt_nranks = 1, c->pt_px = 64;
    for (int i = 0; i < 3; i++) c->tex_org[i] = 0;
    return VR_OK;
}

extern "C" int vr_partition_box(const uint32_t grid[3], const uint32_t full_dims[3], uint32_t index, uint32_t org[3],
                                uint32_t dims[3])
{
    if (!grid || !full_dims || !org || !dims) return VR_ERR_INVALID;
    const uint32_t total = grid[0] * grid[1] * grid[2];
    if (total == 0 || index >= total) return VR_ERR_INVALID;
    const uint32_t b[3] = {index % grid[0], (index / grid[0]) % grid[1], index / (grid[0] * grid[1])};
    for (int i = 0; i < 3; i++) {
        if (grid[i] > full_dims[i]) return VR_ERR_INVALID;
        const uint64_t lo = (uint64_t)b[i] * full_dims[i] / grid[i], hi = (uint64_t)(b[i] + 1) * full_dims[i] / grid[i];
        org[i] = (uint32_t)lo, dims[i] = (uint32_t)(hi - lo);
    }
    return VR_OK;
}

namespace
{
uint32_t regular_plane(const uint32_t grid[3], const uint32_t full_dims[3], int a, uint32_t k) { return (uint32_t)((uint64_t)k * full_dims[a] / grid[a]); }
bool margins_ok(const uint32_t grid[3], const uint32_t full_dims[3], const uint32_t margin[3])
{
    for (int a = 0; a < 3; a++) {
        if (grid[a] == 0 || grid[a] > full_dims[a]) return false;
        /* two neighbouring planes may each move by the margin: they must not meet */
        if (margin && grid[a] > 1 && (uint64_t)2 * margin[a] + 1 > full_dims[a] / grid[a]) return false;
    }
    return true;
}
}  // namespace

extern "C" int vr_partition_stored_box(const uint32_t grid[3], const uint32_t full_dims[3], uint32_t index, const uint32_t margin[3],
                                       uint32_t org[3], uint32_t dims[3])
{
    if (!grid || !full_dims || !margin || !org || !dims || !margins_ok(grid, full_dims, margin)) return VR_ERR_INVALID;
    const uint32_t total = grid[0] * grid[1] * grid[2];
    if (index >= total) return VR_ERR_INVALID;
    const uint32_t b[3] = {index % grid[0], (index / grid[0]) % grid[1], index / (grid[0] * grid[1])};
    for (int a = 0; a < 3; a++) {
        const uint32_t lo = regular_plane(grid, full_dims, a, b[a]), hi = regular_plane(grid, full_dims, a, b[a] + 1);
        const uint32_t slo = b[a] > 0 ? lo - margin[a] : lo, shi = b[a] + 1 < grid[a] ? hi + margin[a] : hi;
        org[a] = slo, dims[a] = shi - slo;
    }
    return VR_OK;
}

extern "C" int vr_partition_balance(const uint32_t grid[3], const uint32_t full_dims[3], const float camera_position[3], const uint32_t margin[3],
                                    uint32_t *planes)
{
    if (!grid || !full_dims || !camera_position || !margin || !planes || !margins_ok(grid, full_dims, margin)) return VR_ERR_INVALID;
    /* Samples per unit volume of a perspective frame fall with the squared distance from the eye (a pixel's solid angle
     * is fixed, the steps along its ray are equidistant): density ~ 1 / r^2 over the unit cube, integrated on Q^3 cell
     * centres in double -- the same numbers on every rank.  Per axis the interior planes go to the quantiles k / g of
     * the marginal, no further than `margin` voxels from the regular split. */
    constexpr int Q = 16;
    double marg[3][Q] = {};
    for (int z = 0; z < Q; z++)
        for (int y = 0; y < Q; y++)
            for (int x = 0; x < Q; x++) {
                const double px = (x + 0.5) / Q - 0.5 - (double)camera_position[0], py = (y + 0.5) / Q - 0.5 - (double)camera_position[1],
                             pz = (z + 0.5) / Q - 0.5 - (double)camera_position[2];
                const double r2 = px * px + py * py + pz * pz;
                const double rho = 1.0 / (r2 > 1e-4 ? r2 : 1e-4);
                marg[0][x] += rho, marg[1][y] += rho, marg[2][z] += rho;
            }
    uint32_t *out = planes;
    for (int a = 0; a < 3; a++) {
        double total = 0.0;
        for (int i = 0; i < Q; i++) total += marg[a][i];
        for (uint32_t k = 1; k < grid[a]; k++) {
            const double want = total * (double)k / (double)grid[a];
            double acc = 0.0, pos = 0.5;
            for (int i = 0; i < Q; i++) {
                if (acc + marg[a][i] >= want) {
                    pos = ((double)i + (want - acc) / marg[a][i]) / Q; /* texture coordinate of the quantile */
                    break;
                }
                acc += marg[a][i];
            }
            const int64_t reg = (int64_t)regular_plane(grid, full_dims, a, k);
            int64_t v = (int64_t)std::llround(pos * (double)full_dims[a]);
            v = std::min<int64_t>(std::max<int64_t>(v, reg - (int64_t)margin[a]), reg + (int64_t)margin[a]);
            *out++ = (uint32_t)v;
        }
    }
    return VR_OK;
}

namespace
{
/* plane k of axis a: 0 and grid[a] are the faces of the volume, the others come from `planes` (NULL: the regular split) */
uint32_t plane_at(const uint32_t grid[3], const uint32_t full_dims[3], const uint32_t *planes, int a, uint32_t k)
{
    if (k == 0) return 0;
    if (k >= grid[a]) return full_dims[a];
    if (!planes) return regular_plane(grid, full_dims, a, k);
    uint32_t off = 0;
    for (int i = 0; i < a; i++) off += grid[i] - 1;
    return planes[off + k - 1];
}
}  // namespace

extern "C" int vr_partition_box_planes(const uint32_t grid[3], const uint32_t full_dims[3], const uint32_t *planes, uint32_t index,
                                       uint32_t org[3], uint32_t dims[3])
{
    if (!grid || !full_dims || !org || !dims || !margins_ok(grid, full_dims, nullptr)) return VR_ERR_INVALID;
    const uint32_t total = grid[0] * grid[1] * grid[2];
    if (index >= total) return VR_ERR_INVALID;
    const uint32_t b[3] = {index % grid[0], (index / grid[0]) % grid[1], index / (grid[0] * grid[1])};
    for (int a = 0; a < 3; a++) {
        const uint32_t lo = plane_at(grid, full_dims, planes, a, b[a]), hi = plane_at(grid, full_dims, planes, a, b[a] + 1);
        if (hi <= lo || hi > full_dims[a]) return VR_ERR_INVALID;
        org[a] = lo, dims[a] = hi - lo;
    }
    return VR_OK;
}

extern "C" int vr_partition_order(const uint32_t grid[3], const uint32_t full_dims[3], const float camera_position[3],
                                  uint32_t *out)
{
    return vr_partition_order_planes(grid, full_dims, nullptr, camera_position, out);
}

extern "C" int vr_partition_order_planes(const uint32_t grid[3], const uint32_t full_dims[3], const uint32_t *planes,
                                         const float camera_position[3], uint32_t *out)
{
    if (!grid || !full_dims || !camera_position || !out) return VR_ERR_INVALID;
    const uint32_t total = grid[0] * grid[1] * grid[2];
    if (total == 0) return VR_ERR_INVALID;
    /* per axis: index of the slab holding the camera (-1 / grid when outside the cube).  A ray
     * leaving the camera moves monotonically away from it on every axis, so |slab - camera slab|
     * never decreases along a ray: sorting by the sum of the three distances is a valid
     * front-to-back order for every pixel of the frame. */
    int cs[3];
    for (int a = 0; a < 3; a++) {
        const double t = (double)camera_position[a] + 0.5; /* model -> texture space */
        int s = -1;
        for (uint32_t k = 0; k < grid[a]; k++) {
            const double lo = (double)plane_at(grid, full_dims, planes, a, k) / (double)full_dims[a];
            if (t >= lo) s = (int)k;
        }
        if (t >= 1.0) s = (int)grid[a];
        cs[a] = s;
    }
    std::vector<std::pair<int, uint32_t>> keyed(total);
    for (uint32_t i = 0; i < total; i++) {
        const int b[3] = {(int)(i % grid[0]), (int)((i / grid[0]) % grid[1]), (int)(i / (grid[0] * grid[1]))};
        int key = 0;
        for (int a = 0; a < 3; a++) key += std::abs(b[a] - cs[a]);
        keyed[i] = {key, i};
    }
    std::stable_sort(keyed.begin(), keyed.end(), [](const auto &x, const auto &y) { return x.first < y.first; });
    for (uint32_t i = 0; i < total; i++) out[i] = keyed[i].second;
    return VR_OK;
}

extern "C" int vr_composite(vr_ctx *c, const void *const *parts, uint32_t nparts, uint64_t pixel_begin, uint64_t pixel_end,
                            void *cuda_stream)
{
    return vr_composite_to(c, parts, nparts, pixel_begin, pixel_end, nullptr, cuda_stream);
}

/* test hook: the fp32 columns of inverse(proj*view) the kernel will use */
extern "C" int vr_debug_ray_context(const vr_ctx *c, float out[16])
{
    if (!c || !out || !c->ray_valid) return VR_ERR_INVALID;
    std::memcpy(out, c->rc, sizeof(float) * 16);
    return VR_OK;
}

/* Loader without a context (host only): what vr_load_nrrd would upload.  *data is malloc'd,
 * release with vr_free_host. */
extern "C" int vr_nrrd_read_host(const char *path, int *dtype, uint32_t dims[3], void **data, char *errbuf, size_t errbuf_len)
{
    if (!path || !dtype || !dims || !data) return VR_ERR_INVALID;
    VrNrrd n;
    std::string err;
    const int rc = vr_nrrd_read(path, n, err);
    if (rc != VR_OK) {
        if (errbuf && errbuf_len) std::snprintf(errbuf, errbuf_len, "%s", err.c_str());
        return rc;
    }
    void *p = std::malloc(n.bytes.size());
    if (!p) return VR_ERR_NOMEM;
    std::memcpy(p, n.bytes.data(), n.bytes.size());
    *dtype = n.dtype, dims[0] = n.dims[0], dims[1] = n.dims[1], dims[2] = n.dims[2], *data = p;
    return VR_OK;
}

extern "C" void vr_free_host(void *p) { std::free(p); }

/* The header's per-axis sample distances alone (host only): `spacings`, or the length of `space directions`; NaN where none. */
extern "C" int vr_nrrd_read_spacing(const char *path, double spacing[3], char *errbuf, size_t errbuf_len)
{
    if (!path || !spacing) return VR_ERR_INVALID;
    VrNrrd n;
    VrNrrdPayload pl; /* asking for the payload's place keeps a raw file's data from being read */
    std::string err;
    const int rc = vr_nrrd_read(path, n, err, &pl);
    if (rc != VR_OK) {
        if (errbuf && errbuf_len) std::snprintf(errbuf, errbuf_len, "%s", err.c_str());
        return rc;
    }
    for (int i = 0; i < 3; i++) spacing[i] = n.spacing[i];
    return VR_OK;
}

/* CSV slices (src/data/csv_file_parser.cpp:14-50 through src/data/importer.cpp:30-46) */
extern "C" int vr_csv_read_host(const char *const *paths, uint32_t npaths, uint32_t dims[3], float **data, float *mn, float *mx,
                                char *errbuf, size_t errbuf_len)
{
    if (!paths || !dims || !data || !mn || !mx) return VR_ERR_INVALID;
    VrCsv v;
    std::string err;
    const int rc = vr_csv_read(paths, npaths, v, err);
    if (rc != VR_OK) {
        if (errbuf && errbuf_len) std::snprintf(errbuf, errbuf_len, "%s", err.c_str());
        return rc;
    }
    float *p = static_cast<float *>(std::malloc(std::max<size_t>(v.data.size(), 1) * sizeof(float)));
    if (!p) return VR_ERR_NOMEM;
    if (!v.data.empty()) std::memcpy(p, v.data.data(), v.data.size() * sizeof(float));
    dims[0] = v.dims[0], dims[1] = v.dims[1], dims[2] = v.dims[2], *data = p, *mn = v.min, *mx = v.max;
    return VR_OK;
}

extern "C" int vr_load_csv(vr_ctx *c, const char *const *paths, uint32_t npaths)
{
    if (!c || !paths) return fail(c, VR_ERR_INVALID, "vr_load_csv: null argument");
    VrCsv v;
    std::string err;
    const int rc = vr_csv_read(paths, npaths, v, err);
    if (rc != VR_OK) return fail(c, rc, err);
    if (!v.dims[0] || !v.dims[1] || !v.dims[2] || v.data.size() != (size_t)v.dims[0] * v.dims[1] * v.dims[2])
        return fail(c, VR_ERR_FORMAT, "vr_load_csv: empty dataset"); /* the reference would ask Vulkan for a zero-sized image */
    return vr_set_volume(c, v.data.data(), VR_F32, v.dims[0], v.dims[1], v.dims[2], &v.min, &v.max);
}

/* ============================================== peer memory (one process per GPU) == */

extern "C" int vr_set_image_target(vr_ctx *c, void *dev_rgba8) { return vr_set_image_target_frame(c, 0, dev_rgba8); }

extern "C" int vr_set_image_target_frame(vr_ctx *c, uint32_t frame_index, void *dev_rgba8)
{
    if (!c || frame_index >= VR_MAX_FRAMES_IN_FLIGHT) return fail(c, VR_ERR_INVALID, "vr_set_image_target_frame: bad arguments");
    if (frame_index == 0) c->ext_image = reinterpret_cast<uchar4 *>(dev_rgba8);
    else c->ext_image_slot[frame_index] = reinterpret_cast<uchar4 *>(dev_rgba8);
    c->slot_clear[frame_index].valid = false; /* what a (re-)attached target holds is not known: the next frame clears all of it */
    return VR_OK;
}

extern "C" int vr_get_flags(vr_ctx *c, void **dev_u32)
{
    if (!c || !dev_u32) return VR_ERR_INVALID;
    *dev_u32 = c->d_flags;
    return VR_OK;
}

extern "C" int vr_stream_signal(vr_ctx *c, void *const *flags, uint32_t n, uint32_t value, void *cuda_stream)
{
    if (!c || !flags || n == 0 || n > VR_SIGNAL_MAX) return fail(c, VR_ERR_INVALID, "vr_stream_signal: need 1..32 flags");
    for (uint32_t i = 0; i < n; i++)
        if (!flags[i]) return fail(c, VR_ERR_INVALID, "vr_stream_signal: null flag pointer");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
    VR_CUDA(c, vr_launch_signal(flags, (int)n, value, s));
    return VR_OK;
}

extern "C" int vr_stream_wait(vr_ctx *c, const void *flags, uint32_t n, uint32_t value, uint32_t timeout_ms, void *cuda_stream)
{
    if (!c || !flags || n == 0 || n > 256) return fail(c, VR_ERR_INVALID, "vr_stream_wait: need 1..256 flags");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
    VR_CUDA(c, vr_launch_wait(flags, (int)n, value, timeout_ms ? timeout_ms : 10000u, &c->h_counters[7], s));
    return VR_OK;
}

extern "C" int vr_ipc_export(vr_ctx *c, int which, unsigned char handle[64])
{
    const bool frame_slot = which >= 16 && which < 16 + VR_MAX_FRAMES_IN_FLIGHT;
    const bool raw_slot = which >= 32 && which < 32 + VR_MAX_FRAMES_IN_FLIGHT;
    if (!c || !handle || (which != 0 && which != 1 && which != 2 && !frame_slot && !raw_slot)) return fail(c, VR_ERR_INVALID, "vr_ipc_export: bad arguments");
    if (raw_slot) {
        void *rp = nullptr;
        const int rc = vr_get_raw_frame(c, (uint32_t)(which - 32), &rp);
        if (rc != VR_OK) return rc;
        cudaIpcMemHandle_t h;
        VR_CUDA(c, cudaIpcGetMemHandle(&h, rp));
        std::memcpy(handle, &h, 64);
        return VR_OK;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    if (frame_slot) {
        const int rc = ensure_slot(c, (uint32_t)(which - 16));
        if (rc != VR_OK) return rc;
    }
    void *ptr = which == 0 ? (void *)c->d_image : (which == 1 ? (void *)c->d_raw : (which == 2 ? (void *)c->d_flags : nullptr));
    if (frame_slot) ptr = which == 16 ? (void *)c->d_image : (void *)c->d_image_slot[which - 16];
    if (!ptr) return fail(c, VR_ERR_STATE, "vr_ipc_export: that buffer does not exist (options.write_raw?)");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    VR_CUDA(c, cudaIpcGetMemHandle(&h, ptr));
    std::memcpy(handle, &h, 64);
    return VR_OK;
}

extern "C" int vr_ipc_open(vr_ctx *c, const unsigned char handle[64], void **dev_ptr)
{
    if (!c || !handle || !dev_ptr) return fail(c, VR_ERR_INVALID, "vr_ipc_open: bad arguments");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, 64);
    VR_CUDA(c, cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return VR_OK;
}

extern "C" int vr_ipc_close(vr_ctx *c, void *dev_ptr)
{
    if (!c || !dev_ptr) return VR_ERR_INVALID;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaIpcCloseMemHandle(dev_ptr));
    return VR_OK;
}

extern "C" int vr_composite_to(vr_ctx *c, const void *const *parts, uint32_t nparts, uint64_t pixel_begin,
                               uint64_t pixel_end, void *out_rgba8, void *cuda_stream)
{
    return vr_composite_rects(c, parts, nullptr, nparts, pixel_begin, pixel_end, out_rgba8, cuda_stream);
}

extern "C" int vr_composite_region(vr_ctx *c, const void *const *parts, const uint32_t *rects, uint32_t nparts, const uint32_t region[4],
                                   uint32_t band_index, uint32_t band_count, void *out_rgba8, void *cuda_stream)
{
    if (!c || !parts || !rects || !region || nparts == 0 || band_count == 0 || band_index >= band_count)
        return fail(c, VR_ERR_INVALID, "vr_composite_region: bad arguments");
    if (region[2] > c->W || region[3] > c->H) return fail(c, VR_ERR_INVALID, "vr_composite_region: region outside the frame");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
    c->last_stream = s;
    uchar4 *out = out_rgba8 ? reinterpret_cast<uchar4 *>(out_rgba8) : (c->ext_image ? c->ext_image : c->d_image);
    VR_CUDA(c, vr_launch_composite_region(reinterpret_cast<const float4 *const *>(parts), rects, nparts, region, band_index, band_count, c->W, out, s));
    return VR_OK;
}

/* the region of an image slot a sort-last frame has to merge: the bounding rectangle of what this frame (the union of its
 * sub-box rectangles, stored in `slot_region` for the next time) and the frame before it in the slot can have drawn */
extern "C" void vr_sort_last_region(const uint32_t *rects, uint32_t nparts, uint32_t W, uint32_t H, int first_into_slot, uint32_t slot_region[4],
                                    uint32_t region[4])
{
    uint32_t cur[4] = {W, H, 0, 0};
    for (uint32_t i = 0; i < nparts; i++)
        if (rects[4 * i + 2] > rects[4 * i] && rects[4 * i + 3] > rects[4 * i + 1]) {
            cur[0] = std::min(cur[0], rects[4 * i]), cur[1] = std::min(cur[1], rects[4 * i + 1]);
            cur[2] = std::max(cur[2], rects[4 * i + 2]), cur[3] = std::max(cur[3], rects[4 * i + 3]);
        }
    if (cur[2] <= cur[0] || cur[3] <= cur[1]) cur[0] = cur[1] = cur[2] = cur[3] = 0;
    region[0] = region[1] = 0, region[2] = W, region[3] = H;
    if (!first_into_slot) {
        const uint32_t *pr = slot_region;
        const bool prev_empty = pr[2] <= pr[0] || pr[3] <= pr[1], cur_empty = cur[2] <= cur[0];
        if (prev_empty && cur_empty) region[0] = region[1] = region[2] = region[3] = 0;
        else if (prev_empty) std::memcpy(region, cur, sizeof cur);
        else if (cur_empty) std::memcpy(region, pr, sizeof cur);
        else region[0] = std::min(pr[0], cur[0]), region[1] = std::min(pr[1], cur[1]), region[2] = std::max(pr[2], cur[2]), region[3] = std::max(pr[3], cur[3]);
    }
    std::memcpy(slot_region, cur, sizeof cur);
}

extern "C" int vr_sort_last_frame(vr_ctx *c, const vr_sort_last_wiring *w, uint64_t k, const float cam[3], void *march_stream,
                                  void *merge_stream, void *present_stream)
{
    if (!c || !w || !cam || !w->flags || !w->raw || !w->image || w->world == 0 || w->world > 31 || w->rank >= w->world || w->display >= w->world ||
        w->slots == 0 || w->slots > VR_MAX_FRAMES_IN_FLIGHT)
        return fail(c, VR_ERR_INVALID, "vr_sort_last_frame: bad wiring");
    if (!c->sort_last) return fail(c, VR_ERR_STATE, "vr_sort_last_frame: call vr_set_partition_sort_last first");
    const uint32_t n = w->world, S = w->slots, s = (uint32_t)(k % S), rank = w->rank, disp = w->display;
    const uint32_t READY = 256, READ_DONE = 384, RANGE_DONE = 512, CONSUMED = 640; /* multigpu.py _L_* */
    auto flag = [&](uint32_t r, uint32_t index) { return (void *)(static_cast<char *>(w->flags[r]) + 4 * (size_t)index); };
    if ((uint64_t)w->grid[0] * w->grid[1] * w->grid[2] != n) return fail(c, VR_ERR_INVALID, "vr_sort_last_frame: grid does not match the world size");
    /* this frame's planes, boxes, order and rectangles: pure functions of the camera, the same on every rank */
    uint32_t planes[3 * 32], order[32], rects[4 * 32], box[32][6];
    const bool balanced = w->margin[0] || w->margin[1] || w->margin[2];
    if (balanced && vr_partition_balance(w->grid, w->full, cam, w->margin, planes) != VR_OK) return fail(c, VR_ERR_INVALID, "vr_partition_balance failed");
    for (uint32_t r = 0; r < n; r++)
        if (vr_partition_box_planes(w->grid, w->full, balanced ? planes : nullptr, r, &box[r][0], &box[r][3]) != VR_OK)
            return fail(c, VR_ERR_INVALID, "vr_partition_box_planes failed");
    if (balanced) {
        const int rc = vr_set_owned_box(c, &box[rank][0], &box[rank][3]);
        if (rc != VR_OK) return rc;
    }
    if (vr_partition_order_planes(w->grid, w->full, balanced ? planes : nullptr, cam, order) != VR_OK) return fail(c, VR_ERR_INVALID, "vr_partition_order failed");
    const void *parts[32];
    for (uint32_t i = 0; i < n; i++) {
        subbox_screen_rect(c, w->full, &box[order[i]][0], &box[order[i]][3], &rects[4 * i]);
        parts[i] = w->raw[(size_t)s * n + order[i]];
    }
    const uint32_t frame_no = (uint32_t)(k + 1), freed = (uint32_t)(k + 1 - S);
    int rc;
    /* march stream: my partial slot has been merged by everyone (frame k - S), then K1, then "ready" to all */
    if (k >= S && (rc = vr_stream_wait(c, flag(rank, READ_DONE + s * 32), n, freed, 10000, march_stream)) != VR_OK) return rc;
    if ((rc = vr_render_frame(c, s, march_stream)) != VR_OK) return rc;
    void *sig[33];
    for (uint32_t r = 0; r < n; r++) sig[r] = flag(r, READY + s * 32 + rank);
    if ((rc = vr_stream_signal(c, sig, n, frame_no, march_stream)) != VR_OK) return rc;
    /* merge stream: every partial of frame k complete, image slot s presented (frame k - S), then K2 on my range */
    if ((rc = vr_stream_wait(c, flag(rank, READY + s * 32), n, frame_no, 10000, merge_stream)) != VR_OK) return rc;
    if (k >= S && (rc = vr_stream_wait(c, flag(rank, CONSUMED + s), 1, freed, 10000, merge_stream)) != VR_OK) return rc;
    {
        /* Outside the screen rectangles of the sub-boxes every partial is the identity and the merged pixel the clear
         * colour.  The first frame into an image slot (k < S) merges the whole frame; later ones only the bounding
         * rectangle of what this frame and the frame before it in the slot (k - S) can have drawn -- the same region on
         * every rank, a pure function of the two cameras -- dealt to the ranks in bands of 8 rows. */
        uint32_t region[4];
        vr_sort_last_region(rects, n, c->W, c->H, k < S || !c->tune_sl_region, c->sl_region[s], region);
        if ((rc = vr_composite_region(c, parts, rects, n, region, rank, n, w->image[s], merge_stream)) != VR_OK) return rc;
    }
    for (uint32_t r = 0; r < n; r++) sig[r] = flag(r, READ_DONE + s * 32 + rank);
    sig[n] = flag(disp, RANGE_DONE + s * 32 + rank);
    if ((rc = vr_stream_signal(c, sig, n + 1, frame_no, merge_stream)) != VR_OK) return rc;
    if (rank == disp) {
        /* the frame is complete in image slot s once every range has been merged; then the slot is free again */
        if ((rc = vr_stream_wait(c, flag(rank, RANGE_DONE + s * 32), n, frame_no, 10000, present_stream)) != VR_OK) return rc;
        for (uint32_t r = 0; r < n; r++) sig[r] = flag(r, CONSUMED + s);
        if ((rc = vr_stream_signal(c, sig, n, frame_no, present_stream)) != VR_OK) return rc;
    }
    return VR_OK;
}

extern "C" int vr_partition_screen_rect(const vr_ctx *c, const uint32_t full_dims[3], const uint32_t org[3], const uint32_t dims[3],
                                        uint32_t rect[4])
{
    if (!c || !full_dims || !org || !dims || !rect) return VR_ERR_INVALID;
    for (int i = 0; i < 3; i++)
        if (full_dims[i] == 0 || dims[i] == 0 || org[i] + dims[i] > full_dims[i]) return VR_ERR_INVALID;
    subbox_screen_rect(c, full_dims, org, dims, rect);
    return VR_OK;
}

extern "C" int vr_composite_rects(vr_ctx *c, const void *const *parts, const uint32_t *rects, uint32_t nparts, uint64_t pixel_begin,
                                  uint64_t pixel_end, void *out_rgba8, void *cuda_stream)
{
    if (!c || !parts || nparts == 0) return fail(c, VR_ERR_INVALID, "vr_composite: bad arguments");
    if (pixel_end > (uint64_t)c->W * c->H || pixel_begin > pixel_end) return fail(c, VR_ERR_INVALID, "vr_composite: pixel range outside the frame");
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
    c->last_stream = s;
    uchar4 *out = out_rgba8 ? reinterpret_cast<uchar4 *>(out_rgba8) : (c->ext_image ? c->ext_image : c->d_image);
    VR_CUDA(c, vr_launch_composite(reinterpret_cast<const float4 *const *>(parts), rects, nparts, pixel_begin, pixel_end, c->W, out, nullptr, s));
    return VR_OK;
}
