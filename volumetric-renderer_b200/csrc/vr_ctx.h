/*
 * vr_ctx.h -- internal state of a render context and the kernel parameter blocks.
 * Not part of the ABI (include/vr.h is).
 */
#ifndef VR_CTX_H
#define VR_CTX_H

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/vr.h"

#define VR_KERNEL_VERSION "r2.v11" /* bump when K1 changes: profiles/traffic.json is keyed by it */
#define VR_PREINT_MAX_TEXELS 1024 /* pre-integration table: n x n float4, 16 MiB at most */
#define VR_TILE 16        /* a CTA marches a 16x16 pixel tile: 8 warps of 8x4 pixels */
#define VR_CTA_THREADS 256
#define VR_BATCH 4        /* steps whose fetches are issued back to back before compositing */
#define VR_OCC_MAX_AXIS 128 /* occupancy / distance grid: at most 128^3 cells (2 MiB of distances, read through L1/L2).
                             * Finer cells dilate the occupied region less: C3 0.227 ms at 64, 0.209 at 128, 0.203 at 256
                             * (where the ranges cost 134 MB and the distance transform several ms per LUT change) */
#define VR_CLEAR_RGB 0.11f /* offscreen_pass.cpp:171 */
#define VR_SIGNAL_MAX 32   /* flags one vr_stream_signal writes */

/* Everything K1 needs, passed by value as a __grid_constant__ kernel parameter. */
struct MarchParams {
    /* S2: columns of inverse(proj*view) (fp32), pixel->NDC scale, camera */
    float c0[4], c1[4], c2[4], c3[4];
    float sx, sy;
    float cam[3];
    uint32_t W, H;
    /* S3/S5 uniforms */
    float min_density, max_density, range;
    float min_slice[3], max_slice[3];
    int steps;
    float step_size;
    /* S4: volume */
    cudaTextureObject_t tex_lin; /* normalised coords, linear, border 0, normalised-float read for u8/u16 */
    cudaTextureObject_t tex_pt;  /* texel coords, point, border 0, element read */
    /* paired layout: 2-D layered array, layer l holds the texel pair (v[w = l-1], v[w = l]) along the pair
     * axis w (zero outside); normalised in-plane coords, linear, border 0, normalised-float read.  One
     * bilinear fetch + one lerp = one sample.  The pair axis is chosen per frame (the axis the view
     * direction is closest to: consecutive samples of a ray then fall into different layers and only
     * 1/step of the layers is read from HBM).  When it is not z, the march runs in a permuted sampling
     * space (u, v, w): every per-axis field of this block is stored in that order by the host, and the
     * kernel permutes the ray's entry position and direction once (pair_axis). */
    cudaTextureObject_t tex_pair;
    cudaTextureObject_t tex_pre;     /* pre-integrated classification: n x n half4 (premultiplied rgb, opacity), x = front, y = back */
    cudaTextureObject_t tex_pair_hi; /* layers >= pair_split live in a second array (2048-layer limit of layered arrays) */
    int pair_split;                  /* 0x10000 when one array holds them all */
    int pair_axis;              /* model axis of w: 2 (identity), 1: (u,v,w) = (x,z,y), 0: (u,v,w) = (y,z,x) */
    float pair_a[3], pair_b[3]; /* u, v: local normalised coord = q * a + b;  w: local texel coord + 1.5*2^23 = q * a + b */
    float pair_zc;              /* 0.5 - forg[w]: frac = q_w * dims[w] + (pair_zc - (floor + 1)) */
    float dims[3];      /* GLOBAL grid size as float */
    int org[3];         /* global voxel index of local texel 0 (0 unless sort-last) */
    float forg[3];      /* same as float */
    float inv_ldims[3]; /* 1 / local texture size */
    float dscale;       /* normalised-float texel -> raw value: 255, 65535 or 1 */
    int dtype;          /* vr_dtype of the texels (point fetches read the element type) */
    /* sort-last ownership, texture space, half-open [lo, hi) */
    float own_lo[3], own_hi[3];
    float box_lo[3], box_hi[3]; /* bounds, slices and ownership intersected: where a step is sampled */
    /* S6: LUT, decoded to linear fp32, padded: P[k] = L[clamp(k-1, 0, n-1)], k = 0..n+2 (n+3 entries) */
    const float4 *lut_pad;
    const float4 *lut_dif; /* D[k] = P[k+1] - P[k], k = 0..n+1 (n+2 entries) */
    const uint4 *lut_h;    /* HW sampler: entry k = {P[k] rgba, D[k] rgba} as eight fp16 (n+2 entries, 16 B each) */
    uint32_t nlut;
    float lutA, lutB; /* HW sampler: padded LUT coordinate = texel * lutA + lutB */
    float lutBh;      /* lutB - 1/2 */
    /* empty-space skipping: bit per cell, x fastest */
    const uint32_t *occ;
    int occ_n[3];
    int occ_stride[3]; /* distance-map index = sum c[a] * occ_stride[a] (sampling-space order) */
    int occ_words;
    float occ_scale[3], occ_off[3]; /* cell = int(pos * occ_scale - occ_off) */
    int ess_guard;             /* empty-space jumps a lane may take ahead of one batch */
    float ext[3], inv_ext[3];  /* non-cubic voxels: physical = cube coordinate * ext; (1, 1, 1) unless spacings are honoured */
    float occ_offh[3];         /* occ_off + 1/2 */
    const uint32_t *tile_mask; /* bit per CTA tile (ty * ntx + tx): 0 = no ray of the tile meets an occupied cell; word [mask_words] != 0:
                                * ignore the mask; nullptr = no mask this frame (volume.cu k_tile_mask) */
    uint32_t mask_words;
    const unsigned char *dist; /* 8 maps (one per direction octant) x cells: edge of the largest empty cube with the cell at
                                * its corner, extending along the octant; 0 = occupied (volume.cu k_octant_distance) */
    uint32_t occ_oct[3];       /* offset of the octant's map contributed by a negative direction on sampling axis a */
    float occ_f_hi[3]; /* extent of the volume in fine-cell units (local texture) */
    /* outputs */
    uchar4 *out;
    float4 *raw;
    unsigned long long *counters; /* [4]: covered, n_ref, n_inslice, n_fetch */
    float ert_eps;
    /* screen partition (sort-first): partition tiles of pt_px pixels, round-robin */
    uint32_t pt_px, pt_nx, pt_ny, pt_rank, pt_nranks;
    /* sort-first launch over this rank's tiles only (pt_enum): the partition tiles [pt_x0, pt_x1) x [pt_y0, ..) that
     * meet the bounding box, pt_cnt owned candidates per row of them, pt_sub_x x pt_sub_y CTA tiles in each */
    uint32_t pt_enum, pt_x0, pt_x1, pt_y0, pt_rows, pt_cnt, pt_sub_x, pt_sub_y;
    uint32_t pt_m;   /* partition tile (x, y) belongs to rank (x + y * pt_m) % pt_nranks (vr_sort_first_owner) */
    uint32_t pt_cnt_clear; /* owned candidates per full row of partition tiles */
    /* CTA tiles: ntx x nty over the target; rect = tiles that can be covered (x0,y0,x1,y1, exclusive);
     * the first sp_count blocks spiral out from tile (sp_cx, sp_cy) over the square containing rect */
    uint32_t ntx, nty, sp_count;
    uint32_t nclr_x, nclr_y; /* clear-colour blocks: 64x32 pixel regions from clr_org on (pt_enum: nclr_x blocks, one per owned partition tile of clr_t) */
    uint32_t clr_org[2];     /* pixel origin of the clear regions (multiples of 64, 32) */
    uint32_t clr_t[4];       /* sort-first: partition tiles [x0, x1) x [y0, y1) that need the clear colour */
    int sp_cx, sp_cy, rect[4];
    int carveout;                                             /* shared-memory carve-out hint in percent, -1 = driver default */
    uint32_t warp_w;                                          /* warp footprint width: 4, 8 or 16 pixels */
    uint32_t tile_h;                                          /* CTA tile is VR_TILE x tile_h pixels */
    uint32_t bbox[4];                                         /* x0, y0, x1, y1 (exclusive): pixels that can be covered */
};

struct vr_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;

    uint32_t W = 0, H = 0;
    uchar4 *d_image = nullptr;                                       /* frame 0 */
    uchar4 *d_image_slot[VR_MAX_FRAMES_IN_FLIGHT] = {nullptr, nullptr, nullptr, nullptr}; /* frames 1.. (lazily) */
    float4 *d_raw = nullptr;
    float4 *d_raw_slot[VR_MAX_FRAMES_IN_FLIGHT] = {nullptr, nullptr, nullptr, nullptr}; /* float4 output of frames 1.. (lazily) */
    uchar4 *ext_image = nullptr; /* vr_set_image_target: frame 0 is written there instead (peer memory) */
    uchar4 *ext_image_slot[VR_MAX_FRAMES_IN_FLIGHT] = {nullptr, nullptr, nullptr, nullptr}; /* same for frames 1.. */
    uint32_t *d_flags = nullptr; /* VR_FLAG_COUNT counters for stream-ordered hand-off between ranks (sync.cu) */
    unsigned long long *d_counters = nullptr;
    unsigned long long *h_counters = nullptr; /* pinned */

    /* camera / uniforms (host copy of the UBO, offscreen_pass.h:29-37) */
    float view[16], proj[16], cam_pos[3];
    float min_density = 0.f, max_density = 0.f;
    float min_slice[3] = {0.f, 0.f, 0.f}, max_slice[3] = {1.f, 1.f, 1.f};
    float ray_dist = 1.8f, step_size = 0.005f;
    bool ray_valid = false; /* inverse(proj*view) exists */
    float rc[4][4];         /* columns of the inverse */
    /* non-cubic voxels (vr_set_volume_spacing; off = the reference's unit cube): the volume fills [-ext/2, ext/2].  The
     * march works in the unit cube's coordinates: view = view_in * diag(ext), cam_pos = cam_in / ext (fp32, exact for 1) */
    /* sort-last merge (vr_sort_last_frame): the region of each image slot that the last frame merged into it may have left
     * different from the clear colour -- the next frame into the slot merges the union of that and its own */
    uint32_t sl_region[VR_MAX_FRAMES_IN_FLIGHT][4] = {};
    bool honour_spacing = false;
    float spacing[3] = {1.0f, 1.0f, 1.0f};
    float file_spacing[3] = {__builtin_nanf(""), __builtin_nanf(""), __builtin_nanf("")}; /* of the NRRD loaded last */
    float ext[3] = {1.0f, 1.0f, 1.0f};
    float view_in[16], cam_in[3]; /* the camera as it was set */

    /* volume */
    int dtype = VR_F32;
    uint32_t nx = 0, ny = 0, nz = 0; /* local texture size */
    cudaArray_t array = nullptr;
    cudaTextureObject_t tex_lin = 0, tex_pt = 0;
    int layout_req = VR_LAYOUT_AUTO;  /* vr_set_volume_layout: applies to the next upload */
    int layout_used = VR_LAYOUT_AUTO; /* the request the current volume was uploaded under */
    bool last_pair = false;           /* the last frame sampled the paired layout */
    cudaArray_t pair_array[3] = {nullptr, nullptr, nullptr}; /* paired layouts along x, y, z (see MarchParams::tex_pair), resident next to `array` */
    cudaArray_t pair_array_hi[3] = {nullptr, nullptr, nullptr}; /* their layers >= pair_split[axis], if any */
    cudaTextureObject_t tex_pair[3] = {0, 0, 0}, tex_pair_hi[3] = {0, 0, 0};
    int pair_split[3] = {0x10000, 0x10000, 0x10000};
    int tune_occ_max_axis = VR_OCC_MAX_AXIS; /* VR_OCC_MAX_AXIS: cells per axis of the occupancy / distance grid */
    int tune_pair_split = 2048; /* VR_PAIR_SPLIT: layers per layered array (tests force small values) */
    int last_pair_axis = 2;

    /* sort-last */
    bool sort_last = false;
    uint32_t full[3] = {0, 0, 0}, own_org[3] = {0, 0, 0}, own_dims[3] = {0, 0, 0};
    uint32_t st_org[3] = {0, 0, 0}, st_dims[3] = {0, 0, 0}; /* the box whose voxels this rank holds (>= the owned box: vr_set_owned_box) */
    int tex_org[3] = {0, 0, 0}; /* global voxel index of local texel 0 */

    /* LUT */
    uint32_t nlut = 0;
    uint32_t *d_lut_rgba8 = nullptr; /* the packed R8G8B8A8_SRGB texels as set (or discretised on the device), nlut of them */
    float *d_lut_loc = nullptr;      /* texel centres of a device-side discretize */
    uint32_t *d_lut_nz = nullptr;    /* nz[k] = padded texels before k with alpha != 0 (nlut + 4 entries): k_classify */
    float *d_markers = nullptr;      /* gradient markers of vr_set_gradient */
    uint32_t lut_texels_cap = 0, lut_tables_cap = 0;
    /* pre-integrated classification (vr_set_preintegration): 0 = off, else sub-samples per table entry */
    uint32_t preint_substeps = 0, preint_n = 0;
    bool preint_dirty = true;
    float4 *d_preint = nullptr;        /* the table in fp32 (vr_get_preintegration_table) */
    uint2 *d_preint_h = nullptr;       /* the table as four fp16 per entry, copied into the array below */
    cudaArray_t preint_array = nullptr;
    cudaTextureObject_t preint_tex = 0;
    size_t markers_cap = 0;
    float4 *d_lut_pad = nullptr, *d_lut_dif = nullptr;
    uint4 *d_lut_h = nullptr;
    /* tuning knobs (environment: VR_BATCH, VR_CTA_THREADS, VR_BBOX, VR_WARP_W, VR_CARVEOUT, VR_ESS_MIN_EMPTY) */
    int tune_batch_ess = 4;  /* steps per batch when the occupancy grid is consulted: one look-up per batch (march.cu skip_empty) */
    int tune_ess_guard = 8;  /* jumps per batch at most */
    int tune_batch = 2, tune_threads = 128, tune_bbox = 1, tune_warp_w = 8, tune_carveout = -1;
    int tune_dirty_clear = 1; /* VR_DIRTY_CLEAR=0: restore the clear colour over the whole target every frame */
    int tune_pair_axis = -1;  /* VR_PAIR_AXIS: force the pair axis (experiments); -1 = the axis the camera looks down most */
    int tune_pair_policy = 2; /* VR_PAIR_POLICY: 0 never sample the paired layout, 1 always, 2 per frame (vr_render_frame) */
    size_t tune_pair_max_bytes = (size_t)160 << 30; /* and never more than the free memory minus a reserve */ /* AUTO layout: pair only if the paired copy fits this (VR_PAIR_MAX_MB) */
    float tune_ess_min_empty = 0.25f; /* consult the occupancy grid only if at least this share of cells is empty */

    /* occupancy */
    int occ_n[3] = {0, 0, 0};
    int occ_cell = 0; /* voxels per cell edge (power of two) */
    float2 *d_cell_range = nullptr; /* per cell (min,max) incl. 1-voxel apron and border zeros */
    uint32_t *d_occ = nullptr;
    int occ_words = 0;
    unsigned char *d_dist = nullptr; /* 8 x cells */
    /* dirty-rectangle downloads (vr_download_frame_dirty_async): where each slot's last frame can differ from the
     * clear colour, and what each known host buffer holds */
    /* what each frame slot's target holds outside the tiles its last frame marched (dirty clear, vr_render_frame):
     * valid = every pixel this context owns of `target` (and `raw`) outside `drawn` is the clear colour (zero) */
    struct SlotClear {
        bool valid = false;
        const void *target = nullptr, *raw = nullptr;
        uint32_t W = 0, H = 0, drawn[4] = {0, 0, 0, 0};
        uint32_t pt_rank = 0, pt_nranks = 1, pt_px = 0;
    } slot_clear[VR_MAX_FRAMES_IN_FLIGHT];
    vr_frame_info last_info = {};
    uint32_t slot_bbox[VR_MAX_FRAMES_IN_FLIGHT][4] = {};
    bool slot_bbox_valid[VR_MAX_FRAMES_IN_FLIGHT] = {false, false, false, false};
    struct HostTarget {
        void *ptr = nullptr;
        uint32_t W = 0, H = 0, rect[4] = {0, 0, 0, 0};
        uint64_t stamp = 0;
    } host_targets[8];
    uint64_t host_target_clock = 0;
    bool bench_fetch_only = false;   /* vr_bench_fetch_only: frames run the fetch-only build of K1 */
    uint32_t *d_sat = nullptr;       /* summed-volume table of the occupancy bits, (n+1)^3 */
    uint32_t *d_occ_list = nullptr;  /* indices of the occupied cells, + their count in d_occ_count */
    uint32_t *d_occ_count = nullptr;
    /* per frame slot two buffers of mask_words + 1 words, used in turn: the kernel that fills one zeroes the other */
    uint32_t *d_tile_mask[VR_MAX_FRAMES_IN_FLIGHT] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t tile_mask_cur[VR_MAX_FRAMES_IN_FLIGHT] = {0, 0, 0, 0};
    uint32_t tile_mask_words = 0;
    bool slot_mask_valid[VR_MAX_FRAMES_IN_FLIGHT] = {false, false, false, false};
    uint32_t slot_mask_shape[VR_MAX_FRAMES_IN_FLIGHT][3] = {}; /* tiles in x, tiles in y, tile height of the frame recorded last */
    int tune_sl_region = 1;          /* VR_SL_REGION=0: vr_sort_last_frame merges the whole frame every time */
    int tune_pair_local = 1;         /* VR_PAIR_LOCAL=0: a sort-last rank never overrides the pair axis by its entry face */
    int tune_tile_mask = 1;          /* VR_TILE_MASK=0: no screen-space mask; 2: marked in global memory directly (frames too large to stage) */
    bool occ_dirty = true;
    uint64_t occ_empty_cells = 0;

    vr_options opt;
    uint32_t pt_px = 64, pt_rank = 0, pt_nranks = 1;
    cudaStream_t last_stream = nullptr;
};

/* ---- kernels / device-side stages (each .cu implements its own) ---- */

/* K1 (march.cu) */
cudaError_t vr_launch_composite_region(const float4 *const *parts, const uint32_t *rects, uint32_t nparts, const uint32_t region[4], uint32_t rank,
                                       uint32_t nranks, uint32_t W, uchar4 *out, cudaStream_t s);
cudaError_t vr_launch_march_preint(const MarchParams &p, bool pair, int threads, cudaStream_t s);
cudaError_t vr_launch_march_probe(const MarchParams &p, bool pair, int ub, int threads, cudaStream_t s);
cudaError_t vr_launch_march(const MarchParams &p, bool exact, bool ess, bool stats, bool part, bool pair, int ub,
                            int threads, cudaStream_t s);
size_t vr_march_smem_bytes(const MarchParams &p, bool ess);

/* K0 (volume.cu) */
cudaError_t vr_k0_global_minmax(const void *dev, int dtype, size_t n, float *h_min, float *h_max, cudaStream_t s);
cudaError_t vr_k0_cell_ranges(const void *dev, int dtype, uint32_t nx, uint32_t ny, uint32_t nz, int cell,
                              const int occ_n[3], float2 *d_ranges, cudaStream_t s);
/* lut.cu: the transfer function built on the device */
cudaError_t vr_lut_upload_srgb_table(const float table[256]);
cudaError_t vr_lut_gradient(const float *d_markers, uint32_t nc, uint32_t na, uint32_t count, float *d_loc, uint32_t *d_texels, cudaStream_t s);
cudaError_t vr_lut_preint_table(const float4 *d_pad, uint32_t n, uint32_t substeps, float4 *d_table, uint2 *d_table_h, cudaStream_t s);
cudaError_t vr_lut_tables(const uint32_t *d_texels, uint32_t n, float4 *d_pad, float4 *d_dif, uint4 *d_h, uint32_t *d_nz, cudaStream_t s);
cudaError_t vr_k0_classify(const float2 *d_ranges, const int occ_n[3], const uint32_t *d_lut_nz, uint32_t nlut,
                           float min_density, float max_density, uint32_t *d_occ, int occ_words,
                           unsigned long long *d_empty_count, cudaStream_t s);
cudaError_t vr_k0_distance(const uint32_t *d_occ, const int occ_n[3], unsigned char *d_dist, uint32_t *d_sat,
                           cudaStream_t s);
/* list of the occupied cells (after k_classify); per frame: the CTA tiles whose rays can meet one of them */
cudaError_t vr_k0_list_occupied(const uint32_t *d_occ, const int occ_n[3], uint32_t *d_list, uint32_t *d_count, cudaStream_t s);
cudaError_t vr_k0_tile_mask(const float pv[16], const float lo[3], const float cell[3], const int occ_n[3], const uint32_t *d_list,
                            const uint32_t *d_count, uint32_t max_count, uint32_t W, uint32_t H, uint32_t ntx, uint32_t tile_h,
                            uint32_t *d_mask, uint32_t *d_wipe, uint32_t mask_words, bool staged, cudaStream_t s);
/* fill the layered array paired along `axis` (n_axis + 1 layers) from the linear volume; dtype u8 or u16 */
cudaError_t vr_k0_pack_pairs(const void *dev, int dtype, uint32_t nx, uint32_t ny, uint32_t nz, int axis,
                             cudaArray_t pair_array, cudaArray_t pair_array_hi, int split, cudaStream_t s);
cudaError_t vr_k0_synth(void *dev, int kind, uint32_t seed, int dtype, uint32_t nx, uint32_t ny, uint32_t nz,
                        const int org[3], const uint32_t full[3], cudaStream_t s);

/* stream-ordered flags (sync.cu) */
cudaError_t vr_launch_signal(void *const *flags, int n, uint32_t value, cudaStream_t s);
cudaError_t vr_launch_wait(const void *flags, int n, uint32_t value, uint32_t timeout_ms, unsigned long long *error_word,
                           cudaStream_t s);

/* K2 (composite.cu) */
cudaError_t vr_launch_present(const uchar4 *src, uint32_t sw, uint32_t sh, uchar4 *dst, uint32_t dw, uint32_t dh,
                              size_t dst_pitch_px, cudaStream_t s);
/* sort-first ownership: tiles are dealt along diagonals, rank = (x + y * m) % n with m the smallest number >= 3
 * that has no common factor with n -- rows are shifted against each other, so every rank gets an even share of any
 * screen rectangle whatever the number of tiles per row */
inline uint32_t vr_sort_first_stride(uint32_t n)
{
    for (uint32_t m = 3;; m++) {
        uint32_t a = m, b = n;
        while (b) {
            const uint32_t t = a % b;
            a = b, b = t;
        }
        if (a == 1) return m;
    }
}
cudaError_t vr_launch_composite(const float4 *const *d_parts_table, const uint32_t *rects, uint32_t nparts, uint64_t begin, uint64_t end,
                                uint32_t W, uchar4 *out, float4 *raw_out, cudaStream_t s);

#endif
