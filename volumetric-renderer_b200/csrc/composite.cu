/*
 * composite.cu -- K2: view-ordered `over` of partial images (sort-last), fused with the
 * display blend and RGBA8 store (src/rendering/offscreen_pass.cpp:715-725,:170-173,:293).
 *
 * Not in the reference (it is single-GPU).  A partial is what res/shaders/volume.frag:44-50
 * leaves in out_color for the samples one rank owns: premultiplied rgb and alpha = 1 - T.
 * Front-to-back accumulation is associative, so merging partials in visiting order
 *     C += T * C_k ;  T *= (1 - A_k)
 * equals the single-GPU loop up to fp32 re-association.
 *
 * The partials are read through plain global loads, 16 bytes per pixel per part; the pointers
 * may be local HBM or peer memory mapped over NVLink (cudaIpcOpenMemHandle /
 * cudaDeviceEnablePeerAccess), in which case the exchange and the merge are this one kernel:
 * no staging copy, the NVLink transfer overlaps the arithmetic of the other pixels.
 */
#include <cuda_runtime.h>

#include "vr_ctx.h"

#define VR_MAX_PARTS 64

namespace
{
struct PartTable {
    const float4 *p[VR_MAX_PARTS];
};
/* Pixel rectangle (x0, y0, x1, y1, exclusive) outside which a partial is the identity (0, 0, 0, 0): the screen
 * bounding box of that rank's sub-box.  Such pixels are neither stored by K1 nor read here ("send
 * bounding rectangles, not full frames"). */
struct RectTable {
    uint32_t r[VR_MAX_PARTS][4];
};

__device__ __forceinline__ float4 ld_stream(const float4 *ptr)
{
    /* read-once data: do not allocate in L1 */
    float4 v;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(ptr));
    return v;
}

template <bool RECTS>
__global__ void __launch_bounds__(256) k_composite(const __grid_constant__ PartTable tab, const __grid_constant__ RectTable rects,
                                                   uint32_t nparts, uint64_t begin, uint64_t end, uint32_t W, uchar4 *out, float4 *raw_out)
{
    for (uint64_t i = begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < end;
         i += (uint64_t)gridDim.x * blockDim.x) {
        float cr = 0.f, cg = 0.f, cb = 0.f, T = 1.f;
        uint32_t x = 0, y = 0;
        if (RECTS) y = (uint32_t)(i / W), x = (uint32_t)(i - (uint64_t)y * W);
        /* issue up to 8 loads before the dependent chain */
        for (uint32_t k0 = 0; k0 < nparts; k0 += 8) {
            float4 v[8];
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (k0 + k < nparts) {
                    v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (!RECTS || (x >= rects.r[k0 + k][0] && y >= rects.r[k0 + k][1] && x < rects.r[k0 + k][2] && y < rects.r[k0 + k][3]))
                        v[k] = ld_stream(tab.p[k0 + k] + i);
                }
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (k0 + k < nparts) {
                    cr = fmaf(T, v[k].x, cr);
                    cg = fmaf(T, v[k].y, cg);
                    cb = fmaf(T, v[k].z, cb);
                    T *= (1.0f - v[k].w);
                }
        }
        const float A = 1.0f - T, iA = 1.0f - A;
        if (raw_out) raw_out[i] = make_float4(cr, cg, cb, A);
        const float ov[4] = {cr * A + VR_CLEAR_RGB * iA, cg * A + VR_CLEAR_RGB * iA, cb * A + VR_CLEAR_RGB * iA,
                             A * A + 1.0f * iA};
        unsigned char q[4];
#pragma unroll
        for (int c = 0; c < 4; c++) {
            float x = ov[c];
            x = x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x);
            if (!(x == x)) x = 0.0f;
            q[c] = (unsigned char)(x * 255.0f + 0.5f);
        }
        out[i] = make_uchar4(q[0], q[1], q[2], q[3]);
    }
}
/* The same merge over a pixel REGION dealt to the ranks in bands of 8 rows (band b of the region belongs to rank b % n):
 * what one rank of a sort-last frame merges when only the region that can differ from the clear colour is touched
 * (api.cu vr_sort_last_frame).  The cube sits in the middle of the frame: contiguous row ranges of the whole frame leave
 * the merging to the ranks of the middle, bands of the region give every rank the same share. */
__global__ void __launch_bounds__(256) k_composite_region(const __grid_constant__ PartTable tab, const __grid_constant__ RectTable rects, uint32_t nparts,
                                                          uint32_t rx0, uint32_t ry0, uint32_t rx1, uint32_t ry1, uint32_t rank, uint32_t nranks,
                                                          uint32_t my_bands, uint32_t W, uchar4 *out)
{
    const uint32_t rw = rx1 - rx0;
    const uint64_t total = (uint64_t)my_bands * 8u * rw;
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < total; j += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t line = (uint32_t)(j / rw), col = (uint32_t)(j - (uint64_t)line * rw);
        const uint32_t y = ry0 + ((line >> 3) * nranks + rank) * 8u + (line & 7u), x = rx0 + col;
        if (y >= ry1) continue;
        const uint64_t i = (uint64_t)y * W + x;
        float cr = 0.f, cg = 0.f, cb = 0.f, T = 1.f;
        for (uint32_t k0 = 0; k0 < nparts; k0 += 8) {
            float4 v[8];
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (k0 + k < nparts) {
                    v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (x >= rects.r[k0 + k][0] && y >= rects.r[k0 + k][1] && x < rects.r[k0 + k][2] && y < rects.r[k0 + k][3])
                        v[k] = ld_stream(tab.p[k0 + k] + i);
                }
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (k0 + k < nparts) {
                    cr = fmaf(T, v[k].x, cr);
                    cg = fmaf(T, v[k].y, cg);
                    cb = fmaf(T, v[k].z, cb);
                    T *= (1.0f - v[k].w);
                }
        }
        const float A = 1.0f - T, iA = 1.0f - A;
        const float ov[4] = {cr * A + VR_CLEAR_RGB * iA, cg * A + VR_CLEAR_RGB * iA, cb * A + VR_CLEAR_RGB * iA, A * A + 1.0f * iA};
        unsigned char q[4];
#pragma unroll
        for (int c = 0; c < 4; c++) {
            float v1 = ov[c];
            v1 = v1 < 0.0f ? 0.0f : (v1 > 1.0f ? 1.0f : v1);
            if (!(v1 == v1)) v1 = 0.0f;
            q[c] = (unsigned char)(v1 * 255.0f + 0.5f);
        }
        out[i] = make_uchar4(q[0], q[1], q[2], q[3]);
    }
}
}  // namespace

cudaError_t vr_launch_composite_region(const float4 *const *parts, const uint32_t *rects, uint32_t nparts, const uint32_t region[4], uint32_t rank,
                                       uint32_t nranks, uint32_t W, uchar4 *out, cudaStream_t s)
{
    if (nparts == 0 || nparts > VR_MAX_PARTS || !rects || nranks == 0 || rank >= nranks) return cudaErrorInvalidValue;
    if (region[2] <= region[0] || region[3] <= region[1]) return cudaSuccess;
    const uint32_t bands = (region[3] - region[1] + 7u) / 8u;
    const uint32_t my_bands = rank < bands ? (bands - rank + nranks - 1u) / nranks : 0u;
    if (my_bands == 0) return cudaSuccess;
    PartTable tab;
    RectTable rt = {};
    for (uint32_t k = 0; k < VR_MAX_PARTS; k++) tab.p[k] = k < nparts ? parts[k] : nullptr;
    for (uint32_t k = 0; k < nparts; k++)
        for (int j = 0; j < 4; j++) rt.r[k][j] = rects[4 * k + j];
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint64_t blocks = ((uint64_t)my_bands * 8u * (region[2] - region[0]) + 255) / 256;
    if (blocks > (uint64_t)sms * 8) blocks = (uint64_t)sms * 8;
    k_composite_region<<<(unsigned)blocks, 256, 0, s>>>(tab, rt, nparts, region[0], region[1], region[2], region[3], rank, nranks, my_bands, W, out);
    return cudaGetLastError();
}

cudaError_t vr_launch_composite(const float4 *const *parts, const uint32_t *rects, uint32_t nparts, uint64_t begin, uint64_t end,
                                uint32_t W, uchar4 *out, float4 *raw_out, cudaStream_t s)
{
    if (nparts == 0 || nparts > VR_MAX_PARTS) return cudaErrorInvalidValue;
    if (end <= begin) return cudaSuccess;
    PartTable tab;
    RectTable rt = {};
    for (uint32_t k = 0; k < VR_MAX_PARTS; k++) tab.p[k] = k < nparts ? parts[k] : nullptr;
    if (rects)
        for (uint32_t k = 0; k < nparts; k++)
            for (int j = 0; j < 4; j++) rt.r[k][j] = rects[4 * k + j];
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint64_t blocks = (end - begin + 255) / 256;
    if (blocks > (uint64_t)sms * 8) blocks = (uint64_t)sms * 8;
    if (rects) k_composite<true><<<(unsigned)blocks, 256, 0, s>>>(tab, rt, nparts, begin, end, W, out, raw_out);
    else k_composite<false><<<(unsigned)blocks, 256, 0, s>>>(tab, rt, nparts, begin, end, W, out, raw_out);
    return cudaGetLastError();
}
