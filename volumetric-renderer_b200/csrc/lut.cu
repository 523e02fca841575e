/*
 * lut.cu -- the transfer function built on the device.
 *
 * The reference builds its look-up table on the host -- Gradient::discretize (src/ui/components/gradient.cpp:90-108)
 * fills a vector of packed R8G8B8A8 texels from the editor's markers -- and uploads it as an R8G8B8A8_SRGB 1-D image
 * (src/rendering/offscreen_pass.cpp:279-288, :1076) that the sampler decodes and filters.  Here:
 *
 *   k_gradient_lut   Gradient::discretize on the device: the same fp32 operations in the same order as
 *                    vr_gradient_discretize (hostside.cpp), so the texels are bit-identical;
 *   k_lut_tables     from packed texels (host- or device-made): the sRGB decode, the edge-padded fp32 table of the exact
 *                    sampler, its differences, the packed fp16 table of the hardware sampler, and the running count of
 *                    texels with alpha != 0 that lets k_classify decide a cell with two reads instead of a scan.
 *
 * The sRGB decode is a 256-entry table made once on the host with the arithmetic the oracle uses (double pow): the
 * device pow differs in the last bits and the exact sampler is bit-identical to the oracle.
 * Compiled with -fmad=false like the rest: P + 0.5 * D rounds twice, as the host code it replaces did.
 */
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "vr_ctx.h"

namespace
{
__constant__ float c_srgb[256];

/* sample_markers (gradient.cpp:471-485): clamp, std::lower_bound on the location, lerp a (1 - t) + b t (:13-17) */
template <int STRIDE>
__device__ void sample_markers(const float *m, uint32_t n, float location, float *out)
{
    location = location < 0.0f ? 0.0f : (location > 1.0f ? 1.0f : location);
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = lo + (hi - lo) / 2;
        if (m[mid * STRIDE] < location) lo = mid + 1;
        else hi = mid;
    }
    if (lo == 0 || lo == n) {
        const float *row = m + (lo == 0 ? 0 : (n - 1)) * STRIDE;
        for (int c = 1; c < STRIDE; c++) out[c - 1] = row[c];
        return;
    }
    const float *prev = m + (lo - 1) * STRIDE, *curr = m + lo * STRIDE;
    const float t = (location - prev[0]) / (curr[0] - prev[0]);
    for (int c = 1; c < STRIDE; c++) out[c - 1] = prev[c] * (1.0f - t) + curr[c] * t;
}

__device__ uint32_t unorm8_sat(float v) /* IM_F32_TO_INT8_SAT */
{
    v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
    return (uint32_t)(int)(v * 255.0f + 0.5f);
}

/* one block; markers = nc rows {location, r, g, b} then na rows {location, a} */
__global__ void __launch_bounds__(256) k_gradient_lut(const float *__restrict__ markers, uint32_t nc, uint32_t na, uint32_t count,
                                                      float *__restrict__ loc, uint32_t *__restrict__ out)
{
    /* gradient.cpp:92-95,:106: the texel centres are accumulated by repeated addition -- a serial chain, kept */
    if (threadIdx.x == 0) {
        const float stride = 1.0f / (float)count;
        float location = stride / 2.0f;
        for (uint32_t i = 0; i < count; i++) {
            loc[i] = location;
            location += stride;
        }
    }
    __syncthreads();
    const float *cm = markers, *am = markers + 4 * nc;
    for (uint32_t i = threadIdx.x; i < count; i += blockDim.x) {
        float rgb[3], a;
        sample_markers<4>(cm, nc, loc[i], rgb);
        sample_markers<2>(am, na, loc[i], &a);
        /* ImGui::ColorConvertFloat4ToU32, R in the low byte */
        out[i] = unorm8_sat(rgb[0]) | (unorm8_sat(rgb[1]) << 8) | (unorm8_sat(rgb[2]) << 16) | (unorm8_sat(a) << 24);
    }
}

/* one block */
__global__ void __launch_bounds__(256) k_lut_tables(const uint32_t *__restrict__ texels, uint32_t n, float4 *__restrict__ pad,
                                                    float4 *__restrict__ dif, uint4 *__restrict__ h, uint32_t *__restrict__ nz)
{
    /* P[k] = L[clamp(k - 1)]: the two taps of the CLAMP_TO_EDGE sampler need no clamp */
    for (uint32_t k = threadIdx.x; k < n + 3; k += blockDim.x) {
        long j = (long)k - 1;
        j = j < 0 ? 0 : (j > (long)n - 1 ? (long)n - 1 : j);
        const uint32_t p = texels[j];
        pad[k] = make_float4(c_srgb[p & 0xff], c_srgb[(p >> 8) & 0xff], c_srgb[(p >> 16) & 0xff], (float)((p >> 24) & 0xff) / 255.0f);
    }
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < n + 2; k += blockDim.x) {
        const float4 a = pad[k], b = pad[k + 1];
        const float4 d = make_float4(b.x - a.x, b.y - a.y, b.z - a.z, b.w - a.w);
        dif[k] = d;
        /* hardware sampler: one 16-byte entry {P + D / 2, D} of eight fp16 -> one LDS.128 per sample (march.cu lut_eval) */
        const float e[8] = {a.x + 0.5f * d.x, a.y + 0.5f * d.y, a.z + 0.5f * d.z, a.w + 0.5f * d.w, d.x, d.y, d.z, d.w};
        uint32_t w[4];
        for (int j = 0; j < 4; j++)
            w[j] = (uint32_t)__half_as_ushort(__float2half_rn(e[2 * j])) | ((uint32_t)__half_as_ushort(__float2half_rn(e[2 * j + 1])) << 16);
        h[k] = make_uint4(w[0], w[1], w[2], w[3]);
    }
    /* nz[k] = padded texels before k whose alpha is not exactly 0 */
    if (threadIdx.x == 0) {
        uint32_t acc = 0;
        for (uint32_t k = 0; k < n + 3; k++) {
            nz[k] = acc;
            acc += pad[k].w != 0.0f;
        }
        nz[n + 3] = acc;
    }
}
/* Pre-integrated classification table (include/vr.h, vr_set_preintegration): entry (back j = blockIdx.x, front i) =
 * M sub-samples of the segment whose LUT coordinate runs linearly from i + 1/2 to j + 1/2, each classified like a sample
 * of the march (padded fp32 table, linear), composited front to back with a' = 1 - (1 - a)^(1/M). */
__global__ void __launch_bounds__(256) k_preint_table(const float4 *__restrict__ pad, uint32_t n, uint32_t substeps, float4 *__restrict__ table,
                                                      uint2 *__restrict__ table_h)
{
    const uint32_t j = blockIdx.x;
    const float inv_m = 1.0f / (float)substeps;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        float cr = 0.f, cg = 0.f, cb = 0.f, T = 1.f;
        for (uint32_t m = 0; m < substeps; m++) {
            const float s = ((float)i + 0.5f) + ((float)m + 0.5f) * inv_m * ((float)j - (float)i);
            /* 1-D sampler, CLAMP_TO_EDGE: u = s - 1/2 in [-1, n]; padded texel k = floor(u) + 1 */
            float u = s - 0.5f;
            u = fminf(fmaxf(u, -1.0f), (float)n);
            const float fu = floorf(u);
            const int k = (int)fu + 1;
            const float w = u - fu;
            const float4 a = pad[k], c = pad[k + 1];
            const float sa = fmaf(w, c.w - a.w, a.w);
            const float ac = 1.0f - powf(1.0f - sa, inv_m);
            const float ta = T * ac;
            cr = fmaf(ta, fmaf(w, c.x - a.x, a.x), cr);
            cg = fmaf(ta, fmaf(w, c.y - a.y, a.y), cg);
            cb = fmaf(ta, fmaf(w, c.z - a.z, a.z), cb);
            T = fmaf(-T, ac, T);
        }
        table[(size_t)j * n + i] = make_float4(cr, cg, cb, 1.0f - T);
        /* what the march samples: four fp16 (8-byte texels filter at twice the rate of 16-byte ones; values in [0, 1],
         * error <= 5e-4, the precision of the march's own fp16 transfer-function table) */
        table_h[(size_t)j * n + i] = make_uint2((uint32_t)__half_as_ushort(__float2half_rn(cr)) | ((uint32_t)__half_as_ushort(__float2half_rn(cg)) << 16),
                                                (uint32_t)__half_as_ushort(__float2half_rn(cb)) | ((uint32_t)__half_as_ushort(__float2half_rn(1.0f - T)) << 16));
    }
}
}  // namespace

cudaError_t vr_lut_preint_table(const float4 *d_pad, uint32_t n, uint32_t substeps, float4 *d_table, uint2 *d_table_h, cudaStream_t s)
{
    k_preint_table<<<n, 256, 0, s>>>(d_pad, n, substeps, d_table, d_table_h);
    return cudaGetLastError();
}

cudaError_t vr_lut_upload_srgb_table(const float table[256]) { return cudaMemcpyToSymbol(c_srgb, table, sizeof(float) * 256); }

cudaError_t vr_lut_gradient(const float *d_markers, uint32_t nc, uint32_t na, uint32_t count, float *d_loc, uint32_t *d_texels, cudaStream_t s)
{
    k_gradient_lut<<<1, 256, 0, s>>>(d_markers, nc, na, count, d_loc, d_texels);
    return cudaGetLastError();
}

cudaError_t vr_lut_tables(const uint32_t *d_texels, uint32_t n, float4 *d_pad, float4 *d_dif, uint4 *d_h, uint32_t *d_nz, cudaStream_t s)
{
    k_lut_tables<<<1, 256, 0, s>>>(d_texels, n, d_pad, d_dif, d_h, d_nz);
    return cudaGetLastError();
}
