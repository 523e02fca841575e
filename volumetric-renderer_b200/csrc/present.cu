/*
 * present.cu -- K3: the step after the path.  The reference shows the offscreen image in an ImGui
 * viewport through a sampler with LINEAR mag/min filters and CLAMP_TO_EDGE addressing
 * (src/rendering/offscreen_pass.cpp:361-375, drawn by src/ui/main_window.cpp:157-165); the image is
 * R8G8B8A8_UNORM (:293).  Headless equivalent: resample the RGBA8 frame into a caller's RGBA8
 * buffer of another size with exactly that filter.
 *
 * Per destination pixel (x, y): unnormalised source coordinate u = (x + 1/2) * sw / dw - 1/2 (Vulkan
 * texel coordinates of the pixel centre, shifted for linear filtering), i0 = floor(u), alpha = u - i0,
 * the four texels at (i0, i0 + 1) x (j0, j0 + 1) clamped to the edge, each channel decoded c / 255,
 *     t = (1-a)(1-b) t00 + a(1-b) t10 + (1-a) b t01 + a b t11        (Vulkan 1.0, 15.8.3)
 * evaluated left to right in fp32 with single roundings (this file is built with -fmad=false), and
 * encoded as round(clamp(t) * 255).  Same size in and out is the identity.  HBM-bound: 4 B written
 * and at most 16 B read per pixel (neighbouring pixels share texels through L1/L2).
 */
#include <cuda_runtime.h>

#include "vr_ctx.h"

namespace
{
__global__ void __launch_bounds__(256) k_present_bilinear(const uchar4 *__restrict__ src, uint32_t sw, uint32_t sh, uchar4 *__restrict__ dst,
                                                          uint32_t dw, uint32_t dh, size_t dst_pitch_px, float kx, float ky)
{
    const uint32_t x = blockIdx.x * 32u + (threadIdx.x & 31u), y = blockIdx.y * 8u + (threadIdx.x >> 5);
    if (x >= dw || y >= dh) return;
    const float u = ((float)x + 0.5f) * kx - 0.5f, v = ((float)y + 0.5f) * ky - 0.5f;
    const float fu = floorf(u), fv = floorf(v);
    const float a = u - fu, b = v - fv;
    const int i0 = (int)fu, j0 = (int)fv;
    const int ia = min(max(i0, 0), (int)sw - 1), ib = min(max(i0 + 1, 0), (int)sw - 1);
    const int ja = min(max(j0, 0), (int)sh - 1), jb = min(max(j0 + 1, 0), (int)sh - 1);
    const uchar4 t00 = __ldg(src + (size_t)ja * sw + ia), t10 = __ldg(src + (size_t)ja * sw + ib);
    const uchar4 t01 = __ldg(src + (size_t)jb * sw + ia), t11 = __ldg(src + (size_t)jb * sw + ib);
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    auto filt = [&](unsigned char c00, unsigned char c10, unsigned char c01, unsigned char c11) {
        float t = w00 * ((float)c00 / 255.0f);
        t = t + w10 * ((float)c10 / 255.0f);
        t = t + w01 * ((float)c01 / 255.0f);
        t = t + w11 * ((float)c11 / 255.0f);
        t = t < 0.0f ? 0.0f : (t > 1.0f ? 1.0f : t);
        return (unsigned char)(t * 255.0f + 0.5f);
    };
    dst[(size_t)y * dst_pitch_px + x] = make_uchar4(filt(t00.x, t10.x, t01.x, t11.x), filt(t00.y, t10.y, t01.y, t11.y),
                                                    filt(t00.z, t10.z, t01.z, t11.z), filt(t00.w, t10.w, t01.w, t11.w));
}
}  // namespace

cudaError_t vr_launch_present(const uchar4 *src, uint32_t sw, uint32_t sh, uchar4 *dst, uint32_t dw, uint32_t dh,
                              size_t dst_pitch_px, cudaStream_t s)
{
    if (!dw || !dh) return cudaSuccess;
    if (!sw || !sh || dst_pitch_px < dw) return cudaErrorInvalidValue;
    const dim3 grid((dw + 31u) / 32u, (dh + 7u) / 8u);
    if (grid.y > 65535u) return cudaErrorInvalidConfiguration;
    k_present_bilinear<<<grid, 256, 0, s>>>(src, sw, sh, dst, dw, dh, dst_pitch_px, (float)sw / (float)dw, (float)sh / (float)dh);
    return cudaGetLastError();
}
