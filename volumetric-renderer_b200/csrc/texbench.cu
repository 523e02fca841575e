/*
 * texbench.cu -- the texture-fetch roofline denominator.
 *
 * MEASURED_PEAKS.json has no texture-fetch figure, so the build measures it: a 32^3 volume
 * of the requested texel type (L1-resident after the first touches) is fetched with tex3D
 * from every SM, addresses laid out like the march's (a warp = 8x4 neighbouring rays
 * advancing along z), 8 independent fetches in flight per thread and no other work.
 * The result (fetches/s) is what K1's executed fetch rate is divided by.
 */
#include <cuda_runtime.h>

#include <vector>

#include "vr_ctx.h"

namespace
{
template <bool PAIR>
__global__ void __launch_bounds__(256) k_texbench(cudaTextureObject_t tex, int iters, float step, float *sink)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    /* 8x4 rays one texel apart, each warp at its own offset */
    const float inv = 1.0f / 32.0f;
    float x = ((lane & 7) + 0.37f + (warp & 3) * 6.0f) * inv;
    float y = ((lane >> 3) + 0.61f + (warp >> 2) * 9.0f + (blockIdx.x & 7)) * inv;
    float z = 0.013f * (blockIdx.x & 15);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (PAIR) {
                /* the paired layout's fetch: one bilinear fetch of a two-channel layered texel + the z-lerp */
                const float w = z * 31.0f, fl = floorf(w);
                const float2 t = tex2DLayered<float2>(tex, x, y, (int)fl);
                acc[k] += fmaf(w - fl, t.y - t.x, t.x);
            } else {
                acc[k] += tex3D<float>(tex, x, y, z);
            }
            z += step;
            if (z > 1.0f) z -= 1.0f;
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; k++) s += acc[k];
    if (s == -12345.678f) sink[0] = s; /* never true: keeps the fetches alive */
}

/* DRAM-fed ceiling of the texture path (modes 3, 4): a paired-layout array far larger than L2 is streamed exactly
 * once with the march's mapping -- a CTA of 16x8 threads stands on a 16x8 (x STRIDE) texel tile and walks down the
 * layers, 8 fetches in flight per thread.  STRIDE 1: one bilinear fetch per texel = 4 bytes of new data per fetch
 * for u16 pairs, the ratio K1 has on C2 (4.7 B per fetch).  STRIDE 2: fetches at the corner between four texels,
 * 16 new bytes per fetch: the byte ceiling of the path. */
template <int STRIDE>
__global__ void __launch_bounds__(128) k_texstream(cudaTextureObject_t tex, int tiles_x, int layers, float inv_w, float inv_h, float *sink)
{
    const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int px = tx * 16 + (warp & 1) * 8 + (lane & 7), py = ty * 8 + (warp >> 1) * 4 + (lane >> 3);
    const float x = ((float)(px * STRIDE) + (STRIDE == 1 ? 0.5f : 1.0f)) * inv_w, y = ((float)(py * STRIDE) + (STRIDE == 1 ? 0.5f : 1.0f)) * inv_h;
    float acc = 0.f;
    for (int l = 0; l < layers; l += 8) {
        float2 t[8];
#pragma unroll
        for (int k = 0; k < 8; k++) t[k] = tex2DLayered<float2>(tex, x, y, l + k);
#pragma unroll
        for (int k = 0; k < 8; k++) acc += fmaf(0.37f, t[k].y - t[k].x, t[k].x);
    }
    if (acc == -12345.678f) sink[0] = acc;
}

__global__ void k_texstream_fill(cudaSurfaceObject_t surf, int w, int h, int layers, int esz)
{
    const size_t n = (size_t)w * h * layers;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int x = (int)(i % w), y = (int)((i / w) % h), l = (int)(i / ((size_t)w * h));
        uint32_t v = (uint32_t)i * 2654435761u;
        v ^= v >> 15;
        if (esz == 4) surf2DLayeredwrite(make_ushort2((unsigned short)v, (unsigned short)(v >> 16)), surf, x * 4, y, l);
        else surf2DLayeredwrite(make_uchar2((unsigned char)v, (unsigned char)(v >> 8)), surf, x * 2, y, l);
    }
}

int tex_stream_rate(int device, int dtype, int stride, double *fetches_per_s)
{
    if (dtype == VR_F32) return VR_ERR_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) return VR_ERR_CUDA;
    const int W = 2048, H = 2048, esz = dtype == VR_U8 ? 2 : 4;
    const int layers = (int)(((size_t)1 << 30) / ((size_t)W * H * esz)); /* 1 GiB of texels, 8x the L2 */
    cudaChannelFormatDesc cd = dtype == VR_U8 ? cudaCreateChannelDesc<uchar2>() : cudaCreateChannelDesc<ushort2>();
    cudaArray_t arr = nullptr;
    cudaTextureObject_t tex = 0;
    cudaSurfaceObject_t surf = 0;
    float *sink = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = VR_ERR_CUDA;
    do {
        if (cudaMalloc3DArray(&arr, &cd, make_cudaExtent(W, H, layers), cudaArrayLayered | cudaArraySurfaceLoadStore) != cudaSuccess) break;
        cudaResourceDesc rd = {};
        rd.resType = cudaResourceTypeArray;
        rd.res.array.array = arr;
        if (cudaCreateSurfaceObject(&surf, &rd) != cudaSuccess) break;
        k_texstream_fill<<<148 * 8, 256>>>(surf, W, H, layers, esz);
        cudaTextureDesc td = {};
        td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeBorder;
        td.filterMode = cudaFilterModeLinear;
        td.readMode = cudaReadModeNormalizedFloat;
        td.normalizedCoords = 1;
        if (cudaCreateTextureObject(&tex, &rd, &td, nullptr) != cudaSuccess) break;
        if (cudaMalloc(&sink, sizeof(float)) != cudaSuccess) break;
        cudaEventCreate(&e0), cudaEventCreate(&e1);
        const int tiles_x = W / stride / 16, tiles_y = H / stride / 8;
        auto kern = stride == 1 ? k_texstream<1> : k_texstream<2>;
        kern<<<tiles_x * tiles_y, 128>>>(tex, tiles_x, layers, 1.0f / W, 1.0f / H, sink); /* warm-up */
        double best = 0.0;
        for (int rep = 0; rep < 4; rep++) {
            cudaEventRecord(e0);
            kern<<<tiles_x * tiles_y, 128>>>(tex, tiles_x, layers, 1.0f / W, 1.0f / H, sink);
            cudaEventRecord(e1);
            if (cudaEventSynchronize(e1) != cudaSuccess) break;
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            const double rate = (double)tiles_x * tiles_y * 128.0 * layers / (ms * 1e-3);
            if (rate > best) best = rate;
        }
        if (cudaGetLastError() != cudaSuccess || best == 0.0) break;
        *fetches_per_s = best;
        rc = VR_OK;
    } while (0);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (sink) cudaFree(sink);
    if (tex) cudaDestroyTextureObject(tex);
    if (surf) cudaDestroySurfaceObject(surf);
    if (arr) cudaFreeArray(arr);
    return rc;
}
}  // namespace

extern "C" int vr_bench_tex_peak(int device, int dtype, int hw_filter, double *fetches_per_s)
{
    if (!fetches_per_s || dtype < VR_U8 || dtype > VR_F32 || hw_filter < 0 || hw_filter > 4) return VR_ERR_INVALID;
    if (hw_filter >= 3) return tex_stream_rate(device, dtype, hw_filter == 3 ? 1 : 2, fetches_per_s);
    const bool pair = hw_filter == 2;
    if (pair && dtype == VR_F32) return VR_ERR_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) return VR_ERR_CUDA;
    const int N = 32;
    const size_t esz = (dtype == VR_U8 ? 1 : (dtype == VR_U16 ? 2 : 4)) * (pair ? 2 : 1);
    std::vector<unsigned char> host((size_t)N * N * N * esz);
    uint32_t h = 0x1234567u;
    for (size_t i = 0; i < (size_t)N * N * N * (pair ? 2 : 1); i++) {
        h = h * 1664525u + 1013904223u;
        if (dtype == VR_U8) host[i] = (unsigned char)(h >> 24);
        else if (dtype == VR_U16) reinterpret_cast<unsigned short *>(host.data())[i] = (unsigned short)(h >> 16);
        else reinterpret_cast<float *>(host.data())[i] = (float)(h >> 8) * (1.0f / 16777216.0f);
    }
    cudaChannelFormatDesc cd = dtype == VR_U8    ? cudaCreateChannelDesc<unsigned char>()
                               : dtype == VR_U16 ? cudaCreateChannelDesc<unsigned short>()
                                                 : cudaCreateChannelDesc<float>();
    if (pair) cd = dtype == VR_U8 ? cudaCreateChannelDesc<uchar2>() : cudaCreateChannelDesc<ushort2>();
    cudaArray_t arr = nullptr;
    cudaTextureObject_t tex = 0;
    float *sink = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = VR_ERR_CUDA;
    do {
        if (cudaMalloc3DArray(&arr, &cd, make_cudaExtent(N, N, N), pair ? cudaArrayLayered : 0) != cudaSuccess) break;
        cudaMemcpy3DParms cp = {};
        cp.srcPtr = make_cudaPitchedPtr(host.data(), N * esz, N, N);
        cp.dstArray = arr;
        cp.extent = make_cudaExtent(N, N, N);
        cp.kind = cudaMemcpyHostToDevice;
        if (cudaMemcpy3D(&cp) != cudaSuccess) break;
        cudaResourceDesc rd = {};
        rd.resType = cudaResourceTypeArray;
        rd.res.array.array = arr;
        cudaTextureDesc td = {};
        td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeBorder;
        td.filterMode = hw_filter ? cudaFilterModeLinear : cudaFilterModePoint;
        td.readMode = dtype == VR_F32 ? cudaReadModeElementType : cudaReadModeNormalizedFloat;
        td.normalizedCoords = 1;
        if (cudaCreateTextureObject(&tex, &rd, &td, nullptr) != cudaSuccess) break;
        if (cudaMalloc(&sink, sizeof(float)) != cudaSuccess) break;
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        const int blocks = sms * 8, iters = 2048;
        const float step = 2.56f / 32.0f * 0.125f;
        cudaEventCreate(&e0), cudaEventCreate(&e1);
        auto kern = pair ? k_texbench<true> : k_texbench<false>;
        kern<<<blocks, 256>>>(tex, 64, step, sink); /* warm-up */
        double best = 0.0;
        for (int rep = 0; rep < 5; rep++) {
            cudaEventRecord(e0);
            kern<<<blocks, 256>>>(tex, iters, step, sink);
            cudaEventRecord(e1);
            if (cudaEventSynchronize(e1) != cudaSuccess) break;
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            const double rate = (double)blocks * 256.0 * iters * 8.0 / (ms * 1e-3);
            if (rate > best) best = rate;
        }
        if (cudaGetLastError() != cudaSuccess || best == 0.0) break;
        *fetches_per_s = best;
        rc = VR_OK;
    } while (0);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (sink) cudaFree(sink);
    if (tex) cudaDestroyTextureObject(tex);
    if (arr) cudaFreeArray(arr);
    return rc;
}
