/*
 * tools/exp_texrate.cu -- experiment: filtered-fetch rates of the texture path on sm_100a for the
 * layouts K1 could use.  L1-resident data, all SMs, 8 independent fetches in flight per thread,
 * a warp = 8x4 neighbouring rays advancing along z (same access shape as texbench.cu).
 *   A  3-D array u16,      trilinear          (what K1 does today)
 *   B  2-D layered ushort2, bilinear           (z, z+1 pairs in one texel; z-lerp in ALU)
 *   C  3-D array ushort2,  trilinear
 *   D  2-D layered u16,    bilinear
 *   E  3-D array u8,       trilinear
 *   F  3-D array u16,      point
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/exp_texrate.bin tools/exp_texrate.cu
 */
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e_ = (x);                                                         \
        if (e_ != cudaSuccess) {                                                      \
            std::printf("%s failed: %s\n", #x, cudaGetErrorString(e_));               \
            std::exit(1);                                                             \
        }                                                                             \
    } while (0)

template <int MODE>
__global__ void __launch_bounds__(256) k_rate(cudaTextureObject_t tex, int iters, float step, int nlayer, float *sink)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float inv = 1.0f / 32.0f;
    float x = ((lane & 7) + 0.37f + (warp & 3) * 6.0f) * inv;
    float y = ((lane >> 3) + 0.61f + (warp >> 2) * 9.0f + (blockIdx.x & 7)) * inv;
    float z = 0.013f * (blockIdx.x & 15);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (MODE == 0) {
                acc[k] += tex3D<float>(tex, x, y, z);
            } else if (MODE == 1) {
                const float w = z * (float)(nlayer - 1);
                const float fl = floorf(w);
                const float2 v = tex2DLayered<float2>(tex, x, y, (int)fl);
                acc[k] += fmaf(w - fl, v.y - v.x, v.x);
            } else if (MODE == 2) {
                const float2 v = tex3D<float2>(tex, x, y, z);
                acc[k] += v.x + v.y;
            } else {
                acc[k] += tex2DLayered<float>(tex, x, y, (int)(z * (float)(nlayer - 1)));
            }
            z += step;
            if (z > 1.0f) z -= 1.0f;
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; k++) s += acc[k];
    if (s == -12345.678f) sink[0] = s;
}

static double run(int mode, int kernel_mode, cudaChannelFormatDesc cd, size_t esz, bool layered, bool linear, int N, int NZ)
{
    std::vector<unsigned char> host((size_t)N * N * NZ * esz);
    unsigned h = 0x1234567u;
    for (auto &b : host) {
        h = h * 1664525u + 1013904223u;
        b = (unsigned char)(h >> 24);
    }
    cudaArray_t arr;
    CK(cudaMalloc3DArray(&arr, &cd, make_cudaExtent(N, N, NZ), layered ? cudaArrayLayered : 0));
    cudaMemcpy3DParms cp = {};
    cp.srcPtr = make_cudaPitchedPtr(host.data(), N * esz, N, N);
    cp.dstArray = arr;
    cp.extent = make_cudaExtent(N, N, NZ);
    cp.kind = cudaMemcpyHostToDevice;
    CK(cudaMemcpy3D(&cp));
    cudaResourceDesc rd = {};
    rd.resType = cudaResourceTypeArray;
    rd.res.array.array = arr;
    cudaTextureDesc td = {};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeBorder;
    td.filterMode = linear ? cudaFilterModeLinear : cudaFilterModePoint;
    td.readMode = cudaReadModeNormalizedFloat;
    td.normalizedCoords = 1;
    cudaTextureObject_t tex;
    CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    float *sink;
    CK(cudaMalloc(&sink, 4));
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int blocks = sms * 8, iters = 2048;
    const float step = 2.56f / 32.0f * 0.125f;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    double best = 0;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(e0);
        const int it = rep == 0 ? 64 : iters;
        switch (kernel_mode) {
            case 0: k_rate<0><<<blocks, 256>>>(tex, it, step, NZ, sink); break;
            case 1: k_rate<1><<<blocks, 256>>>(tex, it, step, NZ, sink); break;
            case 2: k_rate<2><<<blocks, 256>>>(tex, it, step, NZ, sink); break;
            default: k_rate<3><<<blocks, 256>>>(tex, it, step, NZ, sink); break;
        }
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep) {
            const double r = (double)blocks * 256.0 * it * 8.0 / (ms * 1e-3);
            if (r > best) best = r;
        }
    }
    CK(cudaGetLastError());
    cudaDestroyTextureObject(tex);
    cudaFreeArray(arr);
    cudaFree(sink);
    (void)mode;
    return best;
}

int main()
{
    int sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double clk = (double)sms * khz * 1e3;
    struct {
        const char *name;
        int km;
        cudaChannelFormatDesc cd;
        size_t esz;
        bool layered, linear;
    } m[] = {
        {"A 3D u16 trilinear", 0, cudaCreateChannelDesc<unsigned short>(), 2, false, true},
        {"B 2DLayered ushort2 bilinear + ALU z-lerp", 1, cudaCreateChannelDesc<ushort2>(), 4, true, true},
        {"C 3D ushort2 trilinear", 2, cudaCreateChannelDesc<ushort2>(), 4, false, true},
        {"D 2DLayered u16 bilinear", 3, cudaCreateChannelDesc<unsigned short>(), 2, true, true},
        {"E 3D u8 trilinear", 0, cudaCreateChannelDesc<unsigned char>(), 1, false, true},
        {"F 3D u16 point", 0, cudaCreateChannelDesc<unsigned short>(), 2, false, false},
        {"G 2DLayered ushort2 point", 1, cudaCreateChannelDesc<ushort2>(), 4, true, false},
        {"H 2DLayered ushort4 bilinear (as float2 read)", 3, cudaCreateChannelDesc<ushort4>(), 8, true, true},
    };
    for (auto &e : m) {
        const double r = run(0, e.km, e.cd, e.esz, e.layered, e.linear, 32, 32);
        std::printf("%-48s %8.1f Gfetch/s  %.3f fetch/clk/SM\n", e.name, r / 1e9, r / clk);
    }
    return 0;
}
