/*
 * tools/exp_stream.cu -- experiment: streaming (DRAM-sourced) sampling throughput of the texture
 * path for the volume layouts K1 could use, on a march-shaped access pattern:
 * RW x RH parallel rays `pitch` voxels apart, `step` voxels per sample, through an N^3 u16 volume.
 *   L0  3-D CUDA array u16, hardware trilinear                          (K1 today)
 *   L1  2-D layered CUDA array ushort2 = (v[z-1], v[z]), bilinear + z-lerp in ALU
 *   L2  pitch-linear 2-D atlas of the same pairs (addresses known), bilinear + z-lerp
 *   L3  L2 + per-lane prefetch.global.L2 of the position A steps ahead
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/exp_stream.bin tools/exp_stream.cu
 * Run:   tools/exp_stream.bin [N=512] [RW=480] [RH=480]
 */
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                   \
    do {                                                                        \
        cudaError_t e_ = (x);                                                   \
        if (e_ != cudaSuccess) {                                                \
            std::printf("%s failed: %s\n", #x, cudaGetErrorString(e_));         \
            std::exit(1);                                                       \
        }                                                                       \
    } while (0)

struct P {
    cudaTextureObject_t tex;
    int N, RW, RH, steps;
    float pitch, ox, oy, oz;    /* ray grid spacing (voxels), origin of ray (0,0) */
    float ux, uy, uz, vx, vy, vz; /* ray grid axes (unit, voxels) */
    float dx, dy, dz;           /* step vector, voxels */
    /* atlas */
    const ushort2 *atlas;
    int tp, th, tiles_x;        /* tile pitch (texels), tile height (rows), tiles per atlas row */
    size_t apitch;              /* atlas row pitch, texels */
    int ahead;
    unsigned long long *count;
    float *sink;
};

__device__ __forceinline__ unsigned hash3(unsigned x, unsigned y, unsigned z)
{
    unsigned h = x * 0x9E3779B1u ^ y * 0x85EBCA77u ^ z * 0xC2B2AE3Du;
    h ^= h >> 15, h *= 0x2C1B3C6Du, h ^= h >> 12;
    return h;
}
__device__ __forceinline__ unsigned short vox(int x, int y, int z, int N)
{
    if (x < 0 || y < 0 || z < 0 || x >= N || y >= N || z >= N) return 0;
    return (unsigned short)(hash3(x, y, z) >> 16);
}

__global__ void k_fill_linear(unsigned short *v, int N)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, n = (size_t)N * N * N;
    if (i < n) v[i] = vox((int)(i % N), (int)((i / N) % N), (int)(i / ((size_t)N * N)), N);
}
__global__ void k_fill_pairs(ushort2 *v, int N) /* (N, N, N+1) */
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, n = (size_t)N * N * (N + 1);
    if (i < n) {
        const int x = (int)(i % N), y = (int)((i / N) % N), l = (int)(i / ((size_t)N * N));
        v[i] = make_ushort2(vox(x, y, l - 1, N), vox(x, y, l, N));
    }
}
__global__ void k_fill_atlas(ushort2 *a, int N, int tp, int th, int tiles_x, size_t apitch, size_t rows)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= apitch * rows) return;
    const size_t row = i / apitch, col = i % apitch;
    const int ty = (int)(row / th), y = (int)(row % th), tx = (int)(col / tp), x = (int)(col % tp);
    const int l = ty * tiles_x + tx;
    ushort2 r = make_ushort2(0, 0);
    if (tx < tiles_x && l <= N && x < N && y < N) r = make_ushort2(vox(x, y, l - 1, N), vox(x, y, l, N));
    a[i] = r;
}

template <int MODE>
__global__ void __launch_bounds__(128) k_stream(const P p)
{
    const int tiles_w = p.RW / 16;
    const int tx = blockIdx.x % tiles_w, ty = blockIdx.x / tiles_w;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int px = tx * 16 + (warp & 1) * 8 + (lane & 7), py = ty * 8 + (warp >> 1) * 4 + (lane >> 3);
    float x = p.ox + p.pitch * (px * p.ux + py * p.vx);
    float y = p.oy + p.pitch * (px * p.uy + py * p.vy);
    float z = p.oz + p.pitch * (px * p.uz + py * p.vz);
    const float fN = (float)p.N, iN = 1.0f / fN;
    float acc = 0.f;
    unsigned n = 0;
    for (int i = 0; i < p.steps; i += 4) {
        float v[4];
        bool ok[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            ok[u] = x >= 0.f && y >= 0.f && z >= 0.f && x <= fN && y <= fN && z <= fN;
            v[u] = 0.f;
            if (ok[u]) {
                if (MODE == 0) {
                    v[u] = tex3D<float>(p.tex, x * iN, y * iN, z * iN);
                } else {
                    const float w = z - 0.5f, fl = floorf(w), fr = w - fl;
                    const int l = (int)fl + 1;
                    float2 t;
                    if (MODE == 1) {
                        t = tex2DLayered<float2>(p.tex, x * iN, y * iN, l);
                    } else {
                        const int ty_ = l / p.tiles_x, tx_ = l - ty_ * p.tiles_x;
                        t = tex2D<float2>(p.tex, x + (float)(tx_ * p.tp), y + (float)(ty_ * p.th));
                        if (MODE == 3) {
                            const float ax = x + p.ahead * p.dx, ay = y + p.ahead * p.dy, az = z + p.ahead * p.dz;
                            if (ax >= 0.f && ay >= 0.f && az >= 0.5f && ax < fN && ay < fN && az < fN) {
                                const int al = (int)(az - 0.5f) + 1, aty = al / p.tiles_x, atx = al - aty * p.tiles_x;
                                const ushort2 *q = p.atlas + (size_t)(aty * p.th + (int)ay) * p.apitch + (atx * p.tp + (int)ax);
                                asm volatile("prefetch.global.L2 [%0];" ::"l"(q));
                            }
                        }
                    }
                    v[u] = fmaf(fr, t.y - t.x, t.x);
                }
            }
            x += p.dx, y += p.dy, z += p.dz;
        }
#pragma unroll
        for (int u = 0; u < 4; u++) acc = fmaf(acc, 0.999f, v[u]), n += ok[u];
    }
    n = __reduce_add_sync(0xffffffffu, n);
    if (lane == 0) atomicAdd(p.count, (unsigned long long)n);
    if (acc == -12345.678f) p.sink[0] = acc;
}

int main(int argc, char **argv)
{
    const int N = argc > 1 ? atoi(argv[1]) : 512, RW = argc > 2 ? atoi(argv[2]) : 480, RH = argc > 3 ? atoi(argv[3]) : 480;
    const float pitch = (float)N / (float)RW * 1.0f, step = 0.005f * N;
    std::printf("N=%d rays %dx%d pitch %.2f vox step %.2f vox\n", N, RW, RH, pitch, step);
    /* volume in the three layouts */
    unsigned short *lin;
    ushort2 *pairs, *atlas;
    const size_t nvox = (size_t)N * N * N, npair = (size_t)N * N * (N + 1);
    CK(cudaMalloc(&lin, nvox * 2));
    CK(cudaMalloc(&pairs, npair * 4));
    k_fill_linear<<<(unsigned)((nvox + 255) / 256), 256>>>(lin, N);
    k_fill_pairs<<<(unsigned)((npair + 255) / 256), 256>>>(pairs, N);
    const int tp = (N + 1 + 3) / 4 * 4, th = N + 1, tiles_x = 16, tiles_y = (N + 1 + tiles_x - 1) / tiles_x;
    size_t apitch = (size_t)tiles_x * tp;
    apitch = (apitch + 127) / 128 * 128; /* 512-byte aligned rows */
    const size_t arows = (size_t)tiles_y * th;
    CK(cudaMalloc(&atlas, apitch * arows * 4));
    k_fill_atlas<<<(unsigned)((apitch * arows + 255) / 256), 256>>>(atlas, N, tp, th, tiles_x, apitch, arows);
    CK(cudaDeviceSynchronize());
    std::printf("atlas %zu x %zu texels (%.0f MiB)\n", apitch, arows, apitch * arows * 4.0 / 1048576.0);

    cudaTextureObject_t tex[3];
    cudaArray_t a0, a1;
    {
        cudaChannelFormatDesc cd = cudaCreateChannelDesc<unsigned short>();
        CK(cudaMalloc3DArray(&a0, &cd, make_cudaExtent(N, N, N)));
        cudaMemcpy3DParms cp = {};
        cp.srcPtr = make_cudaPitchedPtr(lin, (size_t)N * 2, N, N);
        cp.dstArray = a0, cp.extent = make_cudaExtent(N, N, N), cp.kind = cudaMemcpyDeviceToDevice;
        CK(cudaMemcpy3D(&cp));
        cudaChannelFormatDesc cd2 = cudaCreateChannelDesc<ushort2>();
        CK(cudaMalloc3DArray(&a1, &cd2, make_cudaExtent(N, N, N + 1), cudaArrayLayered));
        cp.srcPtr = make_cudaPitchedPtr(pairs, (size_t)N * 4, N, N);
        cp.dstArray = a1, cp.extent = make_cudaExtent(N, N, N + 1);
        CK(cudaMemcpy3D(&cp));
        cudaTextureDesc td = {};
        td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeBorder;
        td.filterMode = cudaFilterModeLinear, td.readMode = cudaReadModeNormalizedFloat, td.normalizedCoords = 1;
        cudaResourceDesc rd = {};
        rd.resType = cudaResourceTypeArray;
        rd.res.array.array = a0;
        CK(cudaCreateTextureObject(&tex[0], &rd, &td, nullptr));
        rd.res.array.array = a1;
        CK(cudaCreateTextureObject(&tex[1], &rd, &td, nullptr));
        cudaResourceDesc rp = {};
        rp.resType = cudaResourceTypePitch2D;
        rp.res.pitch2D.devPtr = atlas, rp.res.pitch2D.desc = cd2, rp.res.pitch2D.width = apitch, rp.res.pitch2D.height = arows;
        rp.res.pitch2D.pitchInBytes = apitch * 4;
        td.normalizedCoords = 0;
        CK(cudaCreateTextureObject(&tex[2], &rp, &td, nullptr));
    }
    cudaFree(lin), cudaFree(pairs);
    unsigned long long *count;
    float *sink;
    CK(cudaMalloc(&count, 8));
    CK(cudaMalloc(&sink, 4));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);

    struct Dir {
        const char *name;
        float d[3], u[3], v[3];
    } dirs[] = {
        {"+z", {0, 0, 1}, {1, 0, 0}, {0, 1, 0}},
        {"+x", {1, 0, 0}, {0, 1, 0}, {0, 0, 1}},
        {"+y", {0, 1, 0}, {1, 0, 0}, {0, 0, 1}},
        {"z-oblique", {0.30f, 0.20f, 0.933f}, {1, 0, 0}, {0, 1, 0}},
        {"x-oblique", {0.933f, 0.20f, 0.30f}, {0, 1, 0}, {0, 0, 1}},
    };
    const char *names[] = {"L0 3D u16 trilinear", "L1 layered pairs", "L2 atlas pairs", "L3 atlas + L2 prefetch"};
    for (auto &D : dirs) {
        for (int mode = 0; mode < 4; mode++) {
            for (int ahead : {4, 8, 16}) {
                if (mode != 3 && ahead != 4) continue;
                P p = {};
                p.tex = tex[mode >= 2 ? 2 : mode];
                p.N = N, p.RW = RW, p.RH = RH, p.steps = (int)(1.0f / 0.005f) + 4, p.pitch = pitch;
                p.ux = D.u[0], p.uy = D.u[1], p.uz = D.u[2], p.vx = D.v[0], p.vy = D.v[1], p.vz = D.v[2];
                p.dx = D.d[0] * step, p.dy = D.d[1] * step, p.dz = D.d[2] * step;
                p.ox = 0.37f * (D.d[0] == 0 || D.d[0] < 0.9f) , p.oy = 0.41f, p.oz = 0.23f;
                p.atlas = atlas, p.tp = tp, p.th = th, p.tiles_x = tiles_x, p.apitch = apitch, p.ahead = ahead;
                p.count = count, p.sink = sink;
                const int blocks = (RW / 16) * (RH / 8);
                float best = 1e30f;
                unsigned long long cnt = 0;
                for (int rep = 0; rep < 5; rep++) {
                    cudaMemset(count, 0, 8);
                    cudaEventRecord(e0);
                    switch (mode) {
                        case 0: k_stream<0><<<blocks, 128>>>(p); break;
                        case 1: k_stream<1><<<blocks, 128>>>(p); break;
                        case 2: k_stream<2><<<blocks, 128>>>(p); break;
                        default: k_stream<3><<<blocks, 128>>>(p); break;
                    }
                    cudaEventRecord(e1);
                    CK(cudaEventSynchronize(e1));
                    float ms;
                    cudaEventElapsedTime(&ms, e0, e1);
                    if (rep && ms < best) best = ms;
                    cudaMemcpy(&cnt, count, 8, cudaMemcpyDeviceToHost);
                }
                std::printf("%-10s %-24s ahead=%2d  %.4f ms  %6.1f Gsample/s  (%llu samples)\n", D.name, names[mode], ahead, best,
                            cnt / (best * 1e-3) / 1e9, cnt);
            }
        }
    }
    return 0;
}
