/*
 * tools/exp_region.cu -- experiment: does it pay to keep the tiles that are resident on one SM next to one another?
 *
 * Model fitted on vr_bench_tex_peak modes 2/3/4 (profiles/r2_exp_texdram.txt): the texture pipe of an SM spends
 * ~1 clock per quad request plus ~0.9 clock per sector that misses L1.  K1 on config 2 misses 26.5 sectors per
 * CTA-step where the area of a 16x8-pixel tile's footprint is 11.6: tiles that are neighbours on the screen run on
 * different SMs (blocks are dealt round-robin), so every tile pays for its whole rim.  If the ~9 tiles resident on
 * an SM formed one compact block marching in step, the rim would be shared in that SM's L1.
 *
 *   A  tiles dealt to SMs as the hardware does (block index -> tile, row-major / spiral does not matter here)
 *   B  persistent blocks: a block asks for its SM id and takes tiles from that SM's own compact region
 *      (RX x RY tiles), next to the tiles the other resident blocks of the SM hold
 *
 * Same rays, same fetches (paired layout, 512^3 u16, 2.56 voxels per step down the pair axis).
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/exp_region.bin tools/exp_region.cu
 */
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

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
    int N, tiles_x, tiles_y, steps;
    int rx, ry, regions_x, regions_y; /* B: region of rx x ry tiles per SM */
    float pitch, ox, oy, oz, dz;
    int *next;                        /* B: per-SM tile counter */
    float *sink;
};

__global__ void k_fill(cudaSurfaceObject_t surf, int N)
{
    const size_t n = (size_t)N * N * (N + 1);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int x = (int)(i % N), y = (int)((i / N) % N), l = (int)(i / ((size_t)N * N));
        unsigned v = (unsigned)i * 2654435761u;
        v ^= v >> 15;
        surf2DLayeredwrite(make_ushort2((unsigned short)v, (unsigned short)(v >> 16)), surf, x * 4, y, l);
    }
}

__device__ __forceinline__ float march_tile(const P &p, int tx, int ty)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int px = tx * 16 + (warp & 1) * 8 + (lane & 7), py = ty * 8 + (warp >> 1) * 4 + (lane >> 3);
    const float iN = 1.0f / (float)p.N;
    const float x = (p.ox + p.pitch * px) * iN, y = (p.oy + p.pitch * py) * iN;
    float z = p.oz, acc = 0.f;
    float2 a[2], b[2];
    float fa[2], fb[2];
#define ISSUE(T, F)                                               \
    _Pragma("unroll") for (int u = 0; u < 2; u++)                 \
    {                                                             \
        const float w = z - 0.5f, fl = floorf(w);                 \
        F[u] = w - fl;                                            \
        T[u] = tex2DLayered<float2>(p.tex, x, y, (int)fl + 1);    \
        z += p.dz;                                                \
    }
#define CONSUME(T, F) \
    _Pragma("unroll") for (int u = 0; u < 2; u++) acc = fmaf(acc, 0.999f, fmaf(F[u], T[u].y - T[u].x, T[u].x));
    ISSUE(a, fa);
    for (int i = 0; i < p.steps; i += 4) {
        ISSUE(b, fb);
        CONSUME(a, fa);
        ISSUE(a, fa);
        CONSUME(b, fb);
    }
    CONSUME(a, fa);
    return acc;
}

__global__ void __launch_bounds__(128) k_dealt(const P p)
{
    const int tx = blockIdx.x % p.tiles_x, ty = blockIdx.x / p.tiles_x;
    const float acc = march_tile(p, tx, ty);
    if (acc == -12345.678f) p.sink[0] = acc;
}

/* SWIZZLE: tiles of a region in 2x2 order so that consecutive tiles are neighbours in both directions */
__global__ void __launch_bounds__(128) k_region(const P p)
{
    __shared__ int s_tile;
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    const int nreg = p.regions_x * p.regions_y;
    if ((int)smid >= nreg) return;
    const int r_x = (int)smid % p.regions_x, r_y = (int)smid / p.regions_x;
    float acc = 0.f;
    while (true) {
        if (threadIdx.x == 0) s_tile = atomicAdd(&p.next[smid], 1);
        __syncthreads();
        const int t = s_tile;
        __syncthreads();
        if (t >= p.rx * p.ry) break;
        const int tx = r_x * p.rx + t % p.rx, ty = r_y * p.ry + t / p.rx;
        acc += march_tile(p, tx, ty);
        if (threadIdx.x == 0) atomicAdd(&p.next[255], 1);
    }
    if (acc == -12345.678f) p.sink[0] = acc;
}

int main(int argc, char **argv)
{
    const int N = argc > 1 ? atoi(argv[1]) : 512;
    cudaChannelFormatDesc cd = cudaCreateChannelDesc<ushort2>();
    cudaArray_t arr;
    CK(cudaMalloc3DArray(&arr, &cd, make_cudaExtent(N, N, N + 1), cudaArrayLayered | cudaArraySurfaceLoadStore));
    cudaResourceDesc rd = {};
    rd.resType = cudaResourceTypeArray;
    rd.res.array.array = arr;
    cudaSurfaceObject_t surf;
    CK(cudaCreateSurfaceObject(&surf, &rd));
    k_fill<<<148 * 8, 256>>>(surf, N);
    CK(cudaDeviceSynchronize());
    cudaTextureDesc td = {};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeBorder;
    td.filterMode = cudaFilterModeLinear, td.readMode = cudaReadModeNormalizedFloat, td.normalizedCoords = 1;
    cudaTextureObject_t tex;
    CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    float *sink;
    int *next;
    CK(cudaMalloc(&sink, 4));
    CK(cudaMalloc(&next, 4 * 256));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    const int only = argc > 2 ? atoi(argv[2]) : -1; /* 0: A only, 1: B only (for ncu) */
    /* 12 x 12 regions of 3 x 6 tiles (48 x 48 pixels) = 576 x 576 rays; also 4 x 5 tiles on 12 x 12 */
    struct Cfg {
        int rx, ry, regions_x, regions_y;
    } cfgs[] = {{3, 6, 12, 12}, {4, 4, 12, 12}, {2, 9, 12, 12}, {6, 3, 12, 12}};
    for (auto &c : cfgs) {
        P p = {};
        p.tex = tex, p.N = N, p.rx = c.rx, p.ry = c.ry, p.regions_x = c.regions_x, p.regions_y = c.regions_y;
        p.tiles_x = c.rx * c.regions_x, p.tiles_y = c.ry * c.regions_y;
        const int rays_x = p.tiles_x * 16, rays_y = p.tiles_y * 8;
        p.pitch = 0.85f, p.steps = 192, p.ox = 0.5f * (N - p.pitch * rays_x) + 0.37f, p.oy = 0.5f * (N - p.pitch * rays_y) + 0.41f;
        p.oz = 0.23f, p.dz = 0.005f * N, p.next = next, p.sink = sink;
        const double samples = (double)rays_x * rays_y * (p.steps + 2);
        for (int mode = 0; mode < 2; mode++) {
            if (only >= 0 && mode != only) continue;
            float best = 1e30f;
            for (int rep = 0; rep < 5; rep++) {
                CK(cudaMemset(next, 0, 4 * 256));
                cudaEventRecord(e0);
                if (mode == 0) k_dealt<<<p.tiles_x * p.tiles_y, 128>>>(p);
                else k_region<<<148 * 10, 128>>>(p);
                cudaEventRecord(e1);
                CK(cudaEventSynchronize(e1));
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                if (rep && ms < best) best = ms;
            }
            int done = 0;
            CK(cudaMemcpy(&done, next + 255, 4, cudaMemcpyDeviceToHost));
            std::printf("rays %dx%d pitch %.2f  %s region %dx%d tiles: %.4f ms  %6.1f Gsample/s  (tiles done %d of %d)\n", rays_x, rays_y, p.pitch,
                        mode == 0 ? "A dealt     " : "B per-SM    ", c.rx, c.ry, best, samples / (best * 1e-3) / 1e9, mode ? done : p.tiles_x * p.tiles_y,
                        p.tiles_x * p.tiles_y);
        }
        if (only >= 0) break;
    }
    return 0;
}
