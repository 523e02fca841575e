/*
 * tools/exp_alias.cu -- experiment: can the memory behind a CUDA array also be
 * reached through a linear pointer, so that K1 could pull the texels its rays will need into L2 with
 * prefetch instructions (which do not go through the texture unit's miss queue)?
 *
 *   1. a 2-D layered ushort2 array is created with CUDA_ARRAY3D_DEFERRED_MAPPING and mapped onto a
 *      cuMemCreate handle (cuMemMapArrayAsync); the same handle is mapped at a virtual address (cuMemMap).
 *   2. layout discovery: every 32-bit word behind the linear alias is set to its own word index; a point
 *      fetch through the texture then tells where texel (x, y, layer) lives.  The block-linear model
 *      (64 B x 8 row GOBs, blocks of `bh` GOBs, layer stride) is fitted and checked on every texel.
 *   3. the march-shaped streaming kernel of tools/exp_stream.cu on that array, without prefetch, with a
 *      per-lane prefetch.global.L2 of the sample `ahead` steps on, and with one
 *      cp.async.bulk.prefetch.L2 per GOB of the warp's footprint.
 *
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/exp_alias.bin tools/exp_alias.cu -lcuda
 * Run:   tools/exp_alias.bin [N=512]
 */
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x)                                                                   \
    do {                                                                        \
        cudaError_t e_ = (x);                                                   \
        if (e_ != cudaSuccess) {                                                \
            std::printf("%s failed: %s\n", #x, cudaGetErrorString(e_));         \
            std::exit(1);                                                       \
        }                                                                       \
    } while (0)
#define CU(x)                                                                   \
    do {                                                                        \
        CUresult r_ = (x);                                                      \
        if (r_ != CUDA_SUCCESS) {                                               \
            const char *s_ = nullptr;                                           \
            cuGetErrorString(r_, &s_);                                          \
            std::printf("%s failed: %d %s\n", #x, (int)r_, s_ ? s_ : "?");      \
            std::exit(1);                                                       \
        }                                                                       \
    } while (0)

static bool cu_try(CUresult r, const char *what)
{
    if (r == CUDA_SUCCESS) return true;
    const char *s = nullptr;
    cuGetErrorString(r, &s);
    std::printf("  %s -> %d %s\n", what, (int)r, s ? s : "?");
    return false;
}

struct Layout {
    unsigned long long layer_stride; /* bytes */
    unsigned bh;                     /* GOBs per block (vertical) */
    unsigned long long row_stride;   /* bytes between rows of blocks */
    int ok;
};

__host__ __device__ inline unsigned long long texel_offset(const Layout &L, unsigned x, unsigned y, unsigned layer)
{
    const unsigned xb = x * 4u; /* bytes */
    const unsigned gx = xb >> 6, gy = y >> 3;
    const unsigned by = gy / L.bh, gyi = gy - by * L.bh;
    unsigned long long off = (unsigned long long)layer * L.layer_stride + (unsigned long long)by * L.row_stride +
                             (unsigned long long)gx * (512ull * L.bh) + (unsigned long long)gyi * 512ull;
    const unsigned xi = xb & 63u, yi = y & 7u;
    off += ((xi >> 5) << 8) + ((yi >> 1) << 6) + (((xi >> 4) & 1u) << 5) + ((yi & 1u) << 4) + (xi & 15u);
    return off;
}

__global__ void k_fill_index(unsigned *w, size_t nwords)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x) w[i] = (unsigned)i;
}
__global__ void k_read_index(cudaTextureObject_t tex, int N, int layers, unsigned *out)
{
    const size_t n = (size_t)N * N * layers;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int x = (int)(i % N), y = (int)((i / N) % N), l = (int)(i / ((size_t)N * N));
        const ushort2 t = tex2DLayered<ushort2>(tex, x + 0.5f, y + 0.5f, l);
        out[i] = (unsigned)t.x | ((unsigned)t.y << 16);
    }
}
__device__ __forceinline__ unsigned hash3(unsigned x, unsigned y, unsigned z)
{
    unsigned h = x * 0x9E3779B1u ^ y * 0x85EBCA77u ^ z * 0xC2B2AE3Du;
    h ^= h >> 15, h *= 0x2C1B3C6Du, h ^= h >> 12;
    return h;
}
__global__ void k_fill_noise(unsigned *w, size_t nwords)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x)
        w[i] = hash3((unsigned)i, (unsigned)(i >> 32), 17u);
}

struct P {
    cudaTextureObject_t tex;
    int N, RW, RH, steps;
    float pitch, ox, oy, oz, ux, uy, uz, vx, vy, vz, dx, dy, dz;
    const char *alias;
    Layout L;
    int ahead;
    unsigned long long *count;
    float *sink;
};

/* MODE 0: no prefetch.  1: per-lane prefetch.global.L2 of the sample `ahead` steps on.
 * 2: per warp and step, lanes 0..3 issue cp.async.bulk.prefetch.L2 of the (up to 2x2) GOBs under the warp's footprint
 *    `ahead` steps on.  3: like 2 but only every 4th step, for 4 layers at once (lanes 0..15). */
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
        if (MODE == 3) {
            /* lanes 0..15: layer group g = lane / 4 (step i + ahead + g), GOB k = lane % 4 of the warp's footprint */
            const int g = (lane >> 2) & 3, k = lane & 3;
            const float s = (float)(p.ahead + g);
            /* footprint corners: lane 0 and lane 31 of the warp */
            const float x0 = __shfl_sync(0xffffffffu, x, 0), y0 = __shfl_sync(0xffffffffu, y, 0), z0 = __shfl_sync(0xffffffffu, z, 0);
            const float x1 = __shfl_sync(0xffffffffu, x, 31), y1 = __shfl_sync(0xffffffffu, y, 31);
            const float ax = ((k & 1) ? fmaxf(x0, x1) + 0.5f : fminf(x0, x1) - 0.5f) + s * p.dx;
            const float ay = ((k & 2) ? fmaxf(y0, y1) + 0.5f : fminf(y0, y1) - 0.5f) + s * p.dy;
            const float az = z0 + s * p.dz;
            if (lane < 16 && ax >= 0.f && ay >= 0.f && az >= 0.5f && ax < fN && ay < fN && az < fN) {
                const unsigned al = (unsigned)(az - 0.5f) + 1u;
                const unsigned long long off = texel_offset(p.L, (unsigned)ax & ~15u, (unsigned)ay & ~7u, al);
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], 512;" ::"l"(p.alias + off));
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            ok[u] = x >= 0.f && y >= 0.f && z >= 0.f && x <= fN && y <= fN && z <= fN;
            v[u] = 0.f;
            if (MODE == 2) {
                const int k = lane & 3;
                const float s = (float)p.ahead;
                const float x0 = __shfl_sync(0xffffffffu, x, 0), y0 = __shfl_sync(0xffffffffu, y, 0), z0 = __shfl_sync(0xffffffffu, z, 0);
                const float x1 = __shfl_sync(0xffffffffu, x, 31), y1 = __shfl_sync(0xffffffffu, y, 31);
                const float ax = ((k & 1) ? fmaxf(x0, x1) + 0.5f : fminf(x0, x1) - 0.5f) + s * p.dx;
                const float ay = ((k & 2) ? fmaxf(y0, y1) + 0.5f : fminf(y0, y1) - 0.5f) + s * p.dy;
                const float az = z0 + s * p.dz;
                if (lane < 4 && ax >= 0.f && ay >= 0.f && az >= 0.5f && ax < fN && ay < fN && az < fN) {
                    const unsigned al = (unsigned)(az - 0.5f) + 1u;
                    const unsigned long long off = texel_offset(p.L, (unsigned)ax & ~15u, (unsigned)ay & ~7u, al);
                    asm volatile("cp.async.bulk.prefetch.L2.global [%0], 512;" ::"l"(p.alias + off));
                }
            }
            if (ok[u]) {
                const float w = z - 0.5f, fl = floorf(w), fr = w - fl;
                const int l = (int)fl + 1;
                const float2 t = tex2DLayered<float2>(p.tex, x * iN, y * iN, l);
                if (MODE == 1) {
                    const float ax = x + p.ahead * p.dx, ay = y + p.ahead * p.dy, az = z + p.ahead * p.dz;
                    if (ax >= 0.f && ay >= 0.f && az >= 0.5f && ax < fN && ay < fN && az < fN) {
                        const unsigned al = (unsigned)(az - 0.5f) + 1u;
                        const unsigned long long off = texel_offset(p.L, (unsigned)ax, (unsigned)ay, al);
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(p.alias + off));
                    }
                }
                v[u] = fmaf(fr, t.y - t.x, t.x);
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
    const int N = argc > 1 ? atoi(argv[1]) : 512, RW = 480, RH = 480;
    const int layers = N + 1;
    CK(cudaSetDevice(0));
    CK(cudaFree(0));
    CUdevice dev;
    CU(cuDeviceGet(&dev, 0));
    int dm = 0, sp = 0, vmm = 0;
    cuDeviceGetAttribute(&dm, CU_DEVICE_ATTRIBUTE_DEFERRED_MAPPING_CUDA_ARRAY_SUPPORTED, dev);
    cuDeviceGetAttribute(&sp, CU_DEVICE_ATTRIBUTE_SPARSE_CUDA_ARRAY_SUPPORTED, dev);
    cuDeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
    std::printf("deferred-mapping arrays %d, sparse arrays %d, vmm %d\n", dm, sp, vmm);
    if (!dm || !vmm) return 2;

    CUDA_ARRAY3D_DESCRIPTOR ad = {};
    ad.Width = N, ad.Height = N, ad.Depth = layers, ad.Format = CU_AD_FORMAT_UNSIGNED_INT16, ad.NumChannels = 2;
    ad.Flags = CUDA_ARRAY3D_LAYERED | CUDA_ARRAY3D_DEFERRED_MAPPING;
    CUarray arr;
    CU(cuArray3DCreate(&arr, &ad));
    CUDA_ARRAY_MEMORY_REQUIREMENTS req = {};
    CU(cuArrayGetMemoryRequirements(&req, arr, dev));
    std::printf("array %d x %d x %d layers ushort2: size %zu (%.1f MiB; texels %.1f MiB) alignment %zu\n", N, N, layers, req.size,
                req.size / 1048576.0, (double)N * N * layers * 4 / 1048576.0, req.alignment);

    /* probe: which tile-pool allocations does this driver accept at all? */
    for (size_t probe : {(size_t)2 << 20, (size_t)64 << 20, (size_t)512 << 20}) {
        for (int gflag = 0; gflag < 2; gflag++) {
            CUmemAllocationProp prop = {};
            prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
            prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
            prop.location.id = 0;
            prop.allocFlags.usage = CU_MEM_CREATE_USAGE_TILE_POOL;
            size_t g = 0;
            CUresult rg = cuMemGetAllocationGranularity(&g, &prop, gflag ? CU_MEM_ALLOC_GRANULARITY_RECOMMENDED : CU_MEM_ALLOC_GRANULARITY_MINIMUM);
            CUmemGenericAllocationHandle hp = 0;
            CUresult rc = cuMemCreate(&hp, probe, &prop, 0);
            std::printf("probe TILE_POOL size %zu: granularity(%s) rc %d = %zu, cuMemCreate rc %d\n", probe, gflag ? "rec" : "min", (int)rg, g, (int)rc);
            if (rc == CUDA_SUCCESS) cuMemRelease(hp);
        }
    }
    {
        /* the runtime's view: does a plain deferred-mapping array report requirements, does a sparse one? */
        CUDA_ARRAY3D_DESCRIPTOR sd = ad;
        sd.Flags = CUDA_ARRAY3D_LAYERED | CUDA_ARRAY3D_SPARSE;
        CUarray sarr;
        CUresult rs = cuArray3DCreate(&sarr, &sd);
        std::printf("sparse layered array create rc %d\n", (int)rs);
        if (rs == CUDA_SUCCESS) {
            CUDA_ARRAY_SPARSE_PROPERTIES spp = {};
            CUresult rp = cuArrayGetSparseProperties(&spp, sarr);
            std::printf("  sparse properties rc %d: tile %u x %u x %u, miptail first level %u size %llu\n", (int)rp, spp.tileExtent.width,
                        spp.tileExtent.height, spp.tileExtent.depth, spp.miptailFirstLevel, spp.miptailSize);
            cuArrayDestroy(sarr);
        }
    }
    CUmemGenericAllocationHandle h = 0;
    CUdeviceptr va = 0;
    size_t size = 0;
    bool mapped = false;
    for (int attempt = 0; attempt < 2 && !mapped; attempt++) {
        CUmemAllocationProp prop = {};
        prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        prop.location.id = 0;
        prop.allocFlags.usage = attempt == 0 ? CU_MEM_CREATE_USAGE_TILE_POOL : 0;
        size_t gran = 0;
        CU(cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
        if (gran < req.alignment) gran = req.alignment;
        size = (req.size + gran - 1) / gran * gran;
        std::printf("attempt %d (usage %s): granularity %zu size %zu\n", attempt, attempt == 0 ? "TILE_POOL" : "0", gran, size);
        if (!cu_try(cuMemCreate(&h, size, &prop, 0), "cuMemCreate")) continue;
        CUarrayMapInfo mi = {};
        mi.resourceType = CU_RESOURCE_TYPE_ARRAY;
        mi.resource.array = arr;
        mi.memOperationType = CU_MEM_OPERATION_TYPE_MAP;
        mi.memHandleType = CU_MEM_HANDLE_TYPE_GENERIC;
        mi.memHandle.memHandle = h;
        mi.offset = 0;
        mi.deviceBitMask = 1;
        if (!cu_try(cuMemMapArrayAsync(&mi, 1, 0), "cuMemMapArrayAsync")) { cuMemRelease(h); continue; }
        CU(cuStreamSynchronize(0));
        if (!cu_try(cuMemAddressReserve(&va, size, gran, 0, 0), "cuMemAddressReserve")) { cuMemRelease(h); continue; }
        if (!cu_try(cuMemMap(va, size, 0, h, 0), "cuMemMap (linear alias of the array's memory)")) {
            cuMemAddressFree(va, size);
            /* unmap the array from this handle before the next attempt */
            mi.memOperationType = CU_MEM_OPERATION_TYPE_UNMAP;
            mi.memHandle.memHandle = 0;
            cu_try(cuMemMapArrayAsync(&mi, 1, 0), "cuMemMapArrayAsync(unmap)");
            cuStreamSynchronize(0);
            cuMemRelease(h);
            continue;
        }
        CUmemAccessDesc acc = {};
        acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE, acc.location.id = 0, acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        if (!cu_try(cuMemSetAccess(va, size, &acc, 1), "cuMemSetAccess")) return 3;
        mapped = true;
    }
    if (!mapped) {
        std::printf("RESULT: no linear alias of array memory on this driver\n");
        return 4;
    }
    std::printf("RESULT: array memory aliased linearly at %p (%zu bytes)\n", (void *)va, size);

    /* texture objects on the array */
    cudaTextureObject_t tex_pt = 0, tex_lin = 0;
    {
        cudaResourceDesc rd = {};
        rd.resType = cudaResourceTypeArray;
        rd.res.array.array = (cudaArray_t)arr;
        cudaTextureDesc td = {};
        td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeBorder;
        td.filterMode = cudaFilterModePoint, td.readMode = cudaReadModeElementType, td.normalizedCoords = 0;
        CK(cudaCreateTextureObject(&tex_pt, &rd, &td, nullptr));
        td.filterMode = cudaFilterModeLinear, td.readMode = cudaReadModeNormalizedFloat, td.normalizedCoords = 1;
        CK(cudaCreateTextureObject(&tex_lin, &rd, &td, nullptr));
    }
    /* ---- layout discovery ---- */
    const size_t nwords = size / 4, ntex = (size_t)N * N * layers;
    k_fill_index<<<148 * 8, 256>>>((unsigned *)va, nwords);
    CK(cudaDeviceSynchronize());
    unsigned *d_idx;
    CK(cudaMalloc(&d_idx, ntex * 4));
    k_read_index<<<148 * 8, 256>>>(tex_pt, N, layers, d_idx);
    CK(cudaDeviceSynchronize());
    std::vector<unsigned> idx(ntex);
    CK(cudaMemcpy(idx.data(), d_idx, ntex * 4, cudaMemcpyDeviceToHost));
    cudaFree(d_idx);
    auto at = [&](int x, int y, int l) { return (unsigned long long)idx[((size_t)l * N + y) * N + x] * 4ull; };
    std::printf("byte offsets, layer 0, y=0, x=0..39:\n ");
    for (int x = 0; x < 40; x++) std::printf(" %llu", at(x, 0, 0));
    std::printf("\nbyte offsets, layer 0, x=0, y=0..71:\n ");
    for (int y = 0; y < 72 && y < N; y++) std::printf(" %llu", at(0, y, 0));
    std::printf("\nlayer 1 origin %llu, layer 2 origin %llu, last layer origin %llu\n", at(0, 0, 1), at(0, 0, 2), at(0, 0, layers - 1));
    Layout L = {};
    L.layer_stride = at(0, 0, 1) - at(0, 0, 0);
    /* block height: first y (multiple of 8) whose offset is not y/8 * 512 */
    L.bh = 1;
    while (L.bh * 8 < (unsigned)N && at(0, L.bh * 8, 0) == (unsigned long long)L.bh * 512ull) L.bh++;
    L.row_stride = (L.bh * 8 < (unsigned)N) ? at(0, L.bh * 8, 0) : 0;
    size_t bad = 0;
    for (int l = 0; l < layers; l += (layers > 8 ? layers / 8 : 1))
        for (int y = 0; y < N; y++)
            for (int x = 0; x < N; x++)
                if (at(x, y, l) != texel_offset(L, x, y, l)) {
                    if (bad < 8) std::printf("  mismatch (%d,%d,%d): array %llu model %llu\n", x, y, l, at(x, y, l), texel_offset(L, x, y, l));
                    bad++;
                }
    L.ok = bad == 0;
    std::printf("block-linear model: layer stride %llu, block height %u GOBs, block-row stride %llu -> %zu mismatches\n", L.layer_stride,
                L.bh, L.row_stride, bad);
    if (!L.ok) {
        /* leave enough to work the layout out by hand */
        FILE *f = std::fopen("gpurun_out/exp_alias_layout.txt", "w");
        if (f) {
            for (int y = 0; y < 128 && y < N; y++) {
                for (int x = 0; x < 128 && x < N; x++) std::fprintf(f, "%llu ", at(x, y, 0));
                std::fprintf(f, "\n");
            }
            std::fclose(f);
        }
    }

    /* ---- streaming with and without prefetch ---- */
    k_fill_noise<<<148 * 8, 256>>>((unsigned *)va, nwords);
    CK(cudaDeviceSynchronize());
    unsigned long long *count;
    float *sink;
    CK(cudaMalloc(&count, 8));
    CK(cudaMalloc(&sink, 4));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    const float pitch = (float)N / (float)RW, step = 0.005f * N;
    struct Dir {
        const char *name;
        float d[3], u[3], v[3];
    } dirs[] = {
        {"+z", {0, 0, 1}, {1, 0, 0}, {0, 1, 0}},
        {"z-oblique", {0.30f, 0.20f, 0.933f}, {1, 0, 0}, {0, 1, 0}},
    };
    const char *names[] = {"no prefetch", "per-lane prefetch.global.L2", "per-step bulk prefetch of GOBs", "4-step bulk prefetch of GOBs"};
    for (auto &D : dirs)
        for (int mode = 0; mode < 4; mode++) {
            if (mode && !L.ok) continue;
            for (int ahead : {4, 8, 16, 32}) {
                if (mode == 0 && ahead != 4) continue;
                P p = {};
                p.tex = tex_lin, p.N = N, p.RW = RW, p.RH = RH, p.steps = (int)(1.0f / 0.005f) + 4, p.pitch = pitch;
                p.ux = D.u[0], p.uy = D.u[1], p.uz = D.u[2], p.vx = D.v[0], p.vy = D.v[1], p.vz = D.v[2];
                p.dx = D.d[0] * step, p.dy = D.d[1] * step, p.dz = D.d[2] * step;
                p.ox = 0.37f, p.oy = 0.41f, p.oz = 0.23f;
                p.alias = (const char *)va, p.L = L, p.ahead = ahead, p.count = count, p.sink = sink;
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
                std::printf("%-10s %-34s ahead=%2d  %.4f ms  %6.1f Gsample/s  (%llu samples)\n", D.name, names[mode], ahead, best,
                            cnt / (best * 1e-3) / 1e9, cnt);
            }
        }
    return 0;
}
