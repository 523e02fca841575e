/*
 * volume.cu -- K0: what happens to a volume between the loader and the march.
 *
 * Replaces, on the device and in the dataset's native type:
 *   - the two O(N) host passes std::min_element / std::max_element of
 *     NrrdFileParser::parse (src/data/nrrd_file_parser.cpp:39-40)        -> k_global_minmax
 *   - (new) per-cell value ranges for empty-space skipping                -> k_cell_ranges
 *   - (new) cell classification against the transfer function            -> k_classify
 *   - (new) the synthetic volumes BASELINE.json names, generated in HBM  -> k_synth
 * The copy into the bricked (block-linear) 3-D CUDA array that replaces
 * vkCmdCopyBufferToImage (src/rendering/offscreen_pass.cpp:1379-1406) is a cudaMemcpy3D
 * issued from api.cu.
 *
 * Compiled with -fmad=false (see march.cu): k_classify evaluates the transfer-function
 * coordinate with the same single-rounding operations as the march, and the fBm generator is
 * bit-identical to the CPU generator the tests use.
 */
#include <cuda_runtime.h>
#include <math_constants.h>

#include <vector>

#include "vr_ctx.h"

namespace
{

template <typename T>
__device__ __forceinline__ float ldval(const void *base, size_t i)
{
    return (float)__ldg(reinterpret_cast<const T *>(base) + i);
}

__device__ __forceinline__ float ld_any(const void *base, int dtype, size_t i)
{
    switch (dtype) {
        case VR_U8: return ldval<unsigned char>(base, i);
        case VR_U16: return ldval<unsigned short>(base, i);
        default: return ldval<float>(base, i);
    }
}

__device__ __forceinline__ void block_minmax(float &mn, float &mx, float2 *out)
{
    __shared__ float smn[32], smx[32];
    for (int o = 16; o; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    if ((threadIdx.x & 31) == 0) smn[w] = mn, smx[w] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < nw; i++) mn = fminf(mn, smn[i]), mx = fmaxf(mx, smx[i]);
        *out = make_float2(mn, mx);
    }
}

/* grid-stride min/max over the whole volume, 16 bytes per load; one partial per block */
template <typename T>
__global__ void __launch_bounds__(256) k_global_minmax(const T *__restrict__ v, size_t n, float2 *partial)
{
    constexpr int PER = 16 / sizeof(T);
    float mn = CUDART_INF_F, mx = -CUDART_INF_F;
    const size_t nvec = n / PER;
    const uint4 *v4 = reinterpret_cast<const uint4 *>(v);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 q = __ldg(v4 + i);
        const T *e = reinterpret_cast<const T *>(&q);
#pragma unroll
        for (int k = 0; k < PER; k++) {
            const float f = (float)e[k];
            mn = fminf(mn, f), mx = fmaxf(mx, f);
        }
    }
    if (blockIdx.x == 0)
        for (size_t i = nvec * PER + threadIdx.x; i < n; i += blockDim.x) {
            const float f = (float)v[i];
            mn = fminf(mn, f), mx = fmaxf(mx, f);
        }
    block_minmax(mn, mx, partial + blockIdx.x);
}

/* One block per occupancy cell: (min,max) over the voxels any sample positioned in the cell can
 * touch = the cell's own `cell`^3 voxels plus a one-voxel apron; voxels outside the local
 * texture read 0, which is what the border sampler returns (offscreen_pass.cpp:1020-1026). */
__global__ void __launch_bounds__(64) k_cell_ranges(const void *__restrict__ v, int dtype, int nx, int ny, int nz,
                                                    int cell, int cnx, int cny, float2 *ranges)
{
    const int c = blockIdx.x;
    const int cx = c % cnx, cy = (c / cnx) % cny, cz = c / (cnx * cny);
    const int x0 = cx * cell - 1, y0 = cy * cell - 1, z0 = cz * cell - 1;
    const int w = cell + 2;
    float mn = CUDART_INF_F, mx = -CUDART_INF_F;
    bool bad = false;
    for (int row = threadIdx.x; row < w * w; row += blockDim.x) {
        const int y = y0 + row % w, z = z0 + row / w;
        if (y < 0 || y >= ny || z < 0 || z >= nz) {
            mn = fminf(mn, 0.0f), mx = fmaxf(mx, 0.0f);
            continue;
        }
        const size_t base = ((size_t)z * ny + y) * nx;
        for (int k = 0; k < w; k++) {
            const int x = x0 + k;
            const float f = (x < 0 || x >= nx) ? 0.0f : ld_any(v, dtype, base + x);
            mn = fminf(mn, f), mx = fmaxf(mx, f);
            /* fminf / fmaxf drop a NaN; a NaN density samples LUT texel 0 and +-Inf the end texels (march.cu), so a
             * cell holding any non-finite voxel gets the widest range: it is empty only if the whole LUT is */
            if (!(fabsf(f) <= 3.402823466e38f)) bad = true;
        }
    }
    if (bad) mn = -CUDART_INF_F, mx = CUDART_INF_F;
    block_minmax(mn, mx, ranges + c);
}

/* padded-LUT tap index of a raw density, computed exactly as the EXACT march path does */
__device__ __forceinline__ int lut_tap(float v, float min_density, float range, float n)
{
    const float tt = (v - min_density) / range;
    float uu = tt * n - 0.5f;
    uu = (uu > -1.0f) ? uu : -1.0f;
    uu = (uu < n) ? uu : n;
    return (int)floorf(uu) + 1;
}

/* bit = 1: some sample in the cell may have alpha != 0 and must be fetched.
 * bit = 0: every density in the cell's range maps between LUT texels whose alpha is exactly 0,
 * so volume.frag:44-45 leaves colour and transmittance unchanged -- skipping is exact.
 * The tap range is widened by one texel each side for the HW path's fused coordinate. */
__global__ void __launch_bounds__(256) k_classify(const float2 *__restrict__ ranges, int ncells,
                                                  const uint32_t *__restrict__ lut_nz, int nlut, float min_density,
                                                  float max_density, uint32_t *occ, unsigned long long *empty_count)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    bool occupied = false;
    if (c < ncells) {
        const float range = max_density - min_density;
        const float2 r = ranges[c];
        occupied = true;
        if (range > 0.0f && r.x == r.x && r.y == r.y) {
            int k0 = lut_tap(r.x, min_density, range, (float)nlut) - 1;
            int k1 = lut_tap(r.y, min_density, range, (float)nlut) + 2;
            k0 = max(k0, 0), k1 = min(k1, nlut + 2);
            /* padded texels k0..k1 with alpha != 0: two reads of the running count (lut.cu k_lut_tables) */
            occupied = lut_nz[k1 + 1] != lut_nz[k0];
        }
    }
    const uint32_t word = __ballot_sync(0xffffffffu, occupied);
    if ((threadIdx.x & 31) == 0) {
        occ[c >> 5] = word;
        const int live = min(32, ncells - c);
        if (live > 0) {
            const uint32_t mask = live == 32 ? 0xffffffffu : ((1u << live) - 1u);
            atomicAdd(empty_count, (unsigned long long)__popc(~word & mask));
        }
    }
}

/* Directional distance maps.  For each of the eight sign octants (sx, sy, sz) of a ray direction and each
 * empty cell c, D = edge length (in cells) of the largest cube of empty cells that has c at its corner and
 * extends towards (sx, sy, sz); cells outside the grid count as empty; occupied cells get 0; capped at
 * VR_DIST_CAP.  A ray in c whose direction lies in that octant cannot meet an occupied cell before it
 * leaves the cube.  Unlike a centred (Chebyshev) distance this ignores what lies behind the ray, so a ray
 * leaving or passing a surface takes long jumps at once.
 *
 * Built from a summed-volume table S of the occupancy bits (S[x][y][z] = occupied cells in
 * [0,x) x [0,y) x [0,z), (n+1)^3 entries): a cube is empty iff its clipped box sum is 0, which is monotone
 * in the edge length -> binary search, 8 table reads per probe. */
#define VR_DIST_CAP 255
__global__ void __launch_bounds__(256) k_sat_init(const uint32_t *__restrict__ occ, uint32_t *__restrict__ S, int nx, int ny, int nz)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int mx = nx + 1, my = ny + 1;
    if (i >= mx * my * (nz + 1)) return;
    const int x = i % mx, y = (i / mx) % my, z = i / (mx * my);
    uint32_t v = 0;
    if (x > 0 && y > 0 && z > 0) {
        const int q = ((z - 1) * ny + (y - 1)) * nx + (x - 1);
        v = (occ[q >> 5] >> (q & 31)) & 1u;
    }
    S[i] = v;
}

/* running sum along one axis of the (mx, my, mz) table, one thread per line */
__global__ void __launch_bounds__(128) k_sat_scan(uint32_t *__restrict__ S, int mx, int my, int mz, int axis)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = axis == 0 ? mx : (axis == 1 ? my : mz);
    const int na = axis == 0 ? my : mx, nb = axis == 2 ? my : mz; /* the two other extents, fastest first */
    if (t >= na * nb) return;
    const int a = t % na, b = t / na;
    size_t base, stride;
    if (axis == 0) base = ((size_t)b * my + a) * mx, stride = 1;
    else if (axis == 1) base = (size_t)b * my * mx + a, stride = (size_t)mx;
    else base = (size_t)b * mx + a, stride = (size_t)mx * my;
    uint32_t acc = 0;
    for (int i = 0; i < n; i++) {
        acc += S[base + i * stride];
        S[base + i * stride] = acc;
    }
}

__global__ void __launch_bounds__(256) k_octant_distance(const uint32_t *__restrict__ occ, const uint32_t *__restrict__ S,
                                                         unsigned char *__restrict__ out, int nx, int ny, int nz)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int ncells = nx * ny * nz;
    if (c >= ncells) return;
    const int oct = blockIdx.y;
    const int x = c % nx, y = (c / nx) % ny, z = c / (nx * ny);
    int d = 0;
    if (!((occ[c >> 5] >> (c & 31)) & 1u)) {
        const int mx = nx + 1, my = ny + 1;
        const bool negx = oct & 1, negy = oct & 2, negz = oct & 4;
        auto at = [&](int xx, int yy, int zz) { return S[((size_t)zz * my + yy) * mx + xx]; };
        auto empty = [&](int s) {
            const int x0 = max(negx ? x - s + 1 : x, 0), x1 = min(negx ? x + 1 : x + s, nx);
            const int y0 = max(negy ? y - s + 1 : y, 0), y1 = min(negy ? y + 1 : y + s, ny);
            const int z0 = max(negz ? z - s + 1 : z, 0), z1 = min(negz ? z + 1 : z + s, nz);
            const uint32_t n = at(x1, y1, z1) - at(x0, y1, z1) - at(x1, y0, z1) - at(x1, y1, z0) + at(x0, y0, z1) +
                               at(x0, y1, z0) + at(x1, y0, z0) - at(x0, y0, z0);
            return n == 0u;
        };
        int lo = 1, hi = VR_DIST_CAP; /* size 1 = the cell itself, known empty */
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (empty(mid)) lo = mid;
            else hi = mid - 1;
        }
        d = lo;
    }
    out[(size_t)oct * ncells + c] = (unsigned char)d;
}

/* Paired layout along axis `ax`: the 2-D layered array has one layer per position l = 0..n_ax on that
 * axis and texel (a, b) of layer l = (v[.., ax = l-1, ..], v[.., ax = l, ..]), zero outside the volume
 * (that is the sampler's transparent-black border, offscreen_pass.cpp:1020-1026, on the pair axis).
 * In-plane axes (a, b): ax = 2 -> (x, y), ax = 1 -> (x, z), ax = 0 -> (y, z).
 * Written through a surface so the block-linear array is filled in place, without a staging copy. */
/* CUDA layered arrays hold at most 2048 layers: layers >= split go to a second array (surf1). */
template <typename T2>
__device__ __forceinline__ void put_pair(T2 t, cudaSurfaceObject_t surf0, cudaSurfaceObject_t surf1, int split, int a, int b, int l)
{
    if (l < split) surf2DLayeredwrite(t, surf0, a * (int)sizeof(T2), b, l);
    else surf2DLayeredwrite(t, surf1, a * (int)sizeof(T2), b, l - split);
}

template <typename T, typename T2>
__global__ void __launch_bounds__(256) k_pack_pairs(const T *__restrict__ v, int nx, int ny, int nz, int ax,
                                                    cudaSurfaceObject_t surf, cudaSurfaceObject_t surf1, int split)
{
    if (ax != 0) {
        /* a = x: consecutive threads read consecutive voxels */
        const int x = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y, l = blockIdx.z;
        if (x >= nx) return;
        T2 t;
        if (ax == 2) {
            t.x = l > 0 ? __ldg(v + ((size_t)(l - 1) * ny + b) * nx + x) : (T)0;
            t.y = l < nz ? __ldg(v + ((size_t)l * ny + b) * nx + x) : (T)0;
        } else {
            t.x = l > 0 ? __ldg(v + ((size_t)b * ny + (l - 1)) * nx + x) : (T)0;
            t.y = l < ny ? __ldg(v + ((size_t)b * ny + l) * nx + x) : (T)0;
        }
        put_pair(t, surf, surf1, split, x, b, l);
    } else {
        /* layers along x, a = y: a 32 (y) x 33 (x) tile of one z-slice goes through shared memory so that
         * both the reads (along x) and the writes (along y) are contiguous */
        __shared__ T tile[32][34];
        const int z = blockIdx.z, y0 = blockIdx.y * 32, l0 = blockIdx.x * 32; /* layers l0 .. l0+31 need x = l0-1 .. l0+31 */
        const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;               /* 32 x 8 threads */
        for (int r = ty; r < 32; r += 8) {
            const int y = y0 + r;
            for (int c = tx; c < 33; c += 32) {
                const int x = l0 - 1 + c;
                tile[r][c] = (y < ny && x >= 0 && x < nx) ? __ldg(v + ((size_t)z * ny + y) * nx + x) : (T)0;
            }
        }
        __syncthreads();
        for (int r = ty; r < 32; r += 8) { /* r: layer within the tile, tx: y within the tile */
            const int l = l0 + r, y = y0 + tx;
            if (l <= nx && y < ny) {
                T2 t;
                t.x = tile[tx][r], t.y = tile[tx][r + 1];
                put_pair(t, surf, surf1, split, y, z, l);
            }
        }
    }
}

/* ------------------------------------------------------------------ synth -- */

__device__ __forceinline__ uint32_t hash_u32(uint32_t h)
{
    h ^= h >> 16;
    h *= 0x7feb352du;
    h ^= h >> 15;
    h *= 0x846ca68bu;
    h ^= h >> 16;
    return h;
}

__device__ __forceinline__ float lattice(uint32_t seed, int ix, int iy, int iz, int oct)
{
    const uint32_t h = seed ^ ((uint32_t)ix * 0x8da6b343u) ^ ((uint32_t)iy * 0xd8163841u) ^
                       ((uint32_t)iz * 0xcb1ab31fu) ^ ((uint32_t)oct * 0x9e3779b9u);
    return (float)(hash_u32(h) >> 8) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ float fade(float t) { return t * t * (3.0f - 2.0f * t); }
__device__ __forceinline__ float mixf(float a, float b, float t) { return a + t * (b - a); }

__device__ float value_noise(uint32_t seed, float x, float y, float z, int oct)
{
    const float fx = floorf(x), fy = floorf(y), fz = floorf(z);
    const int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    const float tx = fade(x - fx), ty = fade(y - fy), tz = fade(z - fz);
    const float c000 = lattice(seed, ix, iy, iz, oct), c100 = lattice(seed, ix + 1, iy, iz, oct);
    const float c010 = lattice(seed, ix, iy + 1, iz, oct), c110 = lattice(seed, ix + 1, iy + 1, iz, oct);
    const float c001 = lattice(seed, ix, iy, iz + 1, oct), c101 = lattice(seed, ix + 1, iy, iz + 1, oct);
    const float c011 = lattice(seed, ix, iy + 1, iz + 1, oct), c111 = lattice(seed, ix + 1, iy + 1, iz + 1, oct);
    return mixf(mixf(mixf(c000, c100, tx), mixf(c010, c110, tx), ty), mixf(mixf(c001, c101, tx), mixf(c011, c111, tx), ty),
                tz);
}

/* 5-octave value-noise fBm, lacunarity 2, gain 0.5, contrast-stretched about 0.5 */
__device__ float fbm(uint32_t seed, float vx, float vy, float vz, float inv_period)
{
    float sum = 0.0f, amp = 1.0f, norm = 0.0f, f = inv_period;
    for (int o = 0; o < 5; o++) {
        sum += amp * value_noise(seed, vx * f, vy * f, vz * f, o);
        norm += amp;
        amp *= 0.5f;
        f *= 2.0f;
    }
    const float v = 0.5f + (sum / norm - 0.5f) * 2.2f;
    return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
}

/* Marschner-Lobb test signal, f_M = 6, alpha = 0.25 */
__device__ float marschner_lobb(float x, float y, float z)
{
    const float PI = 3.14159265358979323846f, fM = 6.0f, al = 0.25f;
    const float r = sqrtf(x * x + y * y);
    const float pr = cosf(2.0f * PI * fM * cosf(PI * r / 2.0f));
    return (1.0f - sinf(PI * z / 2.0f) + al * (1.0f + pr)) / (2.0f * (1.0f + al));
}

#define VR_NBLOBS 64
struct Blobs {
    float cx[VR_NBLOBS], cy[VR_NBLOBS], cz[VR_NBLOBS], sigma[VR_NBLOBS], amp[VR_NBLOBS];
};

/* sum of truncated Gaussians on an exactly-zero background */
__device__ float blobs_at(const Blobs &b, float x, float y, float z)
{
    const float cut = 2.5f, floor_v = 0.0439369336f; /* exp(-cut^2/2) */
    float sum = 0.0f;
    for (int i = 0; i < VR_NBLOBS; i++) {
        const float dx = x - b.cx[i], dy = y - b.cy[i], dz = z - b.cz[i];
        const float d2 = dx * dx + dy * dy + dz * dz;
        const float s2 = b.sigma[i] * b.sigma[i];
        if (d2 < cut * cut * s2) sum += b.amp[i] * (expf(-0.5f * d2 / s2) - floor_v) / (1.0f - floor_v);
    }
    return sum > 1.0f ? 1.0f : sum;
}

struct SynthParams {
    void *out;
    int kind, dtype;
    uint32_t seed;
    uint32_t nx, ny, nz;    /* local extent written */
    int ox, oy, oz;         /* global index of local voxel 0 */
    uint32_t fx, fy, fz;    /* full grid */
    float inv_period;
};

__global__ void __launch_bounds__(256) k_synth(const __grid_constant__ SynthParams s, const __grid_constant__ Blobs blobs)
{
    const size_t total = (size_t)s.nx * s.ny * s.nz;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const uint32_t li = (uint32_t)(idx % s.nx), lj = (uint32_t)((idx / s.nx) % s.ny), lk = (uint32_t)(idx / ((size_t)s.nx * s.ny));
        const int i = (int)li + s.ox, j = (int)lj + s.oy, k = (int)lk + s.oz;
        float v = 0.0f;
        const bool inside = i >= 0 && j >= 0 && k >= 0 && (uint32_t)i < s.fx && (uint32_t)j < s.fy && (uint32_t)k < s.fz;
        if (inside) {
            if (s.kind == VR_SYNTH_MARSCHNER_LOBB) {
                const float x = -1.0f + 2.0f * ((float)i + 0.5f) / (float)s.fx;
                const float y = -1.0f + 2.0f * ((float)j + 0.5f) / (float)s.fy;
                const float z = -1.0f + 2.0f * ((float)k + 0.5f) / (float)s.fz;
                v = marschner_lobb(x, y, z);
            } else if (s.kind == VR_SYNTH_FBM) {
                v = fbm(s.seed, (float)i + 0.5f, (float)j + 0.5f, (float)k + 0.5f, s.inv_period);
            } else {
                v = blobs_at(blobs, (float)i + 0.5f, (float)j + 0.5f, (float)k + 0.5f);
                if (s.kind == VR_SYNTH_BLOBS_FBM && v > 0.0f)
                    v *= 0.25f + 0.75f * fbm(s.seed, (float)i + 0.5f, (float)j + 0.5f, (float)k + 0.5f, s.inv_period);
            }
        }
        switch (s.dtype) {
            case VR_U8: reinterpret_cast<unsigned char *>(s.out)[idx] = (unsigned char)(v * 255.0f + 0.5f); break;
            case VR_U16: reinterpret_cast<unsigned short *>(s.out)[idx] = (unsigned short)(v * 65535.0f + 0.5f); break;
            default: reinterpret_cast<float *>(s.out)[idx] = v; break;
        }
    }
}

inline uint64_t splitmix64(uint64_t *s)
{
    uint64_t z = (*s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
inline float rnd01(uint64_t *s) { return (float)(splitmix64(s) >> 40) * (1.0f / 16777216.0f); }

}  // namespace

cudaError_t vr_k0_global_minmax(const void *dev, int dtype, size_t n, float *h_min, float *h_max, cudaStream_t s)
{
    int dev_id = 0, sms = 148;
    cudaGetDevice(&dev_id);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev_id);
    const int blocks = sms * 8;
    float2 *d_partial = nullptr;
    cudaError_t e = cudaMalloc(&d_partial, sizeof(float2) * blocks);
    if (e != cudaSuccess) return e;
    switch (dtype) {
        case VR_U8: k_global_minmax<unsigned char><<<blocks, 256, 0, s>>>((const unsigned char *)dev, n, d_partial); break;
        case VR_U16: k_global_minmax<unsigned short><<<blocks, 256, 0, s>>>((const unsigned short *)dev, n, d_partial); break;
        default: k_global_minmax<float><<<blocks, 256, 0, s>>>((const float *)dev, n, d_partial); break;
    }
    e = cudaGetLastError();
    std::vector<float2> h(blocks);
    if (e == cudaSuccess) e = cudaMemcpyAsync(h.data(), d_partial, sizeof(float2) * blocks, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d_partial);
    if (e != cudaSuccess) return e;
    float mn = h[0].x, mx = h[0].y;
    for (int i = 1; i < blocks; i++) mn = fminf(mn, h[i].x), mx = fmaxf(mx, h[i].y);
    *h_min = mn, *h_max = mx;
    return cudaSuccess;
}

cudaError_t vr_k0_cell_ranges(const void *dev, int dtype, uint32_t nx, uint32_t ny, uint32_t nz, int cell,
                              const int occ_n[3], float2 *d_ranges, cudaStream_t s)
{
    const int ncells = occ_n[0] * occ_n[1] * occ_n[2];
    k_cell_ranges<<<ncells, 64, 0, s>>>(dev, dtype, (int)nx, (int)ny, (int)nz, cell, occ_n[0], occ_n[1], d_ranges);
    return cudaGetLastError();
}

cudaError_t vr_k0_classify(const float2 *d_ranges, const int occ_n[3], const uint32_t *d_lut_nz, uint32_t nlut,
                           float min_density, float max_density, uint32_t *d_occ, int occ_words,
                           unsigned long long *d_empty_count, cudaStream_t s)
{
    const int ncells = occ_n[0] * occ_n[1] * occ_n[2];
    cudaError_t e = cudaMemsetAsync(d_empty_count, 0, sizeof(unsigned long long), s);
    if (e != cudaSuccess) return e;
    const int threads = occ_words * 32;
    k_classify<<<(threads + 255) / 256, 256, 0, s>>>(d_ranges, ncells, d_lut_nz, (int)nlut, min_density, max_density,
                                                     d_occ, d_empty_count);
    return cudaGetLastError();
}

/* ---- screen-space mask of the tiles whose rays can meet an occupied cell -------------------------------------
 *   With a sparse volume most rays inside the cube's screen box cross empty cells only: their
 * pixels are the clear colour, and the ray set-up (an unprojection, a slab test, three reciprocals, a normalisation:
 * about a third of config 3's instructions) is spent for nothing.  Once per LUT / volume change the occupied cells
 * are gathered into a list; once per frame every listed cell is projected (its eight corners through proj * view)
 * and the CTA tiles under its screen bounding box are marked.  A sample contributes only if it is POSITIONED in an
 * occupied cell (k_classify), and a pixel whose ray passes through a cell lies inside the projection of that cell,
 * so an unmarked tile is the clear colour exactly: K1 stores it without setting a ray up. */
__global__ void __launch_bounds__(256) k_list_occupied(const uint32_t *__restrict__ occ, int ncells, uint32_t *list, uint32_t *count)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const bool on = c < ncells && ((occ[c >> 5] >> (c & 31)) & 1u);
    const uint32_t m = __ballot_sync(0xffffffffu, on);
    uint32_t base = 0;
    if ((threadIdx.x & 31) == 0 && m) base = atomicAdd(count, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (on) list[base + __popc(m & ((1u << (threadIdx.x & 31)) - 1u))] = (uint32_t)c;
}

struct TileMaskParams {
    /* clip-space x, y, w of cell (0,0,0)'s low corner and their increments per cell along each axis: the eight
     * corners of a cell are sums of these (proj * view is affine in the model position) */
    float h0[3], ax[3], ay[3], az[3];
    int cnx, cny;
    uint32_t W, H, ntx, tile_h, words;
    uint32_t *mask; /* words + 1: the bits, then a flag set when a cell cannot be projected (behind the eye): K1 ignores the mask */
    uint32_t *wipe; /* the slot's other buffer, read by the frame before: zeroed here for the frame after */
};

template <bool STAGED>
__global__ void __launch_bounds__(256) k_tile_mask(const __grid_constant__ TileMaskParams p, const uint32_t *__restrict__ list,
                                                   const uint32_t *__restrict__ count)
{
    /* STAGED: the block marks a copy of the mask in shared memory and merges the words it touched into the global one
     * at the end.  The 256 cells of a block are neighbours in the grid, so their marks pile onto a few dozen words: done
     * directly in global memory those atomics queue up on the same L2 lines (measured: 86 % of the stall samples). */
    extern __shared__ uint32_t s_mask[];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    for (uint32_t w = i; w <= p.words; w += gridDim.x * blockDim.x) p.wipe[w] = 0u;
    if (STAGED) {
        for (uint32_t w = threadIdx.x; w < p.words; w += blockDim.x) s_mask[w] = 0u;
        __syncthreads();
    }
    uint32_t *const m = STAGED ? s_mask : p.mask;
    if (i < *count) {
        const uint32_t c = list[i];
        const uint32_t cxy = c / (uint32_t)p.cnx, cx = c - cxy * (uint32_t)p.cnx, cz = cxy / (uint32_t)p.cny, cy = cxy - cz * (uint32_t)p.cny;
        const float fx = (float)cx, fy = (float)cy, fz = (float)cz;
        float b[3];
#pragma unroll
        for (int k = 0; k < 3; k++) b[k] = fmaf(fz, p.az[k], fmaf(fy, p.ay[k], fmaf(fx, p.ax[k], p.h0[k])));
        float x0 = 1e30f, y0 = 1e30f, x1 = -1e30f, y1 = -1e30f, wmin = 1e30f;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const float hx = b[0] + ((k & 1) ? p.ax[0] : 0.f) + ((k & 2) ? p.ay[0] : 0.f) + ((k & 4) ? p.az[0] : 0.f);
            const float hy = b[1] + ((k & 1) ? p.ax[1] : 0.f) + ((k & 2) ? p.ay[1] : 0.f) + ((k & 4) ? p.az[1] : 0.f);
            const float hw = b[2] + ((k & 1) ? p.ax[2] : 0.f) + ((k & 2) ? p.ay[2] : 0.f) + ((k & 4) ? p.az[2] : 0.f);
            const float iw = __fdividef(1.0f, hw); /* 2 ulp: the bounds below carry a pixel and a half of slack */
            const float sx = hx * iw, sy = hy * iw;
            wmin = fminf(wmin, hw);
            x0 = fminf(x0, sx), x1 = fmaxf(x1, sx), y0 = fminf(y0, sy), y1 = fmaxf(y1, sy);
        }
        if (!(wmin > 1e-4f) || !(x0 == x0) || !(x1 == x1) || !(y0 == y0) || !(y1 == y1)) {
            p.mask[p.words] = 1u;
        } else {
            /* NDC -> pixels; pixel centres inside [lo - 1.5, hi + 1.5) */
            const float hwid = 0.5f * (float)p.W, hhgt = 0.5f * (float)p.H;
            const int px0 = max((int)floorf(fmaf(x0, hwid, hwid) - 1.5f), 0), px1 = min((int)ceilf(fmaf(x1, hwid, hwid) + 1.5f), (int)p.W - 1);
            const int py0 = max((int)floorf(fmaf(y0, hhgt, hhgt) - 1.5f), 0), py1 = min((int)ceilf(fmaf(y1, hhgt, hhgt) + 1.5f), (int)p.H - 1);
            if (px1 >= px0 && py1 >= py0) {
                const uint32_t tx0 = (uint32_t)px0 / VR_TILE, tx1 = (uint32_t)px1 / VR_TILE, ty0 = (uint32_t)py0 / p.tile_h, ty1 = (uint32_t)py1 / p.tile_h;
                for (uint32_t ty = ty0; ty <= ty1; ty++) {
                    /* the tiles of one row are consecutive bits: whole words at a time */
                    const uint32_t b0 = ty * p.ntx + tx0, b1 = ty * p.ntx + tx1;
                    for (uint32_t w = b0 >> 5; w <= (b1 >> 5); w++) {
                        const uint32_t lo_bit = w == (b0 >> 5) ? (b0 & 31u) : 0u, hi_bit = w == (b1 >> 5) ? (b1 & 31u) : 31u;
                        const uint32_t bits = (0xffffffffu >> (31u - hi_bit)) & (0xffffffffu << lo_bit);
                        /* most cells find their bits set already (a stale read only costs a redundant atomic) */
                        const uint32_t have = STAGED ? *(volatile uint32_t *)(m + w) : __ldcg(m + w);
                        if ((have & bits) != bits) atomicOr(m + w, bits);
                    }
                }
            }
        }
    }
    if (STAGED) {
        __syncthreads();
        for (uint32_t w = threadIdx.x; w < p.words; w += blockDim.x) {
            const uint32_t bits = s_mask[w];
            if (bits && (__ldcg(p.mask + w) & bits) != bits) atomicOr(p.mask + w, bits);
        }
    }
}

cudaError_t vr_k0_list_occupied(const uint32_t *d_occ, const int occ_n[3], uint32_t *d_list, uint32_t *d_count, cudaStream_t s)
{
    const int n = occ_n[0] * occ_n[1] * occ_n[2];
    cudaError_t e = cudaMemsetAsync(d_count, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    k_list_occupied<<<(n + 255) / 256, 256, 0, s>>>(d_occ, n, d_list, d_count);
    return cudaGetLastError();
}

cudaError_t vr_k0_tile_mask(const float pv[16], const float lo[3], const float cell[3], const int occ_n[3], const uint32_t *d_list,
                            const uint32_t *d_count, uint32_t max_count, uint32_t W, uint32_t H, uint32_t ntx, uint32_t tile_h,
                            uint32_t *d_mask, uint32_t *d_wipe, uint32_t mask_words, bool staged, cudaStream_t s)
{
    TileMaskParams p;
    const int row[3] = {0, 1, 3}; /* clip x, y, w */
    for (int k = 0; k < 3; k++) {
        const int r = row[k];
        p.h0[k] = pv[r] * lo[0] + pv[4 + r] * lo[1] + pv[8 + r] * lo[2] + pv[12 + r];
        p.ax[k] = pv[r] * cell[0], p.ay[k] = pv[4 + r] * cell[1], p.az[k] = pv[8 + r] * cell[2];
    }
    p.cnx = occ_n[0], p.cny = occ_n[1], p.W = W, p.H = H, p.ntx = ntx, p.tile_h = tile_h, p.words = mask_words;
    p.mask = d_mask, p.wipe = d_wipe;
    /* one thread per occupied cell (the host knows how many since the classification), and at least the threads the wipe needs */
    const uint32_t threads = max_count > mask_words + 1 ? max_count : mask_words + 1;
    const size_t smem = sizeof(uint32_t) * (size_t)mask_words;
    if (staged && smem <= 40 * 1024) k_tile_mask<true><<<(threads + 255u) / 256u, 256, smem, s>>>(p, d_list, d_count);
    else k_tile_mask<false><<<(threads + 255u) / 256u, 256, 0, s>>>(p, d_list, d_count);
    return cudaGetLastError();
}

cudaError_t vr_k0_distance(const uint32_t *d_occ, const int occ_n[3], unsigned char *d_dist, uint32_t *d_sat,
                           cudaStream_t s)
{
    const int nx = occ_n[0], ny = occ_n[1], nz = occ_n[2];
    const int mx = nx + 1, my = ny + 1, mz = nz + 1;
    const int n = nx * ny * nz, m = mx * my * mz;
    k_sat_init<<<(m + 255) / 256, 256, 0, s>>>(d_occ, d_sat, nx, ny, nz);
    k_sat_scan<<<(my * mz + 127) / 128, 128, 0, s>>>(d_sat, mx, my, mz, 0);
    k_sat_scan<<<(mx * mz + 127) / 128, 128, 0, s>>>(d_sat, mx, my, mz, 1);
    k_sat_scan<<<(mx * my + 127) / 128, 128, 0, s>>>(d_sat, mx, my, mz, 2);
    k_octant_distance<<<dim3((n + 255) / 256, 8), 256, 0, s>>>(d_occ, d_sat, d_dist, nx, ny, nz);
    return cudaGetLastError();
}

cudaError_t vr_k0_pack_pairs(const void *dev, int dtype, uint32_t nx, uint32_t ny, uint32_t nz, int axis,
                             cudaArray_t pair_array, cudaArray_t pair_array_hi, int split, cudaStream_t s)
{
    if ((dtype != VR_U8 && dtype != VR_U16) || axis < 0 || axis > 2) return cudaErrorInvalidValue;
    if (nx > 65535u * 32u || ny > 65535u || nz > 65535u) return cudaErrorInvalidConfiguration;
    cudaSurfaceObject_t surf[2] = {0, 0};
    cudaArray_t arrs[2] = {pair_array, pair_array_hi};
    cudaError_t e = cudaSuccess;
    for (int k = 0; k < 2 && e == cudaSuccess; k++) {
        if (!arrs[k]) continue;
        cudaResourceDesc rd = {};
        rd.resType = cudaResourceTypeArray;
        rd.res.array.array = arrs[k];
        e = cudaCreateSurfaceObject(&surf[k], &rd);
    }
    if (e == cudaSuccess) {
        /* axis 2: (x blocks, y, layers z); axis 1: (x blocks, z, layers y); axis 0: (layer tiles, y tiles, z) */
        const dim3 grid = axis == 2   ? dim3((nx + 255) / 256, ny, nz + 1)
                          : axis == 1 ? dim3((nx + 255) / 256, nz, ny + 1)
                                      : dim3((nx + 1 + 31) / 32, (ny + 31) / 32, nz);
        if (dtype == VR_U8)
            k_pack_pairs<unsigned char, uchar2><<<grid, 256, 0, s>>>((const unsigned char *)dev, (int)nx, (int)ny, (int)nz, axis,
                                                                     surf[0], surf[1], split);
        else
            k_pack_pairs<unsigned short, ushort2><<<grid, 256, 0, s>>>((const unsigned short *)dev, (int)nx, (int)ny, (int)nz, axis,
                                                                       surf[0], surf[1], split);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    }
    for (int k = 0; k < 2; k++)
        if (surf[k]) cudaDestroySurfaceObject(surf[k]);
    return e;
}

cudaError_t vr_k0_synth(void *dev, int kind, uint32_t seed, int dtype, uint32_t nx, uint32_t ny, uint32_t nz,
                        const int org[3], const uint32_t full[3], cudaStream_t s)
{
    SynthParams sp;
    sp.out = dev, sp.kind = kind, sp.dtype = dtype, sp.seed = seed;
    sp.nx = nx, sp.ny = ny, sp.nz = nz;
    sp.ox = org[0], sp.oy = org[1], sp.oz = org[2];
    sp.fx = full[0], sp.fy = full[1], sp.fz = full[2];
    float period = (float)full[0] / 8.0f;
    if (period < 2.0f) period = 2.0f;
    sp.inv_period = 1.0f / period;
    Blobs b;
    {
        /* blob table: centres, sigmas (8..40 voxels at 1024^3, scaled with the grid), amplitudes */
        uint64_t st = (uint64_t)seed * 0x100000001b3ull + 0x5eedull;
        const float n = (float)full[0], scale = n / 1024.0f;
        for (int i = 0; i < VR_NBLOBS; i++) {
            b.cx[i] = (0.12f + 0.76f * rnd01(&st)) * n;
            b.cy[i] = (0.12f + 0.76f * rnd01(&st)) * n;
            b.cz[i] = (0.12f + 0.76f * rnd01(&st)) * n;
            b.sigma[i] = (8.0f + 32.0f * rnd01(&st)) * scale;
            b.amp[i] = 0.35f + 0.65f * rnd01(&st);
        }
    }
    int dev_id = 0, sms = 148;
    cudaGetDevice(&dev_id);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev_id);
    const size_t total = (size_t)nx * ny * nz;
    size_t blocks = (total + 255) / 256;
    if (blocks > (size_t)sms * 32) blocks = (size_t)sms * 32;
    if (blocks == 0) blocks = 1;
    k_synth<<<(unsigned)blocks, 256, 0, s>>>(sp, b);
    return cudaGetLastError();
}

/* ---- min/max of a device array with NO allocation (vr_minmax_static, include/vr.h) -------------------------------
 * No counterpart in the reference; added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is k_global_minmax with its
 * per-block partials in a file-scope device array instead of a cudaMalloc'd buffer, and the last reduction on the
 * device instead of on the host: two launches on the caller's stream, nothing else -- no cudaMalloc, no cudaFree, no
 * host allocation, no copy, no synchronisation.
 *
 * The array lives in every GPU that loads the library until the process exits: VR_STATIC_MINMAX_BLOCKS float2 = 16 KiB.
 * ALL calls on a device share it, and it keeps the last call's contents: every launch overwrites each partial it reads
 * later (one per block, all `blocks` of them), so nothing has to start at zero -- but two calls that overlap on the GPU
 * (different streams, no ordering) race on it.  Serialise them.
 *
 * In host code `g_minmax_scratch` is a host-side stand-in, not a device address: the kernels get the address
 * cudaGetSymbolAddress returns, taken once per device with that device current. */
#define VR_STATIC_MINMAX_BLOCKS 2048
#define VR_STATIC_MINMAX_MAX_DEVICES 64
static __device__ __align__(256) float2 g_minmax_scratch[VR_STATIC_MINMAX_BLOCKS];

namespace
{
float2 *g_minmax_scratch_ptr[VR_STATIC_MINMAX_MAX_DEVICES]; /* device address of the array on each device, 0 = not taken yet */
int g_minmax_blocks[VR_STATIC_MINMAX_MAX_DEVICES];

__global__ void __launch_bounds__(256) k_minmax_final(const float2 *__restrict__ partial, int n, float *out)
{
    float mn = CUDART_INF_F, mx = -CUDART_INF_F;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float2 p = partial[i];
        mn = fminf(mn, p.x), mx = fmaxf(mx, p.y);
    }
    __shared__ float2 r;
    block_minmax(mn, mx, &r);
    __syncthreads();
    if (threadIdx.x == 0) out[0] = r.x, out[1] = r.y;
}

int static_fail(char *errbuf, size_t len, int status, const char *msg, cudaError_t e = cudaSuccess)
{
    if (errbuf && len) {
        if (e != cudaSuccess) std::snprintf(errbuf, len, "vr_minmax_static: %s: %s", msg, cudaGetErrorString(e));
        else std::snprintf(errbuf, len, "vr_minmax_static: %s", msg);
    }
    return status;
}
}  // namespace

extern "C" uint64_t vr_minmax_static_max_count(void) { return VR_MINMAX_STATIC_MAX_COUNT; }

extern "C" int vr_minmax_static(int device, const void *dev_values, int dtype, uint64_t count, float *dev_min_max, void *cuda_stream,
                                char *errbuf, size_t errbuf_len)
{
    if (!dev_values || !dev_min_max) return static_fail(errbuf, errbuf_len, VR_ERR_INVALID, "null input or output pointer");
    if (dtype < VR_U8 || dtype > VR_F32) return static_fail(errbuf, errbuf_len, VR_ERR_INVALID, "dtype must be VR_U8, VR_U16 or VR_F32");
    if (count == 0) return static_fail(errbuf, errbuf_len, VR_ERR_INVALID, "count must be >= 1");
    if (count > VR_MINMAX_STATIC_MAX_COUNT)
        return static_fail(errbuf, errbuf_len, VR_ERR_INVALID, "count exceeds the maximum of 2^36 values (vr_minmax_static_max_count): split the array");
    if (reinterpret_cast<uintptr_t>(dev_values) & 15u) return static_fail(errbuf, errbuf_len, VR_ERR_INVALID, "dev_values must be 16-byte aligned");
    if (reinterpret_cast<uintptr_t>(dev_min_max) & 3u) return static_fail(errbuf, errbuf_len, VR_ERR_INVALID, "dev_min_max must be 4-byte aligned");
    if (device < 0 || device >= VR_STATIC_MINMAX_MAX_DEVICES) return static_fail(errbuf, errbuf_len, VR_ERR_INVALID, "device index out of range");
    int before = -1;
    cudaError_t e = cudaGetDevice(&before);
    if (e != cudaSuccess) return static_fail(errbuf, errbuf_len, VR_ERR_CUDA, "no CUDA device (this backend has no CPU fallback)", e);
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return static_fail(errbuf, errbuf_len, VR_ERR_CUDA, "cudaSetDevice", e);
    int rc = VR_OK;
    if (!g_minmax_scratch_ptr[device]) {
        /* set-up, once per device, with that device current: the array's address ON THIS DEVICE */
        float2 *p = nullptr;
        int sms = 0;
        e = cudaGetSymbolAddress((void **)&p, g_minmax_scratch);
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess || !p) rc = static_fail(errbuf, errbuf_len, VR_ERR_CUDA, "cudaGetSymbolAddress of the scratch array", e);
        else g_minmax_blocks[device] = std::min(sms * 8, VR_STATIC_MINMAX_BLOCKS), g_minmax_scratch_ptr[device] = p;
    }
    if (rc == VR_OK) {
        cudaStream_t s = (cudaStream_t)cuda_stream;
        float2 *partial = g_minmax_scratch_ptr[device];
        const int blocks = g_minmax_blocks[device];
        const size_t n = (size_t)count;
        switch (dtype) {
            case VR_U8: k_global_minmax<unsigned char><<<blocks, 256, 0, s>>>((const unsigned char *)dev_values, n, partial); break;
            case VR_U16: k_global_minmax<unsigned short><<<blocks, 256, 0, s>>>((const unsigned short *)dev_values, n, partial); break;
            default: k_global_minmax<float><<<blocks, 256, 0, s>>>((const float *)dev_values, n, partial); break;
        }
        k_minmax_final<<<1, 256, 0, s>>>(partial, blocks, dev_min_max);
        e = cudaGetLastError();
        if (e != cudaSuccess) rc = static_fail(errbuf, errbuf_len, VR_ERR_CUDA, "kernel launch", e);
    }
    if (before >= 0 && before != device) cudaSetDevice(before);
    return rc;
}
