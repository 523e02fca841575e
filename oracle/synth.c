/*
 * synth.c -- procedural test volumes of SURVEY.md section 8(d), CPU restatement.
 *
 * TEST INFRASTRUCTURE ONLY (see vr_oracle.h).  The reference ships no datasets; these
 * are the synthetic inputs BASELINE.json names (Marschner-Lobb, fractal noise, sparse
 * blobs).  The CUDA generators in volumetric-renderer_b200/csrc/synth.cu follow the same
 * formulas; the two may differ by 1 LSB where libm and the GPU round sin/cos/exp
 * differently, so GPU parity tests render the volume the GPU generated.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "vr_oracle.h"

static inline uint32_t hash_u32(uint32_t h)
{
    h ^= h >> 16;
    h *= 0x7feb352du;
    h ^= h >> 15;
    h *= 0x846ca68bu;
    h ^= h >> 16;
    return h;
}

static inline float lattice(uint32_t seed, int ix, int iy, int iz, int oct)
{
    uint32_t h = seed ^ ((uint32_t)ix * 0x8da6b343u) ^ ((uint32_t)iy * 0xd8163841u) ^
                 ((uint32_t)iz * 0xcb1ab31fu) ^ ((uint32_t)oct * 0x9e3779b9u);
    return (float)(hash_u32(h) >> 8) * (1.0f / 16777216.0f);
}

static inline float fade(float t) { return t * t * (3.0f - 2.0f * t); }
static inline float mixf(float a, float b, float t) { return a + t * (b - a); }

static inline float value_noise(uint32_t seed, float x, float y, float z, int oct)
{
    float fx = floorf(x), fy = floorf(y), fz = floorf(z);
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    float tx = fade(x - fx), ty = fade(y - fy), tz = fade(z - fz);
    float c000 = lattice(seed, ix, iy, iz, oct), c100 = lattice(seed, ix + 1, iy, iz, oct);
    float c010 = lattice(seed, ix, iy + 1, iz, oct), c110 = lattice(seed, ix + 1, iy + 1, iz, oct);
    float c001 = lattice(seed, ix, iy, iz + 1, oct), c101 = lattice(seed, ix + 1, iy, iz + 1, oct);
    float c011 = lattice(seed, ix, iy + 1, iz + 1, oct), c111 = lattice(seed, ix + 1, iy + 1, iz + 1, oct);
    return mixf(mixf(mixf(c000, c100, tx), mixf(c010, c110, tx), ty),
                mixf(mixf(c001, c101, tx), mixf(c011, c111, tx), ty), tz);
}

/* kind 0: Marschner-Lobb, f_M = 6, alpha = 0.25, x,y,z in [-1,1] at voxel centres */
static inline float marschner_lobb(float x, float y, float z)
{
    const float PI = 3.14159265358979323846f, fM = 6.0f, al = 0.25f;
    float r = sqrtf(x * x + y * y);
    float pr = cosf(2.0f * PI * fM * cosf(PI * r / 2.0f));
    return (1.0f - sinf(PI * z / 2.0f) + al * (1.0f + pr)) / (2.0f * (1.0f + al));
}

/* kind 1: 5-octave value-noise fBm, lacunarity 2, gain 0.5, base period nx/8 voxels,
 * contrast-stretched about 0.5 so the data spans the full range */
static inline float fbm(uint32_t seed, float vx, float vy, float vz, float inv_period)
{
    float sum = 0.0f, amp = 1.0f, norm = 0.0f, f = inv_period;
    for (int o = 0; o < 5; o++) {
        sum += amp * value_noise(seed, vx * f, vy * f, vz * f, o);
        norm += amp;
        amp *= 0.5f;
        f *= 2.0f;
    }
    float v = 0.5f + (sum / norm - 0.5f) * 2.2f;
    return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
}

/* kind 2: 64 truncated Gaussian blobs on an exactly-zero background */
#define VRO_NBLOBS 64
typedef struct blob {
    float cx, cy, cz, sigma, amp;
} blob;

static inline uint64_t splitmix64(uint64_t *s)
{
    uint64_t z = (*s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

static inline float rnd01(uint64_t *s) { return (float)(splitmix64(s) >> 40) * (1.0f / 16777216.0f); }

static void make_blobs(uint32_t seed, float n, blob *b)
{
    uint64_t s = (uint64_t)seed * 0x100000001b3ull + 0x5eedull;
    float scale = n / 1024.0f;
    for (int i = 0; i < VRO_NBLOBS; i++) {
        b[i].cx = (0.12f + 0.76f * rnd01(&s)) * n;
        b[i].cy = (0.12f + 0.76f * rnd01(&s)) * n;
        b[i].cz = (0.12f + 0.76f * rnd01(&s)) * n;
        b[i].sigma = (8.0f + 32.0f * rnd01(&s)) * scale;
        b[i].amp = 0.35f + 0.65f * rnd01(&s);
    }
}

#define VRO_BLOB_CUT 2.5f
static inline float blobs_at(const blob *b, float x, float y, float z)
{
    const float floor_v = 0.0439369336f; /* exp(-2.5^2/2) */
    float sum = 0.0f;
    for (int i = 0; i < VRO_NBLOBS; i++) {
        float dx = x - b[i].cx, dy = y - b[i].cy, dz = z - b[i].cz;
        float d2 = dx * dx + dy * dy + dz * dz;
        float s2 = b[i].sigma * b[i].sigma;
        if (d2 < VRO_BLOB_CUT * VRO_BLOB_CUT * s2)
            sum += b[i].amp * (expf(-0.5f * d2 / s2) - floor_v) / (1.0f - floor_v);
    }
    return sum > 1.0f ? 1.0f : sum;
}

static inline void store(void *out, int dtype, size_t i, float v01)
{
    switch (dtype) {
        case VRO_U8: ((uint8_t *)out)[i] = (uint8_t)(v01 * 255.0f + 0.5f); break;
        case VRO_U16: ((uint16_t *)out)[i] = (uint16_t)(v01 * 65535.0f + 0.5f); break;
        default: ((float *)out)[i] = v01; break;
    }
}

int vro_synth(int kind, uint32_t seed, int dtype, uint32_t nx, uint32_t ny, uint32_t nz, void *out, int nthreads)
{
    if (!out || kind < 0 || kind > 2 || dtype < 0 || dtype > 2) return -1;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_num_procs();
#else
    nthreads = 1;
#endif
    blob blobs[VRO_NBLOBS];
    if (kind == 2) make_blobs(seed, (float)nx, blobs);
    float period = (float)nx / 8.0f;
    if (period < 2.0f) period = 2.0f;
    const float inv_period = 1.0f / period;
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (long k = 0; k < (long)nz; k++)
        for (uint32_t j = 0; j < ny; j++)
            for (uint32_t i = 0; i < nx; i++) {
                size_t idx = ((size_t)k * ny + j) * nx + i;
                float v;
                if (kind == 0) {
                    float x = -1.0f + 2.0f * ((float)i + 0.5f) / (float)nx;
                    float y = -1.0f + 2.0f * ((float)j + 0.5f) / (float)ny;
                    float z = -1.0f + 2.0f * ((float)k + 0.5f) / (float)nz;
                    v = marschner_lobb(x, y, z);
                } else if (kind == 1) {
                    v = fbm(seed, (float)i + 0.5f, (float)j + 0.5f, (float)k + 0.5f, inv_period);
                } else {
                    v = blobs_at(blobs, (float)i + 0.5f, (float)j + 0.5f, (float)k + 0.5f);
                }
                store(out, dtype, idx, v);
            }
    return 0;
}
