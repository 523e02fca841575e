/*
 * exp_tma.cu -- experiment: could config 2 leave the texture path?
 *
 * The alternative to texture fetches is to stage the volume in shared memory with the bulk-copy engine (TMA,
 * cp.async.bulk behind an mbarrier ring) from a K0-bricked LINEAR copy and filter trilinearly in fp32 FMAs.  This
 * measures the two halves of that design at their best, with everything a real march would add left out:
 *
 *   feed     every CTA streams bricks of 18^3 u16 (16^3 interior + a one-voxel apron, 11 664 contiguous bytes) through a
 *            3-stage shared-memory ring and touches one word per thread: the rate the copy engine feeds the SMs at;
 *   sample   the same, and every thread takes SPB / 128 trilinear samples per brick from shared memory along a ray
 *            through the brick (8 LDS.U16 + the weights + 7 FMA each), accumulating the value: no transfer function, no
 *            compositing, no ray bookkeeping between bricks, bricks dealt perfectly, every ray axis-parallel-ish.
 *
 * Config 2 takes 5.1e7 samples per frame from 32^3 bricks of a 512^3 volume: 1556 samples per brick on average; K1
 * delivers 4.8-5.0e11 samples/s including the transfer function and the compositing (36.75 instructions per sample),
 * its DRAM-fed texture ceiling is 7.66e11 fetches/s.
 *
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/exp_tma.bin tools/exp_tma.cu
 * Run:   tools/exp_tma.bin            (prints one line per variant)
 */
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define B 18
#define BRICK_VOX (B * B * B)
#define BRICK_BYTES (BRICK_VOX * 2) /* 11 664, a multiple of 16 */
#define STAGES 3
#define THREADS 128

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            std::fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            std::exit(1);                                                                       \
        }                                                                                       \
    } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

__global__ void k_fill(unsigned short *v, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        v[i] = (unsigned short)((i * 2654435761u) >> 13);
}

/* SPT = samples per thread and brick (0: feed only) */
template <int SPT>
__global__ void __launch_bounds__(THREADS) k_bricks(const unsigned short *__restrict__ bricks, int nbricks, float *out)
{
    extern __shared__ __align__(128) unsigned char smem[];
    auto stage_ptr = [&](int s) { return reinterpret_cast<unsigned short *>(smem + (size_t)s * ((BRICK_BYTES + 127) / 128 * 128)); };
    __shared__ __align__(8) uint64_t full[STAGES];
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int first = blockIdx.x, stride = gridDim.x;
    const int mine = first < nbricks ? (nbricks - first + stride - 1) / stride : 0;
    if (threadIdx.x == 0)
        for (int i = 0; i < STAGES && i < mine; i++) {
            mbar_expect_tx(&full[i], BRICK_BYTES);
            bulk_g2s(stage_ptr(i), bricks + (size_t)(first + i * stride) * BRICK_VOX, BRICK_BYTES, &full[i]);
        }
    /* this thread's ray through a brick: enters at (1 + x, 1 + y, 1) of the interior, direction mostly along z */
    const float ex = 1.0f + (float)(threadIdx.x & 15) * 0.93f, ey = 1.0f + (float)(threadIdx.x >> 4) * 1.9f;
    const float dx = 0.011f * (float)((int)(threadIdx.x & 3) - 1), dy = 0.007f, dz = SPT ? 15.9f / (float)SPT : 0.0f;
    float acc = 0.0f;
    for (int i = 0; i < mine; i++) {
        const int s = i % STAGES;
        mbar_wait(&full[s], (uint32_t)((i / STAGES) & 1));
        const unsigned short *v = stage_ptr(s);
        if (SPT == 0) {
            acc += (float)v[threadIdx.x * 45];
        } else {
            float px = ex, py = ey, pz = 1.0f;
#pragma unroll 4
            for (int k = 0; k < SPT; k++) {
                const float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
                const float wx = px - fx, wy = py - fy, wz = pz - fz;
                const int o = ((int)fz * B + (int)fy) * B + (int)fx;
                const float c000 = (float)v[o], c100 = (float)v[o + 1], c010 = (float)v[o + B], c110 = (float)v[o + B + 1];
                const float c001 = (float)v[o + B * B], c101 = (float)v[o + B * B + 1], c011 = (float)v[o + B * B + B], c111 = (float)v[o + B * B + B + 1];
                const float a0 = fmaf(wx, c100 - c000, c000), a1 = fmaf(wx, c110 - c010, c010), a2 = fmaf(wx, c101 - c001, c001),
                            a3 = fmaf(wx, c111 - c011, c011);
                const float b0 = fmaf(wy, a1 - a0, a0), b1 = fmaf(wy, a3 - a2, a2);
                acc += fmaf(wz, b1 - b0, b0);
                px += dx, py += dy, pz += dz;
            }
        }
        __syncthreads(); /* everyone has read the stage: refill it */
        if (threadIdx.x == 0 && i + STAGES < mine) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&full[s], BRICK_BYTES);
            bulk_g2s(stage_ptr(s), bricks + (size_t)(first + (i + STAGES) * stride) * BRICK_VOX, BRICK_BYTES, &full[s]);
        }
    }
    if (acc == 123.456f) out[0] = acc;
}

template <int SPT>
static void run(const unsigned short *bricks, int nbricks, float *out, int ctas, const char *label)
{
    const size_t smem = (size_t)STAGES * ((BRICK_BYTES + 127) / 128 * 128);
    CK(cudaFuncSetAttribute(k_bricks<SPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; i++) k_bricks<SPT><<<ctas, THREADS, smem>>>(bricks, nbricks, out);
    CK(cudaDeviceSynchronize());
    const int reps = 20;
    CK(cudaEventRecord(e0));
    for (int i = 0; i < reps; i++) k_bricks<SPT><<<ctas, THREADS, smem>>>(bricks, nbricks, out);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    ms /= reps;
    const double bytes = (double)nbricks * BRICK_BYTES, samples = (double)nbricks * THREADS * SPT;
    std::printf("%-34s %d CTAs: %.4f ms per pass over %d bricks = %.0f GB/s fed", label, ctas, ms, nbricks, bytes / (ms * 1e-3) / 1e9);
    if (SPT) std::printf(", %.1f Gsample/s (%d samples per brick)", samples / (ms * 1e-3) / 1e9, THREADS * SPT);
    std::printf("\n");
}

int main()
{
    const int nbricks = 32 * 32 * 32; /* a 512^3 volume in 16^3 bricks */
    unsigned short *bricks;
    float *out;
    CK(cudaMalloc(&bricks, (size_t)nbricks * BRICK_BYTES));
    CK(cudaMalloc(&out, 4));
    k_fill<<<148 * 8, 256>>>(bricks, (size_t)nbricks * BRICK_VOX);
    CK(cudaDeviceSynchronize());
    int sms = 148;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    std::printf("bricks of %d^3 u16 = %d bytes, %d-stage ring per CTA, %d threads; volume copy %.0f MB (1.42x the 512^3 volume)\n", B, BRICK_BYTES, STAGES,
                THREADS, (double)nbricks * BRICK_BYTES / 1e6);
    for (int per_sm : {2, 4, 6}) {
        run<0>(bricks, nbricks, out, sms * per_sm, "feed only");
        run<4>(bricks, nbricks, out, sms * per_sm, "feed + 4 trilinear samples/thread");
        run<12>(bricks, nbricks, out, sms * per_sm, "feed + 12 samples/thread (config 2)");
        run<24>(bricks, nbricks, out, sms * per_sm, "feed + 24 samples/thread");
    }
    return 0;
}
