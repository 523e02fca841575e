/*
 * sync.cu -- stream-ordered frame hand-off between the GPUs of one box.
 *
 * The reference orders its frames in flight with per-frame fences on one device
 * (src/rendering/main_pass.cpp:55-57,:115-117).  With one process per GPU the same ordering
 * crosses devices: "rank r's tiles of frame k are in the display image", "frame k has been consumed,
 * its slot may be overwritten".  These are 32-bit counters in device memory (a context's flag block,
 * mapped into the peers with CUDA IPC) written and awaited by one-warp kernels in stream order, so
 * a frame never costs a host synchronisation or a collective: the host only enqueues.
 *
 *   k_signal  after everything earlier in the stream (e.g. K1's peer stores): fence, then release-store
 *             the value into up to 32 flags, local or peer.
 *   k_wait    spin (acquire loads, nanosleep back-off) until n local flags have reached the value;
 *             gives up after a timeout and raises the context's error word instead of hanging the GPU.
 */
#include <cuda_runtime.h>

#include "vr_ctx.h"

namespace
{
struct FlagList {
    uint32_t *p[VR_SIGNAL_MAX];
};

__global__ void __launch_bounds__(32) k_signal(const __grid_constant__ FlagList f, int n, uint32_t value)
{
    __threadfence_system();
    if ((int)threadIdx.x < n) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f.p[threadIdx.x]), "r"(value) : "memory");
}

__global__ void __launch_bounds__(256) k_wait(const uint32_t *flags, int n, uint32_t value, unsigned long long timeout_ns,
                                              unsigned long long *error_word)
{
    if ((int)threadIdx.x < n) {
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while (true) {
            uint32_t cur;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(cur) : "l"(flags + threadIdx.x) : "memory");
            if ((int32_t)(cur - value) >= 0) break; /* counters only grow; wrap-safe compare */
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > timeout_ns) {
                *error_word = 1ull + threadIdx.x;
                break;
            }
            __nanosleep(100);
        }
    }
    __syncthreads();
    __threadfence_system();
}
}  // namespace

cudaError_t vr_launch_signal(void *const *flags, int n, uint32_t value, cudaStream_t s)
{
    if (n <= 0) return cudaSuccess;
    if (n > VR_SIGNAL_MAX) return cudaErrorInvalidValue;
    FlagList f = {};
    for (int i = 0; i < n; i++) f.p[i] = reinterpret_cast<uint32_t *>(flags[i]);
    k_signal<<<1, 32, 0, s>>>(f, n, value);
    return cudaGetLastError();
}

cudaError_t vr_launch_wait(const void *flags, int n, uint32_t value, uint32_t timeout_ms, unsigned long long *error_word,
                           cudaStream_t s)
{
    if (n <= 0) return cudaSuccess;
    if (n > 256) return cudaErrorInvalidValue;
    k_wait<<<1, (n + 31) / 32 * 32, 0, s>>>(reinterpret_cast<const uint32_t *>(flags), n, value,
                                            (unsigned long long)timeout_ms * 1000000ull, error_word);
    return cudaGetLastError();
}
