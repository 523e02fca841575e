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
#include <cuda.h> /* types and constants of the stream memory operations only: the entry points are looked up at run time */
#include <cuda_runtime.h>

#include <cstdlib>

#include "vr_ctx.h"

namespace
{
/* VR_SYNC_MEMOPS=1 (experiment, off by default): wait and signal with the driver's STREAM MEMORY OPERATIONS
 * (cuStreamWaitValue32 GEQ / cuStreamWriteValue32) instead of the one-warp kernels below -- the wait is then done by the
 * front end of the stream like an event wait, takes no SM and is not scheduled behind the CTAs of a frame.  There is no
 * time-out in this form.  The library does not link libcuda: the two entry points come from cudaGetDriverEntryPoint. */
typedef CUresult (*MemOpFn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
MemOpFn g_wait32 = nullptr, g_write32 = nullptr;
int g_memops = 0; /* 0 not looked at yet, 1 in use, -1 off / unavailable */

bool memops()
{
    if (g_memops == 0) {
        g_memops = -1;
        const char *e = std::getenv("VR_SYNC_MEMOPS");
        if (e && std::atoi(e) == 1) {
            void *w = nullptr, *v = nullptr;
            cudaDriverEntryPointQueryResult q1, q2;
            if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &v, cudaEnableDefault, &q1) == cudaSuccess && q1 == cudaDriverEntryPointSuccess && v &&
                cudaGetDriverEntryPoint("cuStreamWriteValue32", &w, cudaEnableDefault, &q2) == cudaSuccess && q2 == cudaDriverEntryPointSuccess && w) {
                g_wait32 = reinterpret_cast<MemOpFn>(v), g_write32 = reinterpret_cast<MemOpFn>(w);
                g_memops = 1;
            }
        }
    }
    return g_memops == 1;
}

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

extern "C" int vr_sync_uses_memops(void) { return memops() ? 1 : 0; }

cudaError_t vr_launch_signal(void *const *flags, int n, uint32_t value, cudaStream_t s)
{
    if (n <= 0) return cudaSuccess;
    if (n > VR_SIGNAL_MAX) return cudaErrorInvalidValue;
    if (memops()) {
        /* (the write is ordered behind everything earlier in the stream and flushes its memory writes first) */
        for (int i = 0; i < n; i++)
            if (g_write32((CUstream)s, (CUdeviceptr)(uintptr_t)flags[i], value, CU_STREAM_WRITE_VALUE_DEFAULT) != CUDA_SUCCESS) return cudaErrorUnknown;
        return cudaSuccess;
    }
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
    if (memops()) {
        /* GEQ is the cyclic comparison (int32)(*addr - value) >= 0 of k_wait */
        for (int i = 0; i < n; i++)
            if (g_wait32((CUstream)s, (CUdeviceptr)(uintptr_t)(reinterpret_cast<const uint32_t *>(flags) + i), value, CU_STREAM_WAIT_VALUE_GEQ) != CUDA_SUCCESS)
                return cudaErrorUnknown;
        return cudaSuccess;
    }
    k_wait<<<1, (n + 31) / 32 * 32, 0, s>>>(reinterpret_cast<const uint32_t *>(flags), n, value,
                                            (unsigned long long)timeout_ms * 1000000ull, error_word);
    return cudaGetLastError();
}
