/*
 * group.cu -- vr_group: ONE process driving the GPUs of one box (the C++ multi-GPU host of the path).
 *
 * The reference constructs its single pass at src/rendering/vulkan_context.cpp:51 and calls record() once per
 * frame (src/rendering/main_pass.cpp:91); a group stands in the same place and records one frame on N devices.
 * Everything here is host code over the C ABI of vr.h: a context per rank, peer access instead of CUDA IPC (one
 * address space), and CUDA EVENTS instead of the flag kernels of sync.cu for the frame hand-off -- within one
 * process cudaStreamWaitEvent orders streams of different devices, so nothing ever spins on the GPU.
 *
 *   replicas     every rank renders the frame of its own camera (set through vr_group_context)
 *   sort-first   image tiles dealt to the ranks, volume replicated; every rank's K1 stores its tiles into the display
 *                rank's image slot over NVLink; `slots` frames in flight
 *   sort-last    one sub-box per rank (+ ghosts); K1 -> float4 partial (slot), K2 merges pixel range r of all
 *                partials in view order into the display rank's image slot; march and merge on two streams
 *
 * (The one-process-per-GPU variant of the same protocols, for torchrun: volumetric-renderer_b200/multigpu.py.)
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <array>
#include <atomic>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "vr_ctx.h"

struct vr_group {
    int partition = VR_PARTITION_REPLICAS;
    uint32_t n = 0, slots = 2, W = 0, H = 0;
    std::vector<int> dev;
    std::vector<vr_ctx *> ctx;
    std::vector<cudaStream_t> march, merge; /* per rank */
    cudaStream_t present = nullptr;         /* display rank (0) */
    /* events [rank][slot] */
    std::vector<std::vector<cudaEvent_t>> ready, merged;
    std::vector<cudaEvent_t> consumed; /* [slot], display rank */
    std::vector<void *> image_slot;    /* display rank's RGBA8 targets */
    std::vector<std::vector<void *>> raw_slot; /* [rank][slot] */
    uint64_t frame = 0;
    uint32_t full[3] = {0, 0, 0}, grid[3] = {1, 1, 1};
    std::vector<std::array<uint32_t, 6>> box;    /* org, dims each rank OWNS in the frame being recorded */
    std::vector<std::array<uint32_t, 6>> stored; /* org, dims each rank HOLDS */
    std::vector<std::array<uint32_t, 4>> slot_region; /* per image slot: what the frame merged into it last can have drawn */
    float balance = 0.0f;                        /* sort-last: how far the split planes may follow the camera (vr_group_set_balance) */
    uint32_t margin[3] = {0, 0, 0};
    float cam_pos[3] = {0, -3, 0};
    std::string err;
    /* One worker thread per rank enqueues that rank's part of a frame (above two ranks: a frame is ~20 API calls per rank,
     * and one thread issuing them for eight devices takes longer than the GPUs need for the frame).  A frame is two
     * phases with a host barrier between them, because cudaStreamWaitEvent waits for the record made LAST when it is
     * called: every rank's "ready" event must have been recorded (phase 1) before anyone waits on it (phase 2). */
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv;
    std::atomic<uint64_t> ticket{0};
    std::atomic<uint32_t> done{0};
    int phase = 0; /* 1 march, 2 merge, -1 quit */
    struct Job {
        uint64_t k = 0;
        uint32_t s = 0;
        const void *const *parts = nullptr;
        const uint32_t *rects = nullptr;
        uint32_t region[4] = {0, 0, 0, 0};
    } job;
    std::vector<int> wrc;
    std::vector<std::string> werr;
    cudaStream_t join = nullptr;            /* display rank: joins the ranks' "ready" events into all_ready[slot] */
    std::vector<cudaEvent_t> all_ready;
};

namespace
{
thread_local std::string g_group_error;

int gfail(vr_group *g, int status, const std::string &msg)
{
    if (g) g->err = msg;
    else g_group_error = msg;
    return status;
}

#define G_CUDA(g, call)                                                                              \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess) return gfail(g, VR_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); \
    } while (0)
#define G_VR(g, r, call)                                                                              \
    do {                                                                                              \
        const int rc__ = (call);                                                                      \
        if (rc__ != VR_OK) return gfail(g, rc__, std::string("rank ") + std::to_string(r) + ": " + vr_last_error((g)->ctx[r])); \
    } while (0)

/* split the axes as evenly as possible: 8 -> 2x2x2, 4 -> 2x2x1, 6 -> 3x2x1 (multigpu.grid_for_world) */
void grid_for_world(uint32_t world, uint32_t g[3])
{
    g[0] = g[1] = g[2] = 1;
    std::vector<uint32_t> factors;
    uint32_t w = world;
    for (uint32_t f = 2; w > 1; f++)
        while (w % f == 0) factors.push_back(f), w /= f;
    std::sort(factors.rbegin(), factors.rend());
    for (uint32_t f : factors) {
        int a = 0;
        for (int i = 1; i < 3; i++)
            if (g[i] < g[a]) a = i;
        g[a] *= f;
    }
}


int make_partition(vr_group *g, uint32_t nx, uint32_t ny, uint32_t nz)
{
    g->full[0] = nx, g->full[1] = ny, g->full[2] = nz;
    grid_for_world(g->n, g->grid);
    g->box.assign(g->n, {});
    g->stored.assign(g->n, {});
    /* load balance: the planes may follow the camera by `balance` of the edge; every rank holds that margin around its box */
    for (int i = 0; i < 3; i++) {
        const uint32_t room = g->grid[i] > 1 ? (g->full[i] / g->grid[i] - 1) / 2 : 0;
        g->margin[i] = g->grid[i] > 1 ? std::min<uint32_t>((uint32_t)(g->balance * (float)g->full[i]), room) : 0;
    }
    for (uint32_t r = 0; r < g->n; r++) {
        uint32_t org[3], dims[3], sorg[3], sdims[3];
        if (vr_partition_box(g->grid, g->full, r, org, dims) != VR_OK || vr_partition_stored_box(g->grid, g->full, r, g->margin, sorg, sdims) != VR_OK)
            return gfail(g, VR_ERR_INVALID, "vr_group: the volume is too small for one sub-box per rank");
        for (int i = 0; i < 3; i++) g->box[r][i] = org[i], g->box[r][3 + i] = dims[i], g->stored[r][i] = sorg[i], g->stored[r][3 + i] = sdims[i];
        G_VR(g, r, vr_set_partition_sort_last(g->ctx[r], g->full, sorg, sdims));
    }
    return VR_OK;
}

int reduce_density_range(vr_group *g)
{
    float mn = 0.f, mx = 0.f;
    for (uint32_t r = 0; r < g->n; r++) {
        float a, b;
        G_VR(g, r, vr_get_volume_info(g->ctx[r], nullptr, nullptr, &a, &b));
        mn = r ? std::min(mn, a) : a, mx = r ? std::max(mx, b) : b;
    }
    for (uint32_t r = 0; r < g->n; r++) G_VR(g, r, vr_set_density_range(g->ctx[r], mn, mx));
    return VR_OK;
}

/* image / partial slots and their peer pointers; called once the partition and the options are known */
int wire(vr_group *g)
{
    g->image_slot.assign(g->slots, nullptr);
    if (g->partition == VR_PARTITION_REPLICAS) return VR_OK;
    for (uint32_t s = 0; s < g->slots; s++) {
        unsigned char h[64];
        if (s) G_VR(g, 0, vr_ipc_export(g->ctx[0], 16 + (int)s, h)); /* allocates the slot (the handle itself is not needed here) */
        G_VR(g, 0, vr_get_image_frame(g->ctx[0], s, &g->image_slot[s], nullptr));
    }
    if (g->partition == VR_PARTITION_SORT_FIRST) {
        for (uint32_t r = 0; r < g->n; r++) {
            G_VR(g, r, vr_set_partition_sort_first(g->ctx[r], r, g->n, 64));
            for (uint32_t s = 0; s < g->slots; s++)
                if (r) G_VR(g, r, vr_set_image_target_frame(g->ctx[r], s, g->image_slot[s]));
        }
    } else {
        g->raw_slot.assign(g->n, std::vector<void *>(g->slots, nullptr));
        for (uint32_t r = 0; r < g->n; r++) {
            vr_options o;
            G_VR(g, r, vr_get_options(g->ctx[r], &o));
            o.write_raw = 1;
            G_VR(g, r, vr_set_options(g->ctx[r], &o));
            for (uint32_t s = 0; s < g->slots; s++) G_VR(g, r, vr_get_raw_frame(g->ctx[r], s, &g->raw_slot[r][s]));
        }
    }
    return VR_OK;
}
}  // namespace

extern "C" const char *vr_group_last_error(const vr_group *g) { return g ? g->err.c_str() : g_group_error.c_str(); }
extern "C" uint32_t vr_group_size(const vr_group *g) { return g ? g->n : 0; }
extern "C" vr_ctx *vr_group_context(vr_group *g, uint32_t rank) { return (g && rank < g->n) ? g->ctx[rank] : nullptr; }

namespace
{
/* what rank r enqueues in a phase of frame k; returns a status, the message in *err */
int rank_phase(vr_group *g, uint32_t r, int phase, std::string *err)
{
    const uint64_t k = g->job.k;
    const uint32_t S = g->slots, s = g->job.s, n = g->n;
    auto cu = [&](cudaError_t e, const char *what) {
        if (e == cudaSuccess) return true;
        *err = std::string(what) + ": " + cudaGetErrorString(e);
        return false;
    };
    auto vr = [&](int rc) {
        if (rc == VR_OK) return true;
        *err = std::string("rank ") + std::to_string(r) + ": " + vr_last_error(g->ctx[r]);
        return false;
    };
    if (!cu(cudaSetDevice(g->dev[r]), "cudaSetDevice")) return VR_ERR_CUDA;
    if (phase == 1) {
        /* the slot is free again: frame k - S has been presented out of the image slot (sort-first), every rank has merged
         * my partial of it (sort-last: `consumed` is recorded behind all the `merged` events) */
        if (k >= S && !cu(cudaStreamWaitEvent(g->march[r], g->consumed[s], 0), "cudaStreamWaitEvent")) return VR_ERR_CUDA;
        const int rc = vr_render_frame(g->ctx[r], s, g->march[r]);
        if (!vr(rc)) return rc;
        if (!cu(cudaEventRecord(g->ready[r][s], g->march[r]), "cudaEventRecord")) return VR_ERR_CUDA;
        return VR_OK;
    }
    /* phase 2, sort-last: every partial of frame k complete, image slot s presented, then K2 on my bands of the region */
    if (!cu(cudaStreamWaitEvent(g->merge[r], g->all_ready[s], 0), "cudaStreamWaitEvent")) return VR_ERR_CUDA;
    if (k >= S && !cu(cudaStreamWaitEvent(g->merge[r], g->consumed[s], 0), "cudaStreamWaitEvent")) return VR_ERR_CUDA;
    const int rc = vr_composite_region(g->ctx[r], g->job.parts, g->job.rects, n, g->job.region, r, n, g->image_slot[s], g->merge[r]);
    if (!vr(rc)) return rc;
    if (!cu(cudaEventRecord(g->merged[r][s], g->merge[r]), "cudaEventRecord")) return VR_ERR_CUDA;
    return VR_OK;
}

void worker_main(vr_group *g, uint32_t r)
{
    uint64_t seen = 0;
    for (;;) {
        int spins = 0;
        while (g->ticket.load(std::memory_order_acquire) == seen) {
            if (++spins < 4000) {
                std::this_thread::yield();
            } else { /* idle: sleep until the next frame */
                std::unique_lock<std::mutex> lk(g->mu);
                g->cv.wait(lk, [&] { return g->ticket.load(std::memory_order_acquire) != seen; });
            }
        }
        seen = g->ticket.load(std::memory_order_acquire);
        const int phase = g->phase;
        if (phase < 0) return;
        g->wrc[r] = rank_phase(g, r, phase, &g->werr[r]);
        g->done.fetch_add(1, std::memory_order_release);
    }
}

/* run a phase on every rank: on the workers when there are any, else inline; first failure wins */
int run_phase(vr_group *g, int phase)
{
    const uint32_t n = g->n;
    if (g->workers.empty()) {
        for (uint32_t r = 0; r < n; r++) {
            std::string err;
            const int rc = rank_phase(g, r, phase, &err);
            if (rc != VR_OK) return gfail(g, rc, err);
        }
        return VR_OK;
    }
    g->done.store(0, std::memory_order_relaxed);
    {
        std::lock_guard<std::mutex> lk(g->mu);
        g->phase = phase;
        g->ticket.fetch_add(1, std::memory_order_release);
    }
    g->cv.notify_all();
    while (g->done.load(std::memory_order_acquire) < n) std::this_thread::yield();
    for (uint32_t r = 0; r < n; r++)
        if (g->wrc[r] != VR_OK) return gfail(g, g->wrc[r], g->werr[r]);
    return VR_OK;
}

void stop_workers(vr_group *g)
{
    if (g->workers.empty()) return;
    {
        std::lock_guard<std::mutex> lk(g->mu);
        g->phase = -1;
        g->ticket.fetch_add(1, std::memory_order_release);
    }
    g->cv.notify_all();
    for (auto &t : g->workers) t.join();
    g->workers.clear();
}
}  // namespace

extern "C" int vr_group_destroy(vr_group *g)
{
    if (!g) return VR_OK;
    stop_workers(g);
    for (uint32_t r = 0; r < g->ctx.size(); r++) {
        cudaSetDevice(g->dev[r]);
        cudaDeviceSynchronize();
    }
    for (uint32_t r = 0; r < g->ctx.size(); r++) {
        cudaSetDevice(g->dev[r]);
        for (auto &e : g->ready.size() > r ? g->ready[r] : std::vector<cudaEvent_t>()) cudaEventDestroy(e);
        for (auto &e : g->merged.size() > r ? g->merged[r] : std::vector<cudaEvent_t>()) cudaEventDestroy(e);
        if (r < g->march.size() && g->march[r]) cudaStreamDestroy(g->march[r]);
        if (r < g->merge.size() && g->merge[r]) cudaStreamDestroy(g->merge[r]);
    }
    if (!g->dev.empty()) cudaSetDevice(g->dev[0]);
    for (auto &e : g->consumed) cudaEventDestroy(e);
    for (auto &e : g->all_ready) cudaEventDestroy(e);
    if (g->present) cudaStreamDestroy(g->present);
    if (g->join) cudaStreamDestroy(g->join);
    for (auto *c : g->ctx) vr_destroy(c);
    delete g;
    return VR_OK;
}

extern "C" int vr_create_group(uint32_t ndev, const int *devs, int partition, uint32_t width, uint32_t height, uint32_t slots,
                               vr_group **out)
{
    if (!out || !devs || ndev == 0 || ndev > 31 || width == 0 || height == 0 || partition < VR_PARTITION_REPLICAS ||
        partition > VR_PARTITION_SORT_LAST)
        return gfail(nullptr, VR_ERR_INVALID, "vr_create_group: bad arguments");
    *out = nullptr;
    vr_group *g = new (std::nothrow) vr_group;
    if (!g) return gfail(nullptr, VR_ERR_NOMEM, "out of host memory");
    g->partition = partition, g->n = ndev, g->W = width, g->H = height;
    g->slots = std::min<uint32_t>(std::max<uint32_t>(slots ? slots : 2, 1), VR_MAX_FRAMES_IN_FLIGHT);
    g->dev.assign(devs, devs + ndev);
    auto bail = [&](int rc, const std::string &msg) {
        g_group_error = msg;
        vr_group_destroy(g);
        return rc;
    };
    for (uint32_t r = 0; r < ndev; r++) {
        vr_ctx *c = nullptr;
        const int rc = vr_create(devs[r], width, height, &c);
        if (rc != VR_OK) return bail(rc, std::string("vr_create_group: rank ") + std::to_string(r) + ": " + vr_last_error(nullptr));
        g->ctx.push_back(c);
    }
    /* every device reaches every other's memory (one address space: plain pointers, no IPC handles) */
    for (uint32_t a = 0; a < ndev; a++)
        for (uint32_t b = 0; b < ndev; b++) {
            if (devs[a] == devs[b]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devs[a], devs[b]);
            if (!can && partition != VR_PARTITION_REPLICAS) return bail(VR_ERR_CUDA, "vr_create_group: the devices cannot access each other's memory");
            cudaSetDevice(devs[a]);
            const cudaError_t e = cudaDeviceEnablePeerAccess(devs[b], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled && partition != VR_PARTITION_REPLICAS)
                return bail(VR_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
            cudaGetLastError();
        }
    g->march.assign(ndev, nullptr), g->merge.assign(ndev, nullptr);
    g->ready.assign(ndev, std::vector<cudaEvent_t>(g->slots, nullptr));
    g->merged.assign(ndev, std::vector<cudaEvent_t>(g->slots, nullptr));
    for (uint32_t r = 0; r < ndev; r++) {
        if (cudaSetDevice(devs[r]) != cudaSuccess || cudaStreamCreateWithFlags(&g->march[r], cudaStreamNonBlocking) != cudaSuccess ||
            cudaStreamCreateWithFlags(&g->merge[r], cudaStreamNonBlocking) != cudaSuccess)
            return bail(VR_ERR_CUDA, "vr_create_group: cudaStreamCreate failed");
        for (uint32_t s = 0; s < g->slots; s++)
            if (cudaEventCreateWithFlags(&g->ready[r][s], cudaEventDisableTiming) != cudaSuccess ||
                cudaEventCreateWithFlags(&g->merged[r][s], cudaEventDisableTiming) != cudaSuccess)
                return bail(VR_ERR_CUDA, "vr_create_group: cudaEventCreate failed");
    }
    cudaSetDevice(devs[0]);
    if (cudaStreamCreateWithFlags(&g->present, cudaStreamNonBlocking) != cudaSuccess) return bail(VR_ERR_CUDA, "vr_create_group: cudaStreamCreate failed");
    g->consumed.assign(g->slots, nullptr);
    for (uint32_t s = 0; s < g->slots; s++)
        if (cudaEventCreateWithFlags(&g->consumed[s], cudaEventDisableTiming) != cudaSuccess) return bail(VR_ERR_CUDA, "vr_create_group: cudaEventCreate failed");
    if (cudaStreamCreateWithFlags(&g->join, cudaStreamNonBlocking) != cudaSuccess) return bail(VR_ERR_CUDA, "vr_create_group: cudaStreamCreate failed");
    g->all_ready.assign(g->slots, nullptr);
    for (uint32_t s = 0; s < g->slots; s++)
        if (cudaEventCreateWithFlags(&g->all_ready[s], cudaEventDisableTiming) != cudaSuccess) return bail(VR_ERR_CUDA, "vr_create_group: cudaEventCreate failed");
    if (partition == VR_PARTITION_SORT_FIRST) {
        const int rc = wire(g);
        if (rc != VR_OK) return bail(rc, g->err);
    }
    g->wrc.assign(ndev, VR_OK), g->werr.assign(ndev, std::string());
    if (ndev > 2 && partition != VR_PARTITION_REPLICAS) {
        try {
            for (uint32_t r = 0; r < ndev; r++) g->workers.emplace_back(worker_main, g, r);
        } catch (...) {
            return bail(VR_ERR_NOMEM, "vr_create_group: cannot start the worker threads");
        }
    }
    *out = g;
    return VR_OK;
}

/* ---- mutators: the reference's members, applied to every rank ------------------------------------------------- */

extern "C" int vr_group_set_camera(vr_group *g, const float view[16], const float proj[16], const float cam[3])
{
    if (!g || !cam) return gfail(g, VR_ERR_INVALID, "vr_group_set_camera: null argument");
    for (uint32_t r = 0; r < g->n; r++) G_VR(g, r, vr_set_camera(g->ctx[r], view, proj, cam));
    std::memcpy(g->cam_pos, cam, sizeof g->cam_pos);
    return VR_OK;
}

extern "C" int vr_group_set_transfer_function(vr_group *g, const uint32_t *rgba8, uint32_t n)
{
    if (!g) return VR_ERR_INVALID;
    for (uint32_t r = 0; r < g->n; r++) G_VR(g, r, vr_set_transfer_function(g->ctx[r], rgba8, n));
    return VR_OK;
}

extern "C" int vr_group_set_slicing(vr_group *g, const float mn[3], const float mx[3])
{
    if (!g) return VR_ERR_INVALID;
    for (uint32_t r = 0; r < g->n; r++) G_VR(g, r, vr_set_slicing(g->ctx[r], mn, mx));
    return VR_OK;
}

extern "C" int vr_group_set_options(vr_group *g, const vr_options *o)
{
    if (!g || !o) return VR_ERR_INVALID;
    vr_options oo = *o;
    if (g->partition == VR_PARTITION_SORT_LAST) oo.write_raw = 1; /* the partial IS the float4 output */
    for (uint32_t r = 0; r < g->n; r++) G_VR(g, r, vr_set_options(g->ctx[r], &oo));
    return VR_OK;
}

extern "C" int vr_group_set_balance(vr_group *g, float fraction)
{
    if (!g || !(fraction >= 0.0f) || fraction > 0.25f) return gfail(g, VR_ERR_INVALID, "vr_group_set_balance: need 0 <= fraction <= 0.25");
    g->balance = fraction; /* takes effect with the next vr_group_set_volume / vr_group_synth_volume */
    return VR_OK;
}

extern "C" int vr_group_set_volume(vr_group *g, const void *data, int dtype, uint32_t nx, uint32_t ny, uint32_t nz, const float *mn,
                                   const float *mx)
{
    if (!g || !data || dtype < VR_U8 || dtype > VR_F32 || !nx || !ny || !nz) return gfail(g, VR_ERR_INVALID, "vr_group_set_volume: bad arguments");
    if (g->partition != VR_PARTITION_SORT_LAST) {
        for (uint32_t r = 0; r < g->n; r++) G_VR(g, r, vr_set_volume(g->ctx[r], data, dtype, nx, ny, nz, mn, mx)); /* replicated */
        return VR_OK;
    }
    int rc = make_partition(g, nx, ny, nz);
    if (rc != VR_OK) return rc;
    const size_t esz = dtype == VR_U8 ? 1 : (dtype == VR_U16 ? 2 : 4);
    std::vector<unsigned char> sub;
    for (uint32_t r = 0; r < g->n; r++) {
        /* what the rank holds (its sub-box + the balance margin) plus one ghost voxel towards each neighbour, cut out of the host array */
        uint32_t lo[3], l[3];
        for (int i = 0; i < 3; i++) {
            const uint32_t org = g->stored[r][i], dims = g->stored[r][3 + i], end = org + dims;
            lo[i] = org - (org > 0 ? 1 : 0);
            l[i] = dims + (org > 0 ? 1 : 0) + (end < g->full[i] ? 1 : 0);
        }
        sub.resize((size_t)l[0] * l[1] * l[2] * esz);
        for (uint32_t z = 0; z < l[2]; z++)
            for (uint32_t y = 0; y < l[1]; y++)
                std::memcpy(sub.data() + ((size_t)z * l[1] + y) * l[0] * esz,
                            static_cast<const unsigned char *>(data) + (((size_t)(lo[2] + z) * ny + (lo[1] + y)) * nx + lo[0]) * esz, (size_t)l[0] * esz);
        G_VR(g, r, vr_set_volume(g->ctx[r], sub.data(), dtype, l[0], l[1], l[2], mn, mx));
    }
    if (!mn || !mx) {
        rc = reduce_density_range(g);
        if (rc != VR_OK) return rc;
    }
    return wire(g);
}

extern "C" int vr_group_synth_volume(vr_group *g, int kind, uint32_t seed, int dtype, uint32_t nx, uint32_t ny, uint32_t nz)
{
    if (!g) return VR_ERR_INVALID;
    if (g->partition != VR_PARTITION_SORT_LAST) {
        for (uint32_t r = 0; r < g->n; r++) G_VR(g, r, vr_synth_volume(g->ctx[r], kind, seed, dtype, nx, ny, nz));
        return VR_OK;
    }
    int rc = make_partition(g, nx, ny, nz);
    if (rc != VR_OK) return rc;
    for (uint32_t r = 0; r < g->n; r++) G_VR(g, r, vr_synth_subvolume(g->ctx[r], kind, seed, dtype));
    rc = reduce_density_range(g);
    return rc != VR_OK ? rc : wire(g);
}

/* ---- one frame on all ranks ----------------------------------------------------------------------------------- */

extern "C" int vr_group_render(vr_group *g)
{
    if (!g) return VR_ERR_INVALID;
    const uint64_t k = g->frame;
    const uint32_t S = g->slots, s = (uint32_t)(k % S), n = g->n;
    if (g->partition == VR_PARTITION_REPLICAS) {
        for (uint32_t r = 0; r < n; r++) G_VR(g, r, vr_render_frame(g->ctx[r], s, g->march[r]));
        g->frame++;
        return VR_OK;
    }
    g->job.k = k, g->job.s = s;
    if (g->partition == VR_PARTITION_SORT_FIRST) {
        /* every rank: slot free -> K1 (tiles straight into the display rank's slot) -> ready */
        const int rc = run_phase(g, 1);
        if (rc != VR_OK) return rc;
        G_CUDA(g, cudaSetDevice(g->dev[0]));
        for (uint32_t r = 0; r < n; r++) G_CUDA(g, cudaStreamWaitEvent(g->present, g->ready[r][s], 0));
        G_CUDA(g, cudaEventRecord(g->consumed[s], g->present)); /* the frame is complete (a consumer would read it here) */
        g->frame++;
        return VR_OK;
    }
    /* sort-last */
    if (g->raw_slot.empty()) return gfail(g, VR_ERR_STATE, "vr_group_render: no volume (vr_group_set_volume / vr_group_synth_volume)");
    std::vector<uint32_t> order(n), planes(g->grid[0] + g->grid[1] + g->grid[2]);
    const bool balanced = g->margin[0] || g->margin[1] || g->margin[2];
    if (balanced) {
        /* this frame's split planes (a pure function of the camera) and every rank's box between them */
        if (vr_partition_balance(g->grid, g->full, g->cam_pos, g->margin, planes.data()) != VR_OK) return gfail(g, VR_ERR_INVALID, "vr_partition_balance failed");
        for (uint32_t r = 0; r < n; r++) {
            if (vr_partition_box_planes(g->grid, g->full, planes.data(), r, &g->box[r][0], &g->box[r][3]) != VR_OK)
                return gfail(g, VR_ERR_INVALID, "vr_partition_box_planes failed");
            G_VR(g, r, vr_set_owned_box(g->ctx[r], &g->box[r][0], &g->box[r][3]));
        }
    }
    if (vr_partition_order_planes(g->grid, g->full, balanced ? planes.data() : nullptr, g->cam_pos, order.data()) != VR_OK)
        return gfail(g, VR_ERR_INVALID, "vr_partition_order failed");
    std::vector<uint32_t> rects(4 * n);
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t b = order[i];
        if (vr_partition_screen_rect(g->ctx[0], g->full, &g->box[b][0], &g->box[b][3], &rects[4 * i]) != VR_OK)
            return gfail(g, VR_ERR_INVALID, "vr_partition_screen_rect failed");
    }
    /* phase 1, every rank: partial slot free -> K1 -> ready */
    int rc = run_phase(g, 1);
    if (rc != VR_OK) return rc;
    /* the ranks' ready events joined into one (a rank then waits on one event instead of n) */
    G_CUDA(g, cudaSetDevice(g->dev[0]));
    for (uint32_t r = 0; r < n; r++) G_CUDA(g, cudaStreamWaitEvent(g->join, g->ready[r][s], 0));
    G_CUDA(g, cudaEventRecord(g->all_ready[s], g->join));
    /* only the region that can differ from the clear colour is merged, in bands of 8 rows dealt to the ranks */
    if (g->slot_region.size() != S) g->slot_region.assign(S, {});
    vr_sort_last_region(rects.data(), n, g->W, g->H, k < S, g->slot_region[s].data(), g->job.region);
    std::vector<const void *> parts(n);
    for (uint32_t i = 0; i < n; i++) parts[i] = g->raw_slot[order[i]][s];
    g->job.parts = parts.data(), g->job.rects = rects.data();
    /* phase 2, every rank: all partials ready, image slot presented -> K2 -> merged */
    rc = run_phase(g, 2);
    g->job.parts = nullptr, g->job.rects = nullptr;
    if (rc != VR_OK) return rc;
    G_CUDA(g, cudaSetDevice(g->dev[0]));
    for (uint32_t r = 0; r < n; r++) G_CUDA(g, cudaStreamWaitEvent(g->present, g->merged[r][s], 0));
    G_CUDA(g, cudaEventRecord(g->consumed[s], g->present));
    g->frame++;
    return VR_OK;
}

extern "C" int vr_group_sync(vr_group *g)
{
    if (!g) return VR_ERR_INVALID;
    for (uint32_t r = 0; r < g->n; r++) G_VR(g, r, vr_sync(g->ctx[r]));
    return VR_OK;
}

/* The frame recorded last, complete, from the display rank's slot (replicas: rank's own frame). */
extern "C" int vr_group_download(vr_group *g, uint32_t rank, void *host_rgba8)
{
    if (!g || !host_rgba8 || rank >= g->n) return gfail(g, VR_ERR_INVALID, "vr_group_download: bad arguments");
    if (g->frame == 0) return gfail(g, VR_ERR_STATE, "vr_group_download: nothing rendered yet");
    const int rc = vr_group_sync(g);
    if (rc != VR_OK) return rc;
    const uint32_t s = (uint32_t)((g->frame - 1) % g->slots), r = g->partition == VR_PARTITION_REPLICAS ? rank : 0;
    G_VR(g, r, vr_download_frame_async(g->ctx[r], s, host_rgba8, nullptr));
    G_VR(g, r, vr_sync(g->ctx[r]));
    return VR_OK;
}
