// Error state, version, launch counter and the optional in-stream kernel timer of the C ABI.
#include "common.cuh"
#include <atomic>
#include <string.h>

namespace hg {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char* get_error() { return g_err; }

static std::atomic<long long> g_launches{0};
void count_launches(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// CUDA-event pairs recorded around the fused tensor-core kernel on its own stream, so a
// benchmark can report that kernel's duration without a profiler attached.
constexpr int kProfSlots = 4096;
static std::atomic<bool> g_prof_on{false};      // read by every enqueueing host thread
static cudaEvent_t g_prof_ev[2 * kProfSlots];
static bool g_prof_made = false;
static std::atomic<int> g_prof_used{0};
static thread_local int t_prof_slot = -1;          // builds may be enqueued from several host threads

void prof_begin(cudaStream_t st) {
    t_prof_slot = -1;
    if (!g_prof_on.load(std::memory_order_acquire)) return;
    const int slot = g_prof_used.fetch_add(1);
    if (slot >= kProfSlots) return;
    t_prof_slot = slot;
    cudaEventRecord(g_prof_ev[2 * slot], st);
}
void prof_end(cudaStream_t st) {
    if (t_prof_slot < 0) return;
    cudaEventRecord(g_prof_ev[2 * t_prof_slot + 1], st);
    t_prof_slot = -1;
}
}  // namespace hg

extern "C" int hg_version(void) { return 100; }   // 0.1.0
extern "C" const char* hg_last_error(void) { return hg::get_error(); }
extern "C" int64_t hg_launch_count(void) { return hg::g_launches.load(); }

extern "C" int hg_profile_enable(int on) {
    if (on && !hg::g_prof_made) {
        for (int i = 0; i < 2 * hg::kProfSlots; ++i) HG_CHECK_CUDA(cudaEventCreate(&hg::g_prof_ev[i]));
        hg::g_prof_made = true;
    }
    hg::g_prof_used.store(0);
    hg::g_prof_on.store(on != 0, std::memory_order_release);
    return HG_OK;
}

// Sum of the recorded durations (ms) of the fused similarity+top-k kernel since the last
// enable/collect, and how many launches that covers.  Synchronises on the recorded events.
extern "C" int hg_profile_collect(float* ms_total, int* launches) {
    HG_REQUIRE(ms_total && launches, HG_ERR_BAD_ARG, "hg_profile_collect: null");
    float sum = 0.f;
    const int used = hg::g_prof_used.load() < hg::kProfSlots ? hg::g_prof_used.load() : hg::kProfSlots;
    for (int i = 0; i < used; ++i) {
        float ms = 0.f;
        HG_CHECK_CUDA(cudaEventSynchronize(hg::g_prof_ev[2 * i + 1]));
        HG_CHECK_CUDA(cudaEventElapsedTime(&ms, hg::g_prof_ev[2 * i], hg::g_prof_ev[2 * i + 1]));
        sum += ms;
    }
    *ms_total = sum;
    *launches = used;
    hg::g_prof_used.store(0);
    return HG_OK;
}

// Maps a peer process's allocation (the 64-byte cudaIpcMemHandle_t it exported) into THIS process for the CURRENT
// device, with peer access enabled lazily: the returned base is what kernels on this device may load / store.
// (torch rebuilds shared tensors with the EXPORTING device current -- fine for copies, not for kernels running on
// another device, hence this pair.)  One open per exported allocation and process; hg_ipc_close unmaps.
extern "C" int hg_ipc_open(const void* handle64, void** base_out) {
    HG_REQUIRE(handle64 && base_out, HG_ERR_BAD_ARG, "hg_ipc_open: null");
    cudaIpcMemHandle_t h;
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(&h, handle64, sizeof(h));
    const cudaError_t e = cudaIpcOpenMemHandle(base_out, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaGetLastError();                          // do not leave the error for the next launch check to find
        hg::set_error("hg_ipc_open: cudaIpcOpenMemHandle -> %s", cudaGetErrorString(e));
        return HG_ERR_CUDA;
    }
    return HG_OK;
}
// A device buffer that peers can map: plain cudaMalloc (torch's caching allocator may hand out expandable-segment
// memory, whose handles cudaIpcOpenMemHandle does not take) + its IPC handle.  The one place where the library
// allocates: a long-lived exchange buffer the caller asks for explicitly and frees with hg_peer_free.
extern "C" int hg_peer_alloc(size_t bytes, void** ptr_out, void* handle64_out) {
    HG_REQUIRE(bytes > 0 && ptr_out && handle64_out, HG_ERR_BAD_ARG, "hg_peer_alloc: bad argument");
    void* p = nullptr;
    HG_CHECK_CUDA(cudaMalloc(&p, bytes));
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        cudaFree(p);
        hg::set_error("hg_peer_alloc: cudaIpcGetMemHandle -> %s", cudaGetErrorString(e));
        return HG_ERR_CUDA;
    }
    memcpy(handle64_out, &h, sizeof(h));
    *ptr_out = p;
    return HG_OK;
}
extern "C" int hg_peer_free(void* ptr) {
    HG_REQUIRE(ptr, HG_ERR_BAD_ARG, "hg_peer_free: null");
    HG_CHECK_CUDA(cudaFree(ptr));
    return HG_OK;
}
extern "C" int hg_ipc_close(void* base) {
    HG_REQUIRE(base, HG_ERR_BAD_ARG, "hg_ipc_close: null");
    HG_CHECK_CUDA(cudaIpcCloseMemHandle(base));
    return HG_OK;
}

// Lets kernels running on the CURRENT device dereference memory of `peer_device` that this process has mapped (its own
// allocations there, or CUDA-IPC mappings opened in that device's context): the bucket pushes of the partitioned
// symmetric pass (hg_knn_sympart_sweep) store straight into the owning GPU's arrays over NVLink.
extern "C" int hg_enable_peer_access(int peer_device) {
    int dev = 0;
    HG_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev == peer_device) return HG_OK;
    int can = 0;
    HG_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer_device));
    HG_REQUIRE(can != 0, HG_ERR_UNSUPPORTED, "hg_enable_peer_access: device %d cannot access device %d", dev, peer_device);
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return HG_OK; }
    HG_CHECK_CUDA(e);
    return HG_OK;
}
