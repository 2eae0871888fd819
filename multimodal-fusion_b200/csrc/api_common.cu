// Error state, version, launch counter and the optional in-stream kernel timer of the C ABI.
#include "common.cuh"
#include <atomic>

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
