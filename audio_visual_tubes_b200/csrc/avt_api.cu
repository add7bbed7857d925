// Library-level entry points: version, per-thread error text, device check.
#include <stdarg.h>
#include <atomic>
#include "avt_common.cuh"

namespace avt {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---- launch counter + stage timing ---------------------------------------------------
// Debug / measurement aid (bench.py): process-global, not meant for concurrent callers.
static std::atomic<long long> g_launches{0};
static std::atomic<int> g_prof_on{0};
constexpr int kProfMaxPairs = 2048;
struct ProfState {
    cudaEvent_t ev[kProfMaxPairs][2];
    int created = 0;     // event pairs created so far
    int used = 0;        // pairs recorded since the last reset
    bool open = false;
};
static ProfState g_prof[kProfSlots];

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
bool prof_enabled() { return g_prof_on.load(std::memory_order_relaxed) != 0; }

void prof_begin(int slot, cudaStream_t st) {
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    ProfState& p = g_prof[slot];
    if (p.used >= kProfMaxPairs) return;
    if (p.used >= p.created) {
        if (cudaEventCreate(&p.ev[p.created][0]) != cudaSuccess ||
            cudaEventCreate(&p.ev[p.created][1]) != cudaSuccess) { cudaGetLastError(); return; }
        ++p.created;
    }
    cudaEventRecord(p.ev[p.used][0], st);
    p.open = true;
}
void prof_end(int slot, cudaStream_t st) {
    ProfState& p = g_prof[slot];
    if (!p.open) return;
    cudaEventRecord(p.ev[p.used][1], st);
    ++p.used;
    p.open = false;
}

}  // namespace avt

extern "C" long long avt_launch_count(void) { return avt::g_launches.load(); }

extern "C" int avt_profile_enable(int on) {
    for (int s = 0; s < avt::kProfSlots; ++s) { avt::g_prof[s].used = 0; avt::g_prof[s].open = false; }
    avt::g_prof_on.store(on ? 1 : 0);
    return AVT_OK;
}

extern "C" int avt_profile_read(int slot, double* total_ms, int* intervals) {
    AVT_REQUIRE(slot >= 0 && slot < avt::kProfSlots && total_ms && intervals,
                "avt_profile_read: bad slot %d or null pointer", slot);
    avt::ProfState& p = avt::g_prof[slot];
    double t = 0.0;
    for (int i = 0; i < p.used; ++i) {
        AVT_CHECK_CUDA(cudaEventSynchronize(p.ev[i][1]));
        float ms = 0.f;
        AVT_CHECK_CUDA(cudaEventElapsedTime(&ms, p.ev[i][0], p.ev[i][1]));
        t += ms;
    }
    *total_ms = t;
    *intervals = p.used;
    return AVT_OK;
}

extern "C" int avt_version(void) { return 100; }

extern "C" const char* avt_last_error(void) { return avt::g_err; }

extern "C" int avt_device_ok(void) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return 0; }
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return major == 10 ? 1 : 0;
}
