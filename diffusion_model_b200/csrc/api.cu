// Error reporting, version, launch counter and the optional per-kernel CUDA-event profiler of the C ABI.
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

static thread_local char g_err[512] = "";

void egnn_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* egnn_b200_last_error(void) { return g_err; }
extern "C" int egnn_b200_version(void) { return EGNN_B200_ABI_VERSION; }

// ---- launch counter: every kernel launch of the library goes through EGNN_CHECK_LAUNCH -------------------------
static long long g_launches = 0;
void egnn_count_launch() { ++g_launches; }
extern "C" int64_t egnn_launch_count(void) { return g_launches; }

// ---- profiler: when enabled, selected kernel families are bracketed by CUDA events on their own stream --------
namespace {
constexpr int kMaxEv = 8192;
bool g_prof = false;
cudaEvent_t g_ev[2 * kMaxEv];
int g_kind[kMaxEv];
int g_nev = 0, g_created = 0;
}  // namespace

bool egnn_prof_on() { return g_prof; }
void egnn_prof_begin(int kind, cudaStream_t st) {
    if (!g_prof || g_nev >= kMaxEv) return;
    while (g_created <= g_nev) {
        cudaEventCreate(&g_ev[2 * g_created]);
        cudaEventCreate(&g_ev[2 * g_created + 1]);
        ++g_created;
    }
    g_kind[g_nev] = kind;
    cudaEventRecord(g_ev[2 * g_nev], st);
}
void egnn_prof_end(cudaStream_t st) {
    if (!g_prof || g_nev >= kMaxEv) return;
    cudaEventRecord(g_ev[2 * g_nev + 1], st);
    ++g_nev;
}
extern "C" int egnn_profile_enable(int32_t on) {
    g_prof = on != 0;
    g_nev = 0;
    return EGNN_OK;
}
extern "C" int egnn_profile_read(int32_t kind, double* total_ms_host, int64_t* launches_host) {
    EGNN_REQUIRE(total_ms_host && launches_host, "egnn_profile_read: NULL pointer");
    double tot = 0.0;
    long long n = 0;
    for (int k = 0; k < g_nev; ++k) {
        if (g_kind[k] != kind) continue;
        EGNN_CUDA(cudaEventSynchronize(g_ev[2 * k + 1]));
        float ms = 0.f;
        EGNN_CUDA(cudaEventElapsedTime(&ms, g_ev[2 * k], g_ev[2 * k + 1]));
        tot += ms;
        ++n;
    }
    *total_ms_host = tot;
    *launches_host = n;
    return EGNN_OK;
}
