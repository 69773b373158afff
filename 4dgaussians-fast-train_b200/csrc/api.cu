// Library info entry points of the fdgs C ABI (include/fdgs.h).
#include "fdgs_common.cuh"

extern "C" {

int fdgs_abi_version(void) { return FDGS_ABI_VERSION; }

int fdgs_compiled_arch(void) {
#ifdef FDGS_ARCH
    return FDGS_ARCH;
#else
    return 100;
#endif
}

const char* fdgs_error_string(int code) {
    switch (code) {
        case FDGS_OK: return "ok";
        case FDGS_E_BADARG: return "fdgs: bad argument (null pointer, inconsistent shapes or exclusive inputs)";
        case FDGS_E_WORKSPACE: return "fdgs: caller-owned workspace too small";
        case FDGS_E_UNSUPPORTED: return "fdgs: configuration not supported by the sm_100a kernels";
        case FDGS_E_SMEM: return "fdgs: configuration exceeds the shared-memory budget";
        case FDGS_E_TOO_LARGE: return "fdgs: problem size exceeds the fixed maximum of this allocation-free entry point";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "fdgs: unknown error";
}

}  // extern "C"

// ---- per-kernel profiling (include/fdgs.h: fdgs_profile_*) -----------------------------------------
#include "prof.cuh"
#include <map>
#include <string>
#include <vector>
#include <cstdio>
#include <cstring>

namespace {
struct Rec { std::string name; cudaEvent_t a, b; };
bool g_prof_on = false;
std::vector<Rec> g_recs;
std::vector<cudaEvent_t> g_pool;
unsigned long long g_launches = 0;
cudaEvent_t get_event() {
    if (!g_pool.empty()) { cudaEvent_t e = g_pool.back(); g_pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}
}  // namespace

static int g_sm_limit = 0;
int fdgs_sm_limit_value() { return g_sm_limit; }
void fdgs_count_launch() { ++g_launches; }
void fdgs_prof_begin(const char* name, cudaStream_t s) {
    ++g_launches;
    if (!g_prof_on) return;
    Rec r; r.name = name; r.a = get_event(); r.b = get_event();
    cudaEventRecord(r.a, s);
    g_recs.push_back(r);
}
void fdgs_prof_end(cudaStream_t s) {
    if (!g_prof_on || g_recs.empty()) return;
    cudaEventRecord(g_recs.back().b, s);
}

extern "C" {
void fdgs_set_sm_limit(int n) { g_sm_limit = n > 0 ? n : 0; }
void fdgs_profile_enable(int on) { g_prof_on = on != 0; }
unsigned long long fdgs_launch_count(void) { return g_launches; }
// Synchronises the device, then writes "name count total_ms\n" per kernel into buf; clears the records.
int fdgs_profile_collect(char* buf, size_t n) {
    cudaDeviceSynchronize();
    std::map<std::string, std::pair<int, double>> agg;
    for (auto& r : g_recs) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) { agg[r.name].first++; agg[r.name].second += ms; }
        g_pool.push_back(r.a); g_pool.push_back(r.b);
    }
    g_recs.clear();
    std::string out;
    char line[256];
    for (auto& kv : agg) { snprintf(line, sizeof line, "%s %d %.6f\n", kv.first.c_str(), kv.second.first, kv.second.second); out += line; }
    if (buf && n) { strncpy(buf, out.c_str(), n - 1); buf[n - 1] = 0; }
    return (int)agg.size();
}
}
