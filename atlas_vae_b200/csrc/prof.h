// Optional per-kernel CUDA-event timing (bench.py roofline leg): when enabled every launch made through
// PROF(name, stream) is bracketed by two events on the launching stream; oevae_profile_report() aggregates them.
#pragma once
#include <cuda_runtime.h>
#include <map>
#include <string>
#include <vector>

namespace oevae {

struct Profiler {
    bool on = false;
    struct Rec { const char* name; cudaEvent_t a, b; };
    std::vector<Rec> recs;
    std::vector<cudaEvent_t> pool;
    cudaEvent_t get() {
        if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
        cudaEvent_t e; cudaEventCreate(&e); return e;
    }
    void clear() { for (auto& r : recs) { pool.push_back(r.a); pool.push_back(r.b); } recs.clear(); }
};
extern Profiler g_prof;

struct ProfScope {
    cudaStream_t s; int idx = -1;
    ProfScope(const char* name, cudaStream_t st) : s(st) {
        if (!g_prof.on) return;
        Profiler::Rec r{name, g_prof.get(), g_prof.get()};
        cudaEventRecord(r.a, s);
        g_prof.recs.push_back(r); idx = (int)g_prof.recs.size() - 1;
    }
    ~ProfScope() { if (idx >= 0) cudaEventRecord(g_prof.recs[idx].b, s); }
};
#define PROF(name, stream) ProfScope _prof_scope(name, stream)

}  // namespace oevae
