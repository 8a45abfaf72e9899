// prof.cu — per-kernel-family device timing behind nb200_profile_{enable,report}. Events are recorded on the stream
// the kernel is launched on; durations are resolved when the report is requested (after a stream synchronise).
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include "common.cuh"

namespace nb {

thread_local bool g_prof_on = false;
bool g_pdl = getenv("NB200_NO_PDL") == nullptr;

namespace {
thread_local std::string g_dup = getenv("NB200_DUP") ? getenv("NB200_DUP") : "";
}
bool dup_family(const char* family) { return !g_dup.empty() && g_dup == family; }
void set_dup_family(const char* family) { g_dup = family ? family : ""; }

namespace {
struct Rec {
    const char* family;
    double flops, bytes;
    cudaEvent_t a, b;
};
thread_local std::vector<Rec> g_recs;
thread_local std::vector<cudaEvent_t> g_pool;

cudaEvent_t get_event() {
    if (!g_pool.empty()) {
        cudaEvent_t e = g_pool.back();
        g_pool.pop_back();
        return e;
    }
    cudaEvent_t e;
    CUDA_CHECK(cudaEventCreate(&e));
    return e;
}
}  // namespace

void prof_begin(cudaStream_t st, const char* family, double flops, double bytes) {
    Rec r{family, flops, bytes, get_event(), get_event()};
    CUDA_CHECK(cudaEventRecord(r.a, st));
    g_recs.push_back(r);
}

void prof_end(cudaStream_t st) { CUDA_CHECK(cudaEventRecord(g_recs.back().b, st)); }

struct ProfAgg {
    long long launches = 0;
    double ms = 0, flops = 0, bytes = 0;
};

// JSON object: {"family": {"launches": n, "ms": total, "flops": total, "bytes": total}, ...}; clears the records.
std::string prof_report() {
    std::map<std::string, ProfAgg> agg;
    for (Rec& r : g_recs) {
        CUDA_CHECK(cudaEventSynchronize(r.b));
        float ms = 0;
        CUDA_CHECK(cudaEventElapsedTime(&ms, r.a, r.b));
        ProfAgg& a = agg[r.family];
        a.launches++;
        a.ms += ms;
        a.flops += r.flops;
        a.bytes += r.bytes;
        g_pool.push_back(r.a);
        g_pool.push_back(r.b);
    }
    g_recs.clear();
    std::string out = "{";
    bool first = true;
    for (auto& kv : agg) {
        char buf[256];
        snprintf(buf, sizeof buf, "%s\"%s\": {\"launches\": %lld, \"ms\": %.6f, \"flops\": %.6e, \"bytes\": %.6e}", first ? "" : ", ",
                 kv.first.c_str(), kv.second.launches, kv.second.ms, kv.second.flops, kv.second.bytes);
        out += buf;
        first = false;
    }
    return out + "}";
}

}  // namespace nb
