// ABI bookkeeping for libweclip_sm100.so: version, error text, device check, opt-in kernel-class profiler.
#include "common.cuh"

#include <vector>

namespace wc {
thread_local char g_err[512] = "";
bool g_prof_on = false;

struct ProfRec { int kind; double work; cudaEvent_t e0, e1; int tag[4]; };
static std::vector<ProfRec> g_recs;
static std::vector<cudaEvent_t> g_pool;
static size_t g_pool_used = 0;
static bool g_open = false;

static cudaEvent_t pool_event() {
    if (g_pool_used == g_pool.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
        g_pool.push_back(e);
    }
    return g_pool[g_pool_used++];
}

void prof_begin(int kind, double work, cudaStream_t st) {
    if (g_open || g_recs.size() >= (1u << 20)) return;
    ProfRec r{kind, work, pool_event(), pool_event(), {0, 0, 0, 0}};
    if (!r.e0 || !r.e1) return;
    cudaEventRecord(r.e0, st);
    g_recs.push_back(r);
    g_open = true;
}
void prof_tag(int a, int b, int c, int d) {
    if (!g_open) return;
    int* t = g_recs.back().tag;
    t[0] = a; t[1] = b; t[2] = c; t[3] = d;
}
void prof_end(cudaStream_t st) {
    if (!g_open) return;
    cudaEventRecord(g_recs.back().e1, st);
    g_open = false;
}
}  // namespace wc

using namespace wc;

extern "C" int wc_abi_version(void) { return WC_ABI_VERSION; }
extern "C" const char* wc_last_error(void) { return wc::g_err; }
extern "C" int wc_device_ok(void) {
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
    return major == 10;
}

extern "C" int wc_prof_kinds(void) { return PK_COUNT; }
extern "C" const char* wc_prof_kind_name(int kind) {
    static const char* names[PK_COUNT] = {"gemm_tc", "layernorm", "softmax_headmean", "par_affinity", "par_iter",
                                          "affinity_refine", "gradcam_misc", "encoder_misc", "cam_upsample", "eval_fuse",
                                          "attention"};
    return (kind >= 0 && kind < PK_COUNT) ? names[kind] : "";
}
extern "C" int wc_prof_enable(int on) {
    g_prof_on = on != 0;
    return WC_OK;
}
extern "C" int wc_prof_reset(void) {
    g_recs.clear();
    g_pool_used = 0;
    g_open = false;
    return WC_OK;
}
// Synchronises the device, then sums per kind: ms[k] device milliseconds, work[k] algorithmic flops or bytes,
// launches[k] timed launches.  Arrays hold wc_prof_kinds() entries.
extern "C" int wc_prof_summary(double* ms, double* work, long* launches) {
    WC_CHECK_ARG(ms && work && launches);
    WC_CUDA(cudaDeviceSynchronize());
    for (int k = 0; k < PK_COUNT; ++k) { ms[k] = 0; work[k] = 0; launches[k] = 0; }
    for (const ProfRec& r : g_recs) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, r.e0, r.e1) != cudaSuccess) continue;
        ms[r.kind] += t; work[r.kind] += r.work; launches[r.kind] += 1;
    }
    return WC_OK;
}

// Raw records (after wc_prof_summary has synchronised): up to cap entries of {kind, ms, work, tag[4]}; returns the count.
extern "C" long wc_prof_records(int* kind, double* ms, double* work, int* tags, long cap) {
    long n = 0;
    for (const ProfRec& r : g_recs) {
        if (n >= cap) break;
        float t = 0.f;
        if (cudaEventElapsedTime(&t, r.e0, r.e1) != cudaSuccess) continue;
        kind[n] = r.kind; ms[n] = t; work[n] = r.work;
        for (int i = 0; i < 4; ++i) tags[4 * n + i] = r.tag[i];
        ++n;
    }
    return n;
}
