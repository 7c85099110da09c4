// covid-abm_b200/csrc/cabm.cu -- C ABI of include/cabm.h over the sm_100a day pipeline.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -shared -Xcompiler -fPIC
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>
#include <climits>
#include <dlfcn.h>
#include <nccl.h>
#include "cabm_kernels2.cuh"
#include "../../include/cabm_popfile.h"

namespace {

struct Buf { void** slot; size_t bytes; };

} // namespace

// ---- scratch that needs no allocation (cabm_serosurvey_static, include/cabm.h) -----------------------------------------
// A file-scope device array: every GPU that loads this library holds CABM_STATIC_SCRATCH_BYTES of it until the process exits.
// Layout: bytes [0, 256) the 64-bit accumulator of the running call (rest of the block unused), then up to
// CABM_STATIC_MAX_IDS person ids. In HOST code the name g_static_scratch is only a stand-in for the symbol, not a device
// address: the address is asked for once per device, with that device current, in cabm_create (cudaGetSymbolAddress) and
// handed to the kernel as an argument.
#define CABM_STATIC_HEAD_BYTES 256
#define CABM_STATIC_SCRATCH_BYTES (CABM_STATIC_HEAD_BYTES + (size_t) CABM_STATIC_MAX_IDS * sizeof(unsigned int))
static __device__ __align__(256) unsigned char g_static_scratch[CABM_STATIC_SCRATCH_BYTES];
static void* g_static_scratch_addr[64] = { nullptr };        // per device ordinal, filled by cabm_create

// k_serosurvey on the static scratch: ids (when use_ids) and the accumulator live in `scratch`, nothing else differs
// (doSerosurvey(NATURAL, pop, day), Community.cpp:1224-1274)
__global__ void k_serosurvey_static(const Dev D, const int day, unsigned char* scratch, const unsigned int n, const int use_ids) {
    unsigned long long* out = reinterpret_cast<unsigned long long*>(scratch);
    const unsigned int* __restrict__ ids = reinterpret_cast<const unsigned int*>(scratch + CABM_STATIC_HEAD_BYTES);
    unsigned int mine = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        mine += sero_positive(D, D.par, use_ids ? ids[i] : i, day) ? 1u : 0u;
    mine = warp_sum(mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(out, (unsigned long long) mine);
}

struct cabm {
    Dev d{};
    std::string err;
    cudaStream_t stream = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    double acc_ms = 0.0; uint64_t launches = 0;
    int device = 0, sms = 148;
    int gm[9] = { 8, 8, 8, 4, 8, 2, 2, 2, 4 };   // blocks per SM: active, hazard, trials, gate, apply+status, hosp+import, (unused), apply1, (unused) (CABM_GRID_MULT overrides: development knob)
    int pdl = 0;                                 // programmatic dependent launch, one bit per kernel of the day (bit k: kernel k may start before its predecessor
                                                 // has drained): CABM_PDL=<mask>. Measured at 2.7 M agents, graph replay: every edge on 151 us per day, none 141.
    int visit_bps = 2;                           // resident k_visit blocks per SM (from the occupancy calculator, given the hot bitmap's size)
    std::vector<float> risk_sorted;              // distinct household riskiness values, ascending: DayCtx::sd_rank = how many lie below the day's social distancing
    void* scratch = nullptr; size_t scratch_bytes = 0;   // handle-owned device scratch of the small entry points (grows, never shrinks)
    unsigned char* static_scratch = nullptr; std::string static_scratch_err;   // device address of g_static_scratch on this handle's device (cabm_create)
    bool have_par = false, have_pop = false;
    cabm_params hpar{};                       // host copy (pointers nulled)
    std::vector<double> h_ti[3];
    std::vector<void*> allocs;
    std::vector<std::pair<void*, size_t>> dyn;    // dynamic state buffers (snapshot set)
    DayCtx* h_ctx = nullptr; DayCtx* d_ctx = nullptr; int ctx_cap = 0;
    int* d_tmp_int = nullptr;
    cudaGraphExec_t day_graph = nullptr, tick_graph = nullptr, import_graph = nullptr, step_graph = nullptr;   // run_days (tick + importation), cabm_tick, cabm_import, tick + import of one day through the per-day calls
    cabm_params* d_par = nullptr; double* d_pfd = nullptr; double* d_imr = nullptr;
    unsigned long long h_n_log_cap = 0;
    uint64_t launches_per_day = 9, launches_tick = 9, launches_import = 2;
    unsigned int bitmap_words = 1; unsigned int* d_out_new = nullptr; DayCtx* d_today = nullptr; unsigned long long* d_diag = nullptr; unsigned int* d_det = nullptr;
    void* nccl_comm = nullptr; int world = 1; size_t xcap = 0, xcap_alloc = 0; unsigned int x_seen = UINT_MAX; FlagRaw* d_xrecv = nullptr; unsigned int* d_xcounts = nullptr; unsigned int* h_xcounts = nullptr;
    Dev* d_batch = nullptr; size_t batch_cap = 0; cudaGraphExec_t batch_graph = nullptr; size_t batch_graph_n = 0;   // cabm_batch_run_days: the communities of the batch, indexed by blockIdx.y
    int infect_seq = 0;                          // cabm_infect: sequence number the kernel echoes into h_pin[1] after the result
    // cabm_tick only stages its day; the launch happens in the next call into the handle. When that call is cabm_import for the same
    // day (the simulate_epidemic loop: tick(), then seed_epidemic(), simulator.h:338-374) both run as ONE day graph, 9 launches
    // instead of 9 + 2 and one host synchronisation instead of two; any other call launches the staged tick first (settle()).
    bool tick_pending = false; int pending_day = -1;
    bool row_ok = false;                         // the ledger row in h_pin (written by close_day) is still the device's: no call has touched the ledger since
    int* h_pin = nullptr;                        // pinned, device-visible scratch: single-word results are written by the kernel itself (no staged copy)
    const FlagRaw* peer_flags[CABM_MAX_SHARDS] = { nullptr }; int peer_mode = 0;      // exchange A over peer memory (cudaIpc mappings of the other handles' flag lists)
    int* d_hosp_ids = nullptr; int* d_pat_save = nullptr; int* d_pat_buf = nullptr; unsigned int n_hosp = 0; unsigned int n_flags_own = 0;   // sharding
    int profile = 0, n_marks = 0; std::vector<cudaEvent_t> marks; std::vector<int> mark_id; double ktime[32] = {0}; uint64_t kcount[32] = {0};
};

static void drop_graphs(cabm_t* h);

struct cabm_snap {
    std::vector<std::pair<void*, size_t>> bufs;   // device copies, same order as cabm::dyn
    void* log = nullptr; unsigned long long n_log = 0;
};

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { h->err = std::string(#call) + ": " + cudaGetErrorString(e_); return CABM_E_CUDA; } } while (0)

template <typename T>
static int dalloc(cabm_t* h, T** p, size_t n, bool dynamic = false, int fill = 0) {
    const size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
    CK(cudaMalloc((void**) p, bytes));
    CK(cudaMemsetAsync(*p, fill, bytes, h->stream));
    h->allocs.push_back(*p);
    if (dynamic) h->dyn.push_back({ (void*) *p, bytes });
    return CABM_OK;
}
template <typename T>
static int upload(cabm_t* h, T** p, const std::vector<T>& v) {
    int rc = dalloc(h, p, v.size()); if (rc) return rc;
    if (!v.empty()) CK(cudaMemcpyAsync(*p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return CABM_OK;
}

// handle-owned device scratch for the small entry points (dose lists, quarantine lists, gathers): grows, is never freed before
// cabm_destroy, so a failing call leaks nothing and a daily call allocates nothing
static int scratch(cabm_t* h, size_t bytes, void** out) {
    if (bytes > h->scratch_bytes) {
        if (h->scratch) { cudaStreamSynchronize(h->stream); cudaFree(h->scratch); h->scratch = nullptr; h->scratch_bytes = 0; }
        const size_t want = std::max<size_t>(bytes + bytes / 2, 1 << 16);
        CK(cudaMalloc(&h->scratch, want));
        h->scratch_bytes = want;
    }
    *out = h->scratch;
    return CABM_OK;
}
// the infection log is the only record of past infections (Person::infectionHistory): a record that did not fit is an error
static int log_count(cabm_t* h, unsigned long long* n) {
    CK(cudaMemcpy(n, &h->d.ctr->n_log, sizeof(*n), cudaMemcpyDeviceToHost));
    if (*n > h->d.cap_log) {
        h->err = "infection log capacity exceeded (" + std::to_string(*n) + " records, capacity " + std::to_string(h->d.cap_log) + "): histories are incomplete";
        return CABM_E_NOMEM;
    }
    return CABM_OK;
}

// ---- host-side keyed Poisson for the daily importation count (simulator.h:209) ----
static void philox_host(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t) 0xD2511F53u * c0, p1 = (uint64_t) 0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t) (p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t) p1, n2 = (uint32_t) (p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t) p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
static int import_count(uint64_t seed, int day, double mean) {
    int total = 0; uint32_t chunk = 0;
    const uint32_t key[2] = { (uint32_t) seed, (uint32_t) (seed >> 32) };
    while (mean > 0.0) {
        const double m = mean > 16.0 ? 16.0 : mean; mean -= m;
        const double limit = cabm_exp(-m);
        double prod = 1.0; int k = 0; uint32_t idx = 0, buf[4];
        for (;;) {
            if ((idx & 3u) == 0) { const uint32_t ctr[4] = { idx >> 2, chunk, 0u, ((uint32_t) day << 8) | CABM_P_IMPORT_COUNT }; philox_host(ctr, key, buf); }
            prod *= (double) buf[idx & 3u] * (1.0 / 4294967296.0); idx++;
            if (!(prod > limit)) break; ++k;
        }
        total += k; chunk++;
    }
    return total;
}

static unsigned int grid_for(const cabm_t* h, size_t n, int per_sm) {
    const size_t want = (n + TPB - 1) / TPB;
    return (unsigned int) std::max<size_t>(1, std::min<size_t>(want, (size_t) h->sms * per_sm));
}

static void fill_ctx(const cabm_t* h, DayCtx& c, int day, int julian_day, int is_weekday) {
    c.day = day; c.julian_day = julian_day; c.is_weekday = is_weekday;
    c.season = h->hpar.seasonality[julian_day - 1];
    c.school_closure = h->h_ti[0][day]; c.business_closure = h->h_ti[1][day]; c.sd = h->h_ti[2][day];
    float up = (float) c.sd; if ((double) up < c.sd) up = nextafterf(up, INFINITY);
    c.sd_up = up;
    // (double) risk < sd for a household  <=>  the rank of its riskiness among the distinct values < sd_rank
    c.sd_rank = (unsigned int) (std::lower_bound(h->risk_sorted.begin(), h->risk_sorted.end(), c.sd, [](float r, double sd) { return (double) r < sd; }) - h->risk_sorted.begin());
}

// kernel ids for per-kernel timing (cabm_kernel_times)
enum { KID_SCAN = 0, KID_ACTIVE, KID_VISIT, KID_HAZARD, KID_TRIALS, KID_GATE0, KID_APPLY_STATUS, KID_HOSP_IMPORT, KID_GATE1, KID_APPLY1,
       KID_END, KID_SMALL, KID_XA, KID_XB, KID_XC, KID_N };
static const char* const KID_NAMES[KID_N] = { "k_scan", "k_active", "k_visit", "k_hazard", "k_trials", "k_gate", "k_apply_status",
    "k_hosp_import", "(unused)", "k_apply(phase1)", "(unused)", "(unused)", "exchange A (flags)", "exchange B (census)", "exchange C (patients)" };

// LAUNCH records an event in front of every kernel when per-kernel profiling is on (stream path only, never under graph capture)
#define LAUNCH(id, ...) do { if (h->profile) prof_mark(h, id); __VA_ARGS__; h->launches++; } while (0)
static void prof_mark(cabm_t* h, int id) {
    if (h->n_marks < (int) h->marks.size()) { cudaEventRecord(h->marks[h->n_marks], h->stream); h->mark_id[h->n_marks] = id; h->n_marks++; }
}
static void prof_collect(cabm_t* h) {      // call after the stream has been synchronised
    if (!h->profile || h->n_marks == 0) return;
    for (int i = 0; i + 1 < h->n_marks; ++i) {
        float ms = 0; cudaEventElapsedTime(&ms, h->marks[i], h->marks[i + 1]);
        if (h->mark_id[i] >= 0) { h->ktime[h->mark_id[i]] += ms; h->kcount[h->mark_id[i]]++; }
    }
    h->n_marks = 0;
}

// One kernel of the day pipeline. With h->pdl the launch carries the programmatic-stream-serialization attribute: the kernel
// may start while its predecessor in the stream is still running and synchronises with it at its own griddepcontrol.wait
// (every kernel of the pipeline has one in front of its first dependent access). Captured into a graph this becomes a
// programmatic dependency edge. Per-kernel profiling (an event between kernels) launches without it.
template <typename... KArgs, typename... Args>
static void launch_k(cabm_t* h, int edge, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = h->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = (((h->pdl >> edge) & 1) && !h->profile) ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// The launch sequence of one day. tick = Community::tick (Community.cpp:1147-1192), import = the daily
// seed_epidemic (simulator.h:201-223). With both set the importation draws share the hospital pass's gate/apply
// launches: an agent hit by both keeps the hospital exposure (lower order key), exactly what running
// seed_epidemic after tick() yields. Nine launches for tick + importation, two for an importation alone.
// one wave of resident blocks: more, smaller blocks were measured slower (every block ends in a latency-bound tail)
static unsigned int visit_grid(const cabm_t* h, const Dev& D) {
    const unsigned int tiles = (D.own_hi - (D.own_lo & ~(VIS_TILE - 1u)) + VIS_TILE - 1) / VIS_TILE;
    return std::max(1u, std::min<unsigned int>((tiles + VIS_WARPS - 1) / VIS_WARPS, (unsigned int) (h->sms * h->visit_bps)));
}
// k_scan: one thread per 128 agents (four bitmap words) of the owned range
static unsigned int scan_grid(const Dev& D) {
    const unsigned int w_lo = (D.own_lo >> 5) & ~3u, w_hi = (D.own_hi + 31u) >> 5;
    return std::max(1u, (w_hi - w_lo + TPB * 4 - 1) / (TPB * 4));
}
// advance: bit 0 = move on to the next staged day context (run_days), bit 1 = leave the day's ledger row in pinned host memory
// (the per-day calls; a bulk run has no reader for it and saves the PCIe writes)
static void launch_day(cabm_t* h, bool tick, bool import, int advance) {
    const Dev& D = h->d;
    const unsigned int gSM = h->sms;
    if (tick) {
        LAUNCH(KID_SCAN, launch_k(h, 8, k_scan, dim3(scan_grid(D)), dim3(TPB), 0, D));
        LAUNCH(KID_ACTIVE, launch_k(h, 0, k_active, dim3(gSM * h->gm[0]), dim3(TPB), 0, D));
        LAUNCH(KID_VISIT, launch_k(h, 1, k_visit, dim3(visit_grid(h, D)), dim3(TPB), visit_smem_bytes(h->bitmap_words), D, h->bitmap_words));
        LAUNCH(KID_HAZARD, launch_k(h, 2, k_hazard, dim3(gSM * h->gm[1]), dim3(TPB), 0, D));
        LAUNCH(KID_TRIALS, launch_k(h, 3, k_trials, dim3(gSM * h->gm[2]), dim3(TPB), 0, D));
        LAUNCH(KID_GATE0, launch_k(h, 7, k_gate, dim3(gSM * h->gm[3]), dim3(128), 0, D));
        LAUNCH(KID_APPLY_STATUS, launch_k(h, 4, k_apply_status, dim3(gSM * h->gm[4]), dim3(128), 0, D, 0, 0));
    }
    LAUNCH(KID_HOSP_IMPORT, launch_k(h, 5, k_hosp_import, dim3(gSM * h->gm[5]), dim3(TPB), 0, D, tick ? 1 : 0, import ? 1 : 0));
    LAUNCH(KID_APPLY1, launch_k(h, 6, k_apply_status, dim3(gSM * h->gm[7]), dim3(128), 0, D, 1, (tick ? 1 : 0) | ((advance & 1) ? 2 : 0) | ((advance & 2) ? 4 : 0)));
    if (h->profile) prof_mark(h, -1);
}

// kernel arguments (the Dev record) are baked into the captured graphs: whatever changes it drops them
static void drop_graphs(cabm_t* h) {
    for (cudaGraphExec_t* g : { &h->day_graph, &h->tick_graph, &h->import_graph, &h->step_graph }) if (*g) { cudaGraphExecDestroy(*g); *g = nullptr; }
}
// capture one launch sequence of the day on the handle's stream; the launches it holds are returned through n_launches
static int capture_day(cabm_t* h, cudaGraphExec_t* out, bool tick, bool import, int advance, uint64_t* n_launches) {
    cudaGraph_t g = nullptr;
    CK(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
    const uint64_t l0 = h->launches;
    launch_day(h, tick, import, advance);
    if (n_launches) *n_launches = h->launches - l0;
    h->launches = l0;
    CK(cudaStreamEndCapture(h->stream, &g));
    CK(cudaGraphInstantiate(out, g, 0));
    cudaGraphDestroy(g);
    return CABM_OK;
}

static int ready(cabm_t* h, int day) {
    if (!h->have_par || !h->have_pop) { h->err = "population and parameters must be set first"; return CABM_E_STATE; }
    if (day < 0 || day >= h->hpar.run_length) { h->err = "day outside run_length"; return CABM_E_ARG; }
    return CABM_OK;
}
static int check_overflow(cabm_t* h) {
    unsigned int ov = 0;
    CK(cudaMemcpy(&ov, &h->d.ctr->overflow, sizeof(ov), cudaMemcpyDeviceToHost));
    if (ov & 0x80000000u) { h->err = "checked build: a device-side index check failed (mask " + std::to_string(ov) + ")"; return CABM_E_STATE; }
    if (ov & ~1u) { h->err = "device list capacity exceeded (mask " + std::to_string(ov) + ")"; return CABM_E_NOMEM; }
    return CABM_OK;
}
// after a synchronised launch sequence that closed a day (close_day): the overflow mask is already in pinned host memory (no copy)
static int check_overflow_host(cabm_t* h) {
    const unsigned int ov = h->d.host_status[0];
    if (ov & 0x80000000u) { h->err = "checked build: a device-side index check failed (mask " + std::to_string(ov) + ")"; return CABM_E_STATE; }
    if (ov & ~1u) { h->err = "device list capacity exceeded (mask " + std::to_string(ov) + ")"; return CABM_E_NOMEM; }
    return CABM_OK;
}
static int timed_sync(cabm_t* h) {
    CK(cudaEventRecord(h->e1, h->stream));
    CK(cudaEventSynchronize(h->e1));
    float ms = 0; CK(cudaEventElapsedTime(&ms, h->e0, h->e1)); h->acc_ms += ms;
    CK(cudaGetLastError());
    prof_collect(h);
    return CABM_OK;
}

// ---- NCCL, resolved at run time so that the library loads where NCCL is absent (CPU-only checks) ----
namespace {
struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};
NcclApi& nccl_api() {
    static NcclApi a; static bool tried = false;
    if (tried) return a;
    tried = true;
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return a;
    a.GetUniqueId = (decltype(a.GetUniqueId)) dlsym(lib, "ncclGetUniqueId"); a.CommInitRank = (decltype(a.CommInitRank)) dlsym(lib, "ncclCommInitRank");
    a.CommDestroy = (decltype(a.CommDestroy)) dlsym(lib, "ncclCommDestroy"); a.AllGather = (decltype(a.AllGather)) dlsym(lib, "ncclAllGather");
    a.AllReduce = (decltype(a.AllReduce)) dlsym(lib, "ncclAllReduce"); a.GetErrorString = (decltype(a.GetErrorString)) dlsym(lib, "ncclGetErrorString");
    a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllGather && a.AllReduce && a.GetErrorString;
    return a;
}
}
#define PM(id) do { if (h->profile) prof_mark(h, id); } while (0)
static const int SHARD_GROUP = 4;      // days enqueued between two host looks at the flag counts (cabm_shard_run_days)
#define NK(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) { h->err = std::string(#call) + ": " + nccl_api().GetErrorString(r_); return CABM_E_CUDA; } } while (0)


extern "C" {
static int settle(cabm_t* h);

int cabm_create(cabm_t** out) {
    if (!out) return CABM_E_ARG;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { *out = nullptr; return CABM_E_NODEVICE; }
    cabm_t* h = new (std::nothrow) cabm(); if (!h) return CABM_E_NOMEM;
    cudaGetDevice(&h->device);
    cudaDeviceProp prop; if (cudaGetDeviceProperties(&prop, h->device) != cudaSuccess) { delete h; *out = nullptr; return CABM_E_NODEVICE; }
    h->sms = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; *out = nullptr; return CABM_E_CUDA; }
    cudaEventCreate(&h->e0); cudaEventCreate(&h->e1);
    if (cudaHostAlloc((void**) &h->h_pin, 64 * sizeof(int), cudaHostAllocMapped) != cudaSuccess) { delete h; *out = nullptr; return CABM_E_CUDA; }
    std::memset(h->h_pin, 0, 64 * sizeof(int));
    h->d.host_status = reinterpret_cast<unsigned int*>(h->h_pin) + 4;       // written by close_day: [0] overflow mask, [1] new infections of the hospital pass + importation,
                                                                            // [2] the day it closed, [12..43] that day's ledger row (h_pin words 16-47)
    {   // device address of the static scratch: once per device, with that device current (it is: cudaGetDevice above)
        void*& addr = g_static_scratch_addr[h->device & 63];
        if (!addr) {
            void* p = nullptr;
            const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
            if (e == cudaSuccess && p) addr = p;
            else { cudaGetLastError(); h->static_scratch_err = std::string("cudaGetSymbolAddress(g_static_scratch): ") + cudaGetErrorString(e); }
        }
        h->static_scratch = (unsigned char*) addr;          // nullptr: cabm_serosurvey_static reports static_scratch_err; nothing else uses it
    }
    if (const char* e = getenv("CABM_PDL")) h->pdl = atoi(e);
    if (const char* e = getenv("CABM_GRID_MULT")) { int k = 0; for (const char* q = e; *q && k < 9; ++k) { h->gm[k] = std::max(1, atoi(q)); while (*q && *q != ',') ++q; if (*q == ',') ++q; } }
    *out = h;
    return CABM_OK;
}

int cabm_destroy(cabm_t* h) {
    if (!h) return CABM_OK;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    drop_graphs(h);
    if (h->nccl_comm && nccl_api().ok) nccl_api().CommDestroy((ncclComm_t) h->nccl_comm);
    for (int r = 0; r < CABM_MAX_SHARDS; ++r) if (h->peer_flags[r] && r != h->d.shard_rank) cudaIpcCloseMemHandle((void*) h->peer_flags[r]);
    if (h->d_xrecv) cudaFree(h->d_xrecv);
    if (h->h_xcounts) cudaFreeHost(h->h_xcounts);
    if (h->h_pin) cudaFreeHost(h->h_pin);
    if (h->scratch) cudaFree(h->scratch);
    if (h->batch_graph) cudaGraphExecDestroy(h->batch_graph);
    if (h->d_batch) cudaFree(h->d_batch);
    for (void* p : h->allocs) cudaFree(p);
    if (h->h_ctx) cudaFreeHost(h->h_ctx);
    cudaEventDestroy(h->e0); cudaEventDestroy(h->e1); cudaStreamDestroy(h->stream);
    delete h;
    return CABM_OK;
}

const char* cabm_last_error(const cabm_t* h) { return h ? h->err.c_str() : "null handle"; }

int cabm_set_params(cabm_t* h, const cabm_params* par) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (h) h->row_ok = false;                       // the ledger row in pinned memory may no longer be the device's after this call
    if (!h || !par || !par->prob_first_detection || par->run_length <= 0 || par->run_length > 32000) return CABM_E_ARG;
    cudaSetDevice(h->device);
    const size_t R = (size_t) par->run_length;
    if (h->have_par && par->run_length != h->hpar.run_length) { h->err = "run_length cannot change"; return CABM_E_STATE; }
    if (!h->have_par) {
        int rc;
        if ((rc = dalloc(h, &h->d_par, 1))) return rc;
        if ((rc = dalloc(h, &h->d_pfd, R * CABM_N_OUTCOMES))) return rc;
        if ((rc = dalloc(h, &h->d_imr, R))) return rc;
        if ((rc = dalloc(h, &h->d.ledger, (size_t) CABM_N_COUNTERS * R, true))) return rc;
        if ((rc = dalloc(h, &h->d.ctr, 1, true))) return rc;
        if ((rc = dalloc(h, &h->d_tmp_int, 4))) return rc;
        if ((rc = dalloc(h, &h->d_out_new, 1))) return rc;
        if ((rc = dalloc(h, &h->d_today, 1))) return rc;
        h->ctx_cap = (int) R;
        if ((rc = dalloc(h, &h->d_ctx, R))) return rc;
        CK(cudaMallocHost((void**) &h->h_ctx, R * sizeof(DayCtx)));
        for (int t = 0; t < 3; ++t) h->h_ti[t].assign(R, 0.0);
    }
    h->hpar = *par;
    CK(cudaMemcpyAsync(h->d_pfd, par->prob_first_detection, R * CABM_N_OUTCOMES * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (par->icu_mortality_reduction) CK(cudaMemcpyAsync(h->d_imr, par->icu_mortality_reduction, R * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    h->hpar.prob_first_detection = nullptr; h->hpar.icu_mortality_reduction = nullptr;
    CK(cudaMemcpyAsync(h->d_par, &h->hpar, sizeof(cabm_params), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    h->d.par = h->d_par; h->d.pfd = h->d_pfd; h->d.imr = h->d_imr; h->d.n_imr = par->icu_mortality_reduction ? (int) R : 0;
    h->d.R = (int) R; h->d.dayctx = h->d_ctx; h->d.today = h->d_today; h->d.seed = par->seed;
    for (int k = 0; k < 4; ++k) {                // u01 < cdf  <=>  x < ceil(cdf * 2^32) for a 32-bit draw x (exact: power-of-two scaling)
        const double c = std::ceil(par->visit_cdf[k] * 4294967296.0);
        h->d.visit_thr[k] = c >= 4294967295.0 ? 0xFFFFFFFFu : (unsigned int) c;
    }
    drop_graphs(h);    // kernel arguments are baked into the graph
    h->have_par = true;
    return CABM_OK;
}

int cabm_set_timed_intervention(cabm_t* h, int which, const double* v, size_t n) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !h->have_par || which < 0 || which > 2 || !v || n == 0) return CABM_E_ARG;
    const size_t R = (size_t) h->hpar.run_length;
    for (size_t d = 0; d < R; ++d) h->h_ti[which][d] = v[std::min(d, n - 1)];
    return CABM_OK;
}

int cabm_load_population(cabm_t* h, const cabm_population* pop) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (h) h->row_ok = false;                       // the ledger row in pinned memory may no longer be the device's after this call
    if (!h || !pop) return CABM_E_ARG;
    if (!h->have_par) { h->err = "cabm_set_params must precede cabm_load_population"; return CABM_E_STATE; }
    if (h->have_pop) { h->err = "population already loaded"; return CABM_E_STATE; }
    if (!pop->starting_natural_efficacy || !pop->cross_protection_probability || !pop->natural_immunity_duration) {
        h->err = "load-time immunity attributes are required"; return CABM_E_ARG; }
    if (pop->n_patron > 5) { h->err = "at most 5 patronised locations per person (Community.cpp:279-288)"; return CABM_E_ARG; }
    cudaSetDevice(h->device);
    const uint32_t N = pop->n_people, L = pop->n_loc, NP = pop->n_patron;
    Dev& D = h->d; D.N = N; D.L = L; D.n_patron = std::max<uint32_t>(NP, 1);
    D.own_lo = 0; D.own_hi = N; D.shard_rank = 0;
    int rc;
    // ---- static agents ----
    std::vector<int> home(pop->home, pop->home + N), dayl(pop->day, pop->day + N);
    std::vector<unsigned char> age(pop->age, pop->age + N), pflags(N);
    std::vector<AgRec> ag(N);                                                          // zeroed: no infection yet (imm == 0 says so)
    // patronised ("public") locations get a dense index q: the visit pass tests hotness in a shared-memory bitmap over q.
    // The riskiness of a household enters public_activity only through `risk < social_distancing[day]` (Community.cpp:1020),
    // so the patron row carries its RANK among the distinct household values and the day context the matching threshold.
    for (uint32_t p = 0; p < N; ++p)
        if (home[p] < 0 || (uint32_t) home[p] >= L || dayl[p] >= (int) L) { h->err = "home/day location out of range"; return CABM_E_ARG; }
    {   // one value per household, not per person
        std::vector<unsigned char> is_home(L, 0);
        for (uint32_t p = 0; p < N; ++p) is_home[home[p]] = 1;
        h->risk_sorted.clear();
        for (uint32_t l = 0; l < L; ++l) if (is_home[l]) h->risk_sorted.push_back(pop->loc_risk[l]);
    }
    std::sort(h->risk_sorted.begin(), h->risk_sorted.end());
    h->risk_sorted.erase(std::unique(h->risk_sorted.begin(), h->risk_sorted.end()), h->risk_sorted.end());
    h->risk_sorted.shrink_to_fit();
    if (h->risk_sorted.size() >= PROW_MAX_RANK) { h->err = "more than 2^24 distinct household riskiness values: the patron row holds a 24-bit rank"; return CABM_E_ARG; }
    const size_t n_rows = ((size_t) N + VIS_TILE - 1) / VIS_TILE * VIS_TILE;          // k_visit copies whole tiles: padded with empty rows
    std::vector<unsigned int> prow(std::max<size_t>(n_rows, VIS_TILE) * PROW_WORDS, 0u);
    std::vector<int> loc_pub(L, -1), pub_loc;
    std::vector<unsigned int> home_rank(L, 0xFFFFFFFFu);
    for (uint32_t p = 0; p < N; ++p) {
        pflags[p] = (pop->comorbid[p] ? PF_COMORBID : 0) | (pop->ltc[p] ? PF_LTC : 0);
        unsigned int e[5] = { 0, 0, 0, 0, 0 }, np = 0;
        for (uint32_t k = 0; k < NP; ++k) {
            const int loc = pop->patron[(size_t) p * NP + k]; if (loc < 0) break;
            if ((uint32_t) loc >= L) { h->err = "patronised location out of range"; return CABM_E_ARG; }
            if (loc_pub[loc] < 0) { loc_pub[loc] = (int) pub_loc.size(); pub_loc.push_back(loc); }
            if ((unsigned int) loc_pub[loc] >= PROW_MAX_Q) { h->err = "more than 2^19 patronised locations: the patron row holds 19-bit indices"; return CABM_E_ARG; }
            e[np++] = (unsigned int) loc_pub[loc] | (pop->loc_pubrisk[loc] == CABM_PT_HIGH ? 1u << PROW_HIGH_BIT : 0u);
        }
        unsigned int& rank = home_rank[home[p]];
        if (rank == 0xFFFFFFFFu) rank = (unsigned int) (std::lower_bound(h->risk_sorted.begin(), h->risk_sorted.end(), pop->loc_risk[home[p]]) - h->risk_sorted.begin());
        { const PRow r = prow_pack(e, np, rank); std::memcpy(&prow[(size_t) p * PROW_WORDS], r.w, sizeof(r.w)); }
        AgRec& a = ag[p];
        std::memset(&a, 0, sizeof(a));
        a.home = home[p]; a.day_loc = dayl[p]; a.hospital = pop->loc_hospital[home[p]];
        a.pbyte = (unsigned char) np; a.prow = prow16_pack(e, np, rank);
    }
    D.Q = (unsigned int) pub_loc.size();
    h->bitmap_words = std::max<unsigned int>(1, (D.Q + 31) / 32);
    {   // dynamic shared memory of k_visit / k_visit_b: the attribute is process-wide per kernel, so it only ever grows
        static size_t smem_set_dev[64] = { 0 };
        size_t& smem_set = smem_set_dev[h->device & 63];
        const size_t need = visit_smem_bytes(h->bitmap_words);
        if (need > 200 * 1024) { h->err = "too many patronised locations for the shared-memory hot bitmap"; return CABM_E_ARG; }
        if (need > smem_set) {
            CK(cudaFuncSetAttribute(k_visit, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) need));
            CK(cudaFuncSetAttribute(k_visit_b, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) need));
            smem_set = need;
        }
        int bps = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_visit, TPB, need));
        h->visit_bps = std::max(1, bps);
    }
    if (pub_loc.empty()) pub_loc.push_back(0);
    std::vector<double> nat_eff(pop->starting_natural_efficacy, pop->starting_natural_efficacy + N);
    std::vector<float> xprot(pop->cross_protection_probability, pop->cross_protection_probability + N);
    std::vector<int> nat_dur(pop->natural_immunity_duration, pop->natural_immunity_duration + N);
    if ((rc = upload(h, (int**) &D.home, home)) || (rc = upload(h, (int**) &D.day_loc, dayl)) || (rc = upload(h, (unsigned int**) &D.prow, prow)) ||
        (rc = upload(h, (unsigned char**) &D.age, age)) || (rc = upload(h, (unsigned char**) &D.pflags, pflags)) ||
        (rc = upload(h, (double**) &D.nat_eff, nat_eff)) ||
        (rc = upload(h, (float**) &D.xprot, xprot)) || (rc = upload(h, (int**) &D.nat_dur, nat_dur)) ||
        (rc = upload(h, (int**) &D.loc_pub, loc_pub)) || (rc = upload(h, (int**) &D.pub_loc, pub_loc))) return rc;
    // ---- static locations ----
    std::vector<unsigned char> ltype(pop->loc_type, pop->loc_type + L), lflags(L);
    for (uint32_t l = 0; l < L; ++l) lflags[l] = (pop->loc_essential[l] ? LF_ESSENTIAL : 0) | ((pop->loc_pubrisk[l] & 3) << 1);
    std::vector<float> lrisk(pop->loc_risk, pop->loc_risk + L);
    std::vector<int> lhosp(pop->loc_hospital, pop->loc_hospital + L);
    if (pop->roster_off[L] >= (int64_t) 0xFFFFFFFFll || pop->nbr_off[L] >= (int64_t) 0xFFFFFFFFll) { h->err = "CSR too large"; return CABM_E_ARG; }
    std::vector<unsigned int> roff(L + 1), noff(L + 1);
    for (uint32_t l = 0; l <= L; ++l) { roff[l] = (unsigned int) pop->roster_off[l]; noff[l] = (unsigned int) pop->nbr_off[l]; }
    std::vector<int> roster(pop->roster, pop->roster + pop->roster_off[L]), nbr(pop->nbr, pop->nbr + pop->nbr_off[L]);
    if ((rc = upload(h, (unsigned char**) &D.loc_type, ltype)) || (rc = upload(h, (unsigned char**) &D.loc_flags, lflags)) ||
        (rc = upload(h, (float**) &D.loc_risk, lrisk)) || (rc = upload(h, (int**) &D.loc_hospital, lhosp)) ||
        (rc = upload(h, (unsigned int**) &D.roster_off, roff)) || (rc = upload(h, (unsigned int**) &D.nbr_off, noff)) ||
        (rc = upload(h, (int**) &D.roster, roster)) || (rc = upload(h, (int**) &D.nbr, nbr))) return rc;
    {   // the same static words, one 32-byte record per location
        std::vector<LocS> locs(L);
        for (uint32_t l = 0; l < L; ++l) {
            LocS& r = locs[l]; std::memset(&r, 0, sizeof(r));
            r.roster_begin = roff[l]; r.roster_size = roff[l + 1] - roff[l]; r.nbr_begin = noff[l]; r.nbr_deg = noff[l + 1] - noff[l];
            r.risk = lrisk[l]; r.pub = loc_pub[l]; r.hospital = lhosp[l]; r.type = ltype[l]; r.flags = lflags[l];
        }
        if ((rc = upload(h, (LocS**) &D.locs, locs))) return rc;
    }
    {   // hospitals, for the census exchange of a sharded community (cabm_shard_patients)
        std::vector<int> hosp_ids;
        for (uint32_t l = 0; l < L; ++l) if (ltype[l] == CABM_HOSPITAL) hosp_ids.push_back((int) l);
        h->n_hosp = (unsigned int) hosp_ids.size();
        if (hosp_ids.empty()) hosp_ids.push_back(0);
        if ((rc = upload(h, &h->d_hosp_ids, hosp_ids)) || (rc = dalloc(h, &h->d_pat_save, hosp_ids.size())) || (rc = dalloc(h, &h->d_pat_buf, hosp_ids.size()))) return rc;
    }
    // ---- dynamic state ----
    const size_t Q1 = std::max<size_t>(D.Q, 1);
    if ((rc = dalloc(h, &D.ag, N, true)) ||
        (rc = dalloc(h, &D.imm, N, true)) || (rc = dalloc(h, &D.vax, N, true)) || (rc = dalloc(h, &D.vax_days, (size_t) N * CABM_MAX_DOSES, true)) ||
        (rc = dalloc(h, &D.quar, N, true)) ||
        (rc = dalloc(h, &D.cand, N)) || (rc = dalloc(h, &D.inf_place, N, true, 0xFF)) || (rc = dalloc(h, &D.inf_by, N, true, 0xFF)) ||
        (rc = dalloc(h, &D.locd, L)) || (rc = dalloc(h, &D.pat_count, L, true)) ||
        (rc = dalloc(h, &D.nvis, Q1)) || (rc = dalloc(h, &D.pub_T, Q1)) || (rc = dalloc(h, &D.pub_hot, ((size_t) h->bitmap_words + 3) & ~(size_t) 3)) ||
        (rc = dalloc(h, &D.inhosp, n_rows / 32 + 4, true)) || (rc = dalloc(h, &D.actbits, ((size_t) N + 1023) / 1024 * 32, true)) || (rc = dalloc(h, &D.active_list, (size_t) N))) return rc;
    CK(cudaMemcpyAsync(D.ag, ag.data(), (size_t) N * sizeof(AgRec), cudaMemcpyHostToDevice, h->stream));
    // ---- day lists ----
    D.cap_hot = L; D.cap_active = N; D.cap_flags = 3 * N + 1024; D.cap_vis = 2 * N + 1024; D.cap_hits = 2 * N + 4096 * HIT_LISTS; D.hit_sub = D.cap_hits / HIT_LISTS; D.cap_new = N; D.cap_pat = N;
    D.cap_items = (unsigned int) std::min<size_t>(0x7FFFFFFF, roster.size() / CHUNK + roster.size() / SMALL_ROSTER + 1024);
    D.cap_log = (unsigned long long) 3 * N + 1024;
    if (const char* e = getenv("CABM_LOG_CAP")) D.cap_log = std::max<unsigned long long>(1, strtoull(e, nullptr, 10));   // test knob: force the log-overflow error
    D.cap_sitems = (unsigned int) std::min<size_t>(0x7FFFFFFF, (size_t) L + nbr.size() + 1024);
    if ((rc = dalloc(h, &D.hot_list, D.cap_hot)) || (rc = dalloc(h, &D.hot_meta, D.cap_hot)) || (rc = dalloc(h, &D.hot_hosp, D.cap_hot)) || (rc = dalloc(h, &D.hot_off, D.cap_hot)) || (rc = dalloc(h, &D.hot_T, (size_t) 2 * D.cap_hot)) || (rc = dalloc(h, &D.sitems, D.cap_sitems)) || (rc = dalloc(h, &D.flags_raw, D.cap_flags)) || (rc = dalloc(h, &D.flags, D.cap_flags)) ||
        (rc = dalloc(h, &D.visit_rec, D.cap_vis)) ||
        (rc = dalloc(h, &D.items, D.cap_items)) || (rc = dalloc(h, &D.hits, D.cap_hits)) || (rc = dalloc(h, &D.hit_ctr, HIT_LISTS * 32)) || (rc = dalloc(h, &D.newinf_list, D.cap_new)) ||
        (rc = dalloc(h, &D.patient_list, D.cap_pat, true)) || (rc = dalloc(h, &D.log, D.cap_log)) || (rc = dalloc(h, &D.foi, D.cap_hot)) ||
        (rc = dalloc(h, &D.n_foi, 1))) return rc;
#ifdef CABM_CHECKS
    if (getenv("CABM_CHECK_SELFTEST")) D.Q = 1;      // checked build only: the device then believes there is one patronised location, so the
                                                     // first hot visit elsewhere must trip `q < D.Q` (arrays keep their real size: nothing is out of bounds)
#endif
    k_init_state<<<grid_for(h, std::max(N, L), 8), TPB, 0, h->stream>>>(D);
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    h->have_pop = true;
    return CABM_OK;
}

// ---- the built population as one binary file (cabm.h: cabm_population_save / cabm_load_population_file; format in cabm_popfile.h) ----
int cabm_population_save(const cabm_population* pop, const char* path) { return cabm_popfile::save(pop, path); }

int cabm_load_population_file(cabm_t* h, const char* path, const double* starting_natural_efficacy,
                              const float* cross_protection_probability, const int32_t* natural_immunity_duration) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !path) return CABM_E_ARG;
    cabm_popfile::Mapped m;
    int rc = m.open_file(path, &h->err);
    if (rc) return rc;
    m.pop.starting_natural_efficacy = starting_natural_efficacy; m.pop.cross_protection_probability = cross_protection_probability;
    m.pop.natural_immunity_duration = natural_immunity_duration;
    return cabm_load_population(h, &m.pop);                 // the mapping goes away with `m`
}

// the staged day of cabm_tick (h_ctx[0]) as its own launch sequence
static int launch_tick_now(cabm_t* h) {
    int rc;
    cudaSetDevice(h->device);
    CK(cudaMemcpyAsync(h->d_ctx, h->h_ctx, sizeof(DayCtx), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemsetAsync(&h->d.ctr->day_idx, 0, sizeof(unsigned int), h->stream));
    CK(cudaEventRecord(h->e0, h->stream));
    if (h->profile) launch_day(h, true, false, 2);
    else {
        if (!h->tick_graph && (rc = capture_day(h, &h->tick_graph, true, false, 2, &h->launches_tick))) return rc;
        CK(cudaGraphLaunch(h->tick_graph, h->stream)); h->launches += h->launches_tick;
    }
    if ((rc = timed_sync(h))) return rc;
    h->row_ok = true;
    return check_overflow_host(h);
}
// every entry point but cabm_import calls this first: a tick that was staged and not yet launched runs now, and its errors are
// this call's errors
static int settle(cabm_t* h) {
    if (!h || !h->tick_pending) return CABM_OK;
    h->tick_pending = false;
    return launch_tick_now(h);
}

int cabm_tick(cabm_t* h, int sim_day, int julian_day, int is_weekday) {
    if (!h) return CABM_E_ARG; int rc = settle(h); if (rc) return rc;
    if ((rc = ready(h, sim_day))) return rc;
    if (julian_day < 1 || julian_day > 366) return CABM_E_ARG;
    DayCtx& c = h->h_ctx[0]; std::memset(&c, 0, sizeof(c));
    fill_ctx(h, c, sim_day, julian_day, is_weekday);
    if (h->profile || h->d.own_lo != 0 || h->d.own_hi != h->d.N || getenv("CABM_EAGER_TICK")) return launch_tick_now(h);
    h->tick_pending = true; h->pending_day = sim_day;        // launched by the next call (see tick_pending)
    return CABM_OK;
}

int cabm_import(cabm_t* h, int sim_day, double expected, const double strain_w[CABM_N_STRAINS], int32_t* out_infected) {
    if (!h || !strain_w) return CABM_E_ARG;
    int rc;
    if (h->tick_pending && h->pending_day != sim_day && (rc = settle(h))) return rc;
    if ((rc = ready(h, sim_day))) return rc;
    cudaSetDevice(h->device);
    const int n_imports = expected > 0 ? import_count(h->hpar.seed, sim_day, expected) : 0;
    if (h->tick_pending) {
        // tick(d) followed by the importation of day d: one day graph, exactly what cabm_run_days replays
        h->tick_pending = false;
        DayCtx& c = h->h_ctx[0];
        c.n_imports = n_imports;
        for (int s = 0; s < CABM_N_STRAINS; ++s) c.strain_w[s] = strain_w[s];
        CK(cudaMemcpyAsync(h->d_ctx, h->h_ctx, sizeof(DayCtx), cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemsetAsync(&h->d.ctr->day_idx, 0, sizeof(unsigned int), h->stream));
        CK(cudaEventRecord(h->e0, h->stream));
        if (!h->step_graph && (rc = capture_day(h, &h->step_graph, true, true, 3, &h->launches_per_day))) return rc;
        CK(cudaGraphLaunch(h->step_graph, h->stream)); h->launches += h->launches_per_day;
    } else {
        DayCtx& c = h->h_ctx[0]; std::memset(&c, 0, sizeof(c));
        c.day = sim_day; c.julian_day = 1; c.n_imports = n_imports;
        for (int s = 0; s < CABM_N_STRAINS; ++s) c.strain_w[s] = strain_w[s];
        CK(cudaMemcpyAsync(h->d_today, h->h_ctx, sizeof(DayCtx), cudaMemcpyHostToDevice, h->stream));   // no k_active in front: publish the context directly
        CK(cudaEventRecord(h->e0, h->stream));
        if (h->profile) launch_day(h, false, true, 2);
        else {
            if (!h->import_graph && (rc = capture_day(h, &h->import_graph, false, true, 2, &h->launches_import))) return rc;
            CK(cudaGraphLaunch(h->import_graph, h->stream)); h->launches += h->launches_import;
        }
    }
    if ((rc = timed_sync(h))) return rc;
    if (out_infected) *out_infected = (int32_t) h->d.host_status[3];               // winners whose exposure was an import, written by close_day before it zeroes the day counters
    h->row_ok = true;
    return check_overflow_host(h);
}

int cabm_infect(cabm_t* h, uint32_t pid, int strain, int sim_day, int32_t* out_infected) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (h) h->row_ok = false;                       // the ledger row in pinned memory may no longer be the device's after this call
    if (!h) return CABM_E_ARG; int rc = ready(h, sim_day); if (rc) return rc;
    if (pid >= h->d.N || strain < 0 || strain >= CABM_N_STRAINS) return CABM_E_ARG;
    cudaSetDevice(h->device);
    // The result and then a sequence number land in pinned host memory over PCIe; the host watches the sequence word instead of
    // waiting for the driver to report the stream idle (the drivers call this once per imported case, hundreds of times a day).
    const int seq = ++h->infect_seq;
    k_seed_one<<<1, 32, 0, h->stream>>>(h->d, pid, strain, sim_day, h->h_pin, seq);
    CK(cudaGetLastError());
    volatile int* pin = h->h_pin;
    bool seen = false;
    for (unsigned int spins = 0; spins < (1u << 22); ++spins) if (pin[1] == seq) { seen = true; break; }   // ~10 us normally; a few ms at most
    if (!seen) { CK(cudaStreamSynchronize(h->stream)); CK(cudaGetLastError()); if (pin[1] != seq) { h->err = "cabm_infect: the kernel did not report"; return CABM_E_CUDA; } }
    if (out_infected) *out_infected = pin[0];
    return CABM_OK;
}

int cabm_infect_batch(cabm_t* h, const uint32_t* pids, const uint8_t* strains, size_t n, int sim_day, uint8_t* out_infected) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (h) h->row_ok = false;                       // the ledger row in pinned memory may no longer be the device's after this call
    if (!h || ((!pids || !strains) && n)) return CABM_E_ARG; int rc = ready(h, sim_day); if (rc) return rc;
    if (n == 0) return CABM_OK;
    if (n > 0x7FFFFFFFu || h->d.N >= (1u << 29)) { h->err = "cabm_infect_batch: at most 2^31 calls and 2^29 agents"; return CABM_E_ARG; }
    for (size_t k = 0; k < n; ++k) if (pids[k] >= h->d.N || strains[k] >= CABM_N_STRAINS) return CABM_E_ARG;
    cudaSetDevice(h->device);
    // Calls on the same person depend on each other, all others do not: wave w holds the (w+1)-th call of every person, waves
    // run one after the other (importation draws people uniformly, so there is almost always one wave)
    std::vector<unsigned int> order(n), wave(n);
    {
        std::vector<std::pair<unsigned int, unsigned int>> by_pid(n);
        for (size_t k = 0; k < n; ++k) by_pid[k] = { pids[k], (unsigned int) k };
        std::sort(by_pid.begin(), by_pid.end());
        for (size_t i = 0; i < n; ++i) wave[by_pid[i].second] = (i && by_pid[i].first == by_pid[i - 1].first) ? wave[by_pid[i - 1].second] + 1 : 0u;
        for (size_t k = 0; k < n; ++k) order[k] = (unsigned int) k;
        std::stable_sort(order.begin(), order.end(), [&](unsigned int a, unsigned int b) { return wave[a] < wave[b]; });
    }
    std::vector<unsigned int> items(n), seq_off(n);
    for (size_t i = 0; i < n; ++i) { const unsigned int k = order[i]; items[i] = pids[k] | ((unsigned int) strains[k] << 29); seq_off[i] = k; }
    const size_t out_off = 2 * n * sizeof(unsigned int);
    void* buf = nullptr; if ((rc = scratch(h, out_off + n, &buf))) return rc;
    unsigned int* d_items = (unsigned int*) buf; unsigned int* d_seq = d_items + n; unsigned char* d_out = (unsigned char*) buf + out_off;
    CK(cudaMemcpyAsync(d_items, items.data(), n * sizeof(unsigned int), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_seq, seq_off.data(), n * sizeof(unsigned int), cudaMemcpyHostToDevice, h->stream));
    for (size_t i0 = 0; i0 < n;) {
        size_t i1 = i0; while (i1 < n && wave[order[i1]] == wave[order[i0]]) ++i1;
        k_seed_batch<<<grid_for(h, i1 - i0, 4), TPB, 0, h->stream>>>(h->d, d_items + i0, d_seq + i0, (unsigned int) (i1 - i0), sim_day, d_out);
        h->launches++;
        i0 = i1;
    }
    k_seed_bump<<<1, 32, 0, h->stream>>>(h->d, (unsigned int) n);
    std::vector<unsigned char> took(n);
    CK(cudaMemcpyAsync(took.data(), d_out, n, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    if (out_infected) std::memcpy(out_infected, took.data(), n);
    return check_overflow(h);
}

int cabm_run_days(cabm_t* h, int first_day, int n_days, const int32_t* julian_day, const int32_t* is_weekday,
                  const double* expected_imports, const double* strain_w) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (h) h->row_ok = false;                       // a bulk run leaves no ledger row in pinned memory
    if (!h || n_days <= 0 || !julian_day || !is_weekday || !expected_imports || !strain_w) return CABM_E_ARG;
    int rc = ready(h, first_day); if (rc) return rc; if ((rc = ready(h, first_day + n_days - 1))) return rc;
    cudaSetDevice(h->device);
    for (int i = 0; i < n_days; ++i) {
        DayCtx& c = h->h_ctx[i]; std::memset(&c, 0, sizeof(c));
        if (julian_day[i] < 1 || julian_day[i] > 366) return CABM_E_ARG;
        fill_ctx(h, c, first_day + i, julian_day[i], is_weekday[i]);
        c.n_imports = expected_imports[i] > 0 ? import_count(h->hpar.seed, first_day + i, expected_imports[i]) : 0;
        for (int s = 0; s < CABM_N_STRAINS; ++s) c.strain_w[s] = strain_w[(size_t) i * CABM_N_STRAINS + s];
    }
    CK(cudaMemcpyAsync(h->d_ctx, h->h_ctx, (size_t) n_days * sizeof(DayCtx), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemsetAsync(&h->d.ctr->day_idx, 0, sizeof(unsigned int), h->stream));
    CK(cudaEventRecord(h->e0, h->stream));
    if (h->profile) {                // per-kernel timing: plain stream launches, one event in front of every kernel
        for (int i = 0; i < n_days; ++i) {
            launch_day(h, true, true, 1);
            CK(cudaStreamSynchronize(h->stream)); prof_collect(h);
        }
        if ((rc = timed_sync(h))) return rc;
        return check_overflow(h);
    }
    if (!h->day_graph && (rc = capture_day(h, &h->day_graph, true, true, 1, &h->launches_per_day))) return rc;   // one day (tick + importation), captured once, replayed every day
    for (int i = 0; i < n_days; ++i) { CK(cudaGraphLaunch(h->day_graph, h->stream)); h->launches += h->launches_per_day; }
    if ((rc = timed_sync(h))) return rc;
    return check_overflow_host(h);
}

int cabm_batch_run_days(cabm_t* const* hs, size_t n, int first_day, int n_days, const int32_t* julian_day, const int32_t* is_weekday,
                        const double* expected_imports, const double* strain_w) {
    for (size_t i = 0; hs && i < n; ++i) if (hs[i]) { hs[i]->row_ok = false; const int rc_settle = settle(hs[i]); if (rc_settle) return rc_settle; }
    if (!hs || n == 0 || n > 65535 || n_days <= 0 || !julian_day || !is_weekday || !expected_imports || !strain_w) return CABM_E_ARG;
    cabm_t* h = hs[0]; if (!h) return CABM_E_ARG;
    int rc = ready(h, first_day); if (rc) return rc; if ((rc = ready(h, first_day + n_days - 1))) return rc;
    for (size_t r = 0; r < n; ++r) {
        cabm_t* g = hs[r];
        if (!g || (rc = ready(g, first_day + n_days - 1))) { h->err = "batch: a handle is not ready for these days"; return CABM_E_STATE; }
        if (g->device != h->device || g->d.N != h->d.N || g->d.L != h->d.L || g->d.Q != h->d.Q || g->bitmap_words != h->bitmap_words ||
            g->d.own_lo != 0 || g->d.own_hi != g->d.N) { h->err = "batch: handles must hold populations of one shape on one device, unsharded"; return CABM_E_ARG; }
        for (size_t q = 0; q < r; ++q) if (hs[q] == g) { h->err = "batch: a handle appears twice"; return CABM_E_ARG; }
    }
    for (int i = 0; i < n_days; ++i) if (julian_day[i] < 1 || julian_day[i] > 366) return CABM_E_ARG;
    cudaSetDevice(h->device);
    cudaStream_t st = h->stream;
    std::vector<Dev> devs(n);
    for (size_t r = 0; r < n; ++r) {
        cabm_t* g = hs[r];
        if (g != h) CK(cudaStreamSynchronize(g->stream));              // earlier work of that handle; from here on everything is on st
        for (int i = 0; i < n_days; ++i) {
            DayCtx& c = g->h_ctx[i]; std::memset(&c, 0, sizeof(c));
            fill_ctx(g, c, first_day + i, julian_day[i], is_weekday[i]);
            c.n_imports = expected_imports[i] > 0 ? import_count(g->hpar.seed, first_day + i, expected_imports[i]) : 0;
            for (int s = 0; s < CABM_N_STRAINS; ++s) c.strain_w[s] = strain_w[(size_t) i * CABM_N_STRAINS + s];
        }
        CK(cudaMemcpyAsync(g->d_ctx, g->h_ctx, (size_t) n_days * sizeof(DayCtx), cudaMemcpyHostToDevice, st));
        CK(cudaMemsetAsync(&g->d.ctr->day_idx, 0, sizeof(unsigned int), st));
        devs[r] = g->d;
    }
    if (h->batch_cap < n) {
        if (h->batch_graph) { cudaGraphExecDestroy(h->batch_graph); h->batch_graph = nullptr; }     // it holds the old table's address
        if (h->d_batch) cudaFree(h->d_batch);
        CK(cudaMalloc(&h->d_batch, n * sizeof(Dev))); h->batch_cap = n;
    }
    CK(cudaMemcpyAsync(h->d_batch, devs.data(), n * sizeof(Dev), cudaMemcpyHostToDevice, st));
    const Dev* B = h->d_batch; const Dev& D = h->d;
    const unsigned int R = (unsigned int) n, gSM = h->sms;
    auto gx = [&](int k) { return std::max(2u, gSM * (unsigned int) h->gm[k] / R); };             // the SMs are shared by the batch
    const dim3 gVisit(std::min<unsigned int>(visit_grid(h, D), std::max(2u, gSM * (unsigned int) h->visit_bps / R)), R);
    auto one_day = [&]() {
        launch_k(h, 8, k_scan_b, dim3(scan_grid(D), R), dim3(TPB), 0, B);
        launch_k(h, 0, k_active_b, dim3(gx(0), R), dim3(TPB), 0, B);
        launch_k(h, 1, k_visit_b, gVisit, dim3(TPB), visit_smem_bytes(h->bitmap_words), B, h->bitmap_words);
        launch_k(h, 2, k_hazard_b, dim3(gx(1), R), dim3(TPB), 0, B);
        launch_k(h, 3, k_trials_b, dim3(gx(2), R), dim3(TPB), 0, B);
        launch_k(h, 7, k_gate_b, dim3(gx(3), R), dim3(128), 0, B);
        launch_k(h, 4, k_apply_status_b, dim3(gx(4), R), dim3(128), 0, B, 0, 0);
        launch_k(h, 5, k_hosp_import_b, dim3(gx(5), R), dim3(TPB), 0, B, 1, 1);
        launch_k(h, 6, k_apply_status_b, dim3(gx(7), R), dim3(128), 0, B, 1, 3);
    };
    if (!h->batch_graph || h->batch_graph_n != n) {                 // the day of this batch as one graph (the Dev table is read at run time)
        if (h->batch_graph) { cudaGraphExecDestroy(h->batch_graph); h->batch_graph = nullptr; }
        cudaGraph_t g = nullptr;
        CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        one_day();
        CK(cudaStreamEndCapture(st, &g));
        CK(cudaGraphInstantiate(&h->batch_graph, g, 0));
        cudaGraphDestroy(g);
        h->batch_graph_n = n;
    }
    CK(cudaEventRecord(h->e0, st));
    for (int i = 0; i < n_days; ++i) { CK(cudaGraphLaunch(h->batch_graph, st)); h->launches += 9; }
    if ((rc = timed_sync(h))) return rc;
    for (size_t r = 0; r < n; ++r) if ((rc = check_overflow(hs[r]))) { if (hs[r] != h) h->err = hs[r]->err; return rc; }
    return CABM_OK;
}

int cabm_vaccinate(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint8_t* out_ok) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || (!pids && n)) return CABM_E_ARG; int rc = ready(h, sim_day); if (rc) return rc;
    if (!h->hpar.immunity_leaky) { h->err = "vaccination requires leaky immunity (Person.cpp:622)"; return CABM_E_STATE; }
    if (n == 0) return CABM_OK;
    for (size_t i = 0; i < n; ++i) if (pids[i] >= h->d.N) return CABM_E_ARG;
    cudaSetDevice(h->device);
    void* buf = nullptr; int rc2 = scratch(h, n * sizeof(unsigned int) + n, &buf); if (rc2) return rc2;
    unsigned int* d_p = (unsigned int*) buf; unsigned char* d_ok = (unsigned char*) buf + n * sizeof(unsigned int);
    CK(cudaMemcpyAsync(d_p, pids, n * sizeof(unsigned int), cudaMemcpyHostToDevice, h->stream));
    k_vaccinate<<<grid_for(h, n, 4), TPB, 0, h->stream>>>(h->d, d_p, (unsigned int) n, sim_day, d_ok);
    std::vector<unsigned char> ok(n);
    CK(cudaMemcpyAsync(ok.data(), d_ok, n, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (out_ok) std::memcpy(out_ok, ok.data(), n);
    return CABM_OK;
}

int cabm_quarantine(cabm_t* h, const uint32_t* pids, size_t n, int from_day, int until_day) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || (!pids && n)) return CABM_E_ARG;
    if (!h->have_pop) return CABM_E_STATE;
    if (n == 0) return CABM_OK;
    for (size_t i = 0; i < n; ++i) if (pids[i] >= h->d.N) return CABM_E_ARG;
    cudaSetDevice(h->device);
    void* buf = nullptr; int rc2 = scratch(h, n * sizeof(unsigned int), &buf); if (rc2) return rc2;
    unsigned int* d_p = (unsigned int*) buf;
    CK(cudaMemcpyAsync(d_p, pids, n * sizeof(unsigned int), cudaMemcpyHostToDevice, h->stream));
    k_quarantine<<<grid_for(h, n, 4), TPB, 0, h->stream>>>(h->d, d_p, (unsigned int) n, from_day, std::min(until_day, 32766));
    CK(cudaStreamSynchronize(h->stream));
    if (!h->d.any_quarantine) { h->d.any_quarantine = 1; drop_graphs(h); }
    return CABM_OK;
}

int cabm_ledger_series(cabm_t* h, int counter, uint64_t* out, size_t n) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !h->have_par || counter < 0 || counter >= CABM_N_COUNTERS || !out) return CABM_E_ARG;
    cudaSetDevice(h->device);
    const size_t R = (size_t) h->hpar.run_length, m = std::min(n, R);
    CK(cudaMemcpy(out, h->d.ledger + (size_t) counter * R, m * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    for (size_t d = m; d < n; ++d) out[d] = 0;
    return CABM_OK;
}
int cabm_ledger_table(cabm_t* h, uint64_t* out, size_t n) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !h->have_par || !out) return CABM_E_ARG;
    cudaSetDevice(h->device);
    const size_t all = (size_t) h->hpar.run_length * CABM_N_COUNTERS, m = std::min(n, all);
    CK(cudaMemcpy(out, h->d.ledger, m * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    for (size_t i = m; i < n; ++i) out[i] = 0;
    return CABM_OK;
}
int cabm_ledger_row(cabm_t* h, int sim_day, uint64_t out[CABM_N_COUNTERS]) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !out) return CABM_E_ARG; int rc = ready(h, sim_day); if (rc) return rc;
    if (h->row_ok && (int) h->d.host_status[2] == sim_day) {                         // the day that was just closed: its row is already in pinned host memory
        std::memcpy(out, h->d.host_status + 12, CABM_N_COUNTERS * sizeof(uint64_t));
        return CABM_OK;
    }
    cudaSetDevice(h->device);
    CK(cudaMemcpy2D(out, sizeof(uint64_t), h->d.ledger + sim_day, (size_t) h->hpar.run_length * sizeof(uint64_t), sizeof(uint64_t), CABM_N_COUNTERS, cudaMemcpyDeviceToHost));
    return CABM_OK;
}
int cabm_cumulative_outcomes(cabm_t* h, uint64_t out[CABM_N_OUTCOMES]) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !h->have_par || !out) return CABM_E_ARG;
    cudaSetDevice(h->device);
    CK(cudaMemcpy(out, h->d.ctr->cumul, sizeof(uint64_t) * CABM_N_OUTCOMES, cudaMemcpyDeviceToHost));
    return CABM_OK;
}
int cabm_work_counters(cabm_t* h, uint64_t out[CABM_N_WORK]) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !h->have_par || !out) return CABM_E_ARG;
    cudaSetDevice(h->device);
    CK(cudaMemcpy(out, h->d.ctr->work, sizeof(uint64_t) * CABM_N_WORK, cudaMemcpyDeviceToHost));
    return CABM_OK;
}

int cabm_sync_histories(cabm_t* h, cabm_infection* out, size_t cap, size_t* out_n) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !h->have_pop) return CABM_E_ARG;
    cudaSetDevice(h->device);
    unsigned long long n = 0;
    { const int rc = log_count(h, &n); if (rc) return rc; }
    if (out_n) *out_n = (size_t) n;
    if (out && cap) CK(cudaMemcpy(out, h->d.log, std::min<size_t>(cap, (size_t) n) * sizeof(cabm_infection), cudaMemcpyDeviceToHost));
    return CABM_OK;
}

int cabm_sync_histories_from(cabm_t* h, size_t first, cabm_infection* out, size_t cap, size_t* out_n) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !h->have_pop) return CABM_E_ARG;
    cudaSetDevice(h->device);
    unsigned long long n = 0;
    { const int rc = log_count(h, &n); if (rc) return rc; }
    const size_t avail = first < n ? (size_t) n - first : 0;
    if (out_n) *out_n = avail;
    if (out && cap && avail) CK(cudaMemcpy(out, h->d.log + first, std::min(cap, avail) * sizeof(cabm_infection), cudaMemcpyDeviceToHost));
    return CABM_OK;
}

// ---------------------------------------------------------------------------------------------
// One community over several handles; see include/cabm.h for the protocol.
// ---------------------------------------------------------------------------------------------
int cabm_shard_config(cabm_t* h, int rank, int world, uint32_t pid_lo, uint32_t pid_hi) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !h->have_pop || world < 1 || rank < 0 || rank >= world || pid_lo > pid_hi || pid_hi > h->d.N) return CABM_E_ARG;
    cudaSetDevice(h->device);
    h->d.own_lo = pid_lo; h->d.own_hi = pid_hi; h->d.shard_rank = rank;
    drop_graphs(h);      // the captured graph carries the old range
    return CABM_OK;
}

int cabm_shard_day_part(cabm_t* h, int part, int sim_day, int julian_day, int is_weekday, double expected_imports,
                        const double* strain_w, int32_t* out_imported) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (h) h->row_ok = false;                       // the ledger row in pinned memory may no longer be the device's after this call
    if (!h || part < 0 || part > 3 || julian_day < 1 || julian_day > 366) return CABM_E_ARG;
    int rc = ready(h, sim_day); if (rc) return rc;
    cudaSetDevice(h->device);
    const Dev& D = h->d; cudaStream_t st = h->stream; const unsigned int gSM = h->sms;
    DayCtx& c = h->h_ctx[0]; std::memset(&c, 0, sizeof(c));
    fill_ctx(h, c, sim_day, julian_day, is_weekday);
    CK(cudaEventRecord(h->e0, st));
    if (part == 0) {
        CK(cudaMemcpyAsync(h->d_ctx, h->h_ctx, sizeof(DayCtx), cudaMemcpyHostToDevice, st));
        CK(cudaMemsetAsync(&h->d.ctr->day_idx, 0, sizeof(unsigned int), st));
        k_scan<<<scan_grid(D), TPB, 0, st>>>(D);
        k_active<<<gSM * h->gm[0], TPB, 0, st>>>(D);
        CK(cudaMemcpyAsync(&h->n_flags_own, &D.ctr->n_flags, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
        h->launches += 2;
    } else if (part == 1) {
        k_visit<<<visit_grid(h, D), TPB, visit_smem_bytes(h->bitmap_words), st>>>(D, h->bitmap_words);
        h->launches += 1;
    } else if (part == 2) {
        if (h->n_hosp) k_patients<<<1, 256, 0, st>>>(D, h->d_hosp_ids, h->n_hosp, h->d_pat_save, h->d_pat_buf, 0);
        k_hazard<<<gSM * h->gm[1], TPB, 0, st>>>(D);
        k_trials<<<gSM * h->gm[2], TPB, 0, st>>>(D);
        k_gate<<<gSM * h->gm[3], 128, 0, st>>>(D);
        k_apply_status<<<gSM * h->gm[4], 128, 0, st>>>(D, 0, 0);
        if (h->n_hosp) k_patients<<<1, 256, 0, st>>>(D, h->d_hosp_ids, h->n_hosp, h->d_pat_save, h->d_pat_buf, 1);
        h->launches += 6;
    } else {
        c.n_imports = (strain_w && expected_imports > 0) ? import_count(h->hpar.seed, sim_day, expected_imports) : 0;
        if (strain_w) for (int s = 0; s < CABM_N_STRAINS; ++s) c.strain_w[s] = strain_w[s];
        CK(cudaMemcpyAsync(h->d_today, h->h_ctx, sizeof(DayCtx), cudaMemcpyHostToDevice, st));
        if (h->n_hosp) k_patients<<<1, 256, 0, st>>>(D, h->d_hosp_ids, h->n_hosp, h->d_pat_save, h->d_pat_buf, 2);
        k_hosp_import<<<gSM * h->gm[5], TPB, 0, st>>>(D, 1, strain_w ? 1 : 0);
        k_apply_status<<<gSM * h->gm[7], 128, 0, st>>>(D, 1, 1);
        h->launches += 3;
    }
    if ((rc = timed_sync(h))) return rc;
    if (part == 3 && out_imported) *out_imported = (int32_t) h->d.host_status[1];   // new infections of the hospital pass and the importation on this handle (written by close_day)
    return part == 3 ? check_overflow(h) : CABM_OK;
}

int cabm_shard_flags_out(cabm_t* h, void** records, size_t* n) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !records || !n || !h->have_pop) return CABM_E_ARG;
    *records = h->d.flags_raw; *n = std::min<size_t>(h->n_flags_own, h->d.cap_flags);
    return CABM_OK;
}

int cabm_shard_flags_in(cabm_t* h, const void* records, size_t n) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (h) h->row_ok = false;                       // the ledger row in pinned memory may no longer be the device's after this call
    if (!h || !h->have_pop || (!records && n)) return CABM_E_ARG;
    if (n == 0) return CABM_OK;
    cudaSetDevice(h->device);
    k_flags_in<<<grid_for(h, n, 4), TPB, 0, h->stream>>>(h->d, (const FlagRaw*) records, (unsigned int) n);
    h->launches += 1;
    CK(cudaStreamSynchronize(h->stream));                        // the caller may reuse `records` after return
    return CABM_OK;
}

int cabm_shard_census(cabm_t* h, void** buf, size_t* n_u32) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !buf || !n_u32 || !h->have_pop) return CABM_E_ARG;
    *buf = h->d.nvis; *n_u32 = h->d.Q;
    return CABM_OK;
}

int cabm_shard_patients(cabm_t* h, void** buf, size_t* n_i32) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !buf || !n_i32 || !h->have_pop) return CABM_E_ARG;
    *buf = h->d_pat_buf; *n_i32 = h->n_hosp;
    return CABM_OK;
}

int cabm_shard_nccl_id(void* out_id) {
    if (!out_id || !nccl_api().ok) return CABM_E_STATE;
    ncclUniqueId id;
    if (nccl_api().GetUniqueId(&id) != ncclSuccess) return CABM_E_CUDA;
    std::memcpy(out_id, &id, CABM_NCCL_ID_BYTES);
    return CABM_OK;
}

int cabm_shard_nccl_init(cabm_t* h, const void* id, int rank, int world) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !id || world < 1 || world > CABM_MAX_SHARDS || rank < 0 || rank >= world) return CABM_E_ARG;
    if (!h->have_pop) { h->err = "load the population and call cabm_shard_config first"; return CABM_E_STATE; }
    if (!nccl_api().ok) { h->err = "libnccl.so.2 not found"; return CABM_E_STATE; }
    cudaSetDevice(h->device);
    ncclUniqueId uid; std::memcpy(&uid, id, CABM_NCCL_ID_BYTES);
    ncclComm_t comm = nullptr;
    NK(nccl_api().CommInitRank(&comm, world, uid, rank));
    h->nccl_comm = comm; h->world = world;
    int rc;
    if (!h->d_xcounts && (rc = dalloc(h, &h->d_xcounts, CABM_MAX_SHARDS + 16))) return rc;
    if (!h->h_xcounts) CK(cudaMallocHost(&h->h_xcounts, 16 * sizeof(unsigned int)));
    // Peer mapping of the flag lists: every rank publishes the IPC handle of its list, opens the others', and the job uses
    // the peer path only if every rank could map every peer (one all-reduce of a yes/no word).
    h->peer_mode = 0;
    const char* env = getenv("CABM_SHARD_PEER");
    if (world > 1 && !(env && atoi(env) == 0)) {
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
        unsigned char* d_hd = nullptr; std::vector<unsigned char> hd((size_t) 64 * world);
        cudaIpcMemHandle_t mine; int ok = cudaIpcGetMemHandle(&mine, h->d.flags_raw) == cudaSuccess ? 1 : 0;
        if (!ok) { cudaGetLastError(); std::memset(&mine, 0, sizeof(mine)); }
        CK(cudaMalloc(&d_hd, hd.size() + 64));
        CK(cudaMemcpyAsync(d_hd + hd.size(), &mine, 64, cudaMemcpyHostToDevice, h->stream));
        NK(nccl_api().AllGather(d_hd + hd.size(), d_hd, 64, ncclUint8, comm, h->stream));
        CK(cudaMemcpyAsync(hd.data(), d_hd, hd.size(), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        for (int r = 0; r < world && ok; ++r) {
            if (r == rank) { h->peer_flags[r] = h->d.flags_raw; continue; }
            cudaIpcMemHandle_t hr; std::memcpy(&hr, hd.data() + (size_t) 64 * r, 64);
            void* ptr = nullptr;
            if (cudaIpcOpenMemHandle(&ptr, hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
            h->peer_flags[r] = (const FlagRaw*) ptr;
        }
        int* d_ok = (int*) d_hd; CK(cudaMemcpyAsync(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice, h->stream));
        NK(nccl_api().AllReduce(d_ok, d_ok, 1, ncclInt32, ncclMin, comm, h->stream));
        CK(cudaMemcpyAsync(&ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        cudaFree(d_hd);
        h->peer_mode = ok;
        if (!ok) for (int r = 0; r < world; ++r) { if (h->peer_flags[r] && r != rank) cudaIpcCloseMemHandle((void*) h->peer_flags[r]); h->peer_flags[r] = nullptr; }
    }
    return CABM_OK;
}

int cabm_shard_exchange_capacity(cabm_t* h, size_t records_per_rank) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || records_per_rank == 0) return CABM_E_ARG;
    h->xcap = std::min<size_t>(records_per_rank, h->d.cap_flags);
    return CABM_OK;
}

int cabm_shard_exchange_mode(cabm_t* h, int* out_peer_memory) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !out_peer_memory) return CABM_E_ARG;
    *out_peer_memory = h->peer_mode;
    return CABM_OK;
}

int cabm_shard_run_days(cabm_t* h, int first_day, int n_days, const int32_t* julian_day, const int32_t* is_weekday,
                        const double* expected_imports, const double* strain_w) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (h) h->row_ok = false;                       // the ledger row in pinned memory may no longer be the device's after this call
    if (!h || n_days <= 0 || !julian_day || !is_weekday || !expected_imports || !strain_w) return CABM_E_ARG;
    if (!h->nccl_comm) { h->err = "cabm_shard_nccl_init first"; return CABM_E_STATE; }
    int rc = ready(h, first_day); if (rc) return rc; if ((rc = ready(h, first_day + n_days - 1))) return rc;
    if (n_days > h->ctx_cap) { h->err = "too many days in one batch"; return CABM_E_ARG; }
    cudaSetDevice(h->device);
    const Dev& D = h->d; cudaStream_t st = h->stream; const unsigned int gSM = h->sms; ncclComm_t comm = (ncclComm_t) h->nccl_comm; NcclApi& N = nccl_api();
    for (int i = 0; i < n_days; ++i) {
        DayCtx& c = h->h_ctx[i]; std::memset(&c, 0, sizeof(c));
        if (julian_day[i] < 1 || julian_day[i] > 366) return CABM_E_ARG;
        fill_ctx(h, c, first_day + i, julian_day[i], is_weekday[i]);
        c.n_imports = expected_imports[i] > 0 ? import_count(h->hpar.seed, first_day + i, expected_imports[i]) : 0;
        for (int s = 0; s < CABM_N_STRAINS; ++s) c.strain_w[s] = strain_w[(size_t) i * CABM_N_STRAINS + s];
    }
    CK(cudaMemcpyAsync(h->d_ctx, h->h_ctx, (size_t) n_days * sizeof(DayCtx), cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(&h->d.ctr->day_idx, 0, sizeof(unsigned int), st));
    // The gather size must be known on the host. Days are enqueued in groups of SHARD_GROUP with no host round trip inside a
    // group; between groups the host reads the largest per-handle flag count and sizes the next gathers at 8x that (floor 8192
    // records, or what cabm_shard_exchange_capacity asked for), so the all-gather moves what the epidemic needs, not N/4 records.
    const int group = h->peer_mode ? (h->profile ? 8 : n_days) : SHARD_GROUP;     // (profiling: the event pool holds eight days of marks)
    PeerFlags peers; for (int r = 0; r < CABM_MAX_SHARDS; ++r) peers.p[r] = h->peer_flags[r];
    for (int g0 = 0; g0 < n_days; g0 += group) {
        size_t want = h->x_seen == UINT_MAX ? std::max<size_t>(65536, (size_t) D.N / (size_t) (4 * h->world)) : std::max<size_t>(8192, (size_t) h->x_seen * 8);
        want = std::min<size_t>(std::max(want, h->xcap), D.cap_flags);
        if (!h->peer_mode && want > h->xcap_alloc) {                                  // receive blocks of the flag all-gather
            CK(cudaStreamSynchronize(st));
            if (h->d_xrecv) cudaFree(h->d_xrecv);
            CK(cudaMalloc(&h->d_xrecv, want * (size_t) h->world * sizeof(FlagRaw)));
            h->xcap_alloc = want;
        }
        const unsigned int cap = (unsigned int) want;
        CK(cudaMemsetAsync(h->d_xcounts + CABM_MAX_SHARDS, 0, sizeof(unsigned int), st));
        CK(cudaEventRecord(h->e0, st));
        for (int i = g0; i < std::min(n_days, g0 + group); ++i) {
            PM(KID_SCAN); k_scan<<<scan_grid(D), TPB, 0, st>>>(D);
            PM(KID_ACTIVE); k_active<<<gSM * h->gm[0], TPB, 0, st>>>(D);
            // exchange A: counts, then the front `cap` records of every handle's flag list
            PM(KID_XA);
            NK(N.AllGather(&D.ctr->n_flags, h->d_xcounts, 1, ncclUint32, comm, st));
            if (h->peer_mode) k_flags_pull<<<gSM * 2, TPB, 0, st>>>(D, peers, h->d_xcounts, h->world, D.shard_rank);
            else {
                NK(N.AllGather(D.flags_raw, h->d_xrecv, (size_t) cap * sizeof(FlagRaw), ncclUint8, comm, st));
                k_flags_in_all<<<gSM * 2, TPB, 0, st>>>(D, h->d_xrecv, h->d_xcounts, cap, h->world, D.shard_rank);
            }
            PM(KID_VISIT); k_visit<<<visit_grid(h, D), TPB, visit_smem_bytes(h->bitmap_words), st>>>(D, h->bitmap_words);
            PM(KID_XB); NK(N.AllReduce(D.nvis, D.nvis, D.Q, ncclUint32, ncclSum, comm, st));                    // exchange B
            if (h->n_hosp) k_patients<<<1, 256, 0, st>>>(D, h->d_hosp_ids, h->n_hosp, h->d_pat_save, h->d_pat_buf, 0);
            PM(KID_HAZARD); k_hazard<<<gSM * h->gm[1], TPB, 0, st>>>(D);
            PM(KID_TRIALS); k_trials<<<gSM * h->gm[2], TPB, 0, st>>>(D);
            PM(KID_GATE0); k_gate<<<gSM * h->gm[3], 128, 0, st>>>(D);
            PM(KID_APPLY_STATUS); k_apply_status<<<gSM * h->gm[4], 128, 0, st>>>(D, 0, 0);
            if (h->n_hosp) {
                PM(KID_XC);
                k_patients<<<1, 256, 0, st>>>(D, h->d_hosp_ids, h->n_hosp, h->d_pat_save, h->d_pat_buf, 1);
                NK(N.AllReduce(h->d_pat_buf, h->d_pat_buf, h->n_hosp, ncclInt32, ncclSum, comm, st));  // exchange C
                k_patients<<<1, 256, 0, st>>>(D, h->d_hosp_ids, h->n_hosp, h->d_pat_save, h->d_pat_buf, 2);
            }
            PM(KID_HOSP_IMPORT); k_hosp_import<<<gSM * h->gm[5], TPB, 0, st>>>(D, 1, 1);
            PM(KID_APPLY1); k_apply_status<<<gSM * h->gm[7], 128, 0, st>>>(D, 1, 3);
            h->launches += 13;
        }
        PM(-1);
        CK(cudaMemcpyAsync(h->h_xcounts, h->d_xcounts + CABM_MAX_SHARDS, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
        if ((rc = timed_sync(h))) return rc;
        h->x_seen = h->h_xcounts[0];
        if (!h->peer_mode && h->x_seen > cap) break;                                  // reported below
    }
    unsigned int ov = 0;
    CK(cudaMemcpy(&ov, &h->d.ctr->overflow, sizeof(ov), cudaMemcpyDeviceToHost));
    if (ov & 512u) { h->err = "flag exchange capacity exceeded: raise it with cabm_shard_exchange_capacity and re-run from a snapshot"; return CABM_E_NOMEM; }
    return check_overflow(h);
}

int cabm_detected_today(cabm_t* h, int sim_day, uint32_t* out_pids, size_t cap, size_t* out_n) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !out_n) return CABM_E_ARG; int rc = ready(h, sim_day); if (rc) return rc;
    cudaSetDevice(h->device);
    if (!h->d_det) { if ((rc = dalloc(h, &h->d_det, (size_t) h->d.N + 1))) return rc; }       // [0] = count, then ids
    CK(cudaMemsetAsync(h->d_det, 0, sizeof(unsigned int), h->stream));
    k_detected_today<<<grid_for(h, h->d.N, 8), TPB, 0, h->stream>>>(h->d, sim_day, h->d_det + 1, h->d.N, h->d_det);
    unsigned int n = 0;
    CK(cudaMemcpyAsync(&n, h->d_det, sizeof(n), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *out_n = n;
    if (out_pids && cap && n) {
        std::vector<unsigned int> ids(n);
        CK(cudaMemcpy(ids.data(), h->d_det + 1, (size_t) n * sizeof(unsigned int), cudaMemcpyDeviceToHost));
        std::sort(ids.begin(), ids.end());
        std::memcpy(out_pids, ids.data(), std::min<size_t>(cap, n) * sizeof(unsigned int));
    }
    return CABM_OK;
}

int cabm_diagnostics(cabm_t* h, int sim_day, uint64_t out[CABM_N_DIAG]) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !out) return CABM_E_ARG; int rc = ready(h, sim_day); if (rc) return rc;
    cudaSetDevice(h->device);
    if (!h->d_diag) { int rc2 = dalloc(h, &h->d_diag, CABM_N_DIAG); if (rc2) return rc2; }
    CK(cudaMemsetAsync(h->d_diag, 0, CABM_N_DIAG * sizeof(unsigned long long), h->stream));
    k_diagnostics<<<grid_for(h, h->d.N, 8), 256, 0, h->stream>>>(h->d, sim_day, h->d_diag);
    CK(cudaMemcpyAsync(out, h->d_diag, CABM_N_DIAG * sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return CABM_OK;
}

int cabm_serosurvey(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint64_t* out_seropos) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !out_seropos) return CABM_E_ARG; int rc = ready(h, sim_day); if (rc) return rc;
    if (!pids) n = h->d.N;
    *out_seropos = 0;
    if (n == 0) return CABM_OK;
    if (pids) for (size_t i = 0; i < n; ++i) if (pids[i] >= h->d.N) return CABM_E_ARG;
    cudaSetDevice(h->device);
    if (!h->d_diag) { int rc2 = dalloc(h, &h->d_diag, CABM_N_DIAG); if (rc2) return rc2; }
    unsigned int* d_p = nullptr;
    if (pids) {
        void* buf = nullptr; int rc2 = scratch(h, n * sizeof(unsigned int), &buf); if (rc2) return rc2;
        d_p = (unsigned int*) buf;
        CK(cudaMemcpyAsync(d_p, pids, n * sizeof(unsigned int), cudaMemcpyHostToDevice, h->stream));
    }
    CK(cudaMemsetAsync(h->d_diag, 0, sizeof(unsigned long long), h->stream));
    k_serosurvey<<<grid_for(h, n, 8), TPB, 0, h->stream>>>(h->d, sim_day, d_p, (unsigned int) n, h->d_diag);
    unsigned long long r = 0;
    CK(cudaMemcpyAsync(&r, h->d_diag, sizeof(r), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *out_seropos = r;
    return CABM_OK;
}

int cabm_eligibility(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint8_t* out_flags) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !out_flags) return CABM_E_ARG; int rc = ready(h, sim_day); if (rc) return rc;
    if (!pids) n = h->d.N;
    if (n == 0) return CABM_OK;
    if (pids) for (size_t i = 0; i < n; ++i) if (pids[i] >= h->d.N) return CABM_E_ARG;
    cudaSetDevice(h->device);
    const size_t out_bytes = (n + 15) & ~(size_t) 15;
    void* buf = nullptr; if ((rc = scratch(h, out_bytes + (pids ? n * sizeof(unsigned int) : 0), &buf))) return rc;
    unsigned char* d_o = (unsigned char*) buf; unsigned int* d_p = pids ? (unsigned int*) ((unsigned char*) buf + out_bytes) : nullptr;
    if (pids) CK(cudaMemcpyAsync(d_p, pids, n * sizeof(unsigned int), cudaMemcpyHostToDevice, h->stream));
    k_eligibility<<<grid_for(h, n, 8), TPB, 0, h->stream>>>(h->d, sim_day, d_p, (unsigned int) n, d_o);
    CK(cudaMemcpyAsync(out_flags, d_o, n, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return CABM_OK;
}

int cabm_reduce(cabm_t* h, int what, int sim_day, const uint32_t* pids, size_t n, double* out_count) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !out_count || what < 0 || what >= CABM_N_REDUCE) return CABM_E_ARG; int rc = ready(h, sim_day); if (rc) return rc;
    if (!pids) n = h->d.N;
    *out_count = 0.0;
    if (n == 0) return CABM_OK;
    if (pids) for (size_t i = 0; i < n; ++i) if (pids[i] >= h->d.N) return CABM_E_ARG;
    cudaSetDevice(h->device);
    if (!h->d_diag) { int rc2 = dalloc(h, &h->d_diag, CABM_N_DIAG); if (rc2) return rc2; }
    unsigned int* d_p = nullptr;
    if (pids) {
        void* buf = nullptr; if ((rc = scratch(h, n * sizeof(unsigned int), &buf))) return rc;
        d_p = (unsigned int*) buf;
        CK(cudaMemcpyAsync(d_p, pids, n * sizeof(unsigned int), cudaMemcpyHostToDevice, h->stream));
    }
    CK(cudaMemsetAsync(h->d_diag, 0, sizeof(unsigned long long), h->stream));
    k_reduce<<<grid_for(h, n, 8), TPB, 0, h->stream>>>(h->d, what, sim_day, d_p, (unsigned int) n, h->d_diag);
    unsigned long long r = 0;
    CK(cudaMemcpyAsync(&r, h->d_diag, sizeof(r), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *out_count = (double) r;
    return CABM_OK;
}

// cabm_serosurvey without any allocation (include/cabm.h): ids and accumulator in the file-scope device array, the result
// comes back through the handle's pinned word. No cudaMalloc / cudaFree / allocating library call on this path.
int cabm_serosurvey_static(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint64_t* out_seropos) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !out_seropos) return CABM_E_ARG; int rc = ready(h, sim_day); if (rc) return rc;
    if (!pids) n = h->d.N;
    *out_seropos = 0;
    if (pids && n > (size_t) CABM_STATIC_MAX_IDS) {
        h->err = "cabm_serosurvey_static: " + std::to_string(n) + " ids exceed CABM_STATIC_MAX_IDS (" + std::to_string((size_t) CABM_STATIC_MAX_IDS) +
                 "): split the list, pass NULL for everybody, or call cabm_serosurvey";
        return CABM_E_ARG;
    }
    if (n == 0) return CABM_OK;
    if (pids) for (size_t i = 0; i < n; ++i) if (pids[i] >= h->d.N) return CABM_E_ARG;
    cudaSetDevice(h->device);
    if (!h->static_scratch) { h->err = h->static_scratch_err.empty() ? "static scratch address unknown" : h->static_scratch_err; return CABM_E_CUDA; }
    unsigned char* sc = h->static_scratch;
    // the array keeps the previous call's contents: the accumulator must start at zero, the ids are overwritten
    CK(cudaMemsetAsync(sc, 0, sizeof(unsigned long long), h->stream));
    if (pids) CK(cudaMemcpyAsync(sc + CABM_STATIC_HEAD_BYTES, pids, n * sizeof(unsigned int), cudaMemcpyHostToDevice, h->stream));
    k_serosurvey_static<<<grid_for(h, n, 8), TPB, 0, h->stream>>>(h->d, sim_day, sc, (unsigned int) n, pids ? 1 : 0);
    unsigned long long* pin = reinterpret_cast<unsigned long long*>(h->h_pin + 8);        // pinned at cabm_create; words 8-9 are this call's
    CK(cudaMemcpyAsync(pin, sc, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    *out_seropos = *pin;
    return CABM_OK;
}

int cabm_get_infections(cabm_t* h, const uint32_t* pids, size_t n, cabm_infection* out) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !h->have_pop || !out) return CABM_E_ARG;
    if (n == 0) return CABM_OK;
    if (pids) { for (size_t i = 0; i < n; ++i) if (pids[i] >= h->d.N) return CABM_E_ARG; } else if (n > h->d.N) return CABM_E_ARG;
    cudaSetDevice(h->device);
    const size_t out_bytes = (n * sizeof(cabm_infection) + 15) & ~(size_t) 15;
    void* buf = nullptr; int rc2 = scratch(h, out_bytes + (pids ? n * sizeof(unsigned int) : 0), &buf); if (rc2) return rc2;
    cabm_infection* d_o = (cabm_infection*) buf; unsigned int* d_p = pids ? (unsigned int*) ((unsigned char*) buf + out_bytes) : nullptr;
    if (pids) CK(cudaMemcpyAsync(d_p, pids, n * sizeof(unsigned int), cudaMemcpyHostToDevice, h->stream));
    k_gather_infections<<<grid_for(h, n, 4), TPB, 0, h->stream>>>(h->d, d_p, (unsigned int) n, d_o);
    CK(cudaMemcpyAsync(out, d_o, n * sizeof(cabm_infection), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return CABM_OK;
}

int cabm_snapshot(cabm_t* h, cabm_snap_t** out) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !out || !h->have_pop) return CABM_E_ARG;
    cudaSetDevice(h->device);
    cabm_snap_t* s = new cabm_snap();
    auto fail = [&](cudaError_t e, const char* what) { h->err = std::string(what) + ": " + cudaGetErrorString(e); cabm_snap_free(s); return CABM_E_CUDA; };
    cudaError_t e;
    for (auto& b : h->dyn) {
        void* p = nullptr;
        if ((e = cudaMalloc(&p, b.second)) != cudaSuccess) return fail(e, "snapshot: cudaMalloc");
        s->bufs.push_back({ p, b.second });
        if ((e = cudaMemcpyAsync(p, b.first, b.second, cudaMemcpyDeviceToDevice, h->stream)) != cudaSuccess) return fail(e, "snapshot: copy");
    }
    { const int rc = log_count(h, &s->n_log); if (rc) { cabm_snap_free(s); return rc; } }
    if (s->n_log) {
        if ((e = cudaMalloc(&s->log, s->n_log * sizeof(cabm_infection))) != cudaSuccess) return fail(e, "snapshot: cudaMalloc(log)");
        if ((e = cudaMemcpyAsync(s->log, h->d.log, s->n_log * sizeof(cabm_infection), cudaMemcpyDeviceToDevice, h->stream)) != cudaSuccess) return fail(e, "snapshot: copy(log)");
    }
    if ((e = cudaStreamSynchronize(h->stream)) != cudaSuccess) return fail(e, "snapshot: sync");
    *out = s;
    return CABM_OK;
}
int cabm_restore(cabm_t* h, const cabm_snap_t* s) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (h) h->row_ok = false;                       // the ledger row in pinned memory may no longer be the device's after this call
    if (!h || !s || s->bufs.size() != h->dyn.size()) return CABM_E_ARG;
    cudaSetDevice(h->device);
    for (size_t i = 0; i < s->bufs.size(); ++i) CK(cudaMemcpyAsync(h->dyn[i].first, s->bufs[i].first, s->bufs[i].second, cudaMemcpyDeviceToDevice, h->stream));
    if (s->n_log) CK(cudaMemcpyAsync(h->d.log, s->log, s->n_log * sizeof(cabm_infection), cudaMemcpyDeviceToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    h->x_seen = UINT_MAX;                   // flag counts of the restored day are unknown to the host
    return CABM_OK;
}
int cabm_snap_free(cabm_snap_t* s) {
    if (!s) return CABM_OK;
    for (auto& b : s->bufs) cudaFree(b.first);
    if (s->log) cudaFree(s->log);
    delete s;
    return CABM_OK;
}

int cabm_dump_foi(cabm_t* h, cabm_foi_row* out, size_t cap, size_t* out_n) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h || !h->have_pop) return CABM_E_ARG;
    cudaSetDevice(h->device);
    if (!h->d.foi_enabled) {       // first call switches the (otherwise skipped) row writes on; rows exist from the next tick
        h->d.foi_enabled = 1; drop_graphs(h);
        if (out_n) *out_n = 0; return CABM_OK;
    }
    unsigned int n = 0;
    CK(cudaMemcpy(&n, h->d.n_foi, sizeof(n), cudaMemcpyDeviceToHost));
    if (out_n) *out_n = n;
    if (out && cap) CK(cudaMemcpy(out, h->d.foi, std::min<size_t>(cap, n) * sizeof(cabm_foi_row), cudaMemcpyDeviceToHost));
    return CABM_OK;
}

int cabm_device_time_ms(cabm_t* h, double* out_ms, uint64_t* out_launches) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h) return CABM_E_ARG;
    if (out_ms) *out_ms = h->acc_ms; if (out_launches) *out_launches = h->launches;
    h->acc_ms = 0.0; h->launches = 0;
    return CABM_OK;
}

// per-kernel device time of the day pipeline; switching it on makes tick/import/run_days use plain stream launches
int cabm_set_profile(cabm_t* h, int on) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h) return CABM_E_ARG;
    cudaSetDevice(h->device);
    if (on && h->marks.empty()) { h->marks.resize(160); h->mark_id.resize(160); for (auto& e : h->marks) cudaEventCreate(&e); }
    h->profile = on ? 1 : 0; h->n_marks = 0;
    return CABM_OK;
}
int cabm_kernel_times(cabm_t* h, double* out_ms, uint64_t* out_count, const char** out_names, size_t cap, size_t* out_n) {
    { const int rc_settle = settle(h); if (rc_settle) return rc_settle; }
    if (!h) return CABM_E_ARG;
    if (out_n) *out_n = KID_N;
    for (size_t k = 0; k < std::min<size_t>(cap, KID_N); ++k) {
        if (out_ms) out_ms[k] = h->ktime[k]; if (out_count) out_count[k] = h->kcount[k]; if (out_names) out_names[k] = KID_NAMES[k];
        h->ktime[k] = 0; h->kcount[k] = 0;
    }
    return CABM_OK;
}

} // extern "C"
