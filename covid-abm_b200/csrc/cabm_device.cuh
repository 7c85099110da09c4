// covid-abm_b200/csrc/cabm_device.cuh -- device-side state layout, Philox stream and the
// per-exposure state-transition sampler (Person::infect, Person.cpp:320-494) for sm_100a.
//
// Layout (all structure-of-arrays in HBM, one element per agent / location):
//   agents   : ag (one 64-byte record: latest Infection, home, day location, hospital, status byte, patron row -- what the
//              passes over the active agents read, in one line), imm (infections per strain),
//              vax (doses, last dose day, naive flag), quar,
//              cand (64-bit winner key of today's exposure resolution), prow (20-byte packed patron row, streamed by k_visit),
//              inhosp (one bit per agent: in hospital today), actbits (one bit per agent: inside an infection window)
//   locations: locd (one 64-byte record: press[4] = u64 fixed-point infectious pressure by strain in 2^-32 units, flags today,
//              hot-list index), nvis, pat_count; static: locs (one 32-byte record) + type/flags/risk/hospital arrays,
//              roster CSR, neighbour CSR
// Day-scoped lists (hot locations, active agents, visits landing on hot locations, exposure work
// items, newly infected agents) are appended with atomics and consumed by grid-stride kernels, so
// a whole day runs without returning to the host.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/cabm.h"
#include "../../include/cabm_det_math.h"

#define CABM_I16_MAX 32767
#define CABM_I16_MIN (-32768)
#define CABM_CAND_EMPTY 0xFFFFFFFFFFFFFFFFull

// exposure passes in tick() order (Community.cpp:1151-1171); the winner key sorts by (pass, source, role)
enum { PASS_HH = 3, PASS_NB = 4, PASS_WORK = 5, PASS_SCHOOL = 7, PASS_NH = 8, PASS_HOSP = 9, PASS_IMPORT = 11, PASS_SEED = 12 };

struct __align__(16) Tl0 { short infected, infectious_begin, infectious_end, symptom_begin, symptom_end, severe_begin, severe_end, hosp_begin; };
struct __align__(16) Tl1 { short critical_begin, critical_end, icu_begin, death, report_day; unsigned char strain, detected; float rel_inf; };

// One 64-byte record per agent: everything the passes over the ACTIVE agents need (k_active, the status pass, the hospital
// pass), so that an active agent costs one line of HBM instead of a sector from each of six arrays. Scattered 32-byte reads
// run at 40-90 G sectors/s on this GPU (profiles/r2_ubench_gather_rate.txt: 2 M agents out of 21.5 M, six arrays 131 us, one
// 64-byte record 57 us) and the heavy-epidemic day at 21.5 M agents was bound by exactly that rate. The static members are
// copies (home, day_loc and the 20-byte patron rows stay as arrays for the kernels that stream them).
struct __align__(64) AgRec {
    Tl0 t0; Tl1 t1;                                           // latest Infection (dynamic)
    int home, day_loc, hospital;                              // static; hospital = loc_hospital[home]
    unsigned char pbyte, hosp_count; unsigned short pad;      // dynamic: patron count | flags (PB_*), admissions not yet discharged
    uint4 prow;                                               // static: the patron row in 16 bytes (prow16_*)
};
static_assert(sizeof(AgRec) == 64, "AgRec is one 64-byte line");

// Locations the same way. LocS: the static words the flagging, hazard and reset passes need about a location, one 32-byte
// sector (the separate arrays stay for the passes that stream or rarely touch them). LocD: a location's day state -- pressure by
// strain, flags today (0 = not hot), index in the hot list -- one 64-byte line that a flag reads and writes once.
struct __align__(32) LocS {
    unsigned int roster_begin, roster_size, nbr_begin, nbr_deg;
    float risk; int pub; int hospital;                       // pub: index among the patronised locations or -1
    unsigned char type, flags; unsigned short pad;
};
static_assert(sizeof(LocS) == 32, "LocS is one sector");
struct __align__(64) LocD {
    unsigned long long press[CABM_N_STRAINS];                // u64 fixed point, 2^-32 units
    unsigned int cnt; int hot; unsigned int pad[6];
};
static_assert(sizeof(LocD) == 64 && CABM_N_STRAINS == 4, "LocD is one 64-byte line");
__device__ __forceinline__ void locd_reset(LocD* d) {           // updateHotLocations for one location: 48 bytes, three stores
    uint4* q = reinterpret_cast<uint4*>(d);
    q[0] = make_uint4(0u, 0u, 0u, 0u); q[1] = make_uint4(0u, 0u, 0u, 0u); q[2] = make_uint4(0u, 0u, 0u, 0u);
}

struct __align__(16) Item { int src; int begin; int count; int purpose; double T; double pad; };   // a 256-member chunk of a big roster
struct __align__(16) SItem { int src; int begin; unsigned int count_purpose; unsigned int t_idx; };   // a small roster: count << 8 | purpose, T = hot_T[t_idx]
struct __align__(16) Hit  { unsigned int pid; unsigned int src; unsigned int purpose; unsigned int sub; }; // a trial with U < T, to be gated
// one (infectious agent, location) flag of the day (Community::flagInfectedLocation, Community.cpp:788-791);
// c = pressure contribution in 2^-32 units with the strain in the top byte. k_active appends FlagRaw records;
// k_trials scatters them into one contiguous segment per hot location (Flag), which k_apply_status searches.
struct __align__(8)  FlagRaw { int loc; unsigned int slot; unsigned int pid; unsigned int pad; unsigned long long c; };
struct __align__(16) Flag { unsigned int pid; unsigned int pad; unsigned long long c; };
#define HIT_LISTS 64u
#define HOT_NOPUB 0x7FFFFFFFu
#define HOT_IS_HOSP 0x80000000u
#define FLAG_C(f) ((f).c & 0x00FFFFFFFFFFFFFFull)
#define FLAG_STRAIN(f) ((int) ((f).c >> 56))

struct DayCtx {                     // per-day scalars, staged by the host for every day of a run_days() batch
    int day, julian_day, is_weekday, n_imports;
    double strain_w[CABM_N_STRAINS];
    double season, sd, school_closure, business_closure;
    float sd_up;                    // smallest float >= sd: for a float r, (double) r < sd  <=>  r < sd_up
    unsigned int sd_rank;           // number of distinct household riskiness values below sd: a household keeps away from
                                    // high-risk places today (Community.cpp:1020) <=> its riskiness rank < sd_rank
};

// counters living in device memory
struct Counters {
    // ---- day-scoped: zeroed by the block that closes the day (k_apply_status, phase 1) ----
    unsigned int n_active, n_hot, n_hot_hosp, n_flags, n_vis, n_items, n_seg, n_sitems;  // n_active: agents inside an infection window today (length of active_list)
    unsigned int n_new[2];                                   // [phase]: 0 = house..nursing home passes, 1 = hospital pass + importation
    unsigned int n_hits_unused;                              // (the hit list has its own counters: Dev::hit_ctr)
    unsigned int n_imported;                                 // winners of the day whose exposure was an importation draw (cabm_import reports them)
    unsigned int day_pad[4];
    // ---- persistent ----
    unsigned int n_pat;                                      // admitted-patient list length (entries are never removed, see k_status)
    unsigned int overflow;                                   // sticky: a list ran out of capacity
    unsigned int day_idx;                                    // index into the staged DayCtx array
    unsigned int seed_calls;
    unsigned int end_ticket;                                 // blocks of the closing kernel that have finished (the last one closes the day)
    unsigned int pad2[5];
    unsigned long long n_log;
    unsigned long long work[CABM_N_WORK];
    unsigned long long cumul[CABM_N_OUTCOMES];
};
#define CABM_DAY_SCOPED_WORDS 16

struct Dev {
    // sizes
    unsigned int N, L, Q, n_patron; int R;
    // one community over several handles (cabm_shard_config): this handle owns agents [own_lo, own_hi); one handle: [0, N)
    unsigned int own_lo, own_hi; int shard_rank;
    // static agents
    const int* home; const int* day_loc; const unsigned int* prow; const unsigned char* age; const unsigned char* pflags;
    const double* nat_eff; const float* xprot; const int* nat_dur;
    // static locations
    const unsigned char* loc_type; const unsigned char* loc_flags; const float* loc_risk; const int* loc_hospital;
    const unsigned int* roster_off; const int* roster; const unsigned int* nbr_off; const int* nbr;
    const int* loc_pub;                  // location -> index among patronised ("public") locations, or -1
    const int* pub_loc;                  // [Q] inverse
    const LocS* locs;                    // the static words of a location in one sector (copies of the arrays above)
    // dynamic agents
    AgRec* ag; unsigned int* imm; unsigned int* vax; short* vax_days; unsigned int* quar;
    unsigned long long* cand; int* inf_place; int* inf_by; unsigned int* n_caused;
    unsigned int* inhosp;                // [ceil(N / 128) * 4] one bit per agent: in hospital today (k_active keeps it, k_visit reads it)
    // The active set: one bit per agent, set when an infection is committed (infect_event), cleared by k_active on the last day
    // of the agent's activity window. k_active scans it every day (N / 8 bytes: 0.34 MB at 2.7 M agents, 2.7 MB at 21.5 M, i.e.
    // L2-resident; the per-agent window words a scan used to read were 32 times that) and so meets the active agents in id
    // order: their state gathers walk memory in one direction. (A persistent list kept in infection order made every gather of
    // k_active and of the status pass random over a multi-GB footprint: +130 and +107 us per day at 21.5 M agents.)
    unsigned int* actbits;               // [ceil(N / 1024) * 32]
    unsigned int* active_list;           // today's active agents, written by k_active in chunks of ascending ids, read by the status pass
    // dynamic locations
    LocD* locd; unsigned int* pat_count;                          // locd: pressure, flags today, index in hot_list (day-scoped)
    unsigned int* nvis; double* pub_T; unsigned int* pub_hot;     // [Q], [Q], [ceil(Q/32)] public-location visitor census, T, hot bitmap
    // day lists
    int* hot_list; int* hot_hosp; unsigned int* hot_off; double* hot_T; FlagRaw* flags_raw; Flag* flags; SItem* sitems;
    // Successful trials waiting for the gate: HIT_LISTS sub-lists of hit_sub records each, one counter per sub-list on its own
    // 128-byte line (hit_ctr[k * 32]); a block of k_trials appends to sub-list blockIdx.x % HIT_LISTS (in a grid of fewer blocks than lists: by warp). One list with one counter
    // took a quarter of k_trials' stall samples in a large epidemic (a quarter of a million returning atomics on one address a day).
    unsigned int* hit_ctr; unsigned int hit_sub;
    unsigned int* hot_meta;              // [i]: public index of hot_list[i] (HOT_NOPUB if none) | HOT_IS_HOSP, for the end-of-day reset
    uint2* visit_rec; Item* items; Hit* hits; unsigned int* newinf_list; unsigned int* patient_list;
    unsigned int cap_hot, cap_active, cap_flags, cap_vis, cap_items, cap_sitems, cap_hits, cap_new, cap_pat; unsigned long long cap_log;
    // outputs
    unsigned long long* ledger;          // [CABM_N_COUNTERS][R]
    cabm_infection* log; cabm_foi_row* foi; unsigned int* n_foi; int foi_enabled;
    Counters* ctr; const DayCtx* dayctx; DayCtx* today;
    unsigned int* host_status;           // pinned host memory: [0] overflow mask, [1] n_new[1] of the day, stored by the block that closes the day
    // parameters
    const cabm_params* par; const double* pfd; const double* imr; int n_imr;
    int any_quarantine;
    unsigned int visit_thr[4];           // ceil(visit_cdf[k] * 2^32), k < 4 (Poisson(1) draw of public_activity in integer form)
    unsigned long long seed;
};

// pflags bits
#define PF_COMORBID 1
#define PF_LTC 2
// pbyte: low 3 bits patron count, bit 7 in hospital today, bit 5 on the admitted-patient list
#define PB_NPAT 7
#define PB_ONLIST 0x20
#define PB_INHOSP 0x80
// loc_flags: bit0 essential, bits 1-2 public transmission risk
#define LF_ESSENTIAL 1
#define LF_PUBRISK(f) (((f) >> 1) & 3)

// The patron row of an agent (Person::patronizedLocations, Person.h:357, plus what public_activity needs to know about the
// agent's household, Community.cpp:1020): five 32-bit words, 20 bytes, so a uniform pick among the agent's places is ONE
// 32-bit load at word j of the row (rows are staged in shared memory by k_visit; stride 5 words = conflict-free banks).
//   word k, bits 0-19 : entry k = index q of the k-th patronised location among the patronised ("public") locations
//                       (19 bits) | HIGH-risk bit 19
//   word 0, bits 20-22: npat = number of patronised places (0..5)
//   words 1-2, bits 20-31: rank (low, high 12 bits) of the household's riskiness among the distinct household riskiness
//                       values of the population, compared with DayCtx::sd_rank
#define PROW_WORDS 5
#define PROW_ENTRY_MASK 0xFFFFFu
#define PROW_Q_MASK 0x7FFFFu
#define PROW_HIGH_BIT 19
#define PROW_MAX_Q (1u << 19)
#define PROW_MAX_RANK (1u << 24)
struct PRow { unsigned int w[PROW_WORDS]; };
__host__ __device__ __forceinline__ unsigned int prow_npat(const unsigned int w0) { return w0 >> 20; }
__host__ __device__ __forceinline__ unsigned int prow_rank(const unsigned int w1, const unsigned int w2) { return (w1 >> 20) | ((w2 >> 20) << 12); }
__host__ __device__ __forceinline__ unsigned int prow_entry(const PRow& r, const unsigned int j) {
    return (j == 0 ? r.w[0] : j == 1 ? r.w[1] : j == 2 ? r.w[2] : j == 3 ? r.w[3] : r.w[4]) & PROW_ENTRY_MASK;
}
__host__ __forceinline__ PRow prow_pack(const unsigned int e[5], const unsigned int npat, const unsigned int rank) {
    PRow r;
    r.w[0] = e[0] | (npat << 20);
    r.w[1] = e[1] | ((rank & 0xFFFu) << 20);
    r.w[2] = e[2] | ((rank >> 12) << 20);
    r.w[3] = e[3];
    r.w[4] = e[4];
    return r;
}
__device__ __forceinline__ PRow prow_load(const unsigned int* __restrict__ rows, const unsigned int p) {
    PRow r; const unsigned int* q = rows + (size_t) p * PROW_WORDS;
#pragma unroll
    for (int k = 0; k < PROW_WORDS; ++k) r.w[k] = q[k];
    return r;
}

// The same row in 16 bytes for the agent record (k_active draws a handful of rows a day, so the unpacking may cost a few
// instructions more): a 128-bit little-endian word with entry k at bits 20 k .. 20 k + 19, the riskiness rank at bits 100-123
// and npat at bits 124-126.
__host__ __forceinline__ uint4 prow16_pack(const unsigned int e[5], const unsigned int npat, const unsigned int rank) {
    uint4 r;
    r.x = e[0] | (e[1] << 20);
    r.y = (e[1] >> 12) | (e[2] << 8) | (e[3] << 28);
    r.z = (e[3] >> 4) | (e[4] << 16);
    r.w = (e[4] >> 16) | (rank << 4) | (npat << 28);
    return r;
}
__host__ __device__ __forceinline__ unsigned int prow16_npat(const uint4& r) { return (r.w >> 28) & 7u; }
__host__ __device__ __forceinline__ unsigned int prow16_rank(const uint4& r) { return (r.w >> 4) & 0xFFFFFFu; }
__host__ __device__ __forceinline__ unsigned int prow16_entry(const uint4& r, const unsigned int j) {
    const unsigned int v = j == 0 ? r.x : j == 1 ? ((r.x >> 20) | (r.y << 12)) : j == 2 ? (r.y >> 8) : j == 3 ? ((r.y >> 28) | (r.z << 4)) : ((r.z >> 16) | (r.w << 16));
    return v & PROW_ENTRY_MASK;
}
// the whole record with four 16-byte loads of one line
__device__ __forceinline__ AgRec ag_load(const AgRec* ag, const unsigned int p) {     // (no __restrict__: the record is written in the same kernels)
    const uint4* q = reinterpret_cast<const uint4*>(ag + p);
    union { AgRec r; uint4 v[4]; } u;
    u.v[0] = q[0]; u.v[1] = q[1]; u.v[2] = q[2]; u.v[3] = q[3];
    return u.r;
}

// Programmatic dependent launch (sm_90+): every kernel of the day pipeline lets its successor start launching when its own
// main loop is over (not earlier: successors released at the top of a kernel sat resident on the SMs, waiting, and the day got
// slower than without the attribute, 152 vs 142 us at 2.7 M agents), and waits for its predecessor's memory before it touches
// anything the predecessor wrote: launch latency and prologues overlap the predecessor's tail. Both are no-ops when the
// kernel was launched without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------
// Philox4x32-10, counter = (sub<<16 | block, a, b, day<<8 | purpose), key = seed
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        // separate high and low products: written as one 64-bit multiply the trial kernel ran 45 % slower (64.5 vs 44.4 us per day
        // at 2.7 M agents, day 250)
        const unsigned int hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const unsigned int hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}

// Out-of-line copy for the event streams (Person::infect draws ~25 numbers at as many call sites): the kernels that walk a
// disease course are latency-bound on one warp per SM and half of their stalls were instruction fetch (k_apply_status was
// 106 KB of code with Philox inlined at every draw). The streaming kernels (k_visit, k_trials) keep the inline version.
__device__ __noinline__ uint4 philox4x32_10_call(uint4 c, uint2 k) { return philox4x32_10(c, k); }

struct Stream {
    uint2 key; unsigned int a, b, dp, sub, idx; uint4 buf;
    __device__ __forceinline__ void begin(unsigned long long seed, unsigned int day, unsigned int purpose, unsigned int a_, unsigned int b_, unsigned int sub_ = 0) {
        key = make_uint2((unsigned int) seed, (unsigned int) (seed >> 32)); a = a_; b = b_; dp = (day << 8) | (purpose & 0xFFu); sub = sub_; idx = 0;
    }
    __device__ __forceinline__ unsigned int next() {
        const unsigned int lane = idx & 3u;
        if (lane == 0) buf = philox4x32_10_call(make_uint4((sub << 16) | (idx >> 2), a, b, dp), key);
        idx++;
        return lane == 0 ? buf.x : lane == 1 ? buf.y : lane == 2 ? buf.z : buf.w;
    }
    __device__ __forceinline__ double u01() { return (double) next() * (1.0 / 4294967296.0); }
};

// Inverse-CDF draw on an integer-support table: the smallest k with u < cdf[k], at most n - 1. Four entries per step: their
// addresses do not depend on one another, so a step is one round trip instead of four. These draws sit in the disease-course
// sampler (Person::infect, ~10 tables per infection), which runs on a handful of lanes per SM at the end of the day: its
// duration is the sum of its dependent loads. Out of line: the sampler's code size matters more than the call.
__device__ __noinline__ unsigned int table_draw(const double* __restrict__ cdf, int n, double u) {
    int k = 0;
    while (k + 4 <= n - 1) {
        const double a = cdf[k], b = cdf[k + 1], c = cdf[k + 2], d = cdf[k + 3];
        if (u < a) return (unsigned int) k;
        if (u < b) return (unsigned int) (k + 1);
        if (u < c) return (unsigned int) (k + 2);
        if (u < d) return (unsigned int) (k + 3);
        k += 4;
    }
#pragma unroll 1
    while (k < n - 1 && !(u < cdf[k])) ++k;
    return (unsigned int) k;
}
// Parameters::sampler (Parameters.h:433-436): the smallest i with !(cdf[i] < r), at most n
__device__ __noinline__ int ref_sampler(const double* __restrict__ cdf, int n, double r) {
    int i = 0;
    while (i + 4 <= n) {
        const double a = cdf[i], b = cdf[i + 1], c = cdf[i + 2], d = cdf[i + 3];
        if (!(a < r)) return i;
        if (!(b < r)) return i + 1;
        if (!(c < r)) return i + 2;
        if (!(d < r)) return i + 3;
        i += 4;
    }
#pragma unroll 1
    while (i < n && cdf[i] < r) ++i;
    return i;
}

// strain of the traced infector (Community.cpp:795-825, :982): the draw x picks a point t in [0, total pressure) and the
// strain whose share contains it; *t_res is the offset inside that strain's share (used by pick_infector)
__device__ __forceinline__ int pick_strain(const unsigned long long* __restrict__ w, unsigned int x, unsigned long long* t_res = nullptr) {
    const unsigned long long w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3];
    const unsigned long long total = w0 + w1 + w2 + w3;
    unsigned long long t = (unsigned long long) x * (total >> 32) + (((unsigned long long) x * (total & 0xffffffffull)) >> 32);
    int s = 3;
    if (t < w0) s = 0; else { t -= w0; if (t < w1) s = 1; else { t -= w1; if (t < w2) s = 2; else t -= w2; } }
    if (t_res) *t_res = t;
    return s;
}

// Parameters::remainingEfficacy (Parameters.h:611-625)
__device__ __noinline__ double remaining_efficacy(const cabm_params* __restrict__ par, double start, double time_delta) {
    if (!par->immunity_wanes) return start;
    const double slope = 0.2, intcpt = 2.4603816;
    const double month_delta = CABM_DIV(time_delta, 30.0);
    const double est = CABM_DIV(CABM_ADD(cabm_log(CABM_SUB(CABM_DIV(1.0, start), 1.0)), intcpt), slope);
    const double l = CABM_SUB(intcpt, CABM_MUL(slope, CABM_ADD(est, month_delta)));
    return cabm_logistic(l);
}

__device__ __forceinline__ bool tl_in(short b, short e, int now) { return b <= now && now < e; }
// last day on which an agent with this infection record needs per-day work (running infection, pending report): the agent's
// activity window is [infected, tl_until]; win[] holds the same two numbers (written by infect_event)
__device__ __forceinline__ int tl_until(const Tl0& t0, const Tl1& t1) {
    int until = max(max((int) t0.infectious_end, (int) t0.symptom_end), max((int) t0.severe_end, (int) t1.critical_end));
    if (t1.death != CABM_I16_MAX) until = max(until, (int) t1.death);
    if (t1.report_day != CABM_I16_MAX) until = max(until, (int) t1.report_day);
    return min(until, CABM_I16_MAX - 1);
}

// Checked build (nvcc -DCABM_CHECKS, covid-abm_b200/csrc/libcabm_checked.so): every index the pipeline computes on the device
// is range-checked before it is used; a violation sets bit 31 of the sticky overflow word, which fails the running call with
// the mask in the error text. Stands in for compute-sanitizer memcheck, which is closed on this GPU pool. No code in the
// product build.
#ifdef CABM_CHECKS
#define CABM_CHECK(D, cond) do { if (!(cond)) atomicOr(&(D).ctr->overflow, 0x80000000u); } while (0)
#else
#define CABM_CHECK(D, cond) do { } while (0)
#endif

__device__ __forceinline__ unsigned long long make_key(int pass, unsigned int src, unsigned int role) {
    return ((unsigned long long) pass << 40) | ((unsigned long long) src << 8) | (unsigned long long) (role & 0xFFu);
}

__device__ __forceinline__ void ledger_add(const Dev& D, int counter, long day) {
    if (day >= 0 && day < (long) D.R) atomicAdd(&D.ledger[(size_t) counter * D.R + day], 1ull);
}
