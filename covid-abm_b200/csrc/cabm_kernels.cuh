// covid-abm_b200/csrc/cabm_kernels.cuh -- first half of the day pipeline (sm_100a): who is infectious where,
// public activity, per-location hazard, exposure trials.
//
//   k_scan     active bitmap   : one bit per agent (set when an infection is committed, cleared on the last day of its window;
//                                N / 8 bytes, L2-resident) -> today's active agents as a list in id order
//   k_active   active agents   : one thread per list entry. Person.cpp:475-486 derived from the stored timeline instead of the
//                                reference's _isHot calendar: pressure at home / day location / hospital / visited locations
//                                (Community.cpp:1031-1035), per-location flag chains, hot lists, in-hospital bit
//   k_visit    all agents      : public_activity (Community.cpp:1006-1039): the 16-byte patron rows of a warp's tile arrive in
//                                shared memory by bulk copy (TMA, double-buffered per warp), Poisson(1) visits from one Philox
//                                block (branch-free body), hot test against a shared-memory bitmap of patronised locations,
//                                visitor census, visit records staged per lane; k >= 4 draws on full warps at the end
//   k_hazard   hot locations   : _tally_infectiousness + the hazard of every pass (Community.cpp:833-969); one thread
//                                per location turns pressure into T and emits work items: small rosters and
//                                neighbour households (resolved by k_trials), 256-member chunks of big rosters
//   k_trials   items + visits  : _transmission's `U < T` (Community.cpp:976) for rosters and visitors; also scatters
//                                the day's flag records into one contiguous segment per hot location. Successful trials
//                                go to the hit list; the susceptibility gate runs in k_gate (cabm_kernels2.cuh), so this
//                                streaming kernel stays at 32 registers
//
// All kernels read their sizes from device-resident counters, so a whole day is one CUDA-graph replay; every kernel
// releases its successor at once (programmatic dependent launch) and waits for its predecessor before the first dependent
// access.
#pragma once
#include <cooperative_groups.h>
#include "cabm_infect.cuh"

namespace cg = cooperative_groups;

#define TPB 256
#define CHUNK 256
#define SMALL_ROSTER 64

__device__ __forceinline__ const DayCtx& day_ctx(const Dev& D) { return *D.today; }

// A batch kernel (one launch, community = blockIdx.y) keeps its community's Dev in shared memory: copied into registers
// (`const Dev D = devs[blockIdx.y]`) it cost up to twice the registers of the single-community kernel, whose Dev is a kernel
// parameter in the constant bank (k_trials 78 against 32, k_active 116 against 60).
#define BATCH_DEV(D, devs)                                                                                                    \
    __shared__ Dev s_dev_;                                                                                                    \
    {                                                                                                                         \
        const unsigned int* src_ = reinterpret_cast<const unsigned int*>(devs + blockIdx.y);                                  \
        unsigned int* dst_ = reinterpret_cast<unsigned int*>(&s_dev_);                                                        \
        for (unsigned int i_ = threadIdx.x; i_ < sizeof(Dev) / 4; i_ += blockDim.x) dst_[i_] = src_[i_];                      \
    }                                                                                                                         \
    __syncthreads();                                                                                                          \
    const Dev& D = s_dev_;

// append n entries to a device list whose length lives at `counter`; one atomic per coalesced group of callers
__device__ __forceinline__ unsigned int list_reserve(unsigned int* counter) {
    cg::coalesced_group g = cg::coalesced_threads();
    unsigned int base = 0;
    if (g.thread_rank() == 0) base = atomicAdd(counter, g.size());
    return g.shfl(base, 0) + g.thread_rank();
}

// reserve n entries (n may differ per thread, 0 allowed) of a device list for every thread of a coalesced group: one atomic
__device__ __forceinline__ unsigned int group_reserve(const cg::coalesced_group& g, unsigned int* counter, const unsigned int n) {
    unsigned int incl = n;
    for (unsigned int o = 1; o < g.size(); o <<= 1) { const unsigned int t = g.shfl_up(incl, o); if (g.thread_rank() >= o) incl += t; }
    unsigned int base = 0;
    if (g.thread_rank() == g.size() - 1 && incl) base = atomicAdd(counter, incl);
    return g.shfl(base, g.size() - 1) + incl - n;
}

// three list reservations for every calling thread with one round trip: the scans first, then the (up to three) atomics
// back to back by the group's last thread, then the broadcasts
__device__ __forceinline__ void list_reserve3(unsigned int* c0, const unsigned int n0, unsigned int* c1, const unsigned int n1, unsigned int* c2, const unsigned int n2,
                                              unsigned int& b0, unsigned int& b1, unsigned int& b2) {
    cg::coalesced_group g = cg::coalesced_threads();
    unsigned int i0 = n0, i1 = n1, i2 = n2;
    for (unsigned int o = 1; o < g.size(); o <<= 1) {
        const unsigned int t0 = g.shfl_up(i0, o), t1 = g.shfl_up(i1, o), t2 = g.shfl_up(i2, o);
        if (g.thread_rank() >= o) { i0 += t0; i1 += t1; i2 += t2; }
    }
    unsigned int a0 = 0, a1 = 0, a2 = 0;
    if (g.thread_rank() == g.size() - 1) {
        if (i0) a0 = atomicAdd(c0, i0);
        if (i1) a1 = atomicAdd(c1, i1);
        if (i2) a2 = atomicAdd(c2, i2);
    }
    b0 = g.shfl(a0, g.size() - 1) + i0 - n0; b1 = g.shfl(a1, g.size() - 1) + i1 - n1; b2 = g.shfl(a2, g.size() - 1) + i2 - n2;
}

// The same for a whole block (every thread calls it, n* may be 0): warp scans, one thread adds the block's three totals to
// the counters, two __syncthreads. One returning atomic per counter and BLOCK instead of per warp: the list counters share a
// line, and the L2 serves returning atomics on one line at 1.5 G/s (profiles/r2_ubench_atomic_rate.txt).
__device__ __forceinline__ void block_reserve3(unsigned int* c0, const unsigned int n0, unsigned int* c1, const unsigned int n1, unsigned int* c2, const unsigned int n2,
                                               unsigned int& b0, unsigned int& b1, unsigned int& b2) {
    __shared__ unsigned int s_w[TPB / 32][3], s_b[3];
    const unsigned int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned int i0 = n0, i1 = n1, i2 = n2;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t0 = __shfl_up_sync(0xffffffffu, i0, o), t1 = __shfl_up_sync(0xffffffffu, i1, o), t2 = __shfl_up_sync(0xffffffffu, i2, o);
        if (lane >= (unsigned int) o) { i0 += t0; i1 += t1; i2 += t2; }
    }
    if (lane == 31) { s_w[wid][0] = i0; s_w[wid][1] = i1; s_w[wid][2] = i2; }
    __syncthreads();
    if (threadIdx.x < 3) {
        unsigned int tot = 0;
        for (unsigned int k = 0; k < blockDim.x / 32; ++k) { const unsigned int c = s_w[k][threadIdx.x]; s_w[k][threadIdx.x] = tot; tot += c; }
        unsigned int* const ctr = threadIdx.x == 0 ? c0 : threadIdx.x == 1 ? c1 : c2;
        s_b[threadIdx.x] = tot ? atomicAdd(ctr, tot) : 0u;
    }
    __syncthreads();
    b0 = s_b[0] + s_w[wid][0] + i0 - n0; b1 = s_b[1] + s_w[wid][1] + i1 - n1; b2 = s_b[2] + s_w[wid][2] + i2 - n2;
    __syncthreads();                                        // the arrays are reused by the block's next call
}

__device__ __forceinline__ unsigned int warp_sum(unsigned int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide sums of up to NC per-thread counts, one global atomic per counter per block
template <int NC>
__device__ __forceinline__ void block_work_add(const Dev& D, const int* ids, const unsigned int* vals) {
    __shared__ unsigned int s_acc[NC];
    if (threadIdx.x < NC) s_acc[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int c = 0; c < NC; ++c) { const unsigned int v = warp_sum(vals[c]); if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s_acc[c], v); }
    __syncthreads();
    if (threadIdx.x < NC && s_acc[threadIdx.x]) atomicAdd(&D.ctr->work[ids[threadIdx.x]], (unsigned long long) s_acc[threadIdx.x]);
}

// Community::flagInfectedLocation (Community.cpp:788-791) on device, second half: the flag record, and for the first
// flag of a location the hot-list entry. `slot` is what atomicAdd(&loc_cnt[loc], 1) returned, f the record index.
__device__ __forceinline__ void flag_location(const Dev& D, int loc, int strain, unsigned long long c, unsigned int pid, unsigned int slot, unsigned int f) {
    if (f < D.cap_flags) { FlagRaw r; r.loc = loc; r.slot = slot; r.pid = pid; r.pad = 0; r.c = c | ((unsigned long long) strain << 56); D.flags_raw[f] = r; }
    else atomicOr(&D.ctr->overflow, 128u);
    if (slot == 0) {
        const unsigned int i = list_reserve(&D.ctr->n_hot);
        if (i < D.cap_hot) {
            const int q = D.loc_pub[loc]; const bool hosp = D.loc_type[loc] == CABM_HOSPITAL;
            D.hot_list[i] = loc; D.locd[loc].hot = (int) i; D.hot_meta[i] = (q >= 0 ? (unsigned int) q : HOT_NOPUB) | (hosp ? HOT_IS_HOSP : 0u);
            if (hosp) { const unsigned int j = list_reserve(&D.ctr->n_hot_hosp); if (j < D.cap_hot) D.hot_hosp[j] = (int) i; }
        } else atomicOr(&D.ctr->overflow, 2u);
        const int q = D.loc_pub[loc];
        if (q >= 0) atomicOr(&D.pub_hot[q >> 5], 1u << (q & 31));
    }
}

// ---------------------------------------------------------------------------------------------
// k_scan: the active bitmap -> today's list of active agents in ascending id order (a block's run is contiguous, blocks append
// in any order). 128 agents (one uint4 of the bitmap) per thread; blocks without an active agent return after one
// __syncthreads_or. Also publishes today's DayCtx for the rest of the pipeline.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void k_scan_body(const Dev& D) {
    __shared__ unsigned int s_warp[TPB / 32], s_base;
    pdl_wait();
    Counters* const ctr = D.ctr;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *D.today = D.dayctx[ctr->day_idx];
        atomicAdd(&ctr->work[CABM_W_AGENTS], (unsigned long long) (D.own_hi - D.own_lo));
    }
    const unsigned int w_lo = (D.own_lo >> 5) & ~3u, w_hi = (D.own_hi + 31u) >> 5;       // bitmap words that hold owned agents (from a uint4 boundary)
    const unsigned int wb = w_lo + (blockIdx.x * TPB + threadIdx.x) * 4u;
    unsigned int w[4] = { 0u, 0u, 0u, 0u };
    if (wb < w_hi) { const uint4 v = *reinterpret_cast<const uint4*>(D.actbits + wb); w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w; }
    unsigned int cnt = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {                                                       // agents of other handles are not this handle's to walk
        const unsigned int a0 = (wb + k) * 32u;
        if (a0 < D.own_lo) w[k] &= (D.own_lo - a0 >= 32u) ? 0u : (0xFFFFFFFFu << (D.own_lo - a0));
        if (a0 + 32u > D.own_hi) w[k] &= (a0 >= D.own_hi) ? 0u : (0xFFFFFFFFu >> (a0 + 32u - D.own_hi));
        cnt += __popc(w[k]);
    }
    pdl_launch_dependents();
    if (__syncthreads_or((int) cnt) == 0) return;
    const unsigned int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (unsigned int) o) incl += t; }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int acc = 0;
#pragma unroll
        for (int k = 0; k < TPB / 32; ++k) { const unsigned int c = s_warp[k]; s_warp[k] = acc; acc += c; }
        s_base = atomicAdd(&ctr->n_active, acc);
    }
    __syncthreads();
    unsigned int o = s_base + s_warp[wid] + incl - cnt;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        for (unsigned int m = w[k]; m; m &= m - 1u, ++o) {
            if (o < D.cap_active) D.active_list[o] = (wb + k) * 32u + (unsigned int) __ffs(m) - 1u; else atomicOr(&ctr->overflow, 4u);
        }
}
__global__ void __launch_bounds__(TPB) k_scan(const Dev D) { k_scan_body(D); }
__global__ void __launch_bounds__(TPB) k_scan_b(const Dev* __restrict__ devs) { BATCH_DEV(D, devs) k_scan_body(D); }   // one launch, replicate = blockIdx.y

// ---------------------------------------------------------------------------------------------
// k_active: today's active agents in the id order k_scan left them in, a block takes 256 of them at a time, in two halves.
//   agents : one thread per agent: timeline, status byte, home, day location and patron row requested at once; the activity
//            window, admissions / discharges, and where the agent is infectious today (Person.cpp:475-486,
//            Community.cpp:1031-1035): home or hospital, day location, up to five distinct patronised places. Every place
//            becomes one 16-byte entry of the block's shared-memory list (position from a shared-memory counter).
//   places : one thread per entry: Community::flagInfectedLocation (Community.cpp:788-791). The location's static words, the
//            pressure RED, the slot atomic and the flag-record reservation are issued back to back (one round trip); only a
//            location's first flag of the day needs a second one (hot list, hot hospitals).
// An agent's places used to be walked by its own thread through arrays in local memory: every slot atomic had to return before
// the next place could be issued (the store of the result waited for it), a dependent chain of up to 2 x 7 round trips per
// thread and 23 % of the kernel's stall samples in a large epidemic (profiles/r2_v6_*). Here every place is its own thread and
// nothing is kept in an indexed array.
// ---------------------------------------------------------------------------------------------
#define ACT_MAXL 7            // home or hospital, day location, five distinct patronised places (Community.cpp:279-288)
struct __align__(16) ActEntry { int loc; unsigned int pid; unsigned long long c; };   // loc < 0: ~(public index); c = pressure | strain << 56

// Community::public_activity draws for one agent (Community.cpp:1012-1026): k ~ Poisson(1) from the first word of the agent's
// Philox block, then k uniform picks among its patronised locations. Kept in registers: the distinct places (public-location
// indices) in qs[0..4] in order of first appearance, their multiplicities six bits each in `hours` (k <= 31). Returns the number
// of distinct places. `row` is the agent's 16-byte patron row.
__device__ __forceinline__ int gen_visits_r(const Dev& D, const DayCtx& cx, const unsigned int pid, const int npat, const uint4& row, unsigned int (&qs)[5], unsigned int& hours) {
    const uint2 key = make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32));
    const unsigned int dp = ((unsigned int) cx.day << 8) | CABM_P_VISIT;
    uint4 b = philox4x32_10(make_uint4(0u, pid, 0u, dp), key);
    hours = 0u;
    int k;
    if (b.x < D.visit_thr[0]) return 0;
    else if (b.x < D.visit_thr[1]) k = 1;
    else if (b.x < D.visit_thr[2]) k = 2;
    else if (b.x < D.visit_thr[3]) k = 3;
    else k = (int) table_draw(D.par->visit_cdf, CABM_T_POISSON, (double) b.x * (1.0 / 4294967296.0));
    int nv = 0;
    const bool avoiding = prow16_rank(row) < cx.sd_rank;
    for (int i = 0; i < k; ++i) {
        const unsigned int idx1 = (unsigned int) i + 1u;              // draw index within the stream (0 was the Poisson draw)
        if ((idx1 & 3u) == 0) b = philox4x32_10(make_uint4(idx1 >> 2, pid, 0u, dp), key);
        const unsigned int w = idx1 & 3u;
        const unsigned int x = w == 0 ? b.x : w == 1 ? b.y : w == 2 ? b.z : b.w;
        const unsigned int e = prow16_entry(row, __umulhi(x, (unsigned int) npat));
        if (((e >> PROW_HIGH_BIT) & 1u) && avoiding) continue;
        const unsigned int q = e & PROW_Q_MASK;
        int t = nv;
#pragma unroll
        for (int u = 4; u >= 0; --u) if (u < nv && qs[u] == q) t = u;
        if (t == nv) {
#pragma unroll
            for (int u = 0; u < 5; ++u) if (u == nv) qs[u] = q;
            ++nv;
        }
        hours += 1u << (6 * t);
    }
    return nv;
}

__device__ __forceinline__ void act_emit(ActEntry* s_e, unsigned int* s_n, const int loc, const unsigned int pid, const unsigned long long c) {
    const unsigned int k = atomicAdd(s_n, 1u);
    ActEntry e; e.loc = loc; e.pid = pid; e.c = c;
    s_e[k] = e;
}

__device__ __forceinline__ void k_active_body(const Dev& D) {
    __shared__ ActEntry s_e[ACT_MAXL * TPB];
    __shared__ unsigned int s_slot[ACT_MAXL * TPB];        // what the slot atomic of entry j returned
    __shared__ unsigned short s_first[ACT_MAXL * TPB];     // entries that were their location's first flag today
    __shared__ unsigned int s_n[2], s_nf[2], s_base[2];    // entries / first flags of the current round and of the next one (reset while idle); list bases
    pdl_wait();
    Counters* const ctr = D.ctr;
    const DayCtx& cx = day_ctx(D); const int day = cx.day;
    const unsigned int n = min(ctr->n_active, D.cap_active);
    unsigned int nflag = 0, ntoday = 0, round = 0;
    if (threadIdx.x < 2) { s_n[threadIdx.x] = 0; s_nf[threadIdx.x] = 0; }
    __syncthreads();
    for (unsigned int base = blockIdx.x * TPB; base < n; base += gridDim.x * TPB, round ^= 1u) {
        unsigned int* const cnt = &s_n[round];
        const unsigned int i = base + threadIdx.x;
        // ---- agents ----
        if (i < n) do {
            const unsigned int p = D.active_list[i];
            CABM_CHECK(D, p < D.N && p >= D.own_lo && p < D.own_hi);
            // the agent's record: one 64-byte line; its activity window is a function of the timeline (tl_until)
            const AgRec ar = ag_load(D.ag, p);
            const Tl0 t0 = ar.t0; const Tl1 t1 = ar.t1;
            const unsigned char pb = ar.pbyte;
            const int hm = ar.home, dl = ar.day_loc;
            const int from = t0.infected, until = tl_until(t0, t1);
            CABM_CHECK(D, from <= day && day <= until && D.imm[p] != 0);                   // a set bit means: inside the window today
            if (until <= day) atomicAnd(&D.actbits[p >> 5], ~(1u << (p & 31)));             // last day of the window: out of the set from tomorrow on
            if (day < from || day > until) break;                                           // (days were skipped by the caller: nothing to do any more)
            ntoday++;
            const bool inhosp = tl_in(t0.hosp_begin, t0.severe_end, day);
            const unsigned char npb = (unsigned char) ((pb & ~PB_INHOSP) | (inhosp ? PB_INHOSP : 0));
            if (npb != pb) {                                                       // admissions and discharges: the status byte and the bit k_visit reads
                D.ag[p].pbyte = npb;
                if (inhosp) atomicOr(&D.inhosp[p >> 5], 1u << (p & 31)); else atomicAnd(&D.inhosp[p >> 5], ~(1u << (p & 31)));
            }
            if (!tl_in(t0.infectious_begin, t0.infectious_end, day)) break;
            CABM_CHECK(D, t1.strain < CABM_N_STRAINS);
            const unsigned long long tag = (unsigned long long) t1.strain << 56;
            const unsigned long long c = (unsigned long long) __double2ll_rn(CABM_MUL((double) t1.rel_inf, 4294967296.0));
            if (inhosp) { const int h = ar.hospital; if (h >= 0 && c) act_emit(s_e, cnt, h, p, c | tag); break; }
            if (c) act_emit(s_e, cnt, hm, p, c | tag);
            if (dl >= 0 && c && !(tl_in(t0.severe_begin, t0.severe_end, day) || tl_in(t1.critical_begin, t1.critical_end, day))) act_emit(s_e, cnt, dl, p, c | tag);
            const int npat = pb & PB_NPAT;
            if (npat) {
                unsigned int qs[5], hours;
                const int nv = gen_visits_r(D, cx, p, npat, ar.prow, qs, hours);
#pragma unroll
                for (int u = 0; u < 5; ++u) if (u < nv) {
                    const int hrs = (int) ((hours >> (6 * u)) & 63u);
                    const unsigned long long cv = (unsigned long long) __double2ll_rn(CABM_MUL(CABM_MUL((double) t1.rel_inf, (double) hrs), 536870912.0));
                    if (cv) act_emit(s_e, cnt, ~(int) qs[u], p, cv | tag);
                }
            }
        } while (false);
        __syncthreads();
        const unsigned int total = *cnt;
        CABM_CHECK(D, total <= ACT_MAXL * TPB);
        // The block's flag records and hot-list entries are reserved by ONE thread, once per round (the list counters share a
        // line; reserved per warp they were a hundred thousand returning atomics on it per day in a large epidemic, and the
        // wait for them half of this kernel's stall samples). Both reservations are in flight while the block works.
        unsigned int fbase = 0, hbase = 0;
        if (threadIdx.x == 0) {
            s_n[round ^ 1u] = 0; s_nf[round ^ 1u] = 0;                                      // nobody touches the other counters until the next round
            if (total) fbase = atomicAdd(&ctr->n_flags, total);
        }
        // ---- places, the location's line: pressure, slot ----
        for (unsigned int j = threadIdx.x; j < total; j += TPB) {
            const ActEntry e = s_e[j];
            int loc = e.loc;
            if (loc < 0) { CABM_CHECK(D, (unsigned int) ~loc < D.Q); loc = D.pub_loc[~loc]; s_e[j].loc = loc; }
            CABM_CHECK(D, loc >= 0 && (unsigned int) loc < D.L);
            const unsigned int strain = (unsigned int) (e.c >> 56);
            atomicAdd(&D.locd[loc].press[strain], e.c & 0x00FFFFFFFFFFFFFFull);
            const unsigned int slot = atomicAdd(&D.locd[loc].cnt, 1u);
            s_slot[j] = slot;
            if (slot == 0) s_first[atomicAdd(&s_nf[round], 1u)] = (unsigned short) j;       // the location's first flag today: hot from now on
            nflag++;
        }
        if (threadIdx.x == 0) s_base[0] = fbase;
        __syncthreads();
        const unsigned int nfirst = s_nf[round];
        if (threadIdx.x == 0 && nfirst) hbase = atomicAdd(&ctr->n_hot, nfirst);
        // ---- the flag records, coalesced ----
        for (unsigned int j = threadIdx.x; j < total; j += TPB) {
            const ActEntry e = s_e[j];
            const unsigned int f = s_base[0] + j;
            if (f < D.cap_flags) { FlagRaw r; r.loc = e.loc; r.slot = s_slot[j]; r.pid = e.pid; r.pad = 0; r.c = e.c; D.flags_raw[f] = r; }
            else atomicOr(&ctr->overflow, 128u);
        }
        if (threadIdx.x == 0) s_base[1] = hbase;
        __syncthreads();
        // ---- the locations that became hot ----
        for (unsigned int k = threadIdx.x; k < nfirst; k += TPB) {
            const int loc = s_e[s_first[k]].loc;
            const LocS ls = D.locs[loc];                                                    // (only a first flag needs the static record)
            const unsigned int hi = s_base[1] + k;
            if (hi < D.cap_hot) {
                D.hot_list[hi] = loc; D.locd[loc].hot = (int) hi;
                D.hot_meta[hi] = (ls.pub >= 0 ? (unsigned int) ls.pub : HOT_NOPUB) | (ls.type == CABM_HOSPITAL ? HOT_IS_HOSP : 0u);   // for the end-of-day reset
                if (ls.type == CABM_HOSPITAL) { const unsigned int hh = list_reserve(&ctr->n_hot_hosp); if (hh < D.cap_hot) D.hot_hosp[hh] = (int) hi; }
            } else atomicOr(&ctr->overflow, 2u);
            if (ls.pub >= 0) atomicOr(&D.pub_hot[ls.pub >> 5], 1u << (ls.pub & 31));
        }
        __syncthreads();                                                                    // the lists are rewritten by the next round
    }
    pdl_launch_dependents();                             // the successor may start launching: only this block's counters are left
    nflag = warp_sum(nflag); ntoday = warp_sum(ntoday);
    if ((threadIdx.x & 31) == 0) {
        if (nflag) atomicAdd(&ctr->work[CABM_W_FLAGS], (unsigned long long) nflag);
        if (ntoday) atomicAdd(&ctr->work[CABM_W_ACTIVE], (unsigned long long) ntoday);
    }
}
__global__ void __launch_bounds__(TPB) k_active(const Dev D) { k_active_body(D); }
__global__ void __launch_bounds__(TPB) k_active_b(const Dev* __restrict__ devs) { BATCH_DEV(D, devs) k_active_body(D); }   // one launch, replicate = blockIdx.y

// ---------------------------------------------------------------------------------------------
// k_visit: every agent with patronised locations that is not in hospital draws its visits (Community.cpp:1006-1039).
//
// A warp takes tiles of 128 consecutive agents, four per lane (lane l of slot a handles agent tile + 32 a + l). The only
// per-agent memory on the path is the agent's 20-byte patron row: the rows of a tile are one contiguous 2.5 KB block that an
// elected lane fetches into the warp's shared-memory buffer with one bulk copy (cp.async.bulk, completion on an mbarrier),
// one tile ahead of the tile being processed, so the warp never waits for HBM and spends no instructions on addresses. The
// in-hospital flags of a tile are four words of a bitmap (one 16-byte load per lane, same address). The Philox block needs
// no memory. Hotness is tested against a shared-memory bitmap over the Q patronised locations (dynamic shared memory,
// fetched with one bulk copy per block). A visit that lands on a hot location becomes one word of the warp's dense staging
// buffer (positions from a warp scan per tile, no atomics); the warp flushes when the next tile might not fit or after 64
// tiles: one reservation, coalesced record stores, one census RED per record. Agents that draw k >= 4 (1.9 %) are queued per block and handled on full warps
// by the generic loop at the end, so the common path carries no divergent tail.
// There is no block-level synchronisation in the loop.
// ---------------------------------------------------------------------------------------------
#define VIS_A 4               // agents per lane and tile
#define VIS_TILE (32 * VIS_A)
#define VIS_STAGE 768         // records a warp can stage (a tile adds at most 3 * VIS_TILE = 384)
#define VIS_FLUSH_AT (VIS_STAGE - 3 * VIS_TILE)   // flush after a tile once more than this is staged
#define VIS_WINDOW 64         // tiles between two flushes at most: a staged record names its tile by 6 bits
#define VIS_SLOWQ (2 * TPB)   // k >= 4 agents a block queues for its final full-warp pass (beyond that: handled in place)
#define VIS_WARPS (TPB / 32)
#define VIS_TILE_BYTES (VIS_TILE * PROW_WORDS * 4u)
#define VIS_MINB 3            // resident blocks per SM the register allocation aims at

struct __align__(16) VisitSmem {
    unsigned int rows[VIS_WARPS][2][VIS_TILE * PROW_WORDS];  // patron rows of the current and the next tile of every warp (bulk-copy destinations)
    unsigned int stage[VIS_WARPS][VIS_STAGE];               // hot visits, dense per warp: public index | slot a << 19 | lane << 21 | tile of the window << 26
    unsigned long long thr[CABM_T_POISSON];                 // ceil(visit_cdf[k] * 2^32)
    unsigned long long bar_rows[VIS_WARPS][2];              // mbarriers of the row buffers
    unsigned long long bar_hot;                             // mbarrier of the hot-bitmap copy
    unsigned int slowq[VIS_SLOWQ];                          // agent | patron count << 29
    unsigned int nslow;
    alignas(16) unsigned int hot[4];                        // hot bitmap over the patronised locations, bitmap_words rounded up to 4 (dynamic); 16-byte aligned: bulk-copy destination
};
static_assert(offsetof(VisitSmem, hot) % 16 == 0 && offsetof(VisitSmem, rows) % 16 == 0 && offsetof(VisitSmem, bar_rows) % 8 == 0 && sizeof(VisitSmem) == offsetof(VisitSmem, hot) + 16, "VisitSmem layout");
static inline size_t visit_smem_bytes(unsigned int bitmap_words) { return sizeof(VisitSmem) - 16 + (size_t) ((bitmap_words + 3u) & ~3u) * 4u; }

__device__ __forceinline__ unsigned int smem_u32(const void* p) { return (unsigned int) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned int bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned int parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}"
                 :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (TMA, 1-D): bytes is a multiple of 16, both addresses 16-byte aligned; completes on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned int bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// the generic public_activity loop for one agent that drew k >= 4 (Community.cpp:1012-1026), one lane per agent.
// thr[j] = ceil(visit_cdf[j] * 2^32): u01 >= cdf  <=>  x >= thr (64-bit compare, exact). A hot visit goes straight to the
// global record list (these agents are 1.9 % of the drawing ones and run on full warps at the end of the block).
__device__ __forceinline__ void visit_slow(const Dev& D, const unsigned int* s_hot, const unsigned long long* s_thr, const unsigned int p, const unsigned int npat,
                                           const unsigned int dp, const uint2 key, const unsigned int sd_rank, unsigned int& w_v0, unsigned int& w_v) {
    const PRow row = prow_load(D.prow, p);
    uint4 bb = philox4x32_10(make_uint4(0u, p, 0u, dp), key);
    unsigned int kk = 0;
    while (kk < CABM_T_POISSON - 1 && (unsigned long long) bb.x >= s_thr[kk]) ++kk;          // table_draw on shared-memory thresholds
    const bool avoiding = prow_rank(row.w[1], row.w[2]) < sd_rank;
    unsigned int qs[5]; unsigned int nv = 0;
    for (unsigned int i = 1; i <= kk; ++i) {
        if ((i & 3u) == 0) bb = philox4x32_10(make_uint4(i >> 2, p, 0u, dp), key);
        const unsigned int x = (i & 3u) == 0 ? bb.x : (i & 3u) == 1 ? bb.y : (i & 3u) == 2 ? bb.z : bb.w;
        const unsigned int e = prow_entry(row, __umulhi(x, npat));
        if (avoiding && ((e >> PROW_HIGH_BIT) & 1u)) continue;
        const unsigned int q = e & PROW_Q_MASK;
        bool seen = false;
#pragma unroll
        for (int t = 0; t < 5; ++t) if (t < (int) nv && qs[t] == q) seen = true;
        if (seen) continue;
#pragma unroll
        for (int t = 0; t < 5; ++t) if (t == (int) nv) qs[t] = q;
        nv++; w_v0++;
        if (s_hot[q >> 5] & (1u << (q & 31))) {
            atomicAdd(&D.nvis[q], 1u);
            const unsigned int vi = list_reserve(&D.ctr->n_vis);
            if (vi < D.cap_vis) D.visit_rec[vi] = make_uint2(p, q); else atomicOr(&D.ctr->overflow, 8u);
            w_v++;
        }
    }
}

// flush a warp's staged records: one reservation, then the dense buffer goes out 32 records at a time (coalesced 8-byte
// stores, one census RED per record). A staged word names its agent by (tile of the window, slot a, lane):
// agent = win_t0 + tile * tile_stride + 32 a + lane.
__device__ __forceinline__ void visit_flush(const Dev& D, const unsigned int* stage, unsigned int& staged, const unsigned int lane,
                                            const unsigned int win_t0, const unsigned int tile_stride, unsigned int& w_v) {
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(&D.ctr->n_vis, staged);
    base = __shfl_sync(0xffffffffu, base, 0);
    for (unsigned int i = lane; i < staged; i += 32) {
        const unsigned int e = stage[i];
        const unsigned int q = e & PROW_Q_MASK, p = win_t0 + (e >> 26) * tile_stride + ((e >> 19) & 3u) * 32u + ((e >> 21) & 31u);
        CABM_CHECK(D, q < D.Q && p >= D.own_lo && p < D.own_hi && staged <= VIS_STAGE);
        atomicAdd(&D.nvis[q], 1u);
        if (base + i < D.cap_vis) D.visit_rec[base + i] = make_uint2(p, q); else atomicOr(&D.ctr->overflow, 8u);
    }
    __syncwarp();                                          // the buffer is written again by the next tile
    if (lane == 0) w_v += staged;
    staged = 0;
}

__device__ __forceinline__ void k_visit_body(const Dev& D, const unsigned int bitmap_words) {
    extern __shared__ __align__(16) unsigned char vis_smem_raw[];
    VisitSmem& S = *reinterpret_cast<VisitSmem*>(vis_smem_raw);
    const unsigned int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const unsigned int own_lo = D.own_lo, own_hi = D.own_hi, own_n = own_hi - own_lo;
    const unsigned int t_base = own_lo & ~(VIS_TILE - 1u);                          // tiles cover the owned agents only
    const unsigned int g_warp = blockIdx.x * VIS_WARPS + wid, tile_stride = gridDim.x * VIS_WARPS * VIS_TILE;
    unsigned int t0 = t_base + g_warp * VIS_TILE;
    // ---- prologue, before the dependency on k_active: barriers, thresholds, the first tile's rows (static data) ----
    if (lane == 0) { mbar_init(&S.bar_rows[wid][0], 1); mbar_init(&S.bar_rows[wid][1], 1); }
    if (threadIdx.x == 0) { mbar_init(&S.bar_hot, 1); S.nslow = 0; }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    if (threadIdx.x < CABM_T_POISSON) S.thr[threadIdx.x] = __double2ull_ru(D.par->visit_cdf[threadIdx.x] * 4294967296.0);
    __syncthreads();
    if (lane == 0 && t0 < own_hi) { mbar_expect_tx(&S.bar_rows[wid][0], VIS_TILE_BYTES); bulk_g2s(&S.rows[wid][0][0], D.prow + (size_t) t0 * PROW_WORDS, VIS_TILE_BYTES, &S.bar_rows[wid][0]); }
    const uint2 key = make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32));
    const unsigned int thr0 = D.visit_thr[0], thr1 = D.visit_thr[1], thr2 = D.visit_thr[2], thr3 = D.visit_thr[3];
    // ---- everything below reads what k_active wrote today ----
    pdl_wait();
    const unsigned int hot_bytes = ((bitmap_words + 3u) & ~3u) * 4u;
    if (threadIdx.x == 0) { mbar_expect_tx(&S.bar_hot, hot_bytes); bulk_g2s(&S.hot[0], D.pub_hot, hot_bytes, &S.bar_hot); }
    const DayCtx& cx = day_ctx(D);
    const unsigned int dp = ((unsigned int) cx.day << 8) | CABM_P_VISIT;
    const unsigned int sd_rank = cx.sd_rank;
    const uint4* __restrict__ hosp4 = reinterpret_cast<const uint4*>(D.inhosp);
    uint4 hw_next = make_uint4(0u, 0u, 0u, 0u);
    if (t0 < own_hi) hw_next = hosp4[t0 / VIS_TILE];
    mbar_wait(&S.bar_hot, 0);
    const unsigned int* __restrict__ s_hot = S.hot;
    unsigned int* const stage = S.stage[wid];
    unsigned int staged = 0, w_pat = 0, w_v0 = 0, w_v = 0, phase = 0;              // staged: records in the warp's buffer (same in every lane); phase: bit b = parity to wait for on row buffer b
    unsigned int win_t0 = t0, win_n = 0;                                            // first tile and number of tiles of the staging window
    for (unsigned int it = 0; t0 < own_hi; ++it, t0 += tile_stride) {
        const unsigned int b = it & 1u;
        const uint4 hw = hw_next;
        const unsigned int t1 = t0 + tile_stride;
        if (t1 < own_hi) {                                                          // next tile: rows by bulk copy, in-hospital words by load
            if (lane == 0) { mbar_expect_tx(&S.bar_rows[wid][b ^ 1u], VIS_TILE_BYTES); bulk_g2s(&S.rows[wid][b ^ 1u][0], D.prow + (size_t) t1 * PROW_WORDS, VIS_TILE_BYTES, &S.bar_rows[wid][b ^ 1u]); }
            hw_next = hosp4[t1 / VIS_TILE];
        }
        // k ~ Poisson(1) and the first three picks of every agent come from one Philox block (Community.cpp:1013-1017). The four
        // agents of a lane are independent: their Philox rounds are interleaved so a warp always has eight multiplies in flight,
        // and the whole computation sits in front of the wait for the tile's rows.
        uint4 x[VIS_A];
        {
            uint2 k = key;
#pragma unroll
            for (int a = 0; a < VIS_A; ++a) x[a] = make_uint4(0u, t0 + a * 32 + lane, 0u, dp);
#pragma unroll
            for (int r = 0; r < 10; ++r) {
#pragma unroll
                for (int a = 0; a < VIS_A; ++a) {
                    const unsigned int hi0 = __umulhi(0xD2511F53u, x[a].x), lo0 = 0xD2511F53u * x[a].x;
                    const unsigned int hi1 = __umulhi(0xCD9E8D57u, x[a].z), lo1 = 0xCD9E8D57u * x[a].z;
                    x[a] = make_uint4(hi1 ^ x[a].y ^ k.x, lo1, hi0 ^ x[a].w ^ k.y, lo0);
                }
                k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
            }
        }
        mbar_wait(&S.bar_rows[wid][b], (phase >> b) & 1u);
        phase ^= 1u << b;
        const unsigned int* __restrict__ rows = S.rows[wid][b];
        unsigned int recq[3 * VIS_A]; unsigned int recmask = 0, slow = 0;          // this lane's hot visits of the tile (k <= 3)
#pragma unroll
        for (int a = 0; a < VIS_A; ++a) {
            const unsigned int* __restrict__ rw = rows + (a * 32 + lane) * PROW_WORDS;    // this agent's row: 5 words, bank = (5 lane + j) mod 32
            const unsigned int p = t0 + a * 32 + lane;
            const unsigned int hword = a == 0 ? hw.x : a == 1 ? hw.y : a == 2 ? hw.z : hw.w;
            const unsigned int w0 = rw[0], w1 = rw[1], w2 = rw[2];
            const unsigned int npat = prow_npat(w0);
            // drawing today: has patronised places, is not in hospital (Community.cpp:1009), and this handle owns the agent
            const unsigned int act = (unsigned int) (npat != 0) & (((hword >> lane) & 1u) ^ 1u) & (unsigned int) ((p - own_lo) < own_n);
            const unsigned int c3 = act & (unsigned int) (x[a].x >= thr3);             // k >= 4: the generic loop at the end of the block
            const unsigned int live = act ^ c3;
            unsigned int v0 = live & (unsigned int) (x[a].x >= thr0), v1 = live & (unsigned int) (x[a].x >= thr1), v2 = live & (unsigned int) (x[a].x >= thr2);
            const unsigned int e0 = rw[__umulhi(x[a].y, npat)], e1 = rw[__umulhi(x[a].z, npat)], e2 = rw[__umulhi(x[a].w, npat)];   // npat == 0: word 0, ignored (act == 0)
            const unsigned int avoiding = (unsigned int) (prow_rank(w1, w2) < sd_rank);   // Community.cpp:1020: compliant households skip high-risk places
            v0 &= (avoiding & (e0 >> PROW_HIGH_BIT)) ^ 1u; v1 &= (avoiding & (e1 >> PROW_HIGH_BIT)) ^ 1u; v2 &= (avoiding & (e2 >> PROW_HIGH_BIT)) ^ 1u;
            const unsigned int q0 = e0 & PROW_Q_MASK, q1 = e1 & PROW_Q_MASK, q2 = e2 & PROW_Q_MASK;
            v2 &= ((v0 & (unsigned int) (q2 == q0)) | (v1 & (unsigned int) (q2 == q1))) ^ 1u;   // visit_map (Community.cpp:1014): one entry per distinct place
            v1 &= (v0 & (unsigned int) (q1 == q0)) ^ 1u;
            slow |= c3 << a;
            w_pat += act;
            w_v0 += v0 + v1 + v2;
            const unsigned int h0 = v0 & __funnelshift_r(s_hot[q0 >> 5], 0u, q0), h1 = v1 & __funnelshift_r(s_hot[q1 >> 5], 0u, q1), h2 = v2 & __funnelshift_r(s_hot[q2 >> 5], 0u, q2);
            recq[a * 3 + 0] = q0; recq[a * 3 + 1] = q1; recq[a * 3 + 2] = q2;
            recmask |= (h0 | (h1 << 1) | (h2 << 2)) << (a * 3);
        }
        // the tile's records go to the warp's dense buffer: positions from a warp scan of the per-lane counts
        if (__any_sync(0xffffffffu, recmask != 0u)) {
            const unsigned int nrec = __popc(recmask);
            unsigned int incl = nrec;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (unsigned int) o) incl += t; }
            unsigned int o = staged + incl - nrec;
            const unsigned int tag = (lane << 21) | (win_n << 26);
#pragma unroll
            for (int r = 0; r < 3 * VIS_A; ++r) if (recmask & (1u << r)) { stage[o] = recq[r] | ((unsigned int) (r / 3) << 19) | tag; ++o; }
            staged += __shfl_sync(0xffffffffu, incl, 31);
        }
        // k >= 4 (1.9 % of the drawing agents): queued for the block's final pass, so the generic loop runs on full warps
        while (slow) {
            const unsigned int a = (unsigned int) __ffs(slow) - 1u; slow &= slow - 1u;
            const unsigned int pos = atomicAdd(&S.nslow, 1u), p = t0 + a * 32 + lane, npat = prow_npat(rows[(a * 32 + lane) * PROW_WORDS]);
            if (pos < VIS_SLOWQ) S.slowq[pos] = p | (npat << 29); else visit_slow(D, s_hot, S.thr, p, npat, dp, key, sd_rank, w_v0, w_v);
        }
        __syncwarp();                                                               // every lane has read its rows: the buffer may be refilled
        ++win_n;
        if (win_n == VIS_WINDOW || staged > VIS_FLUSH_AT) { visit_flush(D, stage, staged, lane, win_t0, tile_stride, w_v); win_t0 = t0 + tile_stride; win_n = 0; }
    }
    if (staged) visit_flush(D, stage, staged, lane, win_t0, tile_stride, w_v);
    pdl_launch_dependents();                             // the tiles are done; the k >= 4 pass and the counters remain
    // the block's k >= 4 agents on full warps
    __syncthreads();
    {
        const unsigned int n = min(S.nslow, (unsigned int) VIS_SLOWQ);
        for (unsigned int i = threadIdx.x; i < n; i += blockDim.x) { const unsigned int e = S.slowq[i]; visit_slow(D, s_hot, S.thr, e & 0x1FFFFFFFu, e >> 29, dp, key, sd_rank, w_v0, w_v); }
    }
    const int ids[3] = { CABM_W_PATRONS, CABM_W_VISITS, CABM_W_HOT_VISITS }; const unsigned int vals[3] = { w_pat, w_v0, w_v };
    block_work_add<3>(D, ids, vals);
}
__global__ void __launch_bounds__(TPB, VIS_MINB) k_visit(const Dev D, const unsigned int bitmap_words) { k_visit_body(D, bitmap_words); }
__global__ void __launch_bounds__(TPB, VIS_MINB) k_visit_b(const Dev* __restrict__ devs, const unsigned int bitmap_words) { BATCH_DEV(D, devs) k_visit_body(D, bitmap_words); }   // one launch, replicate = blockIdx.y

__device__ __forceinline__ double hazard_to_T(double hazard) { return CABM_SUB(1.0, cabm_exp(-hazard)); }

// the susceptibility gate of one successful trial (cabm_kernels2.cuh)
__device__ __forceinline__ void gate_hit(const Dev& D, const int phase, const unsigned int pid, const unsigned int src, const unsigned int purpose, const unsigned int sub);

// one exposure trial: the quarantine skip and `gsl_rng_uniform(RNG) < T` of Community::_transmission (Community.cpp:975-976).
// A success of the house..nursing-home passes (phase 0) is appended to the hit list for the susceptibility gate (k_gate): the
// gate is rare, divergent and register-hungry (Person::infect's front half), and kept out of this kernel so that the streaming
// loop runs at 32 registers and full occupancy (fused in, k_trials needed 70 registers and took 575 instead of 336 us per day
// in the heavy-epidemic regime at 21.5 M agents). The hospital pass (phase 1: a handful of hits) gates in place.
__device__ __forceinline__ void trial(const Dev& D, const int day, const int phase, const unsigned int pid, const int src, const int purpose,
                                      const unsigned int sub, const double T, unsigned int& trials) {
    CABM_CHECK(D, pid < D.N && src >= 0 && (unsigned int) src < D.L);
    if (pid < D.own_lo || pid >= D.own_hi) return;                                 // the owner of an agent draws its exposures
    if (D.any_quarantine) {
        const unsigned int q = D.quar[pid]; const int qf = (short) (q & 0xFFFF), qu = (short) (q >> 16);
        if (day >= qf && day < qu && src != D.home[pid]) return;
    }
    trials++;
    const uint4 b = philox4x32_10(make_uint4(sub << 16, pid, (unsigned int) src, ((unsigned int) day << 8) | (unsigned int) purpose),
                                  make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32)));
    if ((double) b.x * (1.0 / 4294967296.0) < T) {
        if (phase) { gate_hit(D, 1, pid, (unsigned int) src, (unsigned int) purpose, sub); return; }
        const unsigned int k = (gridDim.x >= HIT_LISTS ? blockIdx.x : blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) & (HIT_LISTS - 1u);   // this block's sub-list (in a small grid, as in the batch launches: this warp's), its counter has a line of its own
        const unsigned int h = list_reserve(&D.hit_ctr[k * 32u]);
        if (h < D.hit_sub) { Hit r; r.pid = pid; r.src = (unsigned int) src; r.purpose = (unsigned int) purpose; r.sub = sub; D.hits[(size_t) k * D.hit_sub + h] = r; }
        else atomicOr(&D.ctr->overflow, 256u);
    }
}

__device__ __forceinline__ void write_foi(const Dev& D, unsigned int i, int loc, int type, const unsigned long long* pw, unsigned int size, double T, double Ts) {
    cabm_foi_row r; r.loc = (unsigned int) loc; r.type = (unsigned char) type; r.pad[0] = r.pad[1] = r.pad[2] = 0;
#pragma unroll
    for (int s = 0; s < CABM_N_STRAINS; ++s) r.w[s] = CABM_MUL((double) pw[s], 1.0 / 4294967296.0);
    r.size = size; r.T = T; r.T_social = Ts;
    D.foi[i] = r;
}

// reserve n consecutive entries of a device list for every calling thread (n may differ per thread): one atomic per
// coalesced group of callers
__device__ __forceinline__ unsigned int list_reserve_n(unsigned int* counter, const unsigned int n) {
    cg::coalesced_group g = cg::coalesced_threads();
    unsigned int incl = n;
    for (unsigned int o = 1; o < g.size(); o <<= 1) { const unsigned int t = g.shfl_up(incl, o); if (g.thread_rank() >= o) incl += t; }
    unsigned int base = 0;
    if (g.thread_rank() == g.size() - 1) base = atomicAdd(counter, incl);
    return g.shfl(base, g.size() - 1) + incl - n;
}

// ---------------------------------------------------------------------------------------------
// k_hazard: one thread per hot location (hospitals are left to k_hosp_import, after the status pass): hazard -> T of
// the location's pass(es) (Community.cpp:833-969) and the trial work it implies, as descriptors for k_trials:
//   small items : a roster (or a neighbour's roster) of at most SMALL_ROSTER members, T by reference into hot_T
//   big items   : 256-member chunks of larger rosters
// Every hot location's dependent chain (hot list -> type -> pressure -> roster/neighbour offsets) runs in its own thread,
// so the chains of all hot locations overlap; the trials themselves are spread over lanes in k_trials.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void k_hazard_body(const Dev& D) {
    pdl_wait();
    const DayCtx& cx = day_ctx(D);
    const cabm_params* __restrict__ par = D.par;
    const unsigned int n_hot = min(D.ctr->n_hot, D.cap_hot);
    unsigned int edges = 0;
    for (unsigned int i0 = blockIdx.x * blockDim.x; i0 < n_hot; i0 += gridDim.x * blockDim.x) {       // block-uniform: the reservation below is the block's
        const unsigned int i = i0 + threadIdx.x;
        const bool have = i < n_hot;
        const int loc = D.hot_list[have ? i : i0];                  // (idle threads of the last round read the round's first location and reserve nothing)
        CABM_CHECK(D, loc >= 0 && (unsigned int) loc < D.L && (!have || D.locd[loc].hot == (int) i));
        // everything that depends on `loc` only is loaded in one go (one round trip)
        const LocS ls = D.locs[loc];                          // one sector
        const unsigned int cnt = D.locd[loc].cnt;             // one line
        const ulonglong2* pw = reinterpret_cast<const ulonglong2*>(D.locd[loc].press);
        const ulonglong2 pa = pw[0], pb = pw[1];
        const int type = ls.type;
        const unsigned int rb = ls.roster_begin, size = ls.roster_size;
        const unsigned int nb0 = ls.nbr_begin, deg = ls.nbr_deg;
        const float lrisk = ls.risk;
        const unsigned char lf = ls.flags;
        const int q = ls.pub;
        const unsigned long long pws[4] = { pa.x, pa.y, pb.x, pb.y };
        const double w = CABM_MUL((double) (pws[0] + pws[1] + pws[2] + pws[3]), 1.0 / 4294967296.0);
        double T = 0.0, Ts = 0.0; unsigned int denom = size; int purpose = CABM_P_EXPOSE_HH;
        unsigned int n_nb = 0;                                // small items for the neighbours (between-household pass)
        if (type == CABM_HOUSE) {
            T = hazard_to_T(CABM_MUL(CABM_MUL(par->haz_mult[CABM_HOUSE], cx.season), w));
            const bool risky = (double) lrisk > cx.sd;
            if (risky) { Ts = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(par->haz_mult[CABM_N_LOCTYPES], cx.season), w), (double) (int) size)); if (have) edges += deg; }
            if (have && risky && Ts > 0.0) n_nb = deg;
        } else if (type == CABM_WORK) {
            denom = size + (q >= 0 ? D.nvis[q] : 0u);
            const bool closed = !(lf & LF_ESSENTIAL) && cx.business_closure != 0.0;
            if (!(denom < 2 || closed) && w > 0.0) {
                const double f = LF_PUBRISK(lf) == CABM_PT_HIGH ? 4.0 : 0.25;
                T = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(CABM_MUL(par->haz_mult[CABM_WORK], f), cx.season), w), CABM_SUB((double) (int) denom, 1.0)));
            }
            purpose = CABM_P_EXPOSE_WORK;
        } else if (type == CABM_SCHOOL) {
            const bool open = cx.is_weekday && cx.school_closure < 1;
            const double coef = CABM_MUL(CABM_MUL(CABM_SUB(1.0, cx.school_closure), par->haz_mult[CABM_SCHOOL]), cx.season);
            if (open && coef != 0.0 && size >= 2) T = hazard_to_T(CABM_DIV(CABM_MUL(coef, w), CABM_SUB((double) (int) size, 1.0)));
            purpose = CABM_P_EXPOSE_SCHOOL;
        } else if (type == CABM_NURSINGHOME) {
            if (size >= 2) T = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(par->haz_mult[CABM_NURSINGHOME], cx.season), w), CABM_SUB((double) (int) size, 1.0)));
            purpose = CABM_P_EXPOSE_NH;
        }                                                     // hospitals are left to k_hosp_import: T stays 0 here, no items
        const bool own = have && type != CABM_HOSPITAL && T > 0.0;
        const unsigned int n_small = n_nb + (own && size <= SMALL_ROSTER ? 1u : 0u);
        const unsigned int n_big = own && size > SMALL_ROSTER ? (size + CHUNK - 1) / CHUNK : 0u;
        // this location's segment of the sorted flag array, its small items and its big items: one round trip
        unsigned int seg, sbase, ibase;
        block_reserve3(&D.ctr->n_seg, have ? cnt : 0u, &D.ctr->n_sitems, n_small, &D.ctr->n_items, n_big, seg, sbase, ibase);
        if (!have) continue;
        D.hot_off[i] = seg;
        CABM_CHECK(D, cnt >= 1 && (unsigned long long) seg + cnt <= D.cap_flags && (q < 0 || (unsigned int) q < D.Q));
        if (type == CABM_HOSPITAL) continue;
        D.hot_T[2 * i] = T;
        if (type == CABM_HOUSE) D.hot_T[2 * i + 1] = Ts;
        if (type == CABM_WORK && q >= 0) D.pub_T[q] = T;
        if (n_small) {
            if (sbase + n_small <= D.cap_sitems) {
                // between_household_transmission (Community.cpp:853-875): one small item per neighbour, resolved by k_trials
                // (begin = position in the neighbour array; a neighbour below the social-distancing threshold yields no trials)
                for (unsigned int e = 0; e < n_nb; ++e) { SItem it; it.src = loc; it.begin = (int) (nb0 + e); it.count_purpose = CABM_P_EXPOSE_NB; it.t_idx = 2 * i + 1; D.sitems[sbase + e] = it; }
                if (n_small > n_nb) { SItem it; it.src = loc; it.begin = (int) rb; it.count_purpose = (size << 8) | (unsigned int) purpose; it.t_idx = 2 * i; D.sitems[sbase + n_nb] = it; }
            } else atomicOr(&D.ctr->overflow, 16u);
        }
        if (n_big) {
            if (ibase + n_big <= D.cap_items) {
                for (unsigned int k = 0; k < n_big; ++k) {
                    Item it; it.src = loc; it.begin = (int) (rb + k * CHUNK); it.count = (int) min((unsigned int) CHUNK, size - k * CHUNK); it.purpose = purpose; it.T = T; it.pad = 0.0;
                    D.items[ibase + k] = it;
                }
            } else atomicOr(&D.ctr->overflow, 16u);
        }
        if (D.foi_enabled) write_foi(D, i, loc, type, pws, denom, T, Ts);
    }
    pdl_launch_dependents();
    edges = warp_sum(edges);
    if ((threadIdx.x & 31) == 0 && edges && D.shard_rank == 0) atomicAdd(&D.ctr->work[CABM_W_NBR_EDGES], (unsigned long long) edges);
}
__global__ void __launch_bounds__(TPB) k_hazard(const Dev D) { k_hazard_body(D); }
__global__ void __launch_bounds__(TPB) k_hazard_b(const Dev* __restrict__ devs) { BATCH_DEV(D, devs) k_hazard_body(D); }   // one launch, replicate = blockIdx.y

// ---------------------------------------------------------------------------------------------
// k_trials: block-sized units: the 256-member items of big rosters, the visit records, the flag scatter, the small items.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void k_trials_body(const Dev& D) {
    __shared__ unsigned char s_owner[TPB / 32][32];
    const unsigned int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    pdl_wait();
    const DayCtx& cx = day_ctx(D); const int day = cx.day;
    const unsigned int n_items = min(D.ctr->n_items, D.cap_items), n_vis = min(D.ctr->n_vis, D.cap_vis);
    const unsigned int n_flags = min(D.ctr->n_flags, D.cap_flags);
    const unsigned int n_sitems = min(D.ctr->n_sitems, D.cap_sitems);
    const unsigned int u_vis = n_items + (n_vis + TPB - 1) / TPB;
    const unsigned int u_flag = u_vis + (n_flags + TPB - 1) / TPB;
    const unsigned int units = u_flag + (n_sitems + TPB - 1) / TPB;
    unsigned int trials = 0;
    for (unsigned int u = blockIdx.x; u < units; u += gridDim.x) {
        if (u >= u_flag) {
            // small items, 32 per warp: one lane resolves one item (descriptor, neighbour, roster bounds, T: every item's chain
            // in flight at once), a warp scan lays the items' members end to end, and the trials run one per lane over that
            // range, 32 at a time (a household has two or three members: with eight lanes per item, as before, 70 % of the lanes
            // sat idle through the Philox blocks of the others). Warp-synchronous: no block barrier on the path.
            const unsigned int si = (u - u_flag) * TPB + threadIdx.x;
            unsigned int cnt = 0; int begin = 0, src = 0, purpose = 0; double T = 0.0;
            if (si < n_sitems) {
                const SItem it = D.sitems[si];
                cnt = it.count_purpose >> 8; begin = it.begin; src = it.src;
                purpose = (int) (it.count_purpose & 0xFFu);
                if (purpose == CABM_P_EXPOSE_NB) {                   // a neighbour household: its roster, unless it keeps its distance today
                    CABM_CHECK(D, begin >= 0 && (unsigned int) begin < D.nbr_off[D.L]);
                    const int nb = D.nbr[begin];
                    CABM_CHECK(D, nb >= 0 && (unsigned int) nb < D.L);
                    const LocS ls = D.locs[nb];
                    cnt = (double) ls.risk > cx.sd ? ls.roster_size : 0u; begin = (int) ls.roster_begin;
                }
                CABM_CHECK(D, it.t_idx < 2u * D.cap_hot && (unsigned int) begin + cnt <= D.roster_off[D.L] && (unsigned int) src < D.L);
                T = D.hot_T[it.t_idx];
            }
            unsigned int incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (unsigned int) o) incl += t; }
            const unsigned int off = incl - cnt, total = __shfl_sync(0xffffffffu, incl, 31);
            for (unsigned int cb = 0; cb < total; cb += 32) {
                const unsigned int lo = max(off, cb), hi = min(off + cnt, cb + 32);
                for (unsigned int j = lo; j < hi; ++j) s_owner[wid][j - cb] = (unsigned char) lane;
                __syncwarp();
                const unsigned int j = cb + lane;
                const unsigned int t = j < total ? s_owner[wid][lane] : 0u;
                const int t_begin = __shfl_sync(0xffffffffu, begin, t), t_src = __shfl_sync(0xffffffffu, src, t), t_purpose = __shfl_sync(0xffffffffu, purpose, t);
                const unsigned int t_off = __shfl_sync(0xffffffffu, off, t);
                const double t_T = __shfl_sync(0xffffffffu, T, t);
                if (j < total) trial(D, day, 0, (unsigned int) D.roster[t_begin + (int) (j - t_off)], t_src, t_purpose, 0, t_T, trials);
                __syncwarp();
            }
        } else if (u >= u_vis) {                                    // scatter the day's flags into per-location segments (offsets from k_hazard)
            const unsigned int f = (u - u_vis) * TPB + threadIdx.x;
            if (f < n_flags) {
                const FlagRaw r = D.flags_raw[f];
                CABM_CHECK(D, r.loc >= 0 && (unsigned int) r.loc < D.L && r.pid < D.N && (unsigned int) D.locd[r.loc].hot < min(D.ctr->n_hot, D.cap_hot) &&
                              r.slot < D.locd[r.loc].cnt && (unsigned long long) D.hot_off[D.locd[r.loc].hot] + r.slot < D.cap_flags);
                Flag o; o.pid = r.pid; o.pad = 0; o.c = r.c;
                D.flags[D.hot_off[D.locd[r.loc].hot] + r.slot] = o;
            }
        } else if (u < n_items) {
            const Item it = D.items[u];
            CABM_CHECK(D, it.count >= 1 && it.count <= CHUNK && (unsigned int) it.begin + (unsigned int) it.count <= D.roster_off[D.L] && (unsigned int) it.src < D.L);
            if ((int) threadIdx.x < it.count) trial(D, day, 0, (unsigned int) D.roster[it.begin + threadIdx.x], it.src, it.purpose, 0, it.T, trials);
        } else {
            const unsigned int v = (u - n_items) * TPB + threadIdx.x;
            if (v < n_vis) {
                const uint2 rec = D.visit_rec[v];
                CABM_CHECK(D, rec.x < D.N && rec.y < D.Q);
                const double T = D.pub_T[rec.y]; const int loc = D.pub_loc[rec.y];          // both requested before either is needed
                if (T > 0.0) trial(D, day, 0, rec.x, loc, CABM_P_EXPOSE_WORK_VIS, 0, T, trials);
            }
        }
    }
    pdl_launch_dependents();
    const int ids[1] = { CABM_W_TRIALS }; const unsigned int vals[1] = { trials };
    block_work_add<1>(D, ids, vals);
}
__global__ void __launch_bounds__(TPB, 8) k_trials(const Dev D) { k_trials_body(D); }
__global__ void __launch_bounds__(TPB) k_trials_b(const Dev* __restrict__ devs) { BATCH_DEV(D, devs) k_trials_body(D); }   // one launch, replicate = blockIdx.y
