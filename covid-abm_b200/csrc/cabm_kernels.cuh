// covid-abm_b200/csrc/cabm_kernels.cuh -- first half of the day pipeline (sm_100a): who is infectious where,
// public activity, per-location hazard, exposure trials.
//
//   k_scan     all agents      : streams the 4-byte active-window word (16 agents per thread), block-aggregated
//                                compaction of today's active agents (replaces walking every Person, Community.cpp:684)
//   k_active   active agents   : Person.cpp:475-486 derived from the stored timeline instead of the reference's
//                                _isHot calendar: pressure at home / day location / hospital / visited locations
//                                (Community.cpp:1031-1035), per-location flag chains, hot lists, in-hospital bit
//   k_visit    all agents      : public_activity (Community.cpp:1006-1039): Poisson(1) visits from one Philox block
//                                (branch-free body), hot test against a shared-memory bitmap of patronised locations,
//                                visitor census, visit records staged per warp; k >= 4 draws on full warps at the end
//   k_hazard   hot locations   : _tally_infectiousness + the hazard of every pass (Community.cpp:833-969); one thread
//                                per location turns pressure into T and emits work items: small rosters and
//                                neighbour households (resolved by k_trials), 256-member chunks of big rosters
//   k_trials   items + visits  : _transmission's `U < T` (Community.cpp:976) for rosters and visitors; also scatters
//                                the day's flag records into one contiguous segment per hot location
//
// A trial that succeeds is appended to the hit list; the (rare, heavy) susceptibility gate and the state
// transition run in k_gate / k_apply (cabm_kernels2.cuh) so the streaming kernels stay convergent.
// All kernels read their sizes from device-resident counters, so a whole day is one CUDA-graph replay.
#pragma once
#include <cooperative_groups.h>
#include "cabm_infect.cuh"

namespace cg = cooperative_groups;

#define TPB 256
#define CHUNK 256
#define SMALL_ROSTER 64
#define MAXV 8
#define PATRON_ROW 8          // ints per agent: 5 patronised public-location indices, household riskiness (float bits), 2 pad
#define PATRON_RISK 5

__device__ __forceinline__ const DayCtx& day_ctx(const Dev& D) { return *D.today; }

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

// ---------------------------------------------------------------------------------------------
// k_scan: active-window scan. Also publishes today's DayCtx for the rest of the pipeline.
// ---------------------------------------------------------------------------------------------
#define SCAN_V 4               // uint4 loads per thread in k_scan: a block covers TPB * 4 * SCAN_V agents
__device__ __forceinline__ void k_scan_body(const Dev& D) {
    __shared__ unsigned int s_warp[TPB / 32], s_base;
    const DayCtx& cx = D.dayctx[D.ctr->day_idx];
    const int day = cx.day;
    if (blockIdx.x == 0 && threadIdx.x == 0) { *D.today = cx; atomicAdd(&D.ctr->work[CABM_W_AGENTS], (unsigned long long) (D.own_hi - D.own_lo)); }
    const unsigned int b0 = (D.own_lo & ~3u) + blockIdx.x * (TPB * 4u * SCAN_V);            // agents of other handles are never active here
    uint4 w[SCAN_V];
#pragma unroll
    for (int j = 0; j < SCAN_V; ++j) {                                                       // all loads in flight before the first use
        const unsigned int p0 = b0 + (j * TPB + threadIdx.x) * 4u;
        w[j] = make_uint4(0x80007FFFu, 0x80007FFFu, 0x80007FFFu, 0x80007FFFu);
        if (p0 + 3 < D.N) w[j] = *reinterpret_cast<const uint4*>(D.win + p0);
        else { if (p0 < D.N) w[j].x = D.win[p0]; if (p0 + 1 < D.N) w[j].y = D.win[p0 + 1]; if (p0 + 2 < D.N) w[j].z = D.win[p0 + 2]; }
    }
    unsigned int mask = 0;
#pragma unroll
    for (int j = 0; j < SCAN_V; ++j) {
        const unsigned int ww[4] = { w[j].x, w[j].y, w[j].z, w[j].w };
#pragma unroll
        for (int k = 0; k < 4; ++k) { const int from = (short) (ww[k] & 0xFFFF), until = (short) (ww[k] >> 16); if (day >= from && day <= until) mask |= 1u << (j * 4 + k); }
    }
    const unsigned int cnt = __popc(mask), lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (__syncthreads_or((int) mask) == 0) return;                                           // most blocks outside an epidemic wave: nothing to append
    unsigned int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int acc = 0;
#pragma unroll
        for (int k = 0; k < TPB / 32; ++k) { const unsigned int c = s_warp[k]; s_warp[k] = acc; acc += c; }
        s_base = atomicAdd(&D.ctr->n_active, acc);
        atomicAdd(&D.ctr->work[CABM_W_ACTIVE], (unsigned long long) acc);
    }
    __syncthreads();
    unsigned int o = s_base + s_warp[wid] + incl - cnt;
    while (mask) {
        const unsigned int bit = (unsigned int) __ffs(mask) - 1u; mask &= mask - 1u;
        if (o < D.cap_active) D.active_list[o] = b0 + ((bit >> 2) * TPB + threadIdx.x) * 4u + (bit & 3u); else atomicOr(&D.ctr->overflow, 4u);
        ++o;
    }
}
__global__ void __launch_bounds__(TPB) k_scan(const Dev D) { k_scan_body(D); }
__global__ void __launch_bounds__(TPB) k_scan_b(const Dev* __restrict__ devs) { const Dev D = devs[blockIdx.y]; k_scan_body(D); }   // one launch, replicate = blockIdx.y

// Community::flagInfectedLocation (Community.cpp:788-791) on device, second half: the flag record, and for the first
// flag of a location the hot-list entry. `slot` is what atomicAdd(&loc_cnt[loc], 1) returned, f the record index.
__device__ __forceinline__ void flag_location(const Dev& D, int loc, int strain, unsigned long long c, unsigned int pid, unsigned int slot, unsigned int f) {
    if (f < D.cap_flags) { FlagRaw r; r.loc = loc; r.slot = slot; r.pid = pid; r.pad = 0; r.c = c | ((unsigned long long) strain << 56); D.flags_raw[f] = r; }
    else atomicOr(&D.ctr->overflow, 128u);
    if (slot == 0) {
        const unsigned int i = list_reserve(&D.ctr->n_hot);
        if (i < D.cap_hot) {
            D.hot_list[i] = loc; D.loc_hot[loc] = (int) i;
            if (D.loc_type[loc] == CABM_HOSPITAL) { const unsigned int j = list_reserve(&D.ctr->n_hot_hosp); if (j < D.cap_hot) D.hot_hosp[j] = (int) i; }
        } else atomicOr(&D.ctr->overflow, 2u);
        const int q = D.loc_pub[loc];
        if (q >= 0) atomicOr(&D.pub_hot[q >> 5], 1u << (q & 31));
    }
}

// Community::public_activity draws for one agent (Community.cpp:1012-1026): k ~ Poisson(1) from the first word of
// the agent's Philox block, then k uniform picks among its patronised locations. Returns the number of distinct
// locations; qs[] holds public-location indices, hours[] the multiplicities.
// The agent's 32-byte patron row arrives in (ra, rb): the caller loads it together with the rest of the agent's state.
__device__ __forceinline__ int gen_visits(const Dev& D, const DayCtx& cx, unsigned int pid, int npat, const uint4 ra, const uint4 rb, int* qs, int* hours) {
    uint4 b = philox4x32_10(make_uint4(0u, pid, 0u, ((unsigned int) cx.day << 8) | CABM_P_VISIT), make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32)));
    int k;
    if (b.x < D.visit_thr[0]) return 0;
    else if (b.x < D.visit_thr[1]) k = 1;
    else if (b.x < D.visit_thr[2]) k = 2;
    else if (b.x < D.visit_thr[3]) k = 3;
    else k = (int) table_draw(D.par->visit_cdf, CABM_T_POISSON, (double) b.x * (1.0 / 4294967296.0));
    int nv = 0;
    const bool avoiding = (double) __uint_as_float(rb.y) < cx.sd;      // row[PATRON_RISK]
    for (int i = 0; i < k; ++i) {
        const unsigned int idx1 = (unsigned int) i + 1u;              // draw index within the stream (0 was the Poisson draw)
        if ((idx1 & 3u) == 0) b = philox4x32_10(make_uint4(idx1 >> 2, pid, 0u, ((unsigned int) cx.day << 8) | CABM_P_VISIT), make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32)));
        const unsigned int lane = idx1 & 3u;
        const unsigned int x = lane == 0 ? b.x : lane == 1 ? b.y : lane == 2 ? b.z : b.w;
        const unsigned int j = __umulhi(x, (unsigned int) npat);
        const int e = (int) (j == 0 ? ra.x : j == 1 ? ra.y : j == 2 ? ra.z : j == 3 ? ra.w : rb.x);
        if ((e & PATRON_HIGH) && avoiding) continue;
        const int q = e & 0x3FFFFFFF;
        int t = 0; while (t < nv && qs[t] != q) ++t;
        if (t == nv) { qs[nv] = q; hours[nv] = 1; ++nv; } else hours[t]++;
    }
    return nv;
}

// ---------------------------------------------------------------------------------------------
// k_active: one thread per active agent.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void k_active_body(const Dev& D) {
    const DayCtx& cx = day_ctx(D); const int day = cx.day;
    const unsigned int n = min(D.ctr->n_active, D.cap_active);
    unsigned int nflag = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned int p = D.active_list[i];
        const Tl0 t0 = D.tl0[p]; const Tl1 t1 = D.tl1[p];
        const unsigned char pb = D.pbyte[p];
        const int hm = D.home[p], dl = D.day_loc[p];
        const uint4* __restrict__ row4 = reinterpret_cast<const uint4*>(D.patron + (size_t) p * PATRON_ROW);
        const uint4 ra = row4[0], rb = row4[1];                                    // the patron row, in flight with the rest of the agent's state
        const bool inhosp = tl_in(t0.hosp_begin, t0.severe_end, day);
        const unsigned char npb = (unsigned char) ((pb & ~PB_INHOSP) | (inhosp ? PB_INHOSP : 0));
        if (npb != pb) D.pbyte[p] = npb;
        if (!tl_in(t0.infectious_begin, t0.infectious_end, day)) continue;
        // where this agent is infectious today (Person.cpp:475-486, Community.cpp:1031-1035): up to 2 + MAXV places
        int locs[2 + MAXV]; unsigned long long cs[2 + MAXV]; int nl = 0;
        const unsigned long long c = (unsigned long long) __double2ll_rn(CABM_MUL((double) t1.rel_inf, 4294967296.0));
        if (inhosp) { const int h = D.loc_hospital[hm]; if (h >= 0 && c) { locs[0] = h; cs[0] = c; nl = 1; } }
        else {
            if (c) { locs[0] = hm; cs[0] = c; nl = 1; }
            if (dl >= 0 && c && !(tl_in(t0.severe_begin, t0.severe_end, day) || tl_in(t1.critical_begin, t1.critical_end, day))) { locs[nl] = dl; cs[nl] = c; nl++; }
            const int npat = pb & PB_NPAT;
            if (npat) {
                int qs[MAXV], hours[MAXV];
                const int nv = gen_visits(D, cx, p, npat, ra, rb, qs, hours);
                for (int j = 0; j < nv; ++j) {
                    const unsigned long long cv = (unsigned long long) __double2ll_rn(CABM_MUL(CABM_MUL((double) t1.rel_inf, (double) hours[j]), 536870912.0));
                    if (cv) { locs[nl] = D.pub_loc[qs[j]]; cs[nl] = cv; nl++; }
                }
            }
        }
        if (nl == 0) continue;
        // Community::flagInfectedLocation (Community.cpp:788-791) for all places of the agent at once: the static loads, the
        // pressure REDs, the per-location slot atomics and the flag-list reservation are issued back to back (one round trip),
        // then one reservation covers the hot-list entries of every location this agent was the first to flag
        unsigned int slots[2 + MAXV]; int ltype[2 + MAXV], lpub[2 + MAXV];
        cg::coalesced_group g = cg::coalesced_threads();
        const unsigned int mine = (unsigned int) nl;
        for (int k = 0; k < nl; ++k) { ltype[k] = D.loc_type[locs[k]]; lpub[k] = D.loc_pub[locs[k]]; }
        for (int k = 0; k < nl; ++k) {
            atomicAdd(&D.press[(size_t) locs[k] * CABM_N_STRAINS + t1.strain], cs[k]);
            slots[k] = atomicAdd(&D.loc_cnt[locs[k]], 1u);
        }
        unsigned int nfirst = 0, nfh = 0;
        for (int k = 0; k < nl; ++k) if (slots[k] == 0) { nfirst++; nfh += ltype[k] == CABM_HOSPITAL ? 1u : 0u; }
        // flag records, hot-list entries and hot hospitals reserved together: the three counters are hit by every warp of the
        // grid, and in a large epidemic their round trips (same-address atomics queue up) were a quarter of the kernel
        unsigned int base, hbase, hhbase;
        list_reserve3(&D.ctr->n_flags, mine, &D.ctr->n_hot, nfirst, &D.ctr->n_hot_hosp, nfh, base, hbase, hhbase);
        for (int k = 0; k < nl; ++k) {
            const unsigned int f = base + k;
            if (f < D.cap_flags) { FlagRaw r; r.loc = locs[k]; r.slot = slots[k]; r.pid = p; r.pad = 0; r.c = cs[k] | ((unsigned long long) t1.strain << 56); D.flags_raw[f] = r; }
            else atomicOr(&D.ctr->overflow, 128u);
            if (slots[k] == 0) {
                const unsigned int i = hbase++;
                if (i < D.cap_hot) {
                    D.hot_list[i] = locs[k]; D.loc_hot[locs[k]] = (int) i;
                    if (ltype[k] == CABM_HOSPITAL) { const unsigned int j = hhbase++; if (j < D.cap_hot) D.hot_hosp[j] = (int) i; }
                } else atomicOr(&D.ctr->overflow, 2u);
                if (lpub[k] >= 0) atomicOr(&D.pub_hot[lpub[k] >> 5], 1u << (lpub[k] & 31));
            }
        }
        nflag += mine;
    }
    nflag = warp_sum(nflag);
    if ((threadIdx.x & 31) == 0 && nflag) atomicAdd(&D.ctr->work[CABM_W_FLAGS], (unsigned long long) nflag);
}
__global__ void __launch_bounds__(TPB) k_active(const Dev D) { k_active_body(D); }
__global__ void __launch_bounds__(TPB) k_active_b(const Dev* __restrict__ devs) { const Dev D = devs[blockIdx.y]; k_active_body(D); }   // one launch, replicate = blockIdx.y

// ---------------------------------------------------------------------------------------------
// k_visit: every agent with patronised locations that is not in hospital draws its visits.
// A warp takes tiles of 128 consecutive agents, four per lane. Per agent the only memory on the path is the status byte
// and, when k > 0, the agent's 32-byte patron row (5 public-location indices + the household's riskiness); the Philox
// block needs no memory. Hotness is tested against a shared-memory bitmap over the Q patronised locations (dynamic shared
// memory, ceil(Q/32) words). Visits that land on a hot location are staged in the warp's shared-memory buffer and flushed
// once per tile: one reservation, coalesced record stores, one census RED per record. Agents that draw k >= 4 (1.9 %) are
// queued per warp and handled 32 at a time by the generic loop, so the common path carries no divergent tail.
// There is no block-level synchronisation in the loop.
// ---------------------------------------------------------------------------------------------
#ifndef VIS_A
#define VIS_A 4               // agents per lane and tile: a warp takes tiles of 32 * VIS_A consecutive agents
#endif
#ifndef VIS_MINB
#define VIS_MINB 4            // resident blocks per SM the register allocation aims at
#endif
#define VIS_TILE (32 * VIS_A)
#define VIS_STAGE (4 * VIS_TILE < 512 ? 512 : 4 * VIS_TILE)   // records a warp can stage: a tile adds at most 3 * VIS_TILE, the final k >= 4 pass at most 320
#define VIS_FLUSH_AT (VIS_STAGE - 3 * VIS_TILE < 128 ? VIS_STAGE - 3 * VIS_TILE : 128)   // flush once more than this is staged after a tile
#define VIS_SLOWQ (4 * TPB)   // k >= 4 agents a block queues for its final full-warp pass (beyond that: handled in place)

// a 32-bit load that is issued only when `on` is set (0 otherwise): keeps k_visit's body free of branches
__device__ __forceinline__ unsigned int ldg_if(const int* ptr, unsigned int on) {
    unsigned int v = 0;
    asm("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p ld.global.nc.u32 %0, [%1];\n\t}" : "+r"(v) : "l"(ptr), "r"(on));
    return v;
}

// the generic public_activity loop for one agent that drew k >= 4 (Community.cpp:1012-1026), one lane per agent.
// s_thr[j] = ceil(visit_cdf[j] * 2^32): u01 >= cdf  <=>  x >= s_thr (64-bit compare, exact). The whole patron row comes with two
// 16-byte loads, so the agent's dependent chain is one memory round trip. STAGED: hot visits go to the warp's staging buffer
// (flushed by the caller); otherwise straight to the global record list.
template <bool STAGED>
__device__ __forceinline__ void visit_slow(const Dev& D, const unsigned int* s_hot, const unsigned long long* s_thr, const unsigned int p, const unsigned int npat,
                                           const unsigned int dp, const uint2 key, const float sd_up, uint2* stage, unsigned int* n_stage, unsigned int& w_v0, unsigned int& w_v) {
    const uint4* __restrict__ row4 = reinterpret_cast<const uint4*>(D.patron + (size_t) p * PATRON_ROW);
    const uint4 ra = row4[0], rb = row4[1];
    uint4 bb = philox4x32_10(make_uint4(0u, p, 0u, dp), key);
    unsigned int kk = 0;
    while (kk < CABM_T_POISSON - 1 && (unsigned long long) bb.x >= s_thr[kk]) ++kk;          // table_draw on shared-memory thresholds
    const bool avoiding = __uint_as_float(rb.y) < sd_up;                                  // row[PATRON_RISK]
    unsigned int qs[5]; unsigned int nv = 0;
    for (unsigned int i = 1; i <= kk; ++i) {
        if ((i & 3u) == 0) bb = philox4x32_10(make_uint4(i >> 2, p, 0u, dp), key);
        const unsigned int x = (i & 3u) == 0 ? bb.x : (i & 3u) == 1 ? bb.y : (i & 3u) == 2 ? bb.z : bb.w;
        const unsigned int j = __umulhi(x, npat);
        const unsigned int e = j == 0 ? ra.x : j == 1 ? ra.y : j == 2 ? ra.z : j == 3 ? ra.w : rb.x;
        if (avoiding && (e & PATRON_HIGH)) continue;
        const unsigned int q = e & 0x3FFFFFFFu;
        bool seen = false;
#pragma unroll
        for (int t = 0; t < 5; ++t) if (t < (int) nv && qs[t] == q) seen = true;
        if (seen) continue;
#pragma unroll
        for (int t = 0; t < 5; ++t) if (t == (int) nv) qs[t] = q;
        nv++; w_v0++;
        if (s_hot[q >> 5] & (1u << (q & 31))) {
            if (STAGED) stage[atomicAdd(n_stage, 1u)] = make_uint2(p, q);
            else {
                atomicAdd(&D.nvis[q], 1u);
                const unsigned int vi = list_reserve(&D.ctr->n_vis);
                if (vi < D.cap_vis) D.visit_rec[vi] = make_uint2(p, q); else atomicOr(&D.ctr->overflow, 8u);
                w_v++;
            }
        }
    }
}

// Community::public_activity draws of one agent in 0/1 arithmetic (no branch, predicated loads): k ~ Poisson(1) and the
// first three picks come from one Philox block (Community.cpp:1013-1017). v0..v2 = pick j is a visit (k > j, not avoided,
// not a repeat of an earlier pick), q0..q2 = its public-location index, c3 = the agent drew k >= 4 (generic loop instead).
__device__ __forceinline__ void visit_draw(const Dev& D, const unsigned int p, const unsigned int npat, const unsigned int act, const unsigned int dp, const uint2 key,
                                           const float sd_up, const unsigned int thr0, const unsigned int thr1, const unsigned int thr2, const unsigned int thr3,
                                           unsigned int& q0, unsigned int& q1, unsigned int& q2, unsigned int& v0, unsigned int& v1, unsigned int& v2, unsigned int& c3) {
    const uint4 b = philox4x32_10(make_uint4(0u, p, 0u, dp), key);
    c3 = act & (unsigned int) (b.x >= thr3);
    const unsigned int live = act ^ c3;
    v0 = live & (unsigned int) (b.x >= thr0); v1 = live & (unsigned int) (b.x >= thr1); v2 = live & (unsigned int) (b.x >= thr2);
    const int* __restrict__ row = D.patron + (size_t) p * PATRON_ROW;
    const unsigned int e0 = ldg_if(row + __umulhi(b.y, npat), v0), e1 = ldg_if(row + __umulhi(b.z, npat), v1), e2 = ldg_if(row + __umulhi(b.w, npat), v2);
    const unsigned int avoiding = (unsigned int) (__uint_as_float(ldg_if(row + PATRON_RISK, v0)) < sd_up);   // Community.cpp:1020: compliant households skip high-risk places
    v0 &= (avoiding & (e0 >> 30)) ^ 1u; v1 &= (avoiding & (e1 >> 30)) ^ 1u; v2 &= (avoiding & (e2 >> 30)) ^ 1u;      // bit 30 = PATRON_HIGH
    q0 = e0 & 0x3FFFFFFFu; q1 = e1 & 0x3FFFFFFFu; q2 = e2 & 0x3FFFFFFFu;
    v2 &= ((v0 & (unsigned int) (q2 == q0)) | (v1 & (unsigned int) (q2 == q1))) ^ 1u;   // visit_map (Community.cpp:1014): one entry per distinct place
    v1 &= (v0 & (unsigned int) (q1 == q0)) ^ 1u;
}

// flush a warp's staged records: one reservation, coalesced record stores, one census RED per record
__device__ __forceinline__ void visit_flush(const Dev& D, const uint2* stage, unsigned int* n_stage, const unsigned int total, const unsigned int lane, unsigned int& w_v) {
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(&D.ctr->n_vis, total);
    base = __shfl_sync(0xffffffffu, base, 0);
    for (unsigned int i = lane; i < total; i += 32) {
        const uint2 r = stage[i];
        atomicAdd(&D.nvis[r.y], 1u);
        if (base + i < D.cap_vis) D.visit_rec[base + i] = r; else atomicOr(&D.ctr->overflow, 8u);
    }
    __syncwarp();
    if (lane == 0) { *n_stage = 0; w_v += total; }
    __syncwarp();
}

__device__ __forceinline__ void k_visit_body(const Dev& D, const unsigned int bitmap_words) {
    extern __shared__ unsigned int s_hot[];
    __shared__ uint2 s_stage[TPB / 32][VIS_STAGE];
    __shared__ unsigned int s_slowq[VIS_SLOWQ];                               // agent | patron count << 29
    __shared__ unsigned long long s_thr[CABM_T_POISSON];
    __shared__ unsigned int s_nslow, s_nstage[TPB / 32];
    const DayCtx& cx = day_ctx(D);
    const unsigned int stride = gridDim.x * blockDim.x * VIS_A;
    const unsigned int t_base = D.own_lo & ~(VIS_TILE - 1u);                        // tiles cover the owned agents only
    const unsigned int n_round = t_base + (D.own_hi - t_base + stride - 1) / stride * stride;
    const unsigned int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const unsigned int own_lo = D.own_lo, own_n = D.own_hi - D.own_lo, N = D.N;
    // in slot a lane l handles agent tile + 32 a + l, so the 32-byte patron rows a load instruction touches are
    // consecutive (8 cache lines per request instead of 32)
    const unsigned int t_first = t_base + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * VIS_TILE;
    // the first tile's status bytes and the day's scalars are requested before the block stages the hot bitmap: one memory
    // round trip for the whole prologue instead of three
    unsigned int pb_next = 0;
#pragma unroll
    for (int a = 0; a < VIS_A; ++a) { const unsigned int p = t_first + a * 32 + lane; if (p < N) pb_next |= (unsigned int) D.pbyte[p] << (8 * a); }
    const unsigned int dp = ((unsigned int) cx.day << 8) | CABM_P_VISIT;
    const float sd_up = cx.sd_up;                                                   // (double) risk < sd  <=>  risk < sd_up (fill_ctx)
    for (unsigned int i = threadIdx.x; i < bitmap_words; i += blockDim.x) s_hot[i] = D.pub_hot[i];
    if (threadIdx.x < TPB / 32) s_nstage[threadIdx.x] = 0;
    if (threadIdx.x == 0) s_nslow = 0;
    if (threadIdx.x < CABM_T_POISSON) s_thr[threadIdx.x] = __double2ull_ru(D.par->visit_cdf[threadIdx.x] * 4294967296.0);
    __syncthreads();
    const uint2 key = make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32));
    const unsigned int thr0 = D.visit_thr[0], thr1 = D.visit_thr[1], thr2 = D.visit_thr[2], thr3 = D.visit_thr[3];
    uint2* const stage = s_stage[wid];
    unsigned int w_pat = 0, w_v0 = 0, w_v = 0;
    for (unsigned int t0 = t_first; t0 < n_round; t0 += stride) {
        // the status bytes were loaded one tile ahead (an L2 prefetch of the next tile's patron rows was measured: no gain)
        const unsigned int pb4 = pb_next;
        pb_next = 0;
        if (t0 + stride < n_round) {
#pragma unroll
            for (int a = 0; a < VIS_A; ++a) { const unsigned int p = t0 + stride + a * 32 + lane; if (p < N) pb_next |= (unsigned int) D.pbyte[p] << (8 * a); }
        }
        unsigned int recq[3 * VIS_A]; unsigned int recmask = 0, slow = 0;          // hot visits of this lane's agents (k <= 3)
#pragma unroll
        for (int a = 0; a < VIS_A; ++a) {
            const unsigned int pb = (pb4 >> (8 * a)) & 0xFFu, p = t0 + a * 32 + lane;
            const unsigned int npat = pb & PB_NPAT;
            const unsigned int act = (unsigned int) (npat != 0) & (unsigned int) ((pb & PB_INHOSP) == 0) & (unsigned int) ((p - own_lo) < own_n);   // another handle draws for agents it owns
            unsigned int q0, q1, q2, v0, v1, v2, c3;
            visit_draw(D, p, npat, act, dp, key, sd_up, thr0, thr1, thr2, thr3, q0, q1, q2, v0, v1, v2, c3);
            slow |= c3 << a;
            w_pat += act;
            w_v0 += v0 + v1 + v2;
            const unsigned int h0 = v0 & (s_hot[q0 >> 5] >> (q0 & 31)), h1 = v1 & (s_hot[q1 >> 5] >> (q1 & 31)), h2 = v2 & (s_hot[q2 >> 5] >> (q2 & 31));
            recq[a * 3 + 0] = q0; recq[a * 3 + 1] = q1; recq[a * 3 + 2] = q2;
            recmask |= (h0 | (h1 << 1) | (h2 << 2)) << (a * 3);
        }
        // stage the tile's records in the warp buffer (offset from a shared-memory counter); flush when the next tile might not fit
        const unsigned int nrec = __popc(recmask);
        unsigned int o = 0;
        if (nrec) o = atomicAdd(&s_nstage[wid], nrec);
#pragma unroll
        for (int r = 0; r < 3 * VIS_A; ++r) if (recmask & (1u << r)) { stage[o] = make_uint2(t0 + (r / 3) * 32 + lane, recq[r]); ++o; }
        // k >= 4 (1.9 % of the drawing agents): queued for the block's final pass, so the generic loop runs on full warps
        while (slow) {
            const unsigned int a = (unsigned int) __ffs(slow) - 1u; slow &= slow - 1u;
            const unsigned int pos = atomicAdd(&s_nslow, 1u), p = t0 + a * 32 + lane, npat = (pb4 >> (8 * a)) & PB_NPAT;
            if (pos < VIS_SLOWQ) s_slowq[pos] = p | (npat << 29); else visit_slow<false>(D, s_hot, s_thr, p, npat, dp, key, sd_up, nullptr, nullptr, w_v0, w_v);
        }
        __syncwarp();
        const unsigned int staged = __shfl_sync(0xffffffffu, *(volatile unsigned int*) &s_nstage[wid], 0);   // one value for the whole warp: the flush is collective
        if (staged > VIS_FLUSH_AT) visit_flush(D, stage, &s_nstage[wid], staged, lane, w_v);
    }
    // the block's k >= 4 agents on full warps, their hot visits staged with the warp's leftovers, then one final flush per warp
    __syncthreads();
    {
        const unsigned int n = min(s_nslow, (unsigned int) VIS_SLOWQ);
        for (unsigned int i0 = wid * 32; i0 < n; i0 += blockDim.x) {                  // warp-uniform trip count; a round stages at most 160 records
            if (i0 + lane < n) { const unsigned int e = s_slowq[i0 + lane]; visit_slow<true>(D, s_hot, s_thr, e & 0x1FFFFFFFu, e >> 29, dp, key, sd_up, stage, &s_nstage[wid], w_v0, w_v); }
            __syncwarp();
            const unsigned int staged = __shfl_sync(0xffffffffu, *(volatile unsigned int*) &s_nstage[wid], 0);
            if (staged > VIS_FLUSH_AT) visit_flush(D, stage, &s_nstage[wid], staged, lane, w_v);
        }
    }
    __syncwarp();
    { const unsigned int staged = __shfl_sync(0xffffffffu, *(volatile unsigned int*) &s_nstage[wid], 0); if (staged) visit_flush(D, stage, &s_nstage[wid], staged, lane, w_v); }
    const int ids[3] = { CABM_W_PATRONS, CABM_W_VISITS, CABM_W_HOT_VISITS }; const unsigned int vals[3] = { w_pat, w_v0, w_v };
    block_work_add<3>(D, ids, vals);
}
__global__ void __launch_bounds__(TPB, VIS_MINB) k_visit(const Dev D, const unsigned int bitmap_words) { k_visit_body(D, bitmap_words); }
__global__ void __launch_bounds__(TPB, VIS_MINB) k_visit_b(const Dev* __restrict__ devs, const unsigned int bitmap_words) { const Dev D = devs[blockIdx.y]; k_visit_body(D, bitmap_words); }   // one launch, replicate = blockIdx.y

__device__ __forceinline__ double hazard_to_T(double hazard) { return CABM_SUB(1.0, cabm_exp(-hazard)); }

// the susceptibility gate of one successful trial (cabm_kernels2.cuh)
__device__ __forceinline__ void gate_hit(const Dev& D, const int phase, const unsigned int pid, const unsigned int src, const unsigned int purpose, const unsigned int sub);

// one exposure trial: the quarantine skip and `gsl_rng_uniform(RNG) < T` of Community::_transmission (Community.cpp:975-976).
// A success of the house..nursing-home passes is queued for the susceptibility gate (k_gate); the hospital pass gates in place.
__device__ __forceinline__ void trial(const Dev& D, const int day, const int phase, const unsigned int pid, const int src, const int purpose,
                                      const unsigned int sub, const double T, unsigned int& trials) {
    if (pid < D.own_lo || pid >= D.own_hi) return;                                 // the owner of an agent draws its exposures
    if (D.any_quarantine) {
        const unsigned int q = D.quar[pid]; const int qf = (short) (q & 0xFFFF), qu = (short) (q >> 16);
        if (day >= qf && day < qu && src != D.home[pid]) return;
    }
    trials++;
    const uint4 b = philox4x32_10(make_uint4(sub << 16, pid, (unsigned int) src, ((unsigned int) day << 8) | (unsigned int) purpose),
                                  make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32)));
    if ((double) b.x * (1.0 / 4294967296.0) < T) {
        if (phase) { gate_hit(D, 1, pid, (unsigned int) src, (unsigned int) purpose, sub); return; }   // hospital pass: a handful of hits, gated in place
        const unsigned int h = list_reserve(&D.ctr->n_hits[0]);
        if (h < D.cap_hits) { Hit r; r.pid = pid; r.src = (unsigned int) src; r.purpose = (unsigned int) purpose; r.sub = sub; D.hits[h] = r; }
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
    const DayCtx& cx = day_ctx(D);
    const cabm_params* __restrict__ par = D.par;
    const unsigned int n_hot = min(D.ctr->n_hot, D.cap_hot);
    unsigned int edges = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_hot; i += gridDim.x * blockDim.x) {
        const int loc = D.hot_list[i];
        // everything that depends on `loc` only is loaded in one go (one round trip)
        const int type = D.loc_type[loc];
        const unsigned int cnt = D.loc_cnt[loc];
        const ulonglong2* pw = reinterpret_cast<const ulonglong2*>(&D.press[(size_t) loc * CABM_N_STRAINS]);
        const ulonglong2 pa = pw[0], pb = pw[1];
        const unsigned int rb = D.roster_off[loc], size = D.roster_off[loc + 1] - rb;
        const unsigned int nb0 = D.nbr_off[loc], deg = D.nbr_off[loc + 1] - nb0;
        const float lrisk = D.loc_risk[loc];
        const unsigned char lf = D.loc_flags[loc];
        const int q = D.loc_pub[loc];
        const unsigned long long pws[4] = { pa.x, pa.y, pb.x, pb.y };
        const double w = CABM_MUL((double) (pws[0] + pws[1] + pws[2] + pws[3]), 1.0 / 4294967296.0);
        double T = 0.0, Ts = 0.0; unsigned int denom = size; int purpose = CABM_P_EXPOSE_HH;
        unsigned int n_nb = 0;                                // small items for the neighbours (between-household pass)
        if (type == CABM_HOUSE) {
            T = hazard_to_T(CABM_MUL(CABM_MUL(par->haz_mult[CABM_HOUSE], cx.season), w));
            const bool risky = (double) lrisk > cx.sd;
            if (risky) { Ts = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(par->haz_mult[CABM_N_LOCTYPES], cx.season), w), (double) (int) size)); edges += deg; }
            if (risky && Ts > 0.0) n_nb = deg;
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
        const bool own = type != CABM_HOSPITAL && T > 0.0;
        const unsigned int n_small = n_nb + (own && size <= SMALL_ROSTER ? 1u : 0u);
        const unsigned int n_big = own && size > SMALL_ROSTER ? (size + CHUNK - 1) / CHUNK : 0u;
        // this location's segment of the sorted flag array, its small items and its big items: one round trip
        unsigned int seg, sbase, ibase;
        list_reserve3(&D.ctr->n_seg, cnt, &D.ctr->n_sitems, n_small, &D.ctr->n_items, n_big, seg, sbase, ibase);
        D.hot_off[i] = seg;
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
    edges = warp_sum(edges);
    if ((threadIdx.x & 31) == 0 && edges && D.shard_rank == 0) atomicAdd(&D.ctr->work[CABM_W_NBR_EDGES], (unsigned long long) edges);
}
__global__ void __launch_bounds__(TPB) k_hazard(const Dev D) { k_hazard_body(D); }
__global__ void __launch_bounds__(TPB) k_hazard_b(const Dev* __restrict__ devs) { const Dev D = devs[blockIdx.y]; k_hazard_body(D); }   // one launch, replicate = blockIdx.y

// ---------------------------------------------------------------------------------------------
// k_trials: block-sized units: the 256-member items of big rosters, the visit records, the flag scatter, the small items.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void k_trials_body(const Dev& D) {
    const DayCtx& cx = day_ctx(D); const int day = cx.day;
    const unsigned int n_items = min(D.ctr->n_items, D.cap_items), n_vis = min(D.ctr->n_vis, D.cap_vis);
    const unsigned int n_flags = min(D.ctr->n_flags, D.cap_flags);
    const unsigned int n_sitems = min(D.ctr->n_sitems, D.cap_sitems);
    const unsigned int u_vis = n_items + (n_vis + TPB - 1) / TPB;
    const unsigned int u_flag = u_vis + (n_flags + TPB - 1) / TPB;
    const unsigned int units = u_flag + (n_sitems + TPB / 8 - 1) / (TPB / 8);
    unsigned int trials = 0;
    for (unsigned int u = blockIdx.x; u < units; u += gridDim.x) {
        if (u >= u_flag) {                                   // small items: eight lanes per item
            const unsigned int si = (u - u_flag) * (TPB / 8) + (threadIdx.x >> 3);
            if (si < n_sitems) {
                const SItem it = D.sitems[si];
                unsigned int cnt = it.count_purpose >> 8; int begin = it.begin;
                const int purpose = (int) (it.count_purpose & 0xFFu);
                if (purpose == CABM_P_EXPOSE_NB) {                   // a neighbour household: its roster, unless it keeps its distance today
                    const int nb = D.nbr[begin];
                    const unsigned int b = D.roster_off[nb], e = D.roster_off[nb + 1];
                    cnt = (double) D.loc_risk[nb] > cx.sd ? e - b : 0u; begin = (int) b;
                }
                if (cnt) {
                    const double T = D.hot_T[it.t_idx];
                    for (unsigned int m = threadIdx.x & 7; m < cnt; m += 8) trial(D, day, 0, (unsigned int) D.roster[begin + m], it.src, purpose, 0, T, trials);
                }
            }
        } else if (u >= u_vis) {                                    // scatter the day's flags into per-location segments (offsets from k_hazard)
            const unsigned int f = (u - u_vis) * TPB + threadIdx.x;
            if (f < n_flags) {
                const FlagRaw r = D.flags_raw[f];
                Flag o; o.pid = r.pid; o.pad = 0; o.c = r.c;
                D.flags[D.hot_off[D.loc_hot[r.loc]] + r.slot] = o;
            }
        } else if (u < n_items) {
            const Item it = D.items[u];
            if ((int) threadIdx.x < it.count) trial(D, day, 0, (unsigned int) D.roster[it.begin + threadIdx.x], it.src, it.purpose, 0, it.T, trials);
        } else {
            const unsigned int v = (u - n_items) * TPB + threadIdx.x;
            if (v < n_vis) {
                const uint2 rec = D.visit_rec[v];
                const double T = D.pub_T[rec.y];
                if (T > 0.0) trial(D, day, 0, rec.x, D.pub_loc[rec.y], CABM_P_EXPOSE_WORK_VIS, 0, T, trials);
            }
        }
    }
    const int ids[1] = { CABM_W_TRIALS }; const unsigned int vals[1] = { trials };
    block_work_add<1>(D, ids, vals);
}
__global__ void __launch_bounds__(TPB) k_trials(const Dev D) { k_trials_body(D); }
__global__ void __launch_bounds__(TPB) k_trials_b(const Dev* __restrict__ devs) { const Dev D = devs[blockIdx.y]; k_trials_body(D); }   // one launch, replicate = blockIdx.y
