// covid-abm_b200/csrc/cabm_kernels2.cuh -- second half of the day pipeline: susceptibility gate, winner
// application fused with the status pass, hospital pass + importation, end of day.
//
//   gate_hit        a successful trial : calculate_event_probabilities + the Person::infect bail-out (Person.cpp:322-326)
//                                     against the start-of-phase state; survivors race with atomicMin on the
//                                     agent's 64-bit order key, so the first exposure in reference order wins
//                                     (called by k_trials at the end of each block, by the hospital pass and the
//                                     importation in place)
//   k_apply_status  winners + active: Person::infect for each winner (disease course, detection, ledger writes) and,
//                                     for phase 0, Community::updatePersonStatus (Community.cpp:682-741) over the day's
//                                     active list. The phase-1 launch is the last kernel of the day: its last block
//                                     closes the day (hospital scratch, counters, list swap, next day context)
//   k_hosp_import   hospitals/imports: hospital_transmission (Community.cpp:931-941) after the status pass, and
//                                     the daily importation draws of seed_epidemic (simulator.h:201-223); beside them
//                                     updateHotLocations + clear_public_activity (Community.cpp:999-1003,1042-1046) for
//                                     every hot location that is not a hospital (nothing reads their scratch any more)
#pragma once
#include "cabm_kernels.cuh"

// covid::util::weighted_choice over the strain urn (Utility.h:93-105, simulator.h:215)
__device__ __forceinline__ int weighted_strain(const double* __restrict__ w, double u) {
    const double sum = CABM_ADD(CABM_ADD(CABM_ADD(w[0], w[1]), w[2]), w[3]);
    double ran = CABM_MUL(sum, u);
    for (int k = 0; k < CABM_N_STRAINS; ++k) { if (ran < w[k]) return k; ran = CABM_SUB(ran, w[k]); }
    return CABM_N_STRAINS - 1;
}

// Community::trace_contact (Community.cpp:795-825): infector ~ relInf-weighted among the agents that flagged the source
// location today. The same draw that picked the strain picks the infector: within the strain's share the flags are laid
// out by ascending person id and t is the offset into that share. The location's flags sit in one contiguous segment in
// arrival order, so this is a weighted quickselect on person id (result independent of the arrival order).
// Short segments (households, small workplaces) are searched by the winner's own thread ...
__device__ __forceinline__ int pick_infector_thread(const Flag* __restrict__ seg, const unsigned int n, const int strain, unsigned long long t) {
    unsigned int lo = 0, hi = 0xFFFFFFFFu;
    for (;;) {
        bool have = false; unsigned int pivot = 0; unsigned long long lower = 0, eq = 0;
        for (unsigned int k = 0; k < n; ++k) {
            const Flag e = seg[k];
            if (FLAG_STRAIN(e) != strain || e.pid < lo || e.pid > hi) continue;
            if (!have) { have = true; pivot = e.pid; }
            if (e.pid < pivot) lower += FLAG_C(e); else if (e.pid == pivot) eq += FLAG_C(e);
        }
        if (!have) return -1;
        if (t < lower) hi = pivot - 1;                      // lower > 0 implies pivot > lo >= 0
        else if (t - lower < eq) return (int) pivot;
        else { t -= lower + eq; if (pivot == 0xFFFFFFFFu) return -1; lo = pivot + 1; }
    }
}
__device__ __forceinline__ unsigned long long warp_sum64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// ... long ones (hospitals, big workplaces and schools at the peak) by the whole warp: one strided sweep per pivot,
// the next pivot is any element seen on the side the search continues in. All 32 lanes call with the same arguments.
__device__ __forceinline__ int pick_infector_warp(const Flag* __restrict__ seg, const unsigned int n, const int strain, unsigned long long t) {
    const unsigned int lane = threadIdx.x & 31;
    unsigned int lo = 0, hi = 0xFFFFFFFFu, pivot;
    {
        unsigned int cand = 0; bool h = false;
        for (unsigned int k = lane; k < n && !h; k += 32) { const Flag e = seg[k]; if (FLAG_STRAIN(e) == strain) { cand = e.pid; h = true; } }
        const unsigned int b = __ballot_sync(0xffffffffu, h);
        if (!b) return -1;
        pivot = __shfl_sync(0xffffffffu, cand, __ffs(b) - 1);
    }
    for (;;) {
        unsigned long long lower = 0, eq = 0; unsigned int cl = 0, ch = 0; bool hl = false, hh = false;
        for (unsigned int k = lane; k < n; k += 32) {
            const Flag e = seg[k];
            if (FLAG_STRAIN(e) != strain || e.pid < lo || e.pid > hi) continue;
            if (e.pid < pivot) { lower += FLAG_C(e); cl = e.pid; hl = true; } else if (e.pid == pivot) eq += FLAG_C(e); else { ch = e.pid; hh = true; }
        }
        lower = warp_sum64(lower); eq = warp_sum64(eq);
        if (t < lower) { hi = pivot - 1; const unsigned int b = __ballot_sync(0xffffffffu, hl); pivot = __shfl_sync(0xffffffffu, cl, __ffs(b) - 1); }
        else if (t - lower < eq) return (int) pivot;
        else {
            t -= lower + eq;
            const unsigned int b = __ballot_sync(0xffffffffu, hh);
            if (!b) return -1;
            lo = pivot + 1; pivot = __shfl_sync(0xffffffffu, ch, __ffs(b) - 1);
        }
    }
}
#define INFECTOR_THREAD_MAX 32

// rebuild the event (stream position, strain, keys) of a hit / winner; returns false for malformed input
__device__ __forceinline__ void event_setup(const Dev& D, const DayCtx& cx, const unsigned int pid, const unsigned int src, const unsigned int purpose,
                                            const unsigned int sub, Stream& s, InfectCtx& c, unsigned long long& t_res) {
    c.pid = (int) pid; c.time = cx.day; c.infector = -1; c.purpose = purpose; c.key_sub = sub;
    if (purpose == CABM_P_IMPORT) {
        s.begin(D.seed, (unsigned int) cx.day, CABM_P_IMPORT, src, 0);
        (void) s.next();                                            // the draw that chose the person
        c.strain = weighted_strain(cx.strain_w, s.u01());
        c.source = -1; c.has_source = false; c.key_a = src; c.key_b = 0;
    } else {
        s.begin(D.seed, (unsigned int) cx.day, purpose, pid, src, sub);
        (void) s.next();                                            // the U < T draw
        c.strain = pick_strain(D.locd[src].press, s.next(), &t_res);
        (void) s.next();                                            // reserved (reference: uniform_choice inside the group)
        c.source = (int) src; c.has_source = true; c.key_a = pid; c.key_b = src;
    }
}

__device__ __forceinline__ unsigned long long order_key(const unsigned int purpose, const unsigned int src, const unsigned int sub) {
    int pass = (int) purpose, role = 0;
    if (purpose == CABM_P_EXPOSE_WORK_VIS) { pass = PASS_WORK; role = 0; }
    else if (purpose == CABM_P_EXPOSE_WORK) { pass = PASS_WORK; role = 1; }
    else if (purpose == CABM_P_EXPOSE_HOSP_PAT) { pass = PASS_HOSP; role = 1 + (int) sub; }
    return make_key(pass, src, (unsigned int) role);
}

// calculate_event_probabilities + the Person::infect bail-out (Person.cpp:322-326) for one successful trial or import draw,
// against the start-of-phase state; a survivor races for its agent with atomicMin on the 64-bit order key, and the first
// survivor of an agent puts it on the phase's winner list
__device__ __forceinline__ void gate_hit(const Dev& D, const int phase, const unsigned int pid, const unsigned int src, const unsigned int purpose, const unsigned int sub) {
    const DayCtx& cx = day_ctx(D);
    Stream s; InfectCtx c; unsigned long long t_res = 0;
    event_setup(D, cx, pid, src, purpose, sub, s, c, t_res);
    if (infect_event(D, s, c, false)) {
        const unsigned long long old = atomicMin(&D.cand[pid], order_key(purpose, src, sub));
        if (old == CABM_CAND_EMPTY) {
            const unsigned int nbase = phase ? min(D.ctr->n_new[0], D.cap_new) : 0u;
            const unsigned int ni = nbase + list_reserve(&D.ctr->n_new[phase]);
            if (ni < D.cap_new) D.newinf_list[ni] = pid; else atomicOr(&D.ctr->overflow, 32u);
        }
    }
}

// k_gate: the hits of the house..nursing-home passes (k_trials) pass the susceptibility gate, one thread per hit
__device__ __forceinline__ void k_gate_body(const Dev& D) {
    __shared__ unsigned int s_pref[HIT_LISTS + 1];
    pdl_wait();
    // the sub-lists end to end: prefix sums of their lengths (64 loads, one per thread), then hit i of the concatenation
    if (threadIdx.x < HIT_LISTS) {
        const unsigned int lane = threadIdx.x & 31;
        unsigned int incl = min(D.hit_ctr[threadIdx.x * 32u], D.hit_sub);
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (unsigned int) o) incl += t; }
        s_pref[threadIdx.x + 1] = incl;                     // per warp of 32 sub-lists; the second warp's base is added below
    }
    if (threadIdx.x == 0) s_pref[0] = 0;
    __syncthreads();
    if (threadIdx.x >= 32 && threadIdx.x < HIT_LISTS) { const unsigned int add = s_pref[32]; __syncwarp(); s_pref[threadIdx.x + 1] += add; }
    __syncthreads();
    const unsigned int n = s_pref[HIT_LISTS];
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        unsigned int k = 0;
#pragma unroll
        for (unsigned int step = HIT_LISTS / 2; step; step >>= 1) if (s_pref[k + step] <= i) k += step;
        const Hit h = D.hits[(size_t) k * D.hit_sub + (i - s_pref[k])];
        gate_hit(D, 0, h.pid, h.src, h.purpose, h.sub);
    }
    pdl_launch_dependents();
}
__global__ void __launch_bounds__(128) k_gate(const Dev D) { k_gate_body(D); }
__global__ void __launch_bounds__(128) k_gate_b(const Dev* __restrict__ devs) { BATCH_DEV(D, devs) k_gate_body(D); }   // one launch, replicate = blockIdx.y

// Community::updatePersonStatus (Community.cpp:682-741) for one agent; returns the ledger increments as a bit mask
__device__ __forceinline__ unsigned int status_one(const Dev& D, unsigned int p, int day) {
    const AgRec ar = ag_load(D.ag, p);                                             // timeline, day location, hospital, admission count: one line
    const Tl0 t0 = ar.t0; const Tl1 t1 = ar.t1;
    const int hosp = ar.hospital; const int dl = ar.day_loc;
    const bool inh = tl_in(t0.hosp_begin, t0.severe_end, day);
    if (hosp >= 0 && dl != hosp) {
        if (inh) {
            if (t0.hosp_begin == day) {                                             // goToHospital (Person.h:248)
                D.ag[p].hosp_count = ar.hosp_count + 1; atomicAdd(&D.pat_count[hosp], 1u);
                const unsigned char pb = ar.pbyte;
                if (!(pb & PB_ONLIST)) {                                            // one entry per stay: the patient walk drops it once hosp_count is back to 0
                    D.ag[p].pbyte = pb | PB_ONLIST;
                    const unsigned int pi = list_reserve(&D.ctr->n_pat);
                    if (pi < D.cap_pat) D.patient_list[pi] = p; else atomicOr(&D.ctr->overflow, 64u);
                }
            }
        } else if (day > 0 && tl_in(t0.hosp_begin, t0.severe_end, day - 1)) {     // leaveHospital (Person.h:249, Location.cpp:58-67)
            const unsigned char c = ar.hosp_count;
            if (c > 0) { D.ag[p].hosp_count = c - 1; atomicSub(&D.pat_count[hosp], 1u); }
        }
    }
    unsigned int m = 0;
    if ((int) p < D.par->num_surveilled_people) {
        if (t0.infected == day) m |= (1u << CABM_C_NEW_INFECTED) | (1u << (CABM_C_NEW_INF_STRAIN0 + t1.strain));
        if (t0.symptom_begin == day) { m |= 1u << CABM_C_NEW_SYMPTOMATIC; if (D.vax[p] & 0xFF) m |= 1u << CABM_C_VACCINATED_CASES; }
        if (tl_in(t0.severe_begin, t0.severe_end, day)) m |= 1u << CABM_C_SEVERE_PREV;
        if (inh) {
            m |= 1u << CABM_C_HOSP_PREV;
            if (t0.hosp_begin == day) m |= 1u << CABM_C_HOSP_INC;
            if (tl_in(t1.icu_begin, t1.critical_end, day)) { m |= 1u << CABM_C_ICU_PREV; if (t1.icu_begin == day) m |= 1u << CABM_C_ICU_INC; }
        }
        if (t0.severe_begin == day) m |= 1u << CABM_C_NEW_SEVERE;
        if (t1.critical_begin == day) m |= 1u << CABM_C_NEW_CRITICAL;
        if (t1.death == day) m |= 1u << CABM_C_NEW_DEAD;
    }
    return m;
}

// Closing the day: runs in the last block of the day's last kernel (k_apply_status, phase 1) once every other block has
// finished. tick_done: a whole tick ran (updateHotLocations for the hot hospitals, which the hospital pass and the phase-1
// winners were still reading; list swap); advance: move on to the next staged day context.
__device__ __forceinline__ void close_day(const Dev& D, const int tick_done, const int advance, const int host_row) {
    Counters* c = D.ctr;
    const unsigned int n_hot = min(c->n_hot, D.cap_hot);
    for (unsigned int k = threadIdx.x; k < HIT_LISTS; k += blockDim.x) D.hit_ctr[k * 32u] = 0;   // the hit sub-lists are empty again
    if (tick_done) {
        const unsigned int nh = min(c->n_hot_hosp, D.cap_hot);
        for (unsigned int j = threadIdx.x; j < nh; j += blockDim.x) {
            const int loc = D.hot_list[D.hot_hosp[j]];
            locd_reset(&D.locd[loc]);
            const int q = D.loc_pub[loc];
            if (q >= 0) { D.nvis[q] = 0; D.pub_T[q] = 0.0; D.pub_hot[q >> 5] = 0; }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (tick_done) {
            if (D.shard_rank == 0) c->work[CABM_W_HOT_LOCS] += n_hot;
            if (D.foi_enabled) *D.n_foi = n_hot;
        }
        D.host_status[1] = c->n_new[1]; D.host_status[3] = c->n_imported; D.host_status[0] = c->overflow;             // over PCIe into pinned memory: the host reads them after its sync, no copy
        if (host_row) {   // (per-day calls only) the ledger row of the day that closes: cabm_ledger_row of that day is then a host read. Every writer of the row has
            // finished (the status pass ran in an earlier kernel, this kernel's infections before the ticket)
            const int day = D.today->day;
            unsigned long long* hrow = reinterpret_cast<unsigned long long*>(D.host_status + 12);
            if (day >= 0 && day < D.R) {
#pragma unroll
                for (int k = 0; k < CABM_N_COUNTERS; ++k) hrow[k] = __ldcg(&D.ledger[(size_t) k * D.R + day]);
                D.host_status[2] = (unsigned int) day;
            }
        }
        // (no fence: the host reads the pinned words after it has synchronised with the stream, the next kernel reads the counters
        // behind the kernel boundary; a system-scope fence here waited for the PCIe writes, 1-2 us at the end of every day)
        unsigned int* w = reinterpret_cast<unsigned int*>(c);
#pragma unroll
        for (int k = 0; k < CABM_DAY_SCOPED_WORDS; ++k) w[k] = 0;
        if (advance) c->day_idx++;
        c->end_ticket = 0;
    }
}

// Warp index space: the first n_wu warps take W winners of `phase` each (W = 1 unless the winners outnumber the warps, see
// below); every lane commits its own winner, the whole warp helps with the infector search of long flag segments. Then
// ceil(n_list/32) warps run the status pass (phase 0) over the day's active list for the agents that were not re-infected
// in phase 0: cand[] keeps the winner keys of phase 0 until k_hosp_import so the status half can tell who was.
// Phase 1 (close encodes tick_done | advance << 1 | host_row << 2) is the day's last kernel: winners reset their own candidate word,
// and the last block to finish closes the day.
__device__ __forceinline__ void k_apply_status_body(const Dev& D, const int phase, const int close) {
    __shared__ unsigned int sm[CABM_N_COUNTERS];
    __shared__ unsigned int s_last;
    if (threadIdx.x < CABM_N_COUNTERS) sm[threadIdx.x] = 0;
    __syncthreads();
    pdl_wait();
    const int with_status = phase == 0;
    const DayCtx& cx = day_ctx(D); const int day = cx.day;
    const unsigned int nbase = phase ? min(D.ctr->n_new[0], D.cap_new) : 0u;
    const unsigned int n_w = min(nbase + D.ctr->n_new[phase], D.cap_new) - nbase;
    const unsigned int n_a = with_status ? min(D.ctr->n_active, D.cap_active) : 0u;
    const unsigned int* __restrict__ alist = D.active_list;
    const unsigned int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    // Winners per warp: one while the winners fit in two rounds of the grid's warps (Person::infect is a long dependent chain, a
    // winner alone in its warp finishes soonest, and a second round costs less than two divergent courses in one warp);
    // beyond that W = ceil(winners / (2 * warps)) <= 32 lanes per warp: the lanes walk their disease courses together, every
    // branch arm is issued once per warp, and the loads of all W chains overlap. Measured on the 365-day 21.5 M-agent run:
    // W = 1 always 1148 us per day, packing from one round on 1228, from two rounds on 1078 (three: 1079, four: 1082).
    const unsigned int W = min(32u, max(1u, (n_w + 2 * n_warps - 1) / (2 * n_warps)));
    const unsigned int n_wu = (n_w + W - 1) / W;
    const unsigned int n_units = n_wu + (n_a + 31) / 32;
    for (unsigned int u = warp; u < n_units; u += n_warps) {
        unsigned int m = 0;
        if (u < n_wu) {
            const unsigned int wi = u * W + lane;
            const bool have = lane < W && wi < n_w;
            unsigned int pid = 0; Stream s; InfectCtx c; unsigned long long t_res = 0;
            unsigned int seg_n = 0, seg_off = 0; bool longseg = false;
            if (have) {
                pid = D.newinf_list[nbase + wi];
                CABM_CHECK(D, pid < D.N && D.cand[pid] != CABM_CAND_EMPTY);
                const unsigned long long key = D.cand[pid];
                const int pass = (int) (key >> 40); const unsigned int src = (unsigned int) (key >> 8); const unsigned int role = (unsigned int) (key & 0xFF);
                unsigned int purpose = (unsigned int) pass, sub = 0;
                if (pass == PASS_WORK) purpose = role ? CABM_P_EXPOSE_WORK : CABM_P_EXPOSE_WORK_VIS;
                else if (pass == PASS_HOSP && role) { purpose = CABM_P_EXPOSE_HOSP_PAT; sub = role - 1; }
                event_setup(D, cx, pid, src, purpose, sub, s, c, t_res);
                if (pass == PASS_IMPORT) atomicAdd(&D.ctr->n_imported, 1u);
                if (c.has_source) {
                    CABM_CHECK(D, src < D.L && D.locd[src].cnt >= 1 && (unsigned int) D.locd[src].hot < min(D.ctr->n_hot, D.cap_hot));
                    seg_n = D.locd[src].cnt; seg_off = D.hot_off[D.locd[src].hot];
                    CABM_CHECK(D, (unsigned long long) seg_off + seg_n <= D.cap_flags);
                    if (seg_n <= INFECTOR_THREAD_MAX) c.infector = pick_infector_thread(D.flags + seg_off, seg_n, c.strain, t_res);
                    else longseg = true;
                }
            }
            // long flag segments (hospitals, big workplaces and schools at the peak): the whole warp searches for one winner at a time
            __syncwarp();
            for (unsigned int todo = __ballot_sync(0xffffffffu, longseg); todo; todo &= todo - 1u) {
                const int l = __ffs(todo) - 1;
                const unsigned int n_l = __shfl_sync(0xffffffffu, seg_n, l), off_l = __shfl_sync(0xffffffffu, seg_off, l);
                const int strain_l = __shfl_sync(0xffffffffu, c.strain, l);
                const unsigned long long t_l = __shfl_sync(0xffffffffu, t_res, l);
                const int r = pick_infector_warp(D.flags + off_l, n_l, strain_l, t_l);
                if ((int) lane == l) c.infector = r;
            }
            if (have) {
                infect_event(D, s, c, true);
                if (with_status) m = status_one(D, pid, day);
                else D.cand[pid] = CABM_CAND_EMPTY;                                  // phase 1: nobody reads the key after its winner (phase 0 keys: k_hosp_import)
            }
        } else {
            const unsigned int i = (u - n_wu) * 32 + lane;
            if (i < n_a) {
                const unsigned int p = alist[i];
                CABM_CHECK(D, p < D.N);
                if (D.cand[p] == CABM_CAND_EMPTY) m = status_one(D, p, day);
            }
        }
        __syncwarp();
        if (with_status) {
#pragma unroll
            for (int c = 0; c < CABM_N_COUNTERS; ++c) {
                const unsigned int b = __ballot_sync(0xffffffffu, (m >> c) & 1u);
                if (b && lane == 0) atomicAdd(&sm[c], __popc(b));
            }
        }
    }
    pdl_launch_dependents();
    __syncthreads();
    if (with_status) {
        if (threadIdx.x < CABM_N_COUNTERS && sm[threadIdx.x]) atomicAdd(&D.ledger[(size_t) threadIdx.x * D.R + day], (unsigned long long) sm[threadIdx.x]);
        return;
    }
    // phase 1: the last block to get here closes the day (every other block is done with the day's lists and counters)
    if (threadIdx.x == 0) { __threadfence(); s_last = atomicAdd(&D.ctr->end_ticket, 1u) == gridDim.x - 1 ? 1u : 0u; }
    __syncthreads();
    if (s_last) { __threadfence(); close_day(D, close & 1, (close >> 1) & 1, (close >> 2) & 1); }
}
__global__ void __launch_bounds__(128, 8) k_apply_status(const Dev D, const int phase, const int close) { k_apply_status_body(D, phase, close); }
__global__ void __launch_bounds__(128) k_apply_status_b(const Dev* __restrict__ devs, const int phase, const int close) { BATCH_DEV(D, devs) k_apply_status_body(D, phase, close); }   // one launch, replicate = blockIdx.y

// hospital_transmission (Community.cpp:931-941): census = staff + admitted patients; every consumer computes T itself
__device__ __forceinline__ double hospital_T(const Dev& D, const DayCtx& cx, const int hosp, unsigned int* denom_out, unsigned long long* pws) {
    const unsigned long long* pw = D.locd[hosp].press;
    pws[0] = pw[0]; pws[1] = pw[1]; pws[2] = pw[2]; pws[3] = pw[3];
    const unsigned long long tot = pws[0] + pws[1] + pws[2] + pws[3];
    const unsigned int denom = D.roster_off[hosp + 1] - D.roster_off[hosp] + D.pat_count[hosp];
    *denom_out = denom;
    if (tot == 0 || denom < 2) return 0.0;
    const double w = CABM_MUL((double) tot, 1.0 / 4294967296.0);
    return hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(D.par->haz_mult[CABM_HOSPITAL], cx.season), w), CABM_SUB((double) (int) denom, 1.0)));
}

// do_hosp: warps [0, n_hot_hosp) expose the staff of one hot hospital each, then all threads walk the admitted-patient
// list; do_import: all threads walk today's importation draws (every draw is a "hit": the person is chosen, not exposed)
__device__ __forceinline__ void k_hosp_import_body(const Dev& D, const int do_hosp, const int do_import) {
    pdl_wait();
    const DayCtx& cx = day_ctx(D); const int day = cx.day;
    const unsigned int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    const unsigned int lane = threadIdx.x & 31, warp = tid >> 5, n_warps = nth >> 5;
    unsigned int trials = 0;
    if (do_hosp) {
        // updateHotLocations + clear_public_activity (Community.cpp:999-1003,1042-1046) for today's hot locations other than
        // hospitals: the passes that read their pressure, flags and census are over (the hospital pass reads hospitals only);
        // and the winner keys of phase 0, which the status pass has finished reading. Independent stores issued first: they
        // drain while the dependent chains of the hospital pass wait for memory. Hot hospitals are reset by close_day.
        const unsigned int n_hot = min(D.ctr->n_hot, D.cap_hot);
        for (unsigned int i0 = tid; i0 < n_hot; i0 += 4 * nth) {                      // location and (public index, hospital?) stream in, four per thread in flight
            unsigned int meta[4]; int loc[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) { const unsigned int i = i0 + k * nth; meta[k] = i < n_hot ? D.hot_meta[i] : HOT_IS_HOSP; loc[k] = i < n_hot ? D.hot_list[i] : 0; }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (meta[k] & HOT_IS_HOSP) continue;
                locd_reset(&D.locd[loc[k]]);
                const unsigned int q = meta[k] & HOT_NOPUB;
                if (q != HOT_NOPUB) { D.nvis[q] = 0; D.pub_T[q] = 0.0; D.pub_hot[q >> 5] = 0; }   // the whole bitmap word: every set bit belongs to a hot location
            }
        }
        const unsigned int n_new0 = min(D.ctr->n_new[0], D.cap_new);
        for (unsigned int i = tid; i < n_new0; i += nth) D.cand[D.newinf_list[i]] = CABM_CAND_EMPTY;

        const unsigned int nh = min(D.ctr->n_hot_hosp, D.cap_hot);
        for (unsigned int j = warp; j < nh; j += n_warps) {
            const unsigned int i = (unsigned int) D.hot_hosp[j];                        // index of the hospital in the hot list
            CABM_CHECK(D, i < min(D.ctr->n_hot, D.cap_hot));
            const int hosp = D.hot_list[i];
            CABM_CHECK(D, hosp >= 0 && (unsigned int) hosp < D.L && D.loc_type[hosp] == CABM_HOSPITAL);
            unsigned int denom; unsigned long long pws[4];
            const double T = hospital_T(D, cx, hosp, &denom, pws);
            const unsigned int rb = D.roster_off[hosp], size = D.roster_off[hosp + 1] - rb;
            if (T > 0.0) for (unsigned int m = lane; m < size; m += 32) trial(D, day, 1, (unsigned int) D.roster[rb + m], hosp, CABM_P_EXPOSE_HOSP, 0, T, trials);
            if (D.foi_enabled && lane == 0) write_foi(D, i, hosp, CABM_HOSPITAL, pws, denom, T, 0.0);
        }
        const unsigned int np = min(D.ctr->n_pat, D.cap_pat);
        for (unsigned int i = tid; i < np; i += nth) {
            const unsigned int pid = D.patient_list[i];
            if (pid == 0xFFFFFFFFu) continue;                                         // a discharged patient's entry (below): costs the 4-byte read only
            CABM_CHECK(D, pid < D.N);
            const unsigned int c = D.ag[pid].hosp_count;
            if (c == 0) {                                                             // discharged: the entry becomes a hole and a later admission appends a new one (status_one);
                D.patient_list[i] = 0xFFFFFFFFu; D.ag[pid].pbyte &= (unsigned char) ~PB_ONLIST;   // nobody else writes this agent's status byte in this kernel
                continue;
            }
            const int hosp = D.ag[pid].hospital;
            if (hosp < 0 || D.locd[hosp].cnt == 0) continue;
            unsigned int denom; unsigned long long pws[4];
            const double T = hospital_T(D, cx, hosp, &denom, pws);
            if (T > 0.0) for (unsigned int sub = 0; sub < c; ++sub) trial(D, day, 1, pid, hosp, CABM_P_EXPOSE_HOSP_PAT, sub, T, trials);
        }
    }
    if (do_import) {
        for (unsigned int i = tid; i < (unsigned int) cx.n_imports; i += nth) {
            const uint4 b = philox4x32_10(make_uint4(0u, i, 0u, ((unsigned int) day << 8) | CABM_P_IMPORT), make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32)));
            const unsigned int pid = __umulhi(b.x, D.N);
            if (pid < D.own_lo || pid >= D.own_hi) continue;                          // every handle draws the same people, the owner exposes them
            gate_hit(D, 1, pid, i, CABM_P_IMPORT, 0);
        }
    }
    pdl_launch_dependents();
    trials = warp_sum(trials);
    if (lane == 0 && trials) atomicAdd(&D.ctr->work[CABM_W_TRIALS], (unsigned long long) trials);
}
__global__ void __launch_bounds__(TPB) k_hosp_import(const Dev D, const int do_hosp, const int do_import) { k_hosp_import_body(D, do_hosp, do_import); }
__global__ void __launch_bounds__(TPB) k_hosp_import_b(const Dev* __restrict__ devs, const int do_hosp, const int do_import) { BATCH_DEV(D, devs) k_hosp_import_body(D, do_hosp, do_import); }   // one launch, replicate = blockIdx.y

// Community::infect(id, strain) (Community.cpp:550-553): one agent, gate and commit in one go
__global__ void k_seed_one(const Dev D, const unsigned int pid, const int strain, const int day, int* out, const int echo) {
    if (threadIdx.x || blockIdx.x) return;
    const unsigned int seq = D.ctr->seed_calls++;                                  // every handle counts every call, the owner applies it
    if (pid < D.own_lo || pid >= D.own_hi) { out[0] = 0; __threadfence_system(); *(volatile int*) (out + 1) = echo; return; }
    Stream s; s.begin(D.seed, (unsigned int) day, CABM_P_SEED, pid, seq);
    InfectCtx c; c.pid = (int) pid; c.time = day; c.source = -1; c.has_source = false; c.strain = strain; c.infector = -1;
    c.purpose = CABM_P_SEED; c.key_a = pid; c.key_b = seq; c.key_sub = 0;
    out[0] = infect_event(D, s, c, true) ? 1 : 0;
    __threadfence_system();
    *(volatile int*) (out + 1) = echo;                                              // the host spins on this word (cabm_infect)
}

// n consecutive Community::infect calls (cabm_infect_batch): one thread per call of a WAVE (calls on distinct people, which do
// not see each other); item = person | strain << 29 (N < 2^29 is checked at load time for this entry point), seq_off = position
// of the call in the batch. The sequence counter itself is advanced by k_seed_bump after the last wave.
__global__ void k_seed_batch(const Dev D, const unsigned int* __restrict__ items, const unsigned int* __restrict__ seq_off, const unsigned int n,
                             const int day, unsigned char* out) {
    const unsigned int base = D.ctr->seed_calls;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned int pid = items[i] & 0x1FFFFFFFu; const int strain = (int) (items[i] >> 29);
        const unsigned int k = seq_off[i], seq = base + k;
        unsigned char took = 0;
        if (pid >= D.own_lo && pid < D.own_hi) {                                   // every handle counts every call, the owner applies it
            Stream s; s.begin(D.seed, (unsigned int) day, CABM_P_SEED, pid, seq);
            InfectCtx c; c.pid = (int) pid; c.time = day; c.source = -1; c.has_source = false; c.strain = strain; c.infector = -1;
            c.purpose = CABM_P_SEED; c.key_a = pid; c.key_b = seq; c.key_sub = 0;
            took = infect_event(D, s, c, true) ? 1 : 0;
        }
        out[k] = took;
    }
}
__global__ void k_seed_bump(const Dev D, const unsigned int n) { if (threadIdx.x == 0 && blockIdx.x == 0) D.ctr->seed_calls += n; }

// Person::vaccinate (Person.cpp:611-647), leaky model
__global__ void k_vaccinate(const Dev D, const unsigned int* pids, const unsigned int n, const int day, unsigned char* ok) {
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned int p = pids[i];
        const unsigned int imm = D.imm[p]; const bool had_inf = imm != 0;
        const bool dead = had_inf && D.ag[p].t1.death <= day;
        if (!dead) {
            const unsigned int v = D.vax[p]; const unsigned int nd = v & 0xFF;
            const unsigned int naive = (!had_inf && nd == 0) ? 1u : 0u;
            if (nd < CABM_MAX_DOSES) D.vax_days[(size_t) p * CABM_MAX_DOSES + nd] = (short) day;
            D.vax[p] = min(nd + 1, 255u) | (naive << 8) | ((unsigned int) (unsigned short) (short) day << 16);
        }
        ok[i] = dead ? 0 : 1;
    }
}

__global__ void k_quarantine(const Dev D, const unsigned int* pids, const unsigned int n, const int from, const int until) {
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        D.quar[pids[i]] = ((unsigned int) (unsigned short) (short) from) | ((unsigned int) (unsigned short) (short) until << 16);
}

__global__ void k_gather_infections(const Dev D, const unsigned int* pids, const unsigned int n, cabm_infection* out) {
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned int p = pids ? pids[i] : i;
        cabm_infection r; r.person = p; r.infected_by = -1; r.infected_place = -1;
        if (D.imm[p]) {
            const Tl0 a = D.ag[p].t0; const Tl1 b = D.ag[p].t1;
            r.t[0] = a.infected; r.t[1] = a.infectious_begin; r.t[2] = a.infectious_end; r.t[3] = a.symptom_begin; r.t[4] = a.symptom_end;
            r.t[5] = a.severe_begin; r.t[6] = a.severe_end; r.t[7] = a.hosp_begin; r.t[8] = b.critical_begin; r.t[9] = b.critical_end;
            r.t[10] = b.icu_begin; r.t[11] = b.death; r.report_day = b.report_day; r.strain = b.strain; r.detected_state = b.detected; r.rel_infectiousness = b.rel_inf;
            r.infected_place = D.inf_place[p]; r.infected_by = D.inf_by[p];
        } else {
            for (int k = 0; k < 12; ++k) r.t[k] = CABM_I16_MAX;
            r.t[2] = r.t[4] = r.t[6] = r.t[9] = CABM_I16_MIN;
            r.report_day = CABM_I16_MAX; r.strain = 0; r.detected_state = CABM_N_OUTCOMES; r.rel_infectiousness = 0.0f;
        }
        out[i] = r;
    }
}

// ---------------------------------------------------------------------------------------------
// Caller-side daily diagnostics as one device pass: calculate_vax_stats
// (Community.cpp:1329-1423), doSerosurvey(NATURAL, ...) (:1224-1274), getNumPeopleQuarantining
// (Community.h:437), getNumInfected/Infectious/Symptomatic/Naive (:1196-1222).
// ---------------------------------------------------------------------------------------------
// N-antibody positivity of one person on `day` (Community.cpp:1232-1252)
__device__ __forceinline__ bool sero_positive(const Dev& D, const cabm_params* par, const unsigned int p, const int day) {
    if (D.imm[p] == 0) return false;
    const Tl0 a = D.ag[p].t0;
    const int last_end = max((int) a.infectious_end, (int) a.symptom_end);
    const int tsl = day - last_end;
    const bool converted = (double) a.infected >= par->seroconversion_lag;       // latest infection has the largest infectedBegin
    if (!converted) return false;
    if (par->immunity_leaky) return remaining_efficacy(par, D.nat_eff[p], (double) tsl) > par->seropositivity_threshold;
    return tsl < D.nat_dur[p];
}

__global__ void __launch_bounds__(256) k_diagnostics(const Dev D, const int day, unsigned long long* out) {
    __shared__ unsigned int s_cnt[CABM_N_DIAG];
    if (threadIdx.x < CABM_N_DIAG) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const cabm_params* par = D.par;
    const int lag = par->vaccine_dose_to_protection_lag;
    const unsigned int lane = threadIdx.x & 31;
    for (unsigned int base = blockIdx.x * blockDim.x; base < D.N; base += gridDim.x * blockDim.x) {
        const unsigned int p = base + threadIdx.x;
        unsigned int bits = 0;                                    // one bit per counter this person adds to
        if (p < D.N) {
            const unsigned int v = D.vax[p]; const unsigned int nd = v & 0xFF; const int last_dose = (short) (v >> 16);
            if (nd >= 1 && nd <= 3) bits |= 1u << CABM_D_DOSE1;   // the reference's switch has no case above 3 doses (:1351-1356)
            if (nd == 2 || nd == 3) bits |= 1u << CABM_D_DOSE2;
            if (nd == 3) bits |= 1u << CABM_D_DOSE3;
            const bool full = nd > 1 && (day - last_dose) >= lag;
            if (full) bits |= 1u << CABM_D_FULLY_VAXD;
            const unsigned int imm = D.imm[p];
            if (imm) {
                const Tl0 a = D.ag[p].t0; const Tl1 b = D.ag[p].t1;
                const int o = full ? CABM_D_BT_INF : CABM_D_UV_INF;
                if (a.infected == day) bits |= 1u << o;
                if (a.symptom_begin == day) bits |= 1u << (o + 1);
                if (a.hosp_begin == day) bits |= 1u << (o + 2);
                if (b.death == day) bits |= 1u << (o + 3);
                if (b.detected < CABM_N_OUTCOMES && b.report_day == day) bits |= 1u << (o + 4);
                if (a.infected <= day && day < a.infectious_end) bits |= 1u << CABM_D_INFECTED;
                if (a.infectious_begin <= day && day < a.infectious_end) bits |= 1u << CABM_D_INFECTIOUS;
                if (a.symptom_begin <= day && day < a.symptom_end) bits |= 1u << CABM_D_SYMPTOMATIC;
            } else if (nd == 0) bits |= 1u << CABM_D_NAIVE;
            const unsigned int q = D.quar[p]; const int qf = (short) (q & 0xFFFF), qu = (short) (q >> 16);
            if (q != 0 && day >= qf && day < qu) bits |= 1u << CABM_D_QUARANTINING;
            const bool ped = D.age[p] <= 17;
            if (ped) bits |= 1u << CABM_D_N_PED;
            if (imm && sero_positive(D, par, p, day)) bits |= (1u << CABM_D_SEROPOS_ALL) | (ped ? 1u << CABM_D_SEROPOS_PED : 0u);
        }
#pragma unroll
        for (int c = 0; c < CABM_D_N_PED + 1; ++c) {
            const unsigned int n = __popc(__ballot_sync(0xFFFFFFFFu, (bits >> c) & 1u));
            if (lane == 0 && n) atomicAdd(&s_cnt[c], n);
        }
    }
    __syncthreads();
    if (threadIdx.x < CABM_N_DIAG && s_cnt[threadIdx.x]) atomicAdd(&out[threadIdx.x], (unsigned long long) s_cnt[threadIdx.x]);
}

__global__ void k_serosurvey(const Dev D, const int day, const unsigned int* pids, const unsigned int n, unsigned long long* out) {
    unsigned int mine = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        mine += sero_positive(D, D.par, pids ? pids[i] : i, day) ? 1u : 0u;
    mine = warp_sum(mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(out, (unsigned long long) mine);
}

// one person's predicates on `day`: bits 0-7 as CABM_EL_*, bits 8.. = 1 << (8 + CABM_R_*)
__device__ __forceinline__ unsigned int person_bits(const Dev& D, const cabm_params* par, const unsigned int p, const int day) {
    const unsigned int imm = D.imm[p], v = D.vax[p], nd = v & 0xFF; const int last_dose = (short) (v >> 16);
    unsigned int b = 0; bool alive = true;
    if (imm) {
        const Tl0 a = D.ag[p].t0; const Tl1 t = D.ag[p].t1;
        alive = !(t.death <= day);
        b |= CABM_EL_EVER_INFECTED;
        if (t.detected < CABM_N_OUTCOMES && t.report_day <= day) b |= CABM_EL_PRIOR_CASE;
        if (a.infected <= day && day < a.infectious_end) b |= CABM_EL_INFECTED_NOW | (1u << (8 + CABM_R_INFECTED));
        if (a.infectious_begin <= day && day < a.infectious_end) b |= 1u << (8 + CABM_R_INFECTIOUS);
        if (a.symptom_begin <= day && day < a.symptom_end) b |= 1u << (8 + CABM_R_SYMPTOMATIC);
        if (sero_positive(D, par, p, day)) b |= 1u << (8 + CABM_R_SEROPOSITIVE);
    } else if (nd == 0) b |= 1u << (8 + CABM_R_NAIVE);
    if (alive) b |= CABM_EL_ALIVE | (1u << (8 + CABM_R_ALIVE));
    if (nd) b |= CABM_EL_VACCINATED;
    b |= min(nd, 7u) << CABM_EL_DOSES_SHIFT;
    if (nd > 1 && (day - last_dose) >= par->vaccine_dose_to_protection_lag) b |= 1u << (8 + CABM_R_FULLY_VAXD);
    const unsigned int q = D.quar[p]; const int qf = (short) (q & 0xFFFF), qu = (short) (q >> 16);
    if (q != 0 && day >= qf && day < qu) b |= 1u << (8 + CABM_R_QUARANTINING);
    return b;
}
__global__ void k_eligibility(const Dev D, const int day, const unsigned int* pids, const unsigned int n, unsigned char* out) {
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        out[i] = (unsigned char) (person_bits(D, D.par, pids ? pids[i] : i, day) & 0xFFu);
}
__global__ void k_reduce(const Dev D, const int what, const int day, const unsigned int* pids, const unsigned int n, unsigned long long* out) {
    unsigned int mine = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        mine += (person_bits(D, D.par, pids ? pids[i] : i, day) >> (8 + what)) & 1u;
    mine = warp_sum(mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(out, (unsigned long long) mine);
}

// primary cases of contact tracing (Community.cpp:1062-1070): alive, latest infection detected and reported today
__global__ void k_detected_today(const Dev D, const int day, unsigned int* out, unsigned int cap, unsigned int* n_out) {
    for (unsigned int base = blockIdx.x * blockDim.x; base < D.N; base += gridDim.x * blockDim.x) {
        const unsigned int p = base + threadIdx.x;
        bool hit = false;
        if (p < D.N && D.imm[p]) {
            const Tl1 b = D.ag[p].t1;
            hit = b.detected < CABM_N_OUTCOMES && b.report_day == day && !(b.death <= day);
        }
        if (hit) { const unsigned int i = list_reserve(n_out); if (i < cap) out[i] = p; }
    }
}

// ---------------------------------------------------------------------------------------------
// One community over several handles (include/cabm.h cabm_shard_*): flags of the other handles' agents are applied to
// this handle's replicated location state exactly as k_active applies its own (same pressure sums, own slot numbering);
// the hospital census is kept global by exchanging the day's admissions minus discharges.
// ---------------------------------------------------------------------------------------------
__global__ void k_flags_in(const Dev D, const FlagRaw* __restrict__ in, const unsigned int n) {
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const FlagRaw r = in[i];
        const int strain = FLAG_STRAIN(r); const unsigned long long c = FLAG_C(r);
        atomicAdd(&D.locd[r.loc].press[strain], c);
        const unsigned int slot = atomicAdd(&D.locd[r.loc].cnt, 1u);
        const unsigned int f = list_reserve(&D.ctr->n_flags);
        flag_location(D, r.loc, strain, c, r.pid, slot, f);
    }
}
// exchange A inside the library: recv holds `world` blocks of cap records (ncclAllGather of the front of every handle's
// flags_raw), counts[r] the number of valid records of block r; this handle applies everybody's but its own
__global__ void k_flags_in_all(const Dev D, const FlagRaw* __restrict__ recv, unsigned int* counts, const unsigned int cap,
                               const int world, const int rank) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {          // counts[CABM_MAX_SHARDS]: largest per-handle count since the host last looked (sizes the next gathers)
        unsigned int m = 0;
        for (int r = 0; r < world; ++r) m = max(m, counts[r]);
        atomicMax(&counts[CABM_MAX_SHARDS], m);
    }
    for (int r = 0; r < world; ++r) {
        if (r == rank) continue;
        const unsigned int n = counts[r];
        if (n > cap) { if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(&D.ctr->overflow, 512u); }
        const FlagRaw* in = recv + (size_t) r * cap;
        for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < min(n, cap); i += gridDim.x * blockDim.x) {
            const FlagRaw f = in[i];
            const int strain = FLAG_STRAIN(f); const unsigned long long c = FLAG_C(f);
            atomicAdd(&D.locd[f.loc].press[strain], c);
            const unsigned int slot = atomicAdd(&D.locd[f.loc].cnt, 1u);
            flag_location(D, f.loc, strain, c, f.pid, slot, list_reserve(&D.ctr->n_flags));
        }
    }
}
// Exchange A over peer memory: every handle reads the other handles' flag lists where they lie (NVLink loads through the
// pointers cudaIpcOpenMemHandle gave), exactly counts[r] records each. Ordering comes from the collectives around it: the
// count all-gather in front completes only after every peer's k_active has, and a peer cannot start overwriting its list
// (next day's k_active) before the census all-reduce behind this kernel, which needs this handle's contribution.
struct PeerFlags { const FlagRaw* p[CABM_MAX_SHARDS]; };
__global__ void k_flags_pull(const Dev D, const PeerFlags peers, const unsigned int* __restrict__ counts, const int world, const int rank) {
    for (int r = 0; r < world; ++r) {
        if (r == rank) continue;
        const unsigned int n = min(counts[r], D.cap_flags);
        const unsigned long long* in = reinterpret_cast<const unsigned long long*>(peers.p[r]);
        for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
            const unsigned long long w0 = __ldcv(in + 3 * (size_t) i), w1 = __ldcv(in + 3 * (size_t) i + 1), c_raw = __ldcv(in + 3 * (size_t) i + 2);
            FlagRaw f; f.loc = (int) (unsigned int) w0; f.slot = (unsigned int) (w0 >> 32); f.pid = (unsigned int) w1; f.pad = 0; f.c = c_raw;
            const int strain = FLAG_STRAIN(f); const unsigned long long c = FLAG_C(f);
            atomicAdd(&D.locd[f.loc].press[strain], c);
            const unsigned int slot = atomicAdd(&D.locd[f.loc].cnt, 1u);
            flag_location(D, f.loc, strain, c, f.pid, slot, list_reserve(&D.ctr->n_flags));
        }
    }
}
// mode 0: remember the census before the status pass; 1: buf = own admissions - discharges; 2: census = remembered + summed buf
__global__ void k_patients(const Dev D, const int* __restrict__ hosp_ids, const unsigned int n, int* save, int* buf, const int mode) {
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int loc = hosp_ids[i];
        if (mode == 0) save[i] = (int) D.pat_count[loc];
        else if (mode == 1) buf[i] = (int) D.pat_count[loc] - save[i];
        else D.pat_count[loc] = (unsigned int) (save[i] + buf[i]);
    }
}

__global__ void k_init_state(const Dev D) {
    const unsigned int stride = gridDim.x * blockDim.x;
    for (unsigned int p = blockIdx.x * blockDim.x + threadIdx.x; p < D.N; p += stride) {
        D.cand[p] = CABM_CAND_EMPTY;
    }
}
