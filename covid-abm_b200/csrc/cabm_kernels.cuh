// covid-abm_b200/csrc/cabm_kernels.cuh -- first half of the day pipeline (sm_100a): who is infectious where,
// public activity, per-location hazard, exposure trials.
//
//   k_scan     all agents      : streams the 4-byte active-window word, block-aggregated compaction of today's
//                                active agents (replaces walking every Person, Community.cpp:684)
//   k_active   active agents   : Person.cpp:475-486 derived from the stored timeline instead of the reference's
//                                _isHot calendar: pressure at home / day location / hospital / visited locations
//                                (Community.cpp:1031-1035), per-location flag chains, hot lists, in-hospital bit
//   k_visit    all agents      : public_activity (Community.cpp:1006-1039): Poisson(1) visits from one Philox block,
//                                hot test against a shared-memory bitmap of patronised locations, visitor census,
//                                block-buffered visit records
//   k_hazard   hot locations   : _tally_infectiousness + the hazard of every pass (Community.cpp:833-969); one warp
//                                per location runs the household / neighbour / small-roster trials in place and
//                                splits big rosters into 256-member items
//   k_trials   items + visits  : _transmission's `U < T` (Community.cpp:976) for big rosters and visitors
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
#define SMALL_ROSTER 512
#define MAXV 8
#define VIS_BUF 2048

__device__ __forceinline__ const DayCtx& day_ctx(const Dev& D) { return *D.today; }

// append n entries to a device list whose length lives at `counter`; one atomic per coalesced group of callers
__device__ __forceinline__ unsigned int list_reserve(unsigned int* counter) {
    cg::coalesced_group g = cg::coalesced_threads();
    unsigned int base = 0;
    if (g.thread_rank() == 0) base = atomicAdd(counter, g.size());
    return g.shfl(base, 0) + g.thread_rank();
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
__global__ void __launch_bounds__(TPB) k_scan(const Dev D) {
    __shared__ unsigned int s_warp[TPB / 32], s_base;
    const DayCtx& cx = D.dayctx[D.ctr->day_idx];
    const int day = cx.day;
    if (blockIdx.x == 0 && threadIdx.x == 0) { *D.today = cx; atomicAdd(&D.ctr->work[CABM_W_AGENTS], (unsigned long long) D.N); }
    const unsigned int p0 = (blockIdx.x * TPB + threadIdx.x) * 4u;
    uint4 w = make_uint4(0x80007FFFu, 0x80007FFFu, 0x80007FFFu, 0x80007FFFu);
    if (p0 + 3 < D.N) w = *reinterpret_cast<const uint4*>(D.win + p0);
    else { if (p0 < D.N) w.x = D.win[p0]; if (p0 + 1 < D.N) w.y = D.win[p0 + 1]; if (p0 + 2 < D.N) w.z = D.win[p0 + 2]; }
    const unsigned int ww[4] = { w.x, w.y, w.z, w.w };
    unsigned int mask = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) { const int from = (short) (ww[k] & 0xFFFF), until = (short) (ww[k] >> 16); if (day >= from && day <= until) mask |= 1u << k; }
    const unsigned int cnt = __popc(mask), lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int acc = 0;
#pragma unroll
        for (int k = 0; k < TPB / 32; ++k) { const unsigned int c = s_warp[k]; s_warp[k] = acc; acc += c; }
        s_base = acc ? atomicAdd(&D.ctr->n_active, acc) : 0;
        if (acc) atomicAdd(&D.ctr->work[CABM_W_ACTIVE], (unsigned long long) acc);
    }
    __syncthreads();
    unsigned int o = s_base + s_warp[wid] + incl - cnt;
#pragma unroll
    for (int k = 0; k < 4; ++k) if (mask & (1u << k)) { if (o < D.cap_active) D.active_list[o] = p0 + k; else atomicOr(&D.ctr->overflow, 4u); ++o; }
}

// Community::flagInfectedLocation (Community.cpp:788-791) on device: add to the per-strain pressure and push a flag
// record on the location's chain; whoever finds the chain empty appends the location to the hot list.
__device__ __forceinline__ void add_pressure(const Dev& D, int loc, int strain, unsigned long long c, unsigned int pid) {
    if (loc < 0 || c == 0) return;
    atomicAdd(&D.press[(size_t) loc * CABM_N_STRAINS + strain], c);
    const unsigned int f = list_reserve(&D.ctr->n_flags);
    if (f >= D.cap_flags) { atomicOr(&D.ctr->overflow, 128u); return; }
    const int old = atomicExch(&D.loc_head[loc], (int) f);
    Flag r; r.pid = pid; r.next = old; r.c = c | ((unsigned long long) strain << 56);
    D.flags[f] = r;
    if (old == -1) {
        const unsigned int i = list_reserve(&D.ctr->n_hot);
        if (i < D.cap_hot) D.hot_list[i] = loc; else atomicOr(&D.ctr->overflow, 2u);
        if (D.loc_type[loc] == CABM_HOSPITAL && i < D.cap_hot) { const unsigned int j = list_reserve(&D.ctr->n_hot_hosp); if (j < D.cap_hot) D.hot_hosp[j] = (int) i; }
        const int q = D.loc_pub[loc];
        if (q >= 0) atomicOr(&D.pub_hot[q >> 5], 1u << (q & 31));
    }
}

// Community::public_activity draws for one agent (Community.cpp:1012-1026): k ~ Poisson(1) from the first word of
// the agent's Philox block, then k uniform picks among its patronised locations. Returns the number of distinct
// locations; qs[] holds public-location indices, hours[] the multiplicities.
__device__ __forceinline__ int gen_visits(const Dev& D, const DayCtx& cx, unsigned int pid, int npat, int* qs, int* hours) {
    uint4 b = philox4x32_10(make_uint4(0u, pid, 0u, ((unsigned int) cx.day << 8) | CABM_P_VISIT), make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32)));
    int k;
    if (b.x < D.visit_thr[0]) return 0;
    else if (b.x < D.visit_thr[1]) k = 1;
    else if (b.x < D.visit_thr[2]) k = 2;
    else if (b.x < D.visit_thr[3]) k = 3;
    else k = (int) table_draw(D.par->visit_cdf, CABM_T_POISSON, (double) b.x * (1.0 / 4294967296.0));
    int nv = 0, avoiding = -1;
    const int* __restrict__ row = D.patron + (size_t) pid * D.n_patron;
    for (int i = 0; i < k; ++i) {
        const unsigned int idx1 = (unsigned int) i + 1u;              // draw index within the stream (0 was the Poisson draw)
        if ((idx1 & 3u) == 0) b = philox4x32_10(make_uint4(idx1 >> 2, pid, 0u, ((unsigned int) cx.day << 8) | CABM_P_VISIT), make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32)));
        const unsigned int lane = idx1 & 3u;
        const unsigned int x = lane == 0 ? b.x : lane == 1 ? b.y : lane == 2 ? b.z : b.w;
        const int e = row[__umulhi(x, (unsigned int) npat)];
        if (e & PATRON_HIGH) {
            if (avoiding < 0) avoiding = ((double) D.home_risk[pid] < cx.sd) ? 1 : 0;
            if (avoiding) continue;
        }
        const int q = e & 0x3FFFFFFF;
        int j = 0; while (j < nv && qs[j] != q) ++j;
        if (j == nv) { qs[nv] = q; hours[nv] = 1; ++nv; } else hours[j]++;
    }
    return nv;
}

// ---------------------------------------------------------------------------------------------
// k_active: one thread per active agent.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB) k_active(const Dev D) {
    const DayCtx& cx = day_ctx(D); const int day = cx.day;
    const unsigned int n = min(D.ctr->n_active, D.cap_active);
    unsigned int nflag = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned int p = D.active_list[i];
        const Tl0 t0 = D.tl0[p]; const Tl1 t1 = D.tl1[p];
        const unsigned char pb = D.pbyte[p];
        const int hm = D.home[p], dl = D.day_loc[p];
        const bool inhosp = tl_in(t0.hosp_begin, t0.severe_end, day);
        const unsigned char npb = (unsigned char) ((pb & ~PB_INHOSP) | (inhosp ? PB_INHOSP : 0));
        if (npb != pb) D.pbyte[p] = npb;
        if (!tl_in(t0.infectious_begin, t0.infectious_end, day)) continue;
        const unsigned long long c = (unsigned long long) __double2ll_rn(CABM_MUL((double) t1.rel_inf, 4294967296.0));
        if (inhosp) { add_pressure(D, D.loc_hospital[hm], t1.strain, c, p); nflag++; continue; }
        add_pressure(D, hm, t1.strain, c, p); nflag++;
        if (dl >= 0 && !(tl_in(t0.severe_begin, t0.severe_end, day) || tl_in(t1.critical_begin, t1.critical_end, day))) { add_pressure(D, dl, t1.strain, c, p); nflag++; }
        const int npat = pb & PB_NPAT;
        if (npat) {
            int qs[MAXV], hours[MAXV];
            const int nv = gen_visits(D, cx, p, npat, qs, hours);
            for (int j = 0; j < nv; ++j) {
                const unsigned long long cv = (unsigned long long) __double2ll_rn(CABM_MUL(CABM_MUL((double) t1.rel_inf, (double) hours[j]), 536870912.0));
                if (cv) { add_pressure(D, D.pub_loc[qs[j]], t1.strain, cv, p); nflag++; }
            }
        }
    }
    nflag = warp_sum(nflag);
    if ((threadIdx.x & 31) == 0 && nflag) atomicAdd(&D.ctr->work[CABM_W_FLAGS], (unsigned long long) nflag);
}

// ---------------------------------------------------------------------------------------------
// k_visit: every agent with patronised locations that is not in hospital draws its visits.
// dynamic shared memory: hot bitmap of the Q patronised locations (ceil(Q/32) words)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB) k_visit(const Dev D, const unsigned int bitmap_words) {
    extern __shared__ unsigned int s_hot[];
    __shared__ uint2 s_buf[VIS_BUF];
    __shared__ unsigned int s_cnt, s_gbase;
    const DayCtx& cx = day_ctx(D);
    for (unsigned int i = threadIdx.x; i < bitmap_words; i += blockDim.x) s_hot[i] = D.pub_hot[i];
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int n_round = (D.N + stride - 1) / stride * stride;
    unsigned int w_pat = 0, w_v0 = 0, w_v = 0;
    for (unsigned int p = blockIdx.x * blockDim.x + threadIdx.x; p < n_round; p += stride) {
        if (p < D.N) {
            const unsigned char pb = D.pbyte[p];
            const int npat = pb & PB_NPAT;
            if (npat && !(pb & PB_INHOSP)) {
                w_pat++;
                int qs[MAXV], hours[MAXV];
                const int nv = gen_visits(D, cx, p, npat, qs, hours);
                w_v0 += (unsigned int) nv;
                for (int j = 0; j < nv; ++j) {
                    const int q = qs[j];
                    if (s_hot[q >> 5] & (1u << (q & 31))) {
                        atomicAdd(&D.nvis[q], 1u);
                        s_buf[atomicAdd(&s_cnt, 1u)] = make_uint2(p, (unsigned int) q);
                        w_v++;
                    }
                }
            }
        }
        __syncthreads();
        const unsigned int c = s_cnt;
        __syncthreads();                                     // everyone has read s_cnt before the next round appends to it
        if (c > VIS_BUF - TPB * (MAXV - 1)) {                // not enough room for another round: flush (uniform branch)
            if (threadIdx.x == 0) s_gbase = atomicAdd(&D.ctr->n_vis, c);
            __syncthreads();
            const unsigned int gb = s_gbase;
            if (gb + c <= D.cap_vis) { for (unsigned int i = threadIdx.x; i < c; i += blockDim.x) D.visit_rec[gb + i] = s_buf[i]; }
            else if (threadIdx.x == 0) atomicOr(&D.ctr->overflow, 8u);
            __syncthreads();
            if (threadIdx.x == 0) s_cnt = 0;
            __syncthreads();
        }
    }
    {
        const unsigned int c = s_cnt;
        if (threadIdx.x == 0 && c) s_gbase = atomicAdd(&D.ctr->n_vis, c);
        __syncthreads();
        const unsigned int gb = s_gbase;
        if (c) { if (gb + c <= D.cap_vis) { for (unsigned int i = threadIdx.x; i < c; i += blockDim.x) D.visit_rec[gb + i] = s_buf[i]; } else if (threadIdx.x == 0) atomicOr(&D.ctr->overflow, 8u); }
    }
    const int ids[3] = { CABM_W_PATRONS, CABM_W_VISITS, CABM_W_HOT_VISITS }; const unsigned int vals[3] = { w_pat, w_v0, w_v };
    block_work_add<3>(D, ids, vals);
}

__device__ __forceinline__ double hazard_to_T(double hazard) { return CABM_SUB(1.0, cabm_exp(-hazard)); }

// one exposure trial: the quarantine skip and `gsl_rng_uniform(RNG) < T` of Community::_transmission (Community.cpp:975-976).
// A success is queued for the susceptibility gate (k_gate).
__device__ __forceinline__ void trial(const Dev& D, const int day, const int phase, const unsigned int pid, const int src, const int purpose,
                                      const unsigned int sub, const double T, unsigned int& trials) {
    if (D.any_quarantine) {
        const unsigned int q = D.quar[pid]; const int qf = (short) (q & 0xFFFF), qu = (short) (q >> 16);
        if (day >= qf && day < qu && src != D.home[pid]) return;
    }
    trials++;
    const uint4 b = philox4x32_10(make_uint4(sub << 16, pid, (unsigned int) src, ((unsigned int) day << 8) | (unsigned int) purpose),
                                  make_uint2((unsigned int) D.seed, (unsigned int) (D.seed >> 32)));
    if ((double) b.x * (1.0 / 4294967296.0) < T) {
        const unsigned int base = phase ? min(D.ctr->n_hits[0], D.cap_hits) : 0u;
        const unsigned int h = base + list_reserve(&D.ctr->n_hits[phase]);
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

// ---------------------------------------------------------------------------------------------
// k_hazard: one warp per hot location (hospitals are left to k_hosp, after the status pass).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB) k_hazard(const Dev D) {
    const DayCtx& cx = day_ctx(D); const int day = cx.day;
    const cabm_params* __restrict__ par = D.par;
    const unsigned int n_hot = min(D.ctr->n_hot, D.cap_hot);
    const unsigned int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    unsigned int trials = 0, edges = 0;
    for (unsigned int i = warp; i < n_hot; i += n_warps) {
        const int loc = D.hot_list[i];
        const int type = D.loc_type[loc];
        if (type == CABM_HOSPITAL) continue;
        const unsigned long long* pw = &D.press[(size_t) loc * CABM_N_STRAINS];
        const unsigned long long pws[4] = { pw[0], pw[1], pw[2], pw[3] };
        const double w = CABM_MUL((double) (pws[0] + pws[1] + pws[2] + pws[3]), 1.0 / 4294967296.0);
        const unsigned int rb = D.roster_off[loc], size = D.roster_off[loc + 1] - rb;
        double T = 0.0, Ts = 0.0; unsigned int denom = size; int purpose = 0;
        if (type == CABM_HOUSE) {
            T = hazard_to_T(CABM_MUL(CABM_MUL(par->haz_mult[CABM_HOUSE], cx.season), w));
            const unsigned int nb0 = D.nbr_off[loc], deg = D.nbr_off[loc + 1] - nb0;
            const bool risky = (double) D.loc_risk[loc] > cx.sd;
            if (risky) { Ts = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(par->haz_mult[CABM_N_LOCTYPES], cx.season), w), (double) (int) size)); if (lane == 0) edges += deg; }
            // segment 0 = own household (within_household_transmission), segment s >= 1 = neighbour s-1 (between_household_transmission);
            // 8 lanes per segment, 4 segments per warp round
            const unsigned int n_seg = 1 + ((risky && Ts > 0.0) ? deg : 0);
            for (unsigned int s0 = 0; s0 < n_seg; s0 += 4) {
                const unsigned int s = s0 + (lane >> 3);
                if (s < n_seg) {
                    unsigned int b = rb, cnt = size; double Tt = T; int pp = CABM_P_EXPOSE_HH; bool go = T > 0.0;
                    if (s > 0) {
                        const int nb = D.nbr[nb0 + s - 1];
                        go = (double) D.loc_risk[nb] > cx.sd;
                        b = D.roster_off[nb]; cnt = D.roster_off[nb + 1] - b; Tt = Ts; pp = CABM_P_EXPOSE_NB;
                    }
                    if (go) for (unsigned int m = lane & 7; m < cnt; m += 8) trial(D, day, 0, (unsigned int) D.roster[b + m], loc, pp, 0, Tt, trials);
                }
            }
        } else {
            if (type == CABM_WORK) {
                const unsigned char lf = D.loc_flags[loc];
                const int q = D.loc_pub[loc];
                denom = size + (q >= 0 ? D.nvis[q] : 0u);
                const bool closed = !(lf & LF_ESSENTIAL) && cx.business_closure != 0.0;
                if (!(denom < 2 || closed) && w > 0.0) {
                    const double f = LF_PUBRISK(lf) == CABM_PT_HIGH ? 4.0 : 0.25;
                    T = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(CABM_MUL(par->haz_mult[CABM_WORK], f), cx.season), w), CABM_SUB((double) (int) denom, 1.0)));
                }
                if (q >= 0 && lane == 0) D.pub_T[q] = T;
                purpose = CABM_P_EXPOSE_WORK;
            } else if (type == CABM_SCHOOL) {
                const bool open = cx.is_weekday && cx.school_closure < 1;
                const double coef = CABM_MUL(CABM_MUL(CABM_SUB(1.0, cx.school_closure), par->haz_mult[CABM_SCHOOL]), cx.season);
                if (open && coef != 0.0 && size >= 2) T = hazard_to_T(CABM_DIV(CABM_MUL(coef, w), CABM_SUB((double) (int) size, 1.0)));
                purpose = CABM_P_EXPOSE_SCHOOL;
            } else {  // NURSINGHOME
                if (size >= 2) T = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(par->haz_mult[CABM_NURSINGHOME], cx.season), w), CABM_SUB((double) (int) size, 1.0)));
                purpose = CABM_P_EXPOSE_NH;
            }
            if (T > 0.0) {
                if (size <= SMALL_ROSTER) {
                    for (unsigned int m = lane; m < size; m += 32) trial(D, day, 0, (unsigned int) D.roster[rb + m], loc, purpose, 0, T, trials);
                } else {
                    const unsigned int nch = (size + CHUNK - 1) / CHUNK;
                    unsigned int base = 0;
                    if (lane == 0) base = atomicAdd(&D.ctr->n_items, nch);
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (base + nch <= D.cap_items) {
                        for (unsigned int k = lane; k < nch; k += 32) {
                            Item it; it.src = loc; it.begin = (int) (rb + k * CHUNK); it.count = (int) min((unsigned int) CHUNK, size - k * CHUNK); it.purpose = purpose; it.T = T; it.pad = 0.0;
                            D.items[base + k] = it;
                        }
                    } else if (lane == 0) atomicOr(&D.ctr->overflow, 16u);
                }
            }
        }
        if (D.foi_enabled && lane == 0) write_foi(D, i, loc, type, pws, denom, T, Ts);
    }
    trials = warp_sum(trials);
    if (lane == 0) { if (trials) atomicAdd(&D.ctr->work[CABM_W_TRIALS], (unsigned long long) trials); if (edges) atomicAdd(&D.ctr->work[CABM_W_NBR_EDGES], (unsigned long long) edges); }
}

// ---------------------------------------------------------------------------------------------
// k_trials: 256-trial units: first the roster items of big locations, then the visit records.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB) k_trials(const Dev D) {
    const DayCtx& cx = day_ctx(D); const int day = cx.day;
    const unsigned int n_items = min(D.ctr->n_items, D.cap_items), n_vis = min(D.ctr->n_vis, D.cap_vis);
    const unsigned int units = n_items + (n_vis + TPB - 1) / TPB;
    unsigned int trials = 0;
    for (unsigned int u = blockIdx.x; u < units; u += gridDim.x) {
        if (u < n_items) {
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
