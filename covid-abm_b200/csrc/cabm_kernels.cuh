// covid-abm_b200/csrc/cabm_kernels.cuh -- the day pipeline as sm_100a kernels.
//
//   k_scan            agents        : active-window scan -> infectious pressure at home / day location /
//                                     hospital (Person.cpp:475-486 derived from the stored timeline instead of
//                                     the reference's _isHot calendar), infectious visitors' pressure
//                                     (Community.cpp:1031-1035), active list, in-hospital flag
//   k_visit           agents        : public_activity (Community.cpp:1006-1039): Poisson(1) visits, visitor
//                                     census of hot workplaces, visit records for the exposure pass
//   k_hazard          hot locations : _tally_infectiousness + hazard -> T of every pass (Community.cpp:833-969),
//                                     exposure work items (roster chunks)
//   k_expose*         work items / visit records / patients : _transmission (Community.cpp:971-996)
//   k_apply           winners       : Person::infect for the first successful exposure of each agent
//   k_status          active agents : updatePersonStatus (Community.cpp:682-741)
//   k_import          imports       : seed_epidemic (simulator.h:201-223)
//   k_end / k_*_end                 : updateHotLocations (Community.cpp:999-1003), list resets
//
// All kernels are grid-stride over device-resident counts, so a day needs no host round trip.
#pragma once
#include "cabm_infect.cuh"

#define TPB 256
#define CHUNK 256
#define MAXV 8

__device__ __forceinline__ const DayCtx& day_ctx(const Dev& D) { return D.dayctx[D.ctr->day_idx]; }

template <int NC>
struct BlockTally {            // block-level reduction of per-thread 0/1 increments, one global atomic per block per counter
    unsigned int* sm;
    __device__ __forceinline__ void init(unsigned int* s) { sm = s; if (threadIdx.x < NC) sm[threadIdx.x] = 0; __syncthreads(); }
    __device__ __forceinline__ void add(unsigned int mask_bits) {     // call convergently by the whole warp
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const unsigned int b = __ballot_sync(0xffffffffu, (mask_bits >> c) & 1u);
            if (b && (threadIdx.x & 31) == 0) atomicAdd(&sm[c], __popc(b));
        }
    }
    __device__ __forceinline__ void add_count(int c, unsigned int v) {  // convergent; arbitrary per-thread counts
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (v && (threadIdx.x & 31) == 0) atomicAdd(&sm[c], v);
    }
};

__device__ __forceinline__ void add_pressure(const Dev& D, int loc, int strain, unsigned long long c) {
    if (loc < 0 || c == 0) return;
    atomicAdd(&D.press[(size_t) loc * CABM_N_STRAINS + strain], c);
    if (D.hot_idx[loc] == -1) {
        if (atomicCAS(&D.hot_idx[loc], -1, -2) == -1) {
            const unsigned int i = atomicAdd(&D.ctr->n_hot, 1u);
            if (i < D.cap_hot) { D.hot_list[i] = loc; atomicExch(&D.hot_idx[loc], (int) i); }
            else atomicOr(&D.ctr->overflow, 2u);
        }
    }
}

// Community::public_activity draws for one agent (Community.cpp:1012-1026); returns number of distinct visited locations
__device__ __forceinline__ int gen_visits(const Dev& D, const DayCtx& cx, unsigned int pid, int npat, int* locs, int* hours) {
    Stream s; s.begin(D.par->seed, (unsigned int) cx.day, CABM_P_VISIT, pid, 0);
    const int k = (int) table_draw(D.par->visit_cdf, CABM_T_POISSON, s.u01());
    int nv = 0, avoiding = -1;
    for (int i = 0; i < k; ++i) {
        const unsigned int idx = __umulhi(s.next(), (unsigned int) npat);
        const int e = D.patron[(size_t) pid * D.n_patron + idx];
        if (e & PATRON_HIGH) {
            if (avoiding < 0) avoiding = ((double) D.home_risk[pid] < cx.sd) ? 1 : 0;
            if (avoiding) continue;
        }
        const int loc = e & 0x3FFFFFFF;
        int j = 0; while (j < nv && locs[j] != loc) ++j;
        if (j == nv) { locs[nv] = loc; hours[nv] = 1; ++nv; } else hours[j]++;
    }
    return nv;
}

__global__ void __launch_bounds__(TPB) k_scan(const Dev D) {
    __shared__ unsigned int sm[4]; BlockTally<4> bt; bt.init(sm);
    const DayCtx& cx = day_ctx(D); const int day = cx.day;
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int n_round = (D.N + stride - 1) / stride * stride;
    for (unsigned int p = blockIdx.x * blockDim.x + threadIdx.x; p < n_round; p += stride) {
        unsigned int flags = 0, nflag = 0;
        if (p < D.N) {
            const unsigned int w = D.win[p];
            const int from = (short) (w & 0xFFFF), until = (short) (w >> 16);
            if (day >= from && day <= until) {
                flags = 1;
                const unsigned int ai = atomicAdd(&D.ctr->n_active, 1u);
                if (ai < D.cap_active) D.active_list[ai] = p; else atomicOr(&D.ctr->overflow, 4u);
                const Tl0 t0 = D.tl0[p]; const Tl1 t1 = D.tl1[p];
                const bool inhosp = tl_in(t0.hosp_begin, t0.severe_end, day);
                const unsigned char pb = D.pbyte[p];
                const unsigned char npb = (unsigned char) ((pb & ~(PB_INHOSP | PB_NEW)) | (inhosp ? PB_INHOSP : 0));
                if (npb != pb) D.pbyte[p] = npb;
                if (tl_in(t0.infectious_begin, t0.infectious_end, day)) {
                    const unsigned long long c = (unsigned long long) __double2ll_rn(CABM_MUL((double) t1.rel_inf, 4294967296.0));
                    const int hm = D.home[p];
                    if (inhosp) { add_pressure(D, D.loc_hospital[hm], t1.strain, c); nflag++; }
                    else {
                        add_pressure(D, hm, t1.strain, c); nflag++;
                        const int dl = D.day_loc[p];
                        if (dl >= 0 && !(tl_in(t0.severe_begin, t0.severe_end, day) || tl_in(t1.critical_begin, t1.critical_end, day))) { add_pressure(D, dl, t1.strain, c); nflag++; }
                        const int npat = pb & PB_NPAT;
                        if (npat) {
                            int locs[MAXV], hours[MAXV];
                            const int nv = gen_visits(D, cx, p, npat, locs, hours);
                            for (int j = 0; j < nv; ++j) {
                                const unsigned long long cv = (unsigned long long) __double2ll_rn(CABM_MUL(CABM_MUL((double) t1.rel_inf, (double) hours[j]), 536870912.0));
                                if (cv) { add_pressure(D, locs[j], t1.strain, cv); nflag++; }
                            }
                        }
                    }
                }
            }
        }
        bt.add(flags); bt.add_count(1, nflag);
    }
    __syncthreads();
    if (threadIdx.x == 0) { if (sm[0]) atomicAdd(&D.ctr->work[CABM_W_ACTIVE], (unsigned long long) sm[0]); if (sm[1]) atomicAdd(&D.ctr->work[CABM_W_FLAGS], (unsigned long long) sm[1]); }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&D.ctr->work[CABM_W_AGENTS], (unsigned long long) D.N);
}

__global__ void __launch_bounds__(TPB) k_visit(const Dev D) {
    __shared__ unsigned int sm[4]; BlockTally<4> bt; bt.init(sm);
    const DayCtx& cx = day_ctx(D);
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int n_round = (D.N + stride - 1) / stride * stride;
    for (unsigned int p = blockIdx.x * blockDim.x + threadIdx.x; p < n_round; p += stride) {
        unsigned int isp = 0, v0 = 0, v = 0;
        if (p < D.N) {
            const unsigned char pb = D.pbyte[p];
            const int npat = pb & PB_NPAT;
            if (npat && !(pb & PB_INHOSP)) {
                isp = 1;
                int locs[MAXV], hours[MAXV];
                const int nv = gen_visits(D, cx, p, npat, locs, hours);
                v0 = (unsigned int) nv;
                for (int j = 0; j < nv; ++j) {
                    if (D.hot_idx[locs[j]] != -1) {
                        atomicAdd(&D.nvis[locs[j]], 1u);
                        const unsigned int vi = atomicAdd(&D.ctr->n_vis, 1u);
                        if (vi < D.cap_vis) D.visit_rec[vi] = make_uint2(p, (unsigned int) locs[j]); else atomicOr(&D.ctr->overflow, 8u);
                        v++;
                    }
                }
            }
        }
        bt.add(isp); bt.add_count(1, v0); bt.add_count(2, v);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (sm[0]) atomicAdd(&D.ctr->work[CABM_W_PATRONS], (unsigned long long) sm[0]);
        if (sm[1]) atomicAdd(&D.ctr->work[CABM_W_VISITS], (unsigned long long) sm[1]);
        if (sm[2]) atomicAdd(&D.ctr->work[CABM_W_HOT_VISITS], (unsigned long long) sm[2]);
    }
}

__device__ __forceinline__ double hazard_to_T(double hazard) { return CABM_SUB(1.0, cabm_exp(-hazard)); }

__device__ __forceinline__ void emit_items(const Dev& D, int src, unsigned int begin, unsigned int count, int purpose, double T) {
    if (count == 0) return;
    const unsigned int nch = (count + CHUNK - 1) / CHUNK;
    const unsigned int base = atomicAdd(&D.ctr->n_items, nch);
    if (base + nch > D.cap_items) { atomicOr(&D.ctr->overflow, 16u); return; }
    for (unsigned int k = 0; k < nch; ++k) {
        Item it; it.src = src; it.begin = (int) (begin + k * CHUNK); it.count = (int) min((unsigned int) CHUNK, count - k * CHUNK); it.purpose = purpose; it.T = T;
        D.items[base + k] = it;
    }
}

// phase = 0: HOUSE, WORK, SCHOOL, NURSINGHOME (Community.cpp:1151-1161); phase = 1: HOSPITAL after the status pass (:1171)
__global__ void __launch_bounds__(TPB) k_hazard(const Dev D, const int phase) {
    const DayCtx& cx = day_ctx(D);
    const cabm_params* __restrict__ par = D.par;
    const unsigned int n_hot = min(D.ctr->n_hot, D.cap_hot);
    unsigned int edges = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_hot; i += gridDim.x * blockDim.x) {
        const int loc = D.hot_list[i];
        const int type = D.loc_type[loc];
        if ((type == CABM_HOSPITAL) != (phase == 1)) { if (phase == 0) { D.hot_T[i] = 0.0; D.hot_Ts[i] = 0.0; D.hot_size[i] = 0; } continue; }
        const unsigned long long* pw = &D.press[(size_t) loc * CABM_N_STRAINS];
        const double w = CABM_MUL((double) (pw[0] + pw[1] + pw[2] + pw[3]), 1.0 / 4294967296.0);
        const unsigned int rb = D.roster_off[loc], size = D.roster_off[loc + 1] - rb;
        double T = 0.0, Ts = 0.0; unsigned int denom = size;
        if (type == CABM_HOUSE) {
            T = hazard_to_T(CABM_MUL(CABM_MUL(par->haz_mult[CABM_HOUSE], cx.season), w));
            if (T > 0.0) emit_items(D, loc, rb, size, CABM_P_EXPOSE_HH, T);
            if ((double) D.loc_risk[loc] > cx.sd) {
                Ts = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(par->haz_mult[CABM_N_LOCTYPES], cx.season), w), (double) (int) size));
                for (unsigned int e = D.nbr_off[loc]; e < D.nbr_off[loc + 1]; ++e) {
                    const int nb = D.nbr[e]; edges++;
                    if ((double) D.loc_risk[nb] > cx.sd && Ts > 0.0) emit_items(D, loc, D.roster_off[nb], D.roster_off[nb + 1] - D.roster_off[nb], CABM_P_EXPOSE_NB, Ts);
                }
            }
        } else if (type == CABM_WORK) {
            const unsigned char lf = D.loc_flags[loc];
            denom = size + D.nvis[loc];
            const bool closed = !(lf & LF_ESSENTIAL) && cx.business_closure != 0.0;
            if (!(denom < 2 || closed) && w > 0.0) {
                const double f = LF_PUBRISK(lf) == CABM_PT_HIGH ? 4.0 : 0.25;
                T = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(CABM_MUL(par->haz_mult[CABM_WORK], f), cx.season), w), CABM_SUB((double) (int) denom, 1.0)));
                if (T > 0.0) emit_items(D, loc, rb, size, CABM_P_EXPOSE_WORK, T);
            }
        } else if (type == CABM_SCHOOL) {
            const bool open = cx.is_weekday && cx.school_closure < 1;
            const double coef = CABM_MUL(CABM_MUL(CABM_SUB(1.0, cx.school_closure), par->haz_mult[CABM_SCHOOL]), cx.season);
            if (open && coef != 0.0 && size >= 2) {
                T = hazard_to_T(CABM_DIV(CABM_MUL(coef, w), CABM_SUB((double) (int) size, 1.0)));
                if (T > 0.0) emit_items(D, loc, rb, size, CABM_P_EXPOSE_SCHOOL, T);
            }
        } else if (type == CABM_NURSINGHOME) {
            if (size >= 2) {
                T = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(par->haz_mult[CABM_NURSINGHOME], cx.season), w), CABM_SUB((double) (int) size, 1.0)));
                if (T > 0.0) emit_items(D, loc, rb, size, CABM_P_EXPOSE_NH, T);
            }
        } else {  // HOSPITAL
            denom = size + D.pat_count[loc];
            if (denom >= 2) {
                T = hazard_to_T(CABM_DIV(CABM_MUL(CABM_MUL(par->haz_mult[CABM_HOSPITAL], cx.season), w), CABM_SUB((double) (int) denom, 1.0)));
                if (T > 0.0) emit_items(D, loc, rb, size, CABM_P_EXPOSE_HOSP, T);
            }
        }
        D.hot_T[i] = T; D.hot_Ts[i] = Ts; D.hot_size[i] = denom;
    }
    if (edges) atomicAdd(&D.ctr->work[CABM_W_NBR_EDGES], (unsigned long long) edges);
}

// one exposure trial: Community::_transmission loop body (Community.cpp:973-994), gate only
__device__ __forceinline__ unsigned int try_expose(const Dev& D, const DayCtx& cx, unsigned int pid, int src, int purpose, unsigned int sub, double T) {
    if (D.any_quarantine) {
        const unsigned int q = D.quar[pid]; const int qf = (short) (q & 0xFFFF), qu = (short) (q >> 16);
        if (cx.day >= qf && cx.day < qu && src != D.home[pid]) return 0;
    }
    Stream s; s.begin(D.par->seed, (unsigned int) cx.day, (unsigned int) purpose, pid, (unsigned int) src, sub);
    if (s.u01() < T) {
        const int strain = pick_strain(&D.press[(size_t) src * CABM_N_STRAINS], s.next());
        (void) s.next();
        InfectCtx c; c.pid = (int) pid; c.time = cx.day; c.source = src; c.strain = strain; c.has_source = true;
        c.purpose = (unsigned int) purpose; c.key_a = pid; c.key_b = (unsigned int) src; c.key_sub = sub;
        if (infect_event(D, s, c, false)) {
            int pass = purpose, role = 0;
            if (purpose == CABM_P_EXPOSE_WORK_VIS) { pass = PASS_WORK; role = 0; }
            else if (purpose == CABM_P_EXPOSE_WORK) { pass = PASS_WORK; role = 1; }
            else if (purpose == CABM_P_EXPOSE_HOSP_PAT) { pass = PASS_HOSP; role = 1 + (int) sub; }
            const unsigned long long old = atomicMin(&D.cand[pid], make_key(pass, (unsigned int) src, (unsigned int) role));
            if (old == CABM_CAND_EMPTY) {
                const unsigned int ni = atomicAdd(&D.ctr->n_new, 1u);
                if (ni < D.cap_new) D.newinf_list[ni] = pid; else atomicOr(&D.ctr->overflow, 32u);
            }
        }
    }
    return 1;
}

__global__ void __launch_bounds__(TPB) k_expose_items(const Dev D) {
    const DayCtx& cx = day_ctx(D);
    const unsigned int n_items = min(D.ctr->n_items, D.cap_items);
    const unsigned int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31, n_warps = (gridDim.x * blockDim.x) >> 5;
    unsigned int trials = 0;
    for (unsigned int i = warp; i < n_items; i += n_warps) {
        const Item it = D.items[i];
        for (int j = (int) lane; j < it.count; j += 32) trials += try_expose(D, cx, (unsigned int) D.roster[it.begin + j], it.src, it.purpose, 0, it.T);
    }
    for (int o = 16; o > 0; o >>= 1) trials += __shfl_xor_sync(0xffffffffu, trials, o);
    if (lane == 0 && trials) atomicAdd(&D.ctr->work[CABM_W_TRIALS], (unsigned long long) trials);
}

__global__ void __launch_bounds__(TPB) k_expose_visits(const Dev D) {
    const DayCtx& cx = day_ctx(D);
    const unsigned int n = min(D.ctr->n_vis, D.cap_vis);
    unsigned int trials = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint2 v = D.visit_rec[i];
        const int hi = D.hot_idx[v.y];
        const double T = hi >= 0 ? D.hot_T[hi] : 0.0;
        if (T > 0.0) trials += try_expose(D, cx, v.x, (int) v.y, CABM_P_EXPOSE_WORK_VIS, 0, T);
    }
    if (trials) atomicAdd(&D.ctr->work[CABM_W_TRIALS], (unsigned long long) trials);
}

__global__ void __launch_bounds__(TPB) k_expose_patients(const Dev D) {
    const DayCtx& cx = day_ctx(D);
    const unsigned int n = min(D.ctr->n_pat, D.cap_pat);
    unsigned int trials = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned int pid = D.patient_list[i];
        const unsigned int c = D.hosp_count[pid];
        if (c == 0) continue;
        const int hosp = D.loc_hospital[D.home[pid]];
        const int hi = hosp >= 0 ? D.hot_idx[hosp] : -1;
        const double T = hi >= 0 ? D.hot_T[hi] : 0.0;
        if (T > 0.0) for (unsigned int sub = 0; sub < c; ++sub) trials += try_expose(D, cx, pid, hosp, CABM_P_EXPOSE_HOSP_PAT, sub, T);
    }
    if (trials) atomicAdd(&D.ctr->work[CABM_W_TRIALS], (unsigned long long) trials);
}
