// covid-abm_b200/csrc/cabm_kernels2.cuh -- winner application, status pass, importation, end of day.
#pragma once
#include "cabm_kernels.cuh"

// covid::util::weighted_choice over the strain urn (Utility.h:93-105, simulator.h:215)
__device__ __forceinline__ int weighted_strain(const double* __restrict__ w, double u) {
    const double sum = CABM_ADD(CABM_ADD(CABM_ADD(w[0], w[1]), w[2]), w[3]);
    double ran = CABM_MUL(sum, u);
    for (int k = 0; k < CABM_N_STRAINS; ++k) { if (ran < w[k]) return k; ran = CABM_SUB(ran, w[k]); }
    return CABM_N_STRAINS - 1;
}

// Person::infect for the winning exposure of each newly infected agent (first in reference order)
__global__ void __launch_bounds__(128) k_apply(const Dev D) {
    const DayCtx& cx = day_ctx(D);
    const unsigned int lo = D.ctr->n_new_applied, hi = min(D.ctr->n_new, D.cap_new);
    for (unsigned int i = lo + blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += gridDim.x * blockDim.x) {
        const unsigned int pid = D.newinf_list[i];
        const unsigned long long key = D.cand[pid];
        D.cand[pid] = CABM_CAND_EMPTY;
        const int pass = (int) (key >> 40); const unsigned int src = (unsigned int) (key >> 8); const unsigned int role = (unsigned int) (key & 0xFF);
        Stream s; InfectCtx c; c.pid = (int) pid; c.time = cx.day;
        if (pass == PASS_IMPORT) {
            s.begin(D.par->seed, (unsigned int) cx.day, CABM_P_IMPORT, src, 0);
            (void) s.next();
            c.strain = weighted_strain(cx.strain_w, s.u01());
            c.source = -1; c.has_source = false; c.purpose = CABM_P_IMPORT; c.key_a = src; c.key_b = 0; c.key_sub = 0;
        } else {
            int purpose = pass; unsigned int sub = 0;
            if (pass == PASS_WORK) purpose = role ? CABM_P_EXPOSE_WORK : CABM_P_EXPOSE_WORK_VIS;
            else if (pass == PASS_HOSP && role) { purpose = CABM_P_EXPOSE_HOSP_PAT; sub = role - 1; }
            s.begin(D.par->seed, (unsigned int) cx.day, (unsigned int) purpose, pid, src, sub);
            (void) s.next();
            c.strain = pick_strain(&D.press[(size_t) src * CABM_N_STRAINS], s.next());
            (void) s.next();
            c.source = (int) src; c.has_source = true; c.purpose = (unsigned int) purpose; c.key_a = pid; c.key_b = src; c.key_sub = sub;
        }
        infect_event(D, s, c, true);
    }
}

__global__ void k_phase_end(const Dev D) { if (threadIdx.x == 0 && blockIdx.x == 0) D.ctr->n_new_applied = min(D.ctr->n_new, D.cap_new); }

// Community::updatePersonStatus (Community.cpp:682-741) for one agent; returns the ledger increments as a bit mask
__device__ __forceinline__ unsigned int status_one(const Dev& D, unsigned int p, int day) {
    const Tl0 t0 = D.tl0[p]; const Tl1 t1 = D.tl1[p];
    const int hosp = D.loc_hospital[D.home[p]]; const int dl = D.day_loc[p];
    const bool inh = tl_in(t0.hosp_begin, t0.severe_end, day);
    if (hosp >= 0 && dl != hosp) {
        if (inh) {
            if (t0.hosp_begin == day) {                                             // goToHospital (Person.h:248)
                D.hosp_count[p] = D.hosp_count[p] + 1; atomicAdd(&D.pat_count[hosp], 1u);
                const unsigned char pb = D.pbyte[p];
                if (!(pb & 0x20)) {
                    D.pbyte[p] = pb | 0x20;
                    const unsigned int pi = atomicAdd(&D.ctr->n_pat, 1u);
                    if (pi < D.cap_pat) D.patient_list[pi] = p; else atomicOr(&D.ctr->overflow, 64u);
                }
            }
        } else if (day > 0 && tl_in(t0.hosp_begin, t0.severe_end, day - 1)) {     // leaveHospital (Person.h:249, Location.cpp:58-67)
            const unsigned char c = D.hosp_count[p];
            if (c > 0) { D.hosp_count[p] = c - 1; atomicSub(&D.pat_count[hosp], 1u); }
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

// mode 0: agents active at the start of the day (skipping those re-infected today); mode 1: today's phase-1 infections
__global__ void __launch_bounds__(TPB) k_status(const Dev D, const int mode) {
    __shared__ unsigned int sm[CABM_N_COUNTERS]; BlockTally<CABM_N_COUNTERS> bt; bt.init(sm);
    const int day = day_ctx(D).day;
    const unsigned int n = mode == 0 ? min(D.ctr->n_active, D.cap_active) : min(D.ctr->n_new, D.cap_new);
    const unsigned int* list = mode == 0 ? D.active_list : D.newinf_list;
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int n_round = (n + stride - 1) / stride * stride;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        unsigned int m = 0;
        if (i < n) {
            const unsigned int p = list[i];
            const unsigned char pb = D.pbyte[p];
            if (mode == 0) { if (!(pb & PB_NEW)) m = status_one(D, p, day); }
            else { D.pbyte[p] = pb & ~PB_NEW; m = status_one(D, p, day); }
        }
        bt.add(m);
    }
    __syncthreads();
    if (threadIdx.x < CABM_N_COUNTERS && sm[threadIdx.x]) atomicAdd(&D.ledger[(size_t) threadIdx.x * D.R + day], (unsigned long long) sm[threadIdx.x]);
}

// drop discharged agents from the persistent patient list (single block; the list is tiny)
__global__ void __launch_bounds__(TPB) k_compact_patients(const Dev D) {
    __shared__ unsigned int base;
    const unsigned int n = min(D.ctr->n_pat, D.cap_pat);
    if (threadIdx.x == 0) base = 0;
    __syncthreads();
    for (unsigned int i0 = 0; i0 < n; i0 += blockDim.x) {
        const unsigned int i = i0 + threadIdx.x;
        unsigned int pid = 0; bool keep = false;
        if (i < n) { pid = D.patient_list[i]; keep = D.hosp_count[pid] > 0; if (!keep) D.pbyte[pid] &= ~0x20; }
        const unsigned int b = __ballot_sync(0xffffffffu, keep);
        __shared__ unsigned int wbase[TPB / 32];
        if ((threadIdx.x & 31) == 0) wbase[threadIdx.x >> 5] = __popc(b);
        __syncthreads();
        if (threadIdx.x == 0) { unsigned int acc = base; for (int w = 0; w < TPB / 32; ++w) { const unsigned int c = wbase[w]; wbase[w] = acc; acc += c; } base = acc; }
        __syncthreads();
        if (keep) D.patient_tmp[wbase[threadIdx.x >> 5] + __popc(b & ((1u << (threadIdx.x & 31)) - 1))] = pid;
        __syncthreads();
    }
    for (unsigned int i = threadIdx.x; i < base; i += blockDim.x) D.patient_list[i] = D.patient_tmp[i];
    __syncthreads();
    if (threadIdx.x == 0) D.ctr->n_pat = base;
}

// seed_epidemic(par, community, date, strain_weights) (simulator.h:201-223): gate of every import of the day
__global__ void __launch_bounds__(128) k_import(const Dev D) {
    const DayCtx& cx = day_ctx(D);
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < (unsigned int) cx.n_imports; i += gridDim.x * blockDim.x) {
        Stream s; s.begin(D.par->seed, (unsigned int) cx.day, CABM_P_IMPORT, i, 0);
        const unsigned int pid = __umulhi(s.next(), D.N);
        InfectCtx c; c.pid = (int) pid; c.time = cx.day; c.source = -1; c.has_source = false;
        c.strain = weighted_strain(cx.strain_w, s.u01());
        c.purpose = CABM_P_IMPORT; c.key_a = i; c.key_b = 0; c.key_sub = 0;
        if (infect_event(D, s, c, false)) {
            const unsigned long long old = atomicMin(&D.cand[pid], make_key(PASS_IMPORT, i, 0));
            if (old == CABM_CAND_EMPTY) {
                const unsigned int ni = atomicAdd(&D.ctr->n_new, 1u);
                if (ni < D.cap_new) D.newinf_list[ni] = pid; else atomicOr(&D.ctr->overflow, 32u);
            }
        }
    }
}

// Community::infect(id, strain) (Community.cpp:550-553): one agent, gate and commit in one go
__global__ void k_seed_one(const Dev D, const unsigned int pid, const int strain, const int day, int* out) {
    if (threadIdx.x || blockIdx.x) return;
    const unsigned int seq = D.ctr->seed_calls++;
    Stream s; s.begin(D.par->seed, (unsigned int) day, CABM_P_SEED, pid, seq);
    InfectCtx c; c.pid = (int) pid; c.time = day; c.source = -1; c.has_source = false; c.strain = strain;
    c.purpose = CABM_P_SEED; c.key_a = pid; c.key_b = seq; c.key_sub = 0;
    *out = infect_event(D, s, c, true) ? 1 : 0;
}

// updateHotLocations (Community.cpp:999-1003) + clear_public_activity (:1042-1046): reset what today's hot set touched
__global__ void __launch_bounds__(TPB) k_end(const Dev D) {
    const unsigned int n_hot = min(D.ctr->n_hot, D.cap_hot);
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_hot; i += gridDim.x * blockDim.x) {
        const int loc = D.hot_list[i];
        unsigned long long* pw = &D.press[(size_t) loc * CABM_N_STRAINS];
        if (D.foi_enabled) {
            cabm_foi_row r; r.loc = (unsigned int) loc; r.type = D.loc_type[loc]; r.pad[0] = r.pad[1] = r.pad[2] = 0;
            for (int s = 0; s < CABM_N_STRAINS; ++s) r.w[s] = CABM_MUL((double) pw[s], 1.0 / 4294967296.0);
            r.size = D.hot_size[i]; r.T = D.hot_T[i]; r.T_social = D.hot_Ts[i];
            D.foi[i] = r;
        }
        pw[0] = 0; pw[1] = 0; pw[2] = 0; pw[3] = 0;
        D.hot_idx[loc] = -1; D.nvis[loc] = 0;
    }
}

__global__ void k_tick_end(const Dev D) {
    if (threadIdx.x || blockIdx.x) return;
    Counters* c = D.ctr;
    c->work[CABM_W_HOT_LOCS] += min(c->n_hot, D.cap_hot);
    if (D.foi_enabled) *D.n_foi = min(c->n_hot, D.cap_hot);
    c->n_hot = 0; c->n_active = 0; c->n_vis = 0; c->n_items = 0; c->n_new = 0; c->n_new_applied = 0;
}
__global__ void k_items_reset(const Dev D) { if (threadIdx.x == 0 && blockIdx.x == 0) D.ctr->n_items = 0; }
__global__ void k_import_end(const Dev D, const int advance) {
    if (threadIdx.x || blockIdx.x) return;
    D.ctr->n_new = 0; D.ctr->n_new_applied = 0;
    if (advance) D.ctr->day_idx++;
}

// Person::vaccinate (Person.cpp:611-647), leaky model
__global__ void k_vaccinate(const Dev D, const unsigned int* pids, const unsigned int n, const int day, unsigned char* ok) {
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned int p = pids[i];
        const unsigned int imm = D.imm[p]; const bool had_inf = imm != 0;
        const bool dead = had_inf && D.tl1[p].death <= day;
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
            const Tl0 a = D.tl0[p]; const Tl1 b = D.tl1[p];
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

__global__ void k_init_state(const Dev D) {
    const unsigned int stride = gridDim.x * blockDim.x;
    for (unsigned int p = blockIdx.x * blockDim.x + threadIdx.x; p < D.N; p += stride) {
        D.win[p] = 0x80007FFFu;                       // from = 32767, until = -32768: never active
        D.cand[p] = CABM_CAND_EMPTY;
    }
    for (unsigned int l = blockIdx.x * blockDim.x + threadIdx.x; l < D.L; l += stride) D.hot_idx[l] = -1;
}
