// covid-abm_b200/csrc/cabm_infect.cuh -- the per-exposure state-transition sampler.
//
// infect_event() is Person::infect (Person.cpp:320-494) for one agent: susceptibility gate
// (calculate_event_probabilities :274-316, isCrossProtected :497-535, isVaccineProtected :538-559),
// disease-course sampling through the integer-support CDF tables of cabm_params, processDeath
// (:264-271), the detection cascade with future-dated ledger writes (:419-472, Community.cpp:657-679).
// It is called twice per winning exposure: once with commit=false by the exposure kernels (gate
// only, against the start-of-phase state) and once with commit=true by k_apply for the winner.
// Draws are consumed from the event's Philox stream in the reference's own order.
#pragma once
#include "cabm_device.cuh"

struct InfectCtx { int pid, time, source, strain, infector; bool has_source; unsigned int purpose, key_a, key_b, key_sub; };

__device__ __forceinline__ bool infect_event(const Dev& D, Stream& s, const InfectCtx c, const bool commit) {
    const cabm_params* __restrict__ par = D.par;
    const int pid = c.pid, time = c.time, strain = c.strain;
    CABM_CHECK(D, pid >= 0 && (unsigned int) pid < D.N && strain >= 0 && strain < CABM_N_STRAINS && time >= 0 && time < D.R);
    const int a = D.age[pid]; const unsigned int pf = D.pflags[pid]; const int com = (pf & PF_COMORBID) ? 1 : 0;
    const unsigned int imm = D.imm[pid];
    const int n_inf = (int) ((imm & 0xFF) + ((imm >> 8) & 0xFF) + ((imm >> 16) & 0xFF) + (imm >> 24));
    const unsigned int vax = D.vax[pid];
    const int n_doses = (int) (vax & 0xFF); const bool naive_vax = (vax >> 8) & 1; const int last_dose = (int) (short) (vax >> 16);
    Tl0 o0; o0.infected = CABM_I16_MAX; o0.infectious_end = CABM_I16_MIN; o0.symptom_end = CABM_I16_MIN;
    if (n_inf > 0) o0 = D.ag[pid].t0;

    // ---- isCrossProtected (Person.cpp:497-535) ----
    bool crossProtected = false;
    if (n_inf > 0) {
        const int last_end = max((int) o0.infectious_end, (int) o0.symptom_end);
        const int tsl = time - last_end;
        double rem = remaining_efficacy(par, D.nat_eff[pid], (double) tsl);
        rem = CABM_MUL(rem, CABM_SUB(1.0, par->strain_immune_escape[strain]));
        if (par->immunity_leaky) { if (s.u01() < rem) crossProtected = true; }
        else                     { if (tsl < D.nat_dur[pid]) crossProtected = true; }
        if (!crossProtected) {
            // for (inf : history) if (matrix[inf.strain][strain] and (tsl < 60 or U < 0.85)) -- only the count of matching entries matters
            int m = 0;
#pragma unroll
            for (int k = 0; k < CABM_N_STRAINS; ++k) if (par->cross_protection[k][strain]) m += (int) ((imm >> (8 * k)) & 0xFF);
            for (int k = 0; k < m; ++k) { if (tsl < 60 || s.u01() < 0.85) { crossProtected = true; break; } }
        }
        if (!crossProtected) { if ((double) D.xprot[pid] < par->strain_immune_escape[strain]) crossProtected = true; }
    }
    // ---- isVaccineProtected (Person.cpp:538-559), leaky model ----
    bool vaccineProtected = false;
    const int dose = n_doses > 0 ? min(n_doses - 1, CABM_MAX_DOSES - 1) : 0;
    if (n_doses > 0) {
        const double init_ves = naive_vax ? par->VES_NAIVE[strain][dose] : par->VES[strain][dose];
        const int dsv = time - last_dose;
        double rem = remaining_efficacy(par, init_ves, (double) (dsv - par->vaccine_dose_to_protection_lag));
        rem = CABM_MUL(rem, CABM_SUB(1.0, par->strain_immune_escape[strain]));
        const bool beyond = dsv >= par->vaccine_dose_to_protection_lag;
        const bool remains = s.u01() < rem;
        vaccineProtected = beyond && remains;
    }
    double susceptibility = 1.0;
    if (time < 90) susceptibility = (double) par->susceptibility_by_age[a];
    const double pr_inf = (crossProtected || vaccineProtected) ? 0.0 : susceptibility;

    // ---- bail (Person.cpp:326) ----
    if (n_inf > 0 && tl_in(o0.infected, o0.infectious_end, time)) return false;
    if (s.u01() > pr_inf) return false;
    if (!commit) return true;
    // the infection is certain from here on: its log entry is reserved now, so the round trip of the returning atomic runs beside the
    // disease course below instead of at the end of this chain (a winner's lane is alone in its warp on a light day)
    const unsigned long long li = atomicAdd(&D.ctr->n_log, 1ull);

    // ---- event probabilities (Person.cpp:284-315) ----
    const bool vaccinated = n_doses > 0, prior = n_inf > 0;
    const double VEP = vaccinated ? par->VEP[strain][dose] : 0.0, VEH = vaccinated ? par->VEH[strain][dose] : 0.0;
    const double VEF = vaccinated ? par->VEF[strain][dose] : 0.0, VEI = vaccinated ? par->VEI[strain][dose] : 0.0;
    double pr_symp = CABM_MUL((double) par->pathogenicity_by_age[a], par->strain_rel_pathogenicity[strain]);
    pr_symp = CABM_MUL(pr_symp, prior ? CABM_SUB(1.0, par->IEP) : 1.0);
    pr_symp = CABM_MUL(pr_symp, CABM_SUB(1.0, VEP));
    double pr_sev = CABM_MUL((double) par->prob_severe[com][a], par->strain_rel_severity[strain]);
    pr_sev = CABM_MUL(pr_sev, prior ? CABM_SUB(1.0, par->IEH) : 1.0);
    pr_sev = CABM_MUL(pr_sev, CABM_SUB(1.0, VEH));
    const double pr_hosp = (double) ((pf & PF_LTC) ? 0.2f : 0.45f);
    const double pr_crit = (double) par->prob_critical[com][a];
    double mort = par->strain_rel_mortality[strain];
    mort = CABM_MUL(mort, prior ? CABM_SUB(1.0, par->IEF) : 1.0);
    mort = CABM_MUL(mort, CABM_SUB(1.0, VEF));
    const double pr_nonicu_death = CABM_MUL((double) 0.9f, mort);
    const double icu_coef = CABM_MUL(par->strain_rel_icu_mortality[strain], mort);

    // ---- new infection record ----
    int t_infectious_begin, t_infectious_end, t_symptom_begin = CABM_I16_MAX, t_symptom_end = CABM_I16_MIN;
    int t_severe_begin = CABM_I16_MAX, t_severe_end = CABM_I16_MIN, t_hosp_begin = CABM_I16_MAX;
    int t_critical_begin = CABM_I16_MAX, t_critical_end = CABM_I16_MIN, t_icu_begin = CABM_I16_MAX, t_death = CABM_I16_MAX;
    unsigned long long tally = 1;                                    // bit o set => tallyOutcome(o)
    const int incubation = 1 + (int) table_draw(par->incub_cdf[strain], CABM_INCUB_MAX, s.u01());
    CABM_CHECK(D, incubation >= 1 && incubation <= CABM_INCUB_MAX);
    const int onset = 1 + (int) table_draw(par->onset_cdf[incubation - 1], CABM_INCUB_MAX + 1, s.u01());
    t_infectious_begin = time + onset;
    t_infectious_end = time + incubation + par->strain_infectious_period[c.has_source ? strain : 0];
    double rel = 1.0;
    rel = CABM_MUL(rel, par->strain_rel_infectiousness[strain]);
    rel = CABM_MUL(rel, prior ? CABM_SUB(1.0, par->IEI) : 1.0);
    rel = CABM_MUL(rel, CABM_SUB(1.0, VEI));
    if (s.u01() > pr_symp) {
        rel = CABM_MUL(rel, (s.u01() < par->p_high_asym) ? 4.0 : 0.25);
    } else {
        rel = CABM_MUL(rel, (s.u01() < par->p_high_sym) ? 4.0 : 0.25);
        tally |= 2;
        t_symptom_begin = time + incubation;
        if (!(s.u01() < pr_sev)) {
            t_symptom_end = t_symptom_begin + 1 + (int) table_draw(par->mild_cdf, CABM_T_GAMMA, s.u01());
        } else {
            tally |= 4;
            t_severe_begin = t_symptom_begin + 1 + (int) table_draw(par->presev_cdf, CABM_T_GAMMA, s.u01());
            bool hosp = false;
            if (s.u01() < pr_hosp) { hosp = true; t_hosp_begin = t_severe_begin; }
            if (!(s.u01() < pr_crit)) {
                t_severe_end = t_severe_begin + ref_sampler(par->severe_only_cdf, CABM_N_SEVERE_CDF, s.u01());
                t_symptom_end = t_severe_end;
            } else {
                tally |= 8;
                const int sod = ref_sampler(par->severe_only_cdf, CABM_N_SEVERE_CDF, s.u01());
                const int pre = (sod + 1) / 2, post = sod - pre;
                t_critical_begin = t_severe_begin + pre;
                t_critical_end = t_critical_begin + ref_sampler(par->critical_cdf, CABM_N_CRITICAL_CDF, s.u01());
                t_severe_end = t_critical_end + post;
                t_symptom_end = t_severe_end;
                const double icu_prob = hosp ? 1.0 : 0.5;
                int death_time = CABM_I16_MAX;
                if (s.u01() < icu_prob) {
                    t_icu_begin = t_critical_begin;
                    double imr_v = 0.0;
                    if (D.n_imr > 0) imr_v = t_icu_begin < D.n_imr ? D.imr[t_icu_begin] : D.imr[D.n_imr - 1];
                    const double icu_mort = CABM_MUL(CABM_MUL((double) par->prob_icu_death[com][a], CABM_SUB(1.0, imr_v)), icu_coef);
                    if (!hosp) t_hosp_begin = t_icu_begin;
                    if (s.u01() < icu_mort) death_time = t_critical_begin + (int) table_draw(par->icu_death_cdf, CABM_T_NEGBIN, s.u01());
                } else {
                    if (s.u01() < pr_nonicu_death) death_time = t_critical_begin + 1 + (int) table_draw(par->community_death_cdf, CABM_T_GEOM, s.u01());
                }
                if (death_time != CABM_I16_MAX) {
                    tally |= 16;
                    t_death = death_time;
                    t_infectious_end = min(t_infectious_end, death_time); t_symptom_end = min(t_symptom_end, death_time);
                    t_severe_end = min(t_severe_end, death_time); t_critical_end = min(t_critical_end, death_time);
                }
            }
        }
    }
    const float rel_inf = (float) rel;

    // ---- detection / reporting (Person.cpp:419-472) ----
    int detected_state = CABM_N_OUTCOMES; int report_day = CABM_I16_MAX;
    if (pid < par->num_surveilled_people && time >= 0) {
        Stream r; r.begin(((unsigned long long) s.key.y << 32) | s.key.x, (unsigned int) time, c.purpose | CABM_P_REPORTING, c.key_a, c.key_b, c.key_sub);
        bool detected = false; int state = CABM_N_OUTCOMES; long sample = 0, report = 0;
        const long lag = par->default_reporting_lag;
        const double* __restrict__ pf5 = &D.pfd[(size_t) time * CABM_N_OUTCOMES];
        if (s.u01() < pf5[0]) {
            const unsigned int period = (unsigned int) (t_infectious_end - t_infectious_begin);
            const int tracing_lag = (int) __umulhi(s.next(), period);
            sample = t_infectious_begin + tracing_lag; report = sample + lag; state = 0; detected = true;
        } else if (t_symptom_begin != CABM_I16_MAX && s.u01() < pf5[1]) {
            sample = t_symptom_begin + par->symptom_to_test_lag; report = sample + lag; state = 1; detected = true;
        } else if (t_severe_begin != CABM_I16_MAX && s.u01() < pf5[2]) {
            sample = t_severe_begin; report = sample + lag; state = 2; detected = true;
        } else if (t_critical_begin != CABM_I16_MAX && s.u01() < pf5[3]) {
            sample = t_critical_begin; report = sample + lag; state = 3; detected = true;
        } else if (t_death != CABM_I16_MAX && s.u01() < pf5[4]) {
            sample = t_death; report = sample + (long) table_draw(par->death_lag_cdf, CABM_T_DEATHLAG, r.u01()); state = 4; detected = true;
        }
        if (report >= (long) D.R) detected = false;
        const long death_rep_lag = t_death != CABM_I16_MAX ? (long) table_draw(par->death_lag_cdf, CABM_T_DEATHLAG, r.u01()) : 0;
        if (detected) {
            detected_state = state; report_day = (int) report;
            ledger_add(D, CABM_C_DET_CASES_ONSET, sample);
            if (report < (long) D.R) { ledger_add(D, CABM_C_DET_CASES_REPORT, report); if (t_hosp_begin != CABM_I16_MAX) ledger_add(D, CABM_C_DET_HOSP, report); }
            if (t_death != CABM_I16_MAX) {
                const long me = max((long) t_death + death_rep_lag, report);
                ledger_add(D, CABM_C_DET_DEATHS_ONSET, t_death); ledger_add(D, CABM_C_DET_DEATHS_REPORT, me);
            }
        }
    }

    // ---- commit ----
    Tl0 n0; Tl1 n1;
    n0.infected = (short) time; n0.infectious_begin = (short) t_infectious_begin; n0.infectious_end = (short) t_infectious_end;
    n0.symptom_begin = (short) t_symptom_begin; n0.symptom_end = (short) t_symptom_end; n0.severe_begin = (short) t_severe_begin;
    n0.severe_end = (short) t_severe_end; n0.hosp_begin = (short) t_hosp_begin;
    n1.critical_begin = (short) t_critical_begin; n1.critical_end = (short) t_critical_end; n1.icu_begin = (short) t_icu_begin;
    n1.death = (short) t_death; n1.report_day = (short) report_day; n1.strain = (unsigned char) strain; n1.detected = (unsigned char) detected_state;
    n1.rel_inf = rel_inf;
    D.ag[pid].t0 = n0; D.ag[pid].t1 = n1;
    D.inf_place[pid] = c.has_source ? c.source : -1; D.inf_by[pid] = c.infector;
    const int until = tl_until(n0, n1);
    (void) until;
    atomicOr(&D.actbits[(unsigned int) pid >> 5], 1u << (pid & 31));      // joins the active set (k_active takes it out after day tl_until)
    const unsigned int cnt = (imm >> (8 * strain)) & 0xFF;
    if (cnt < 255) D.imm[pid] = imm + (1u << (8 * strain));
    if (par->retroactive_mature_vaccine && n_doses > 0) D.vax[pid] = vax & ~0x100u;
#pragma unroll
    for (int o = 0; o < CABM_N_OUTCOMES; ++o) if (tally & (1ull << o)) atomicAdd(&D.ctr->cumul[o], 1ull);
    atomicAdd(&D.ctr->work[CABM_W_INFECTIONS], 1ull);
    if (li < D.cap_log) {
        cabm_infection rec;
        rec.person = (unsigned int) pid; rec.infected_by = c.infector; rec.infected_place = c.has_source ? c.source : -1;
        rec.t[0] = n0.infected; rec.t[1] = n0.infectious_begin; rec.t[2] = n0.infectious_end; rec.t[3] = n0.symptom_begin; rec.t[4] = n0.symptom_end;
        rec.t[5] = n0.severe_begin; rec.t[6] = n0.severe_end; rec.t[7] = n0.hosp_begin; rec.t[8] = n1.critical_begin; rec.t[9] = n1.critical_end;
        rec.t[10] = n1.icu_begin; rec.t[11] = n1.death; rec.report_day = n1.report_day; rec.strain = n1.strain; rec.detected_state = n1.detected;
        rec.rel_infectiousness = rel_inf;
        D.log[li] = rec;
    } else { atomicOr(&D.ctr->overflow, 1u); }
    return true;
}
