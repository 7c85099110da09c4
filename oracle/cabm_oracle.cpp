// oracle/cabm_oracle.cpp -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
// Person-level half of the restatement: immunity, event probabilities and Person::infect.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include "cabm_oracle.h"
#include "../include/cabm_det_math.h"

namespace cabm_oracle {

void Stream::begin(uint64_t seed, uint32_t day, uint32_t purpose, uint32_t a, uint32_t b, uint32_t sub_) {
    philox_stream_begin(&s, seed, day, purpose, a, b); sub = sub_;
}
uint32_t Stream::next() {
    const uint32_t lane = s.idx & 3u;
    if (lane == 0) {
        uint32_t ctr[4] = { (sub << 16) | (s.idx >> 2), s.a, s.b, s.dp };
        philox4x32_10_ref(ctr, s.key, s.buf);
    }
    s.idx++;
    return s.buf[lane];
}

// Parameters::sampler (Parameters.h:433-436): first index whose CDF value is >= rand
static size_t ref_sampler(const double* cdf, size_t n, double rand) {
    size_t index = 0;
    while (index < n && cdf[index] < rand) index++;
    return index;
}

// Parameters::remainingEfficacy (Parameters.h:615-625) with _immunityEffectiveStartingTime (:611-613)
double Sim::remaining_efficacy(double starting_efficacy, double time_delta) const {
    if (!par.immunity_wanes) return starting_efficacy;
    const double slope = 0.2, intcpt = 2.4603816;
    const double month_delta = time_delta / 30.0;
    const double est = (cabm_log(1.0 / starting_efficacy - 1.0) + intcpt) / slope;
    const double l = intcpt - slope * (est + month_delta);
    return cabm_logistic(l);
}

// Person::isCrossProtected (Person.cpp:497-535)
bool Sim::is_cross_protected(const Person& p, int pid, int time, int strain, Stream& s) const {
    bool immune = false;
    if (p.hist.size() > 0) {
        const Infection& last = p.hist.back();
        const int last_end = std::max(last.t[T_INFECTIOUS_END], last.t[T_SYMPTOM_END]);          // :503-504
        const int tsl = time - last_end;
        double rem = remaining_efficacy(nat_eff[pid], (double) tsl);
        rem *= 1.0 - par.strain_immune_escape[strain];
        if (par.immunity_leaky) { if (s.u01() < rem) immune = true; }                             // :508
        else                    { if (tsl < nat_dur[pid]) immune = true; }                        // :509
        if (!immune) {
            for (const Infection& inf : p.hist) {                                                 // :515-524
                if (par.cross_protection[inf.strain][strain] && (tsl < 60 || s.u01() < 0.85)) { immune = true; break; }
            }
        }
        if (!immune) {
            if ((double) xprot[pid] < par.strain_immune_escape[strain]) immune = true;            // :529
        }
    }
    return immune;
}

// Person::isVaccineProtected (Person.cpp:538-559); only the leaky model is supported (Person.cpp:622 asserts it)
bool Sim::is_vaccine_protected(const Person& p, int time, int strain, Stream& s) const {
    bool immune = false;
    if (p.vax_days.size() > 0) {
        const size_t dose = std::min<size_t>(p.vax_days.size() - 1, CABM_MAX_DOSES - 1);
        const double init_ves = p.naive_vax ? par.VES_NAIVE[strain][dose] : par.VES[strain][dose];
        const int dsv = time - p.vax_days.back();
        double rem = remaining_efficacy(init_ves, (double) (dsv - par.vaccine_dose_to_protection_lag));
        rem *= 1.0 - par.strain_immune_escape[strain];
        const bool beyond = dsv >= par.vaccine_dose_to_protection_lag;
        const bool remains = s.u01() < rem;                                                       // :553 (leaky)
        immune = beyond && remains;
    }
    return immune;
}

// Person::infect (Person.cpp:320-494) incl. calculate_event_probabilities (:274-316),
// initializeNewInfection (:123-159), processDeath (:264-271) and the detection cascade (:419-472).
// `s` continues the event's RNG stream; the REPORTING_RNG draws come from the sibling stream
// (purpose | CABM_P_REPORTING) of the same key.
bool Sim::infect(int pid, int time, int source_loc, int strain, bool has_source, Stream& s, uint32_t purpose,
                 uint32_t key_a, uint32_t key_b, uint32_t key_sub) {
    Person& p = people[pid];
    const int a = age[pid], com = comorbid[pid];
    last_rep_draws = 0;
    // ---- calculate_event_probabilities ----
    const bool crossProtected   = is_cross_protected(p, pid, time, strain, s);
    const bool vaccineProtected = is_vaccine_protected(p, time, strain, s);
    double susceptibility = 1.0;
    if (time < 90) susceptibility = (double) par.susceptibility_by_age[a];                        // :280-282
    const double pr_inf = (crossProtected || vaccineProtected) ? 0.0 : susceptibility;
    const bool vaccinated = p.vax_days.size() > 0;
    const size_t dose = vaccinated ? std::min<size_t>(p.vax_days.size() - 1, CABM_MAX_DOSES - 1) : 0;
    const double VEP = vaccinated ? par.VEP[strain][dose] : 0.0, VEH = vaccinated ? par.VEH[strain][dose] : 0.0;
    const double VEF = vaccinated ? par.VEF[strain][dose] : 0.0;
    const bool prior = p.hist.size() > 0;
    double pr_symp = (double) par.pathogenicity_by_age[a] * par.strain_rel_pathogenicity[strain];
    pr_symp *= prior ? (1.0 - par.IEP) : 1.0;
    pr_symp *= 1.0 - VEP;
    double pr_sev = (double) par.prob_severe[com][a] * par.strain_rel_severity[strain];
    pr_sev *= prior ? (1.0 - par.IEH) : 1.0;
    pr_sev *= 1.0 - VEH;
    const double pr_hosp = (double) (ltc[pid] ? 0.2f : 0.45f);                                    // Parameters.h:257-258
    const double pr_crit = (double) par.prob_critical[com][a];
    double mort = par.strain_rel_mortality[strain];
    mort *= prior ? (1.0 - par.IEF) : 1.0;
    mort *= 1.0 - VEF;
    const double pr_nonicu_death = (double) 0.9f * mort;                                          // Parameters.h:312
    const double icu_coef = par.strain_rel_icu_mortality[strain] * mort;

    // ---- bail if already infected or not infectable (:326) ----
    if (p.hist.size() > 0 && p.hist.back().isInfected(time)) return false;
    if (s.u01() > pr_inf) return false;

    // ---- new infection record ----
    const int incubation = 1 + (int) table_draw(par.incub_cdf[strain], CABM_INCUB_MAX, s.u01()); // Parameters.h:515-520
    Infection inf;
    inf.t[T_INFECTED] = time; inf.place = has_source ? source_loc : -1; inf.strain = strain; inf.by = has_source ? pending_infector : -1;
    const int onset = 1 + (int) table_draw(par.onset_cdf[incubation - 1], CABM_INCUB_MAX + 1, s.u01()); // Parameters.h:521-524
    inf.t[T_INFECTIOUS_BEGIN] = time + onset;
    const int period_strain = has_source ? strain : 0;                                            // :153,156 (WILDTYPE period for imports)
    inf.t[T_INFECTIOUS_END] = time + incubation + par.strain_infectious_period[period_strain];
    cumul[0]++;                                                                                   // tallyOutcome(ASYMPTOMATIC)
    const bool reinfection = p.hist.size() + 1 > 1;
    const double VEI = vaccinated ? par.VEI[strain][dose] : 0.0;
    double rel = 1.0;
    rel *= par.strain_rel_infectiousness[strain];
    rel *= reinfection ? (1.0 - par.IEI) : 1.0;
    rel *= 1.0 - VEI;

    if (s.u01() > pr_symp) {                                                                      // :350
        rel *= (s.u01() < par.p_high_asym) ? 4.0 : 0.25;
    } else {
        rel *= (s.u01() < par.p_high_sym) ? 4.0 : 0.25;
        cumul[1]++;
        inf.t[T_SYMPTOM_BEGIN] = time + incubation;
        if (!(s.u01() < pr_sev)) {
            inf.t[T_SYMPTOM_END] = inf.t[T_SYMPTOM_BEGIN] + 1 + (int) table_draw(par.mild_cdf, CABM_T_GAMMA, s.u01());
        } else {
            cumul[2]++;
            inf.t[T_SEVERE_BEGIN] = inf.t[T_SYMPTOM_BEGIN] + 1 + (int) table_draw(par.presev_cdf, CABM_T_GAMMA, s.u01());
            bool hosp = false;
            if (s.u01() < pr_hosp) { hosp = true; inf.t[T_HOSP_BEGIN] = inf.t[T_SEVERE_BEGIN]; }
            if (!(s.u01() < pr_crit)) {
                inf.t[T_SEVERE_END] = inf.t[T_SEVERE_BEGIN] + (int) ref_sampler(par.severe_only_cdf, CABM_N_SEVERE_CDF, s.u01());
                inf.t[T_SYMPTOM_END] = inf.t[T_SEVERE_END];
            } else {
                cumul[3]++;
                const int sod = (int) ref_sampler(par.severe_only_cdf, CABM_N_SEVERE_CDF, s.u01());
                const int pre = (sod + 1) / 2;                                                    // round((float) sod / 2) :383
                const int post = sod - pre;
                inf.t[T_CRITICAL_BEGIN] = inf.t[T_SEVERE_BEGIN] + pre;
                inf.t[T_CRITICAL_END] = inf.t[T_CRITICAL_BEGIN] + (int) ref_sampler(par.critical_cdf, CABM_N_CRITICAL_CDF, s.u01());
                inf.t[T_SEVERE_END] = inf.t[T_CRITICAL_END] + post;
                inf.t[T_SYMPTOM_END] = inf.t[T_SEVERE_END];
                const double icu_prob = hosp ? 1.0 : 0.5;                                         // Parameters.h:259-260
                int death_time = INT_MAX;
                if (s.u01() < icu_prob) {
                    inf.t[T_ICU_BEGIN] = inf.t[T_CRITICAL_BEGIN];
                    double imr_v = 0.0;                                                           // Parameters.cpp:207-214
                    if (imr.size() > 0) imr_v = (size_t) inf.t[T_ICU_BEGIN] < imr.size() ? imr[inf.t[T_ICU_BEGIN]] : imr.back();
                    const double icu_mort = ((double) par.prob_icu_death[com][a] * (1.0 - imr_v)) * icu_coef;
                    if (!hosp) inf.t[T_HOSP_BEGIN] = inf.t[T_ICU_BEGIN];
                    if (s.u01() < icu_mort) death_time = inf.t[T_CRITICAL_BEGIN] + (int) table_draw(par.icu_death_cdf, CABM_T_NEGBIN, s.u01());
                } else {
                    if (s.u01() < pr_nonicu_death) death_time = inf.t[T_CRITICAL_BEGIN] + 1 + (int) table_draw(par.community_death_cdf, CABM_T_GEOM, s.u01());
                }
                if (death_time != INT_MAX) {                                                      // processDeath :264-271
                    cumul[4]++;
                    inf.t[T_DEATH] = death_time;
                    inf.t[T_INFECTIOUS_END] = std::min(inf.t[T_INFECTIOUS_END], death_time);
                    inf.t[T_SYMPTOM_END]    = std::min(inf.t[T_SYMPTOM_END], death_time);
                    inf.t[T_SEVERE_END]     = std::min(inf.t[T_SEVERE_END], death_time);
                    inf.t[T_CRITICAL_END]   = std::min(inf.t[T_CRITICAL_END], death_time);
                }
            }
        }
    }
    inf.rel_inf = (float) rel;

    // ---- detection / reporting (:419-472) ----
    if (pid < par.num_surveilled_people && time >= 0) {
        Stream r; r.begin(((uint64_t) s.s.key[1] << 32) | s.s.key[0], (uint32_t) time, purpose | CABM_P_REPORTING, key_a, key_b, key_sub);
        bool detected = false; int state = CABM_N_OUTCOMES; long sample = 0, report = 0;
        const long lag = par.default_reporting_lag;                                               // Parameters.h:539-545 (no rlm)
        const double* pf = &pfd[(size_t) time * CABM_N_OUTCOMES];
        if (s.u01() < pf[0]) {
            const uint32_t period = (uint32_t) (inf.t[T_INFECTIOUS_END] - inf.t[T_INFECTIOUS_BEGIN]);
            const int tracing_lag = (int) (((uint64_t) s.next() * (uint64_t) period) >> 32);      // gsl_rng_uniform_int :428
            sample = inf.t[T_INFECTIOUS_BEGIN] + tracing_lag; report = sample + lag; state = 0; detected = true;
        } else if (inf.t[T_SYMPTOM_BEGIN] != INT_MAX && s.u01() < pf[1]) {
            sample = inf.t[T_SYMPTOM_BEGIN] + par.symptom_to_test_lag; report = sample + lag; state = 1; detected = true;
        } else if (inf.t[T_SEVERE_BEGIN] != INT_MAX && s.u01() < pf[2]) {
            sample = inf.t[T_SEVERE_BEGIN]; report = sample + lag; state = 2; detected = true;
        } else if (inf.t[T_CRITICAL_BEGIN] != INT_MAX && s.u01() < pf[3]) {
            sample = inf.t[T_CRITICAL_BEGIN]; report = sample + lag; state = 3; detected = true;
        } else if (inf.t[T_DEATH] != INT_MAX && s.u01() < pf[4]) {
            sample = inf.t[T_DEATH]; report = sample + (long) table_draw(par.death_lag_cdf, CABM_T_DEATHLAG, r.u01()); state = 4; detected = true;
        }
        if (report >= (long) par.run_length) detected = false;                                    // :457
        const long death_rep_lag = inf.t[T_DEATH] != INT_MAX ? (long) table_draw(par.death_lag_cdf, CABM_T_DEATHLAG, r.u01()) : 0; // :459
        if (detected) {
            inf.detected_state = state; inf.report_day = (int) report;
            const bool hospitalized = inf.t[T_HOSP_BEGIN] != INT_MAX;
            // Community::reportCase (Community.cpp:657-667)
            if (sample < (long) par.run_length) ledger[CABM_C_DET_CASES_ONSET][sample]++;
            if (report < (long) par.run_length) {
                ledger[CABM_C_DET_CASES_REPORT][report]++;
                if (hospitalized) ledger[CABM_C_DET_HOSP][report]++;
            }
            if (inf.t[T_DEATH] != INT_MAX) {                                                      // reportDeath (Community.cpp:674-679)
                const long me = std::max((long) inf.t[T_DEATH] + death_rep_lag, report);
                if (inf.t[T_DEATH] < par.run_length) ledger[CABM_C_DET_DEATHS_ONSET][inf.t[T_DEATH]]++;
                if (me < (long) par.run_length) ledger[CABM_C_DET_DEATHS_REPORT][me]++;
            }
        }
        last_rep_draws = r.s.idx;
    }

    // ---- flag future hot locations (:475-486) ----
    const uint64_t c_fx = (uint64_t) llrint((double) inf.rel_inf * 4294967296.0);
    for (int d = std::max(inf.t[T_INFECTIOUS_BEGIN], 0); d < inf.t[T_INFECTIOUS_END]; d++) {
        if (d >= par.run_length) break;                                                           // Community.cpp:790
        if (inf.inHospital(d)) {
            hot[d].push_back(HotRec{ loc_hospital[home[pid]], pid, c_fx, (uint8_t) strain });
        } else {
            hot[d].push_back(HotRec{ home[pid], pid, c_fx, (uint8_t) strain });
            if (day_loc[pid] >= 0 && !(inf.isSevere(d) || inf.isCritical(d)))
                hot[d].push_back(HotRec{ day_loc[pid], pid, c_fx, (uint8_t) strain });
        }
    }
    if (par.retroactive_mature_vaccine) p.naive_vax = false;                                      // :491
    p.hist.push_back(inf);

    cabm_infection rec{};
    rec.person = (uint32_t) pid; rec.infected_by = inf.by; rec.infected_place = inf.place;
    for (int i = 0; i < T_N; ++i) rec.t[i] = (int16_t) std::max(-32768, std::min(32767, inf.t[i]));
    rec.report_day = (int16_t) std::min(32767, inf.report_day);
    rec.strain = (uint8_t) strain; rec.detected_state = (uint8_t) inf.detected_state; rec.rel_infectiousness = inf.rel_inf;
    log.push_back(rec);
    work[CABM_W_INFECTIONS]++;
    return true;
}

} // namespace cabm_oracle
