// oracle/cabm_oracle_api.cpp -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
// C entry points of the oracle; same signatures as include/cabm.h with the prefix cabm_oracle_.
#include <algorithm>
#include <cstring>
#include <new>
#include "cabm_oracle.h"
#include "cabm_det_math.h"
#include "cabm_popfile.h"

using cabm_oracle::Sim;
struct cabm { Sim sim; };
struct cabm_snap { Sim sim; };

#define H_OK(h) do { if (!(h)) return CABM_E_ARG; } while (0)

extern "C" {

int cabm_oracle_create(cabm_t** out) { if (!out) return CABM_E_ARG; *out = new (std::nothrow) cabm(); return *out ? CABM_OK : CABM_E_NOMEM; }
int cabm_oracle_destroy(cabm_t* h) { delete h; return CABM_OK; }
const char* cabm_oracle_last_error(const cabm_t* h) { return h ? h->sim.err.c_str() : "null handle"; }

// the binary population file (include/cabm_popfile.h): the same host code as the CUDA library's
int cabm_oracle_population_save(const cabm_population* pop, const char* path) { return cabm_popfile::save(pop, path); }
int cabm_oracle_load_population(cabm_t* h, const cabm_population* pop);
int cabm_oracle_load_population_file(cabm_t* h, const char* path, const double* starting_natural_efficacy,
                                     const float* cross_protection_probability, const int32_t* natural_immunity_duration) {
    H_OK(h); if (!path) return CABM_E_ARG;
    cabm_popfile::Mapped m;
    const int rc = m.open_file(path, &h->sim.err);
    if (rc) return rc;
    m.pop.starting_natural_efficacy = starting_natural_efficacy; m.pop.cross_protection_probability = cross_protection_probability;
    m.pop.natural_immunity_duration = natural_immunity_duration;
    return cabm_oracle_load_population(h, &m.pop);
}

int cabm_oracle_load_population(cabm_t* h, const cabm_population* pop) {
    H_OK(h); if (!pop) return CABM_E_ARG;
    Sim& S = h->sim;
    if (!pop->starting_natural_efficacy || !pop->cross_protection_probability || !pop->natural_immunity_duration) {
        S.err = "oracle requires explicit load-time immunity attributes"; return CABM_E_ARG;
    }
    const uint32_t N = pop->n_people, L = pop->n_loc;
    S.n_people = N; S.n_loc = L; S.n_patron = pop->n_patron;
    S.loc_type.assign(pop->loc_type, pop->loc_type + L); S.loc_essential.assign(pop->loc_essential, pop->loc_essential + L);
    S.loc_pubrisk.assign(pop->loc_pubrisk, pop->loc_pubrisk + L); S.loc_risk.assign(pop->loc_risk, pop->loc_risk + L);
    S.loc_hospital.assign(pop->loc_hospital, pop->loc_hospital + L);
    S.nbr_off.assign(pop->nbr_off, pop->nbr_off + L + 1); S.nbr.assign(pop->nbr, pop->nbr + pop->nbr_off[L]);
    S.roster_off.assign(pop->roster_off, pop->roster_off + L + 1); S.roster.assign(pop->roster, pop->roster + pop->roster_off[L]);
    S.home.assign(pop->home, pop->home + N); S.day_loc.assign(pop->day, pop->day + N);
    S.age.assign(pop->age, pop->age + N); S.sex.assign(pop->sex, pop->sex + N);
    S.comorbid.assign(pop->comorbid, pop->comorbid + N); S.ltc.assign(pop->ltc, pop->ltc + N);
    S.patron.assign(pop->patron, pop->patron + (size_t) N * pop->n_patron);
    S.nat_eff.assign(pop->starting_natural_efficacy, pop->starting_natural_efficacy + N);
    S.xprot.assign(pop->cross_protection_probability, pop->cross_protection_probability + N);
    S.nat_dur.assign(pop->natural_immunity_duration, pop->natural_immunity_duration + N);
    S.people.assign(N, cabm_oracle::Person());
    S.patients.clear(); S.log.clear(); S.seed_calls = 0;
    S.own_lo = 0; S.own_hi = N; S.shard_rank = 0; S.shard_world = 1;
    S.census.assign(L, 0u); S.pat_total.assign(L, 0); S.hosp_index.assign(L, -1); S.hosp_ids.clear();
    for (uint32_t l = 0; l < L; ++l) if (S.loc_type[l] == CABM_HOSPITAL) { S.hosp_index[l] = (int32_t) S.hosp_ids.size(); S.hosp_ids.push_back((int32_t) l); }
    S.pat_delta.assign(S.hosp_ids.size(), 0);
    S.have_pop = true;
    return CABM_OK;
}

int cabm_oracle_set_params(cabm_t* h, const cabm_params* par) {
    H_OK(h); if (!par || !par->prob_first_detection || par->run_length <= 0) return CABM_E_ARG;
    Sim& S = h->sim;
    const bool first = !S.have_par;
    S.par = *par;
    const size_t R = (size_t) par->run_length;
    S.pfd.assign(par->prob_first_detection, par->prob_first_detection + R * CABM_N_OUTCOMES);
    if (par->icu_mortality_reduction) S.imr.assign(par->icu_mortality_reduction, par->icu_mortality_reduction + R); else S.imr.clear();
    S.par.prob_first_detection = nullptr; S.par.icu_mortality_reduction = nullptr;
    if (first) {
        S.hot.assign(R, {});
        for (int c = 0; c < CABM_N_COUNTERS; ++c) S.ledger[c].assign(R, 0);
        for (int t = 0; t < 3; ++t) S.ti[t].assign(R, 0.0);
    }
    S.have_par = true;
    return CABM_OK;
}

int cabm_oracle_set_timed_intervention(cabm_t* h, int which, const double* v, size_t n) {
    H_OK(h); Sim& S = h->sim;
    if (!S.have_par || which < 0 || which > 2 || !v || n == 0) return CABM_E_ARG;
    const size_t R = (size_t) S.par.run_length;
    for (size_t d = 0; d < R; ++d) S.ti[which][d] = v[std::min(d, n - 1)];
    return CABM_OK;
}

static int ready(Sim& S, int day) {
    if (!S.have_par || !S.have_pop) { S.err = "population and parameters must be set first"; return CABM_E_STATE; }
    if (day < 0 || day >= S.par.run_length) { S.err = "day outside run_length"; return CABM_E_ARG; }
    return CABM_OK;
}

int cabm_oracle_tick(cabm_t* h, int sim_day, int julian_day, int is_weekday) {
    H_OK(h); int rc = ready(h->sim, sim_day); if (rc) return rc;
    if (julian_day < 1 || julian_day > 366) return CABM_E_ARG;
    h->sim.tick(sim_day, julian_day, is_weekday); return CABM_OK;
}

int cabm_oracle_import(cabm_t* h, int sim_day, double expected, const double strain_w[CABM_N_STRAINS], int32_t* out_infected) {
    H_OK(h); int rc = ready(h->sim, sim_day); if (rc) return rc;
    const int n = h->sim.import_day(sim_day, expected, strain_w);
    if (out_infected) *out_infected = n;
    return CABM_OK;
}

int cabm_oracle_infect(cabm_t* h, uint32_t pid, int strain, int sim_day, int32_t* out_infected) {
    H_OK(h); Sim& S = h->sim; int rc = ready(S, sim_day); if (rc) return rc;
    if (pid >= S.n_people || strain < 0 || strain >= CABM_N_STRAINS) return CABM_E_ARG;
    const uint32_t seq = S.seed_calls++;                              // every instance counts every call, the owner applies it
    if (!S.owns(pid)) { if (out_infected) *out_infected = 0; return CABM_OK; }
    cabm_oracle::Stream s; s.begin(S.par.seed, (uint32_t) sim_day, CABM_P_SEED, pid, seq);
    const bool ok = S.infect((int) pid, sim_day, -1, strain, false, s, CABM_P_SEED, pid, seq, 0);
    if (out_infected) *out_infected = ok ? 1 : 0;
    return CABM_OK;
}

int cabm_oracle_infect_batch(cabm_t* h, const uint32_t* pids, const uint8_t* strains, size_t n, int sim_day, uint8_t* out_infected) {
    H_OK(h); if ((!pids || !strains) && n) return CABM_E_ARG;
    for (size_t k = 0; k < n; ++k) if (pids[k] >= h->sim.n_people || strains[k] >= CABM_N_STRAINS) return CABM_E_ARG;
    for (size_t k = 0; k < n; ++k) {                                  // by definition: the calls one after the other
        int32_t took = 0;
        const int rc = cabm_oracle_infect(h, pids[k], (int) strains[k], sim_day, &took); if (rc) return rc;
        if (out_infected) out_infected[k] = took ? 1 : 0;
    }
    return CABM_OK;
}

int cabm_oracle_run_days(cabm_t* h, int first_day, int n_days, const int32_t* julian_day, const int32_t* is_weekday,
                         const double* expected_imports, const double* strain_w) {
    H_OK(h);
    for (int i = 0; i < n_days; ++i) {
        int rc = cabm_oracle_tick(h, first_day + i, julian_day[i], is_weekday[i]); if (rc) return rc;
        rc = cabm_oracle_import(h, first_day + i, expected_imports[i], strain_w + (size_t) i * CABM_N_STRAINS, nullptr); if (rc) return rc;
    }
    return CABM_OK;
}

// Person::vaccinate (Person.cpp:611-647)
// the batch form of run_days (include/cabm.h): on the CPU the communities simply run one after the other
int cabm_oracle_batch_run_days(cabm_t* const* hs, size_t n, int first_day, int n_days, const int32_t* julian_day, const int32_t* is_weekday,
                               const double* expected_imports, const double* strain_w) {
    if (!hs || n == 0) return CABM_E_ARG;
    for (size_t r = 0; r < n; ++r) {
        if (!hs[r] || hs[r]->sim.n_people != hs[0]->sim.n_people) return CABM_E_ARG;
        const int rc = cabm_oracle_run_days(hs[r], first_day, n_days, julian_day, is_weekday, expected_imports, strain_w);
        if (rc) return rc;
    }
    return CABM_OK;
}

int cabm_oracle_vaccinate(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint8_t* out_ok) {
    H_OK(h); Sim& S = h->sim; int rc = ready(S, sim_day); if (rc) return rc;
    if (!S.par.immunity_leaky) { S.err = "vaccination requires leaky immunity (Person.cpp:622)"; return CABM_E_STATE; }
    for (size_t i = 0; i < n; ++i) {
        if (pids[i] >= S.n_people) return CABM_E_ARG;
        cabm_oracle::Person& p = S.people[pids[i]];
        const bool alive = p.hist.empty() || !p.hist.back().isDead(sim_day);
        if (alive) { p.naive_vax = p.hist.empty() && p.vax_days.empty(); p.vax_days.push_back(sim_day); }
        if (out_ok) out_ok[i] = alive ? 1 : 0;
    }
    return CABM_OK;
}

// Person::scheduleQuarantine (Person.cpp:649-651)
int cabm_oracle_quarantine(cabm_t* h, const uint32_t* pids, size_t n, int from_day, int until_day) {
    H_OK(h); Sim& S = h->sim;
    for (size_t i = 0; i < n; ++i) {
        if (pids[i] >= S.n_people) return CABM_E_ARG;
        cabm_oracle::Person& p = S.people[pids[i]]; p.has_quar = true; p.quar_from = from_day; p.quar_until = until_day;
    }
    return CABM_OK;
}

int cabm_oracle_ledger_series(cabm_t* h, int counter, uint64_t* out, size_t n) {
    H_OK(h); Sim& S = h->sim; if (counter < 0 || counter >= CABM_N_COUNTERS || !out) return CABM_E_ARG;
    for (size_t d = 0; d < n; ++d) out[d] = d < S.ledger[counter].size() ? S.ledger[counter][d] : 0;
    return CABM_OK;
}
int cabm_oracle_ledger_table(cabm_t* h, uint64_t* out, size_t n) {
    H_OK(h); Sim& S = h->sim; if (!out) return CABM_E_ARG;
    const size_t R = (size_t) S.par.run_length;
    for (size_t i = 0; i < n; ++i) { const size_t c = i / R, d = i % R; out[i] = c < CABM_N_COUNTERS && d < S.ledger[c].size() ? S.ledger[c][d] : 0; }
    return CABM_OK;
}
int cabm_oracle_ledger_row(cabm_t* h, int sim_day, uint64_t out[CABM_N_COUNTERS]) {
    H_OK(h); Sim& S = h->sim; int rc = ready(S, sim_day); if (rc) return rc;
    for (int c = 0; c < CABM_N_COUNTERS; ++c) out[c] = S.ledger[c][sim_day];
    return CABM_OK;
}
int cabm_oracle_cumulative_outcomes(cabm_t* h, uint64_t out[CABM_N_OUTCOMES]) { H_OK(h); std::memcpy(out, h->sim.cumul, sizeof(h->sim.cumul)); return CABM_OK; }
int cabm_oracle_work_counters(cabm_t* h, uint64_t out[CABM_N_WORK]) { H_OK(h); std::memcpy(out, h->sim.work, sizeof(h->sim.work)); return CABM_OK; }

int cabm_oracle_sync_histories(cabm_t* h, cabm_infection* out, size_t cap, size_t* out_n) {
    H_OK(h); Sim& S = h->sim;
    if (out_n) *out_n = S.log.size();
    if (out) std::memcpy(out, S.log.data(), std::min(cap, S.log.size()) * sizeof(cabm_infection));
    return CABM_OK;
}

int cabm_oracle_sync_histories_from(cabm_t* h, size_t first, cabm_infection* out, size_t cap, size_t* out_n) {
    H_OK(h); Sim& S = h->sim;
    const size_t avail = first < S.log.size() ? S.log.size() - first : 0;
    if (out_n) *out_n = avail;
    if (out && avail) std::memcpy(out, S.log.data() + first, std::min(cap, avail) * sizeof(cabm_infection));
    return CABM_OK;
}

// Community::doSerosurvey, NATURAL branch, for one person (Community.cpp:1232-1252)
static bool oracle_sero_positive(const Sim& S, uint32_t pid, int day) {
    const cabm_oracle::Person& p = S.people[pid];
    if (p.hist.empty()) return false;
    const cabm_oracle::Infection& last = p.hist.back();
    const int last_end = std::max(last.t[cabm_oracle::T_INFECTIOUS_END], last.t[cabm_oracle::T_SYMPTOM_END]);
    const int tsl = day - last_end;
    bool converted = false;
    for (int i = (int) p.hist.size() - 1; i >= 0; --i) if ((double) p.hist[i].t[cabm_oracle::T_INFECTED] >= S.par.seroconversion_lag) { converted = true; break; }
    if (!converted) return false;
    if (S.par.immunity_leaky) return S.remaining_efficacy(S.nat_eff[pid], (double) tsl) > S.par.seropositivity_threshold;
    return tsl < S.nat_dur[pid];
}

// primary cases of traceForwardContacts (Community.cpp:1062-1070) before the coverage draw
int cabm_oracle_detected_today(cabm_t* h, int sim_day, uint32_t* out_pids, size_t cap, size_t* out_n) {
    H_OK(h); Sim& S = h->sim; if (!out_n) return CABM_E_ARG;
    size_t n = 0;
    for (uint32_t pid = 0; pid < S.n_people; ++pid) {
        const cabm_oracle::Person& p = S.people[pid];
        if (p.hist.empty()) continue;
        const cabm_oracle::Infection& inf = p.hist.back();
        if (inf.isDead(sim_day)) continue;
        if (inf.detected_state < CABM_N_OUTCOMES && inf.report_day == sim_day) { if (out_pids && n < cap) out_pids[n] = pid; ++n; }
    }
    *out_n = n;
    return CABM_OK;
}

// calculate_vax_stats (Community.cpp:1329-1423) and the per-day population counts of simulator.h:402-410,438
int cabm_oracle_diagnostics(cabm_t* h, int sim_day, uint64_t out[CABM_N_DIAG]) {
    H_OK(h); Sim& S = h->sim; if (!out) return CABM_E_ARG;
    for (int k = 0; k < CABM_N_DIAG; ++k) out[k] = 0;
    const int day = sim_day;
    for (uint32_t pid = 0; pid < S.n_people; ++pid) {
        const cabm_oracle::Person& p = S.people[pid];
        switch (p.vax_days.size()) {
            case 3: out[CABM_D_DOSE3]++; [[fallthrough]];
            case 2: out[CABM_D_DOSE2]++; [[fallthrough]];
            case 1: out[CABM_D_DOSE1]++; break;
            default: break;
        }
        const bool full = p.vax_days.size() > 1 && (day - p.vax_days.back()) >= S.par.vaccine_dose_to_protection_lag;
        if (full) out[CABM_D_FULLY_VAXD]++;
        if (!p.hist.empty()) {
            const cabm_oracle::Infection& inf = p.hist.back();
            const int o = full ? CABM_D_BT_INF : CABM_D_UV_INF;
            if (inf.t[cabm_oracle::T_INFECTED] == day) out[o]++;
            if (inf.t[cabm_oracle::T_SYMPTOM_BEGIN] == day) out[o + 1]++;
            if (inf.t[cabm_oracle::T_HOSP_BEGIN] == day) out[o + 2]++;
            if (inf.t[cabm_oracle::T_DEATH] == day) out[o + 3]++;
            if (inf.detected_state < CABM_N_OUTCOMES && inf.report_day == day) out[o + 4]++;
            if (inf.isInfected(day)) out[CABM_D_INFECTED]++;
            if (inf.isInfectious(day)) out[CABM_D_INFECTIOUS]++;
            if (inf.t[cabm_oracle::T_SYMPTOM_BEGIN] <= day && day < inf.t[cabm_oracle::T_SYMPTOM_END]) out[CABM_D_SYMPTOMATIC]++;
        } else if (p.vax_days.empty()) out[CABM_D_NAIVE]++;
        if (S.quarantining(p, day)) out[CABM_D_QUARANTINING]++;
        const bool ped = S.age[pid] <= 17;
        if (ped) out[CABM_D_N_PED]++;
        if (oracle_sero_positive(S, pid, day)) { out[CABM_D_SEROPOS_ALL]++; if (ped) out[CABM_D_SEROPOS_PED]++; }
    }
    return CABM_OK;
}

int cabm_oracle_serosurvey(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint64_t* out_seropos) {
    H_OK(h); Sim& S = h->sim; if (!out_seropos) return CABM_E_ARG;
    if (!pids) n = S.n_people;
    uint64_t c = 0;
    for (size_t i = 0; i < n; ++i) {
        const uint32_t pid = pids ? pids[i] : (uint32_t) i; if (pid >= S.n_people) return CABM_E_ARG;
        c += oracle_sero_positive(S, pid, sim_day) ? 1 : 0;
    }
    *out_seropos = c;
    return CABM_OK;
}

// per-person predicates behind cabm_eligibility / cabm_reduce (include/cabm.h): bits 0-7 CABM_EL_*, bit 8 + k = CABM_R_k
static unsigned oracle_person_bits(const Sim& S, uint32_t pid, int day) {
    const cabm_oracle::Person& p = S.people[pid];
    unsigned b = 0; bool alive = true;
    const unsigned nd = (unsigned) p.vax_days.size();
    if (!p.hist.empty()) {
        const cabm_oracle::Infection& inf = p.hist.back();
        alive = !inf.isDead(day);
        b |= CABM_EL_EVER_INFECTED;
        if (inf.detected_state < CABM_N_OUTCOMES && inf.report_day <= day) b |= CABM_EL_PRIOR_CASE;
        if (inf.isInfected(day)) b |= CABM_EL_INFECTED_NOW | (1u << (8 + CABM_R_INFECTED));
        if (inf.isInfectious(day)) b |= 1u << (8 + CABM_R_INFECTIOUS);
        if (inf.t[cabm_oracle::T_SYMPTOM_BEGIN] <= day && day < inf.t[cabm_oracle::T_SYMPTOM_END]) b |= 1u << (8 + CABM_R_SYMPTOMATIC);
        if (oracle_sero_positive(S, pid, day)) b |= 1u << (8 + CABM_R_SEROPOSITIVE);
    } else if (nd == 0) b |= 1u << (8 + CABM_R_NAIVE);
    if (alive) b |= CABM_EL_ALIVE | (1u << (8 + CABM_R_ALIVE));
    if (nd) b |= CABM_EL_VACCINATED;
    b |= std::min(nd, 7u) << CABM_EL_DOSES_SHIFT;
    if (nd > 1 && (day - p.vax_days.back()) >= S.par.vaccine_dose_to_protection_lag) b |= 1u << (8 + CABM_R_FULLY_VAXD);
    if (S.quarantining(p, day)) b |= 1u << (8 + CABM_R_QUARANTINING);
    return b;
}
int cabm_oracle_eligibility(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint8_t* out_flags) {
    H_OK(h); Sim& S = h->sim; if (!out_flags) return CABM_E_ARG;
    if (!pids) n = S.n_people;
    for (size_t i = 0; i < n; ++i) {
        const uint32_t pid = pids ? pids[i] : (uint32_t) i; if (pid >= S.n_people) return CABM_E_ARG;
        out_flags[i] = (uint8_t) (oracle_person_bits(S, pid, sim_day) & 0xFFu);
    }
    return CABM_OK;
}
int cabm_oracle_reduce(cabm_t* h, int what, int sim_day, const uint32_t* pids, size_t n, double* out_count) {
    H_OK(h); Sim& S = h->sim; if (!out_count || what < 0 || what >= CABM_N_REDUCE) return CABM_E_ARG;
    if (!pids) n = S.n_people;
    uint64_t c = 0;
    for (size_t i = 0; i < n; ++i) {
        const uint32_t pid = pids ? pids[i] : (uint32_t) i; if (pid >= S.n_people) return CABM_E_ARG;
        c += (oracle_person_bits(S, pid, sim_day) >> (8 + what)) & 1u;
    }
    *out_count = (double) c;
    return CABM_OK;
}

int cabm_oracle_get_infections(cabm_t* h, const uint32_t* pids, size_t n, cabm_infection* out) {
    H_OK(h); Sim& S = h->sim;
    for (size_t i = 0; i < n; ++i) {
        const uint32_t pid = pids ? pids[i] : (uint32_t) i; if (pid >= S.n_people) return CABM_E_ARG;
        cabm_infection rec{}; rec.person = pid; rec.infected_by = -1; rec.infected_place = -1;
        const cabm_oracle::Person& p = S.people[pid];
        cabm_oracle::Infection inf; if (!p.hist.empty()) inf = p.hist.back();
        for (int k = 0; k < 12; ++k) rec.t[k] = (int16_t) std::max(-32768, std::min(32767, inf.t[k]));
        rec.report_day = (int16_t) std::min(32767, inf.report_day);
        rec.strain = (uint8_t) inf.strain; rec.detected_state = (uint8_t) inf.detected_state;
        rec.rel_infectiousness = p.hist.empty() ? 0.0f : inf.rel_inf; rec.infected_place = inf.place; rec.infected_by = inf.by;
        out[i] = rec;
    }
    return CABM_OK;
}

int cabm_oracle_snapshot(cabm_t* h, cabm_snap_t** out) { H_OK(h); *out = new cabm_snap{ h->sim }; return CABM_OK; }
int cabm_oracle_restore(cabm_t* h, const cabm_snap_t* s) { H_OK(h); if (!s) return CABM_E_ARG; h->sim = s->sim; return CABM_OK; }
int cabm_oracle_snap_free(cabm_snap_t* s) { delete s; return CABM_OK; }

int cabm_oracle_dump_foi(cabm_t* h, cabm_foi_row* out, size_t cap, size_t* out_n) {
    H_OK(h); Sim& S = h->sim;
    if (out_n) *out_n = S.foi.size();
    if (out) std::memcpy(out, S.foi.data(), std::min(cap, S.foi.size()) * sizeof(cabm_foi_row));
    return CABM_OK;
}
int cabm_oracle_device_time_ms(cabm_t*, double* out_ms, uint64_t* out_launches) { if (out_ms) *out_ms = 0; if (out_launches) *out_launches = 0; return CABM_OK; }


// ---- known-answer hooks (tests/test_reference_kat.py): scripted person state, one Person::infect call ----
// hist: n_hist x {infectedBegin, infectiousEnd, symptomEnd (INT_MIN = none), strain}
int cabm_oracle_kat_set_person(cabm_t* h, uint32_t pid, int n_hist, const int32_t* hist, int n_dose, const int32_t* doses, int naive) {
    H_OK(h); Sim& S = h->sim; if (pid >= S.n_people) return CABM_E_ARG;
    cabm_oracle::Person& p = S.people[pid];
    p = cabm_oracle::Person();
    for (int k = 0; k < n_hist; ++k) {
        cabm_oracle::Infection inf;
        inf.t[cabm_oracle::T_INFECTED] = hist[4 * k]; inf.t[cabm_oracle::T_INFECTIOUS_BEGIN] = hist[4 * k] + 1; inf.t[cabm_oracle::T_INFECTIOUS_END] = hist[4 * k + 1];
        if (hist[4 * k + 2] != INT_MIN) { inf.t[cabm_oracle::T_SYMPTOM_BEGIN] = hist[4 * k] + 1; inf.t[cabm_oracle::T_SYMPTOM_END] = hist[4 * k + 2]; }
        inf.strain = hist[4 * k + 3];
        p.hist.push_back(inf);
    }
    for (int k = 0; k < n_dose; ++k) p.vax_days.push_back(doses[k]);
    p.naive_vax = naive != 0;
    return CABM_OK;
}
// out: ok, t[12], strain, detected_state, report_day, draws_rng, draws_reporting, cumul delta[5], n_flags, then {day, loc} pairs (cap pairs)
int cabm_oracle_kat_infect(cabm_t* h, uint32_t pid, int strain, int day, int has_source, uint32_t key_a, uint32_t key_b,
                           int64_t* out, float* out_rel_inf, int32_t* out_flags, size_t cap_flags) {
    H_OK(h); Sim& S = h->sim; int rc = ready(S, day); if (rc) return rc;
    if (pid >= S.n_people || !out) return CABM_E_ARG;
    uint64_t cum0[CABM_N_OUTCOMES]; for (int k = 0; k < CABM_N_OUTCOMES; ++k) cum0[k] = S.cumul[k];
    for (auto& v : S.hot) v.clear();
    cabm_oracle::Stream s; s.begin(S.par.seed, (uint32_t) day, CABM_P_SEED, key_a, key_b);
    S.pending_infector = -1;
    const bool ok = S.infect((int) pid, day, has_source ? S.home[pid] : -1, strain, has_source != 0, s, CABM_P_SEED, key_a, key_b, 0);
    cabm_oracle::Infection blank; const cabm_oracle::Infection& I = ok ? S.people[pid].hist.back() : blank;
    size_t o = 0; out[o++] = ok ? 1 : 0;
    for (int k = 0; k < cabm_oracle::T_N; ++k) out[o++] = I.t[k];
    out[o++] = ok ? I.strain : 0; out[o++] = I.detected_state; out[o++] = I.report_day; out[o++] = s.s.idx; out[o++] = S.last_rep_draws;
    for (int k = 0; k < CABM_N_OUTCOMES; ++k) out[o++] = (int64_t) (S.cumul[k] - cum0[k]);
    size_t nf = 0;
    for (size_t d = 0; d < S.hot.size(); ++d) for (const cabm_oracle::HotRec& r : S.hot[d]) { if (nf < cap_flags) { out_flags[2 * nf] = (int32_t) d; out_flags[2 * nf + 1] = r.loc; } nf++; }
    out[o++] = (int64_t) nf;
    if (out_rel_inf) *out_rel_inf = ok ? I.rel_inf : 1.0f;
    return CABM_OK;
}
// FNV-style digest + sum of one ledger series, same arithmetic as ref_driver kat's last line
int cabm_oracle_kat_ledger_digest(cabm_t* h, int counter, uint64_t* out_sum, uint64_t* out_hash) {
    H_OK(h); Sim& S = h->sim; if (counter < 0 || counter >= CABM_N_COUNTERS) return CABM_E_ARG;
    uint64_t s = 0, hh = 1469598103934665603ull;
    for (size_t d = 0; d < S.ledger[counter].size(); ++d) { s += S.ledger[counter][d]; hh = (hh ^ (S.ledger[counter][d] + 0x9E37u * d)) * 1099511628211ull; }
    *out_sum = s; *out_hash = hh;
    return CABM_OK;
}

// ---- one community over several instances (include/cabm.h "shard" entry points) ----
int cabm_oracle_shard_config(cabm_t* h, int rank, int world, uint32_t pid_lo, uint32_t pid_hi) {
    H_OK(h); Sim& S = h->sim;
    if (!S.have_pop || world < 1 || rank < 0 || rank >= world || pid_lo > pid_hi || pid_hi > S.n_people) return CABM_E_ARG;
    for (uint32_t p = pid_lo; p < pid_hi; ++p) {                      // household members share an owner
        const int32_t hm = S.home[p];
        for (int64_t e = S.roster_off[hm]; e < S.roster_off[hm + 1]; ++e)
            if (S.home[S.roster[e]] == hm && ((uint32_t) S.roster[e] < pid_lo || (uint32_t) S.roster[e] >= pid_hi)) { S.err = "a household is split between shards"; return CABM_E_ARG; }
    }
    S.shard_rank = rank; S.shard_world = world; S.own_lo = pid_lo; S.own_hi = pid_hi;
    return CABM_OK;
}
int cabm_oracle_shard_day_part(cabm_t* h, int part, int sim_day, int julian_day, int is_weekday, double expected, const double* strain_w, int32_t* out_imported) {
    H_OK(h); Sim& S = h->sim; int rc = ready(S, sim_day); if (rc) return rc;
    if (part < 0 || part > 3 || julian_day < 1 || julian_day > 366) return CABM_E_ARG;
    S.day_part(part, sim_day, julian_day, is_weekday);
    if (part == 3) { const int n = strain_w ? S.import_day(sim_day, expected, strain_w) : 0; if (out_imported) *out_imported = n; }
    return CABM_OK;
}
int cabm_oracle_shard_flags_out(cabm_t* h, void** records, size_t* n) {
    H_OK(h); if (!records || !n) return CABM_E_ARG;
    *records = h->sim.flags_out.data(); *n = h->sim.flags_out.size(); return CABM_OK;
}
int cabm_oracle_shard_flags_in(cabm_t* h, const void* records, size_t n) {
    H_OK(h); if (!records && n) return CABM_E_ARG;
    const cabm_oracle::HotRec* r = (const cabm_oracle::HotRec*) records;
    h->sim.flags_in.insert(h->sim.flags_in.end(), r, r + n); return CABM_OK;
}
int cabm_oracle_shard_census(cabm_t* h, void** buf, size_t* n_u32) {
    H_OK(h); if (!buf || !n_u32) return CABM_E_ARG;
    *buf = h->sim.census.data(); *n_u32 = h->sim.census.size(); return CABM_OK;
}
int cabm_oracle_shard_patients(cabm_t* h, void** buf, size_t* n_i32) {
    H_OK(h); if (!buf || !n_i32) return CABM_E_ARG;
    *buf = h->sim.pat_delta.data(); *n_i32 = h->sim.pat_delta.size(); return CABM_OK;
}

// ---- generator / deterministic-math hooks for tests/test_philox.py and tests/test_det_math.py -------------------------
void cabm_oracle_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) { philox4x32_10_ref(ctr, key, out); }

// the first n words of the event stream keyed (seed; day, purpose, a, b) -- the convention of philox_ref.h
void cabm_oracle_stream_words(uint64_t seed, uint32_t day, uint32_t purpose, uint32_t a, uint32_t b, uint32_t* out, size_t n) {
    philox_stream_ref s; philox_stream_begin(&s, seed, day, purpose, a, b);
    for (size_t i = 0; i < n; ++i) out[i] = philox_stream_next(&s);
}

// kind 0 = cabm_exp, 1 = cabm_log, 2 = cabm_logistic (include/cabm_det_math.h)
void cabm_oracle_det_math(int kind, const double* x, double* y, size_t n) {
    for (size_t i = 0; i < n; ++i) y[i] = kind == 0 ? cabm_exp(x[i]) : kind == 1 ? cabm_log(x[i]) : cabm_logistic(x[i]);
}

} // extern "C"

