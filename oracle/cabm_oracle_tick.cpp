// oracle/cabm_oracle_tick.cpp -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
// Community-level half of the restatement: Community::tick and what it calls, daily importation.
#include <algorithm>
#include <cmath>
#include <cstring>
#include "cabm_oracle.h"
#include "../include/cabm_det_math.h"

namespace cabm_oracle {

// One 32-bit draw picks a point t in [0, total pressure) of the source location: the strain whose share contains t
// is the strain transmitted, and inside that share the flags are laid out by ascending person id, so t also names
// the infector. Stands in for Community::trace_contact (Community.cpp:795-825): weighted_choice over relInf groups,
// uniform inside the group, i.e. infector ~ relInf-weighted; exact integer arithmetic.
static int pick_strain(const uint64_t w_fx[CABM_N_STRAINS], uint32_t x, uint64_t* t_res) {
    const uint64_t total = w_fx[0] + w_fx[1] + w_fx[2] + w_fx[3];
    uint64_t t = (uint64_t) x * (total >> 32) + (((uint64_t) x * (total & 0xffffffffull)) >> 32);
    int s = 0;
    for (; s < CABM_N_STRAINS - 1; ++s) { if (t < w_fx[s]) break; t -= w_fx[s]; }
    *t_res = t;
    return s;
}
static int pick_infector(const LocHot& H, int strain, uint64_t t) {
    std::vector<std::pair<int32_t, uint64_t>> v;
    for (const HotRec& r : H.recs) if (r.strain == strain) v.push_back({ r.pid, r.c_fx });
    std::sort(v.begin(), v.end());
    uint64_t acc = 0;
    for (auto& e : v) { acc += e.second; if (t < acc) return e.first; }
    return -1;
}

// Community::_transmission (Community.cpp:971-996)
void Sim::transmission(int day, int source, const int32_t* at_risk, size_t n, const LocHot& H, double T, uint32_t purpose) {
    for (size_t i = 0; i < n; ++i) {
        const int pid = at_risk[i];
        Person& p = people[pid];
        uint32_t sub = 0;
        if (purpose == CABM_P_EXPOSE_HOSP_PAT) {             // occurrence index of a repeated patient entry
            for (size_t j = 0; j < i; ++j) if (at_risk[j] == pid) sub++;
        }
        if (!owns((uint32_t) pid)) continue;                                               // another instance exposes this agent
        if (quarantining(p, day) && source != home[pid]) continue;                         // :975
        work[CABM_W_TRIALS]++;
        Stream s; s.begin(par.seed, (uint32_t) day, purpose, (uint32_t) pid, (uint32_t) source, sub);
        if (s.u01() < T) {                                                                 // :976
            uint64_t t_res = 0;
            const int strain = pick_strain(H.w_fx, s.next(), &t_res);                      // :980-982
            (void) s.next();                                                               // reserved (reference: uniform_choice inside the group)
            pending_infector = pick_infector(H, strain, t_res);
            infect(pid, day, source, strain, true, s, purpose, (uint32_t) pid, (uint32_t) source, sub);
        }
    }
}

// Community::updatePersonStatus (Community.cpp:682-741)
void Sim::update_person_status(int day) {
    for (uint32_t pid = 0; pid < n_people; ++pid) {
        Person& p = people[pid];
        if (p.hist.empty()) continue;
        const Infection& inf = p.hist.back();
        const int hosp = loc_hospital[home[pid]];
        if (inf.inHospital(day)) {
            if (inf.t[T_HOSP_BEGIN] == day) {
                if (day_loc[pid] != hosp) { patients[hosp].push_back((int32_t) pid); p.hosp_count++; if (hosp >= 0 && hosp_index[hosp] >= 0) pat_delta[hosp_index[hosp]]++; }     // :689
            }
        } else if (day > 0 && inf.inHospital(day - 1)) {
            if (day_loc[pid] != hosp) {                                                                  // :692, Location.cpp:58-67
                std::vector<int32_t>& v = patients[hosp];
                for (size_t i = 0; i < v.size(); ++i) if (v[i] == (int32_t) pid) { v[i] = v.back(); v.pop_back(); p.hosp_count--; if (hosp >= 0 && hosp_index[hosp] >= 0) pat_delta[hosp_index[hosp]]--; break; }
            }
        }
        if ((int) pid < par.num_surveilled_people) {
            if (inf.t[T_INFECTED] == day) { ledger[CABM_C_NEW_INFECTED][day]++; ledger[CABM_C_NEW_INF_STRAIN0 + inf.strain][day]++; }
            if (inf.t[T_SYMPTOM_BEGIN] == day) { ledger[CABM_C_NEW_SYMPTOMATIC][day]++; if (p.vax_days.size() > 0) ledger[CABM_C_VACCINATED_CASES][day]++; }
            if (inf.isSevere(day)) ledger[CABM_C_SEVERE_PREV][day]++;
            if (inf.inHospital(day)) {
                ledger[CABM_C_HOSP_PREV][day]++;
                if (inf.t[T_HOSP_BEGIN] == day) ledger[CABM_C_HOSP_INC][day]++;
                if (inf.inIcu(day)) { ledger[CABM_C_ICU_PREV][day]++; if (inf.t[T_ICU_BEGIN] == day) ledger[CABM_C_ICU_INC][day]++; }
            }
            if (inf.t[T_SEVERE_BEGIN] == day)   ledger[CABM_C_NEW_SEVERE][day]++;
            if (inf.t[T_CRITICAL_BEGIN] == day) ledger[CABM_C_NEW_CRITICAL][day]++;
            if (inf.t[T_DEATH] == day)          ledger[CABM_C_NEW_DEAD][day]++;
        }
    }
}

static double hazard_to_T(double hazard) { return 1.0 - cabm_exp(-hazard); }

// public_activity draws of one person (Community.cpp:1012-1026): Poisson(1) visits, uniform picks among the patronised
// places, high-risk places skipped by compliant households; visit_map keeps one entry per distinct place (hours)
std::map<int32_t, int> Sim::draw_visits(int day, uint32_t pid, int num_locs, bool avoiding) const {
    const int32_t* locs = &patron[(size_t) pid * n_patron];
    Stream s; s.begin(par.seed, (uint32_t) day, CABM_P_VISIT, pid, 0);
    const unsigned visit_hours = table_draw(par.visit_cdf, CABM_T_POISSON, s.u01());             // Poisson(1) :1013
    std::map<int32_t, int> visit_map;
    for (unsigned i = 0; i < visit_hours; ++i) {
        const uint32_t idx = (uint32_t) (((uint64_t) s.next() * (uint64_t) num_locs) >> 32);
        const int32_t loc = locs[idx];
        if (avoiding && loc_pubrisk[loc] == CABM_PT_HIGH) continue;
        visit_map[loc]++;
    }
    return visit_map;
}

// Community::tick (Community.cpp:1147-1192). The day is written in four parts so that several instances, each owning
// a slice of the agents, can advance one community together; between the parts the caller exchanges
//   after part 0: the flag records (every instance needs every infectious (agent, place) pair of the day),
//   after part 1: the visitor census (summed over instances),
//   after part 2: the day's admissions minus discharges per hospital (summed over instances).
// With one instance owning everybody the four parts run back to back and nothing is exchanged.
// Where a person is and whether they are infectious is taken from the infection they carry at the START of the day
// (the reference evaluates public_activity after the household passes, Community.cpp:1151-1154: a person re-infected in
// those passes while still hospitalised for the previous infection would visit; that corner is not reproduced).
void Sim::tick(int day, int julian_day, int is_weekday) {
    for (int part = 0; part < 4; ++part) day_part(part, day, julian_day, is_weekday);
}

void Sim::day_part(int part, int day, int julian_day, int is_weekday) {
    const double season = par.seasonality[julian_day - 1];                                  // Parameters.cpp:252
    const double SD = ti[CABM_TI_SOCIAL_DISTANCING][day];
    auto flag = [&](const HotRec& r) {
        LocHot& H = hotmap[loc_type[r.loc]][r.loc];
        H.w_fx[r.strain] += r.c_fx; H.recs.push_back(r);
    };
    auto add_row = [&](int32_t loc, const LocHot& H, uint32_t size, double T, double Ts) {
        cabm_foi_row row{}; row.loc = (uint32_t) loc; row.type = loc_type[loc]; row.size = size; row.T = T; row.T_social = Ts;
        for (int s = 0; s < CABM_N_STRAINS; ++s) row.w[s] = (double) H.w_fx[s] * (1.0 / 4294967296.0);
        foi.push_back(row);
    };
    auto wsum = [](const LocHot& H) { return (double) H.total() * (1.0 / 4294967296.0); };

    if (part == 0) {
        // today's slice of the calendar for the agents this instance owns (Community.h:133), plus the places its infectious
        // agents visit today (Community.cpp:1031-1035)
        foi.clear(); flags_out.clear(); flags_in.clear(); visitors.clear();
        for (int t = 0; t < CABM_N_LOCTYPES; ++t) hotmap[t].clear();
        for (const HotRec& r : hot[day]) if (r.loc >= 0 && r.c_fx > 0) flags_out.push_back(r);
        for (uint32_t pid = own_lo; pid < own_hi; ++pid) {
            const Person& p = people[pid];
            if (p.hist.empty() || !p.hist.back().isInfectious(day) || p.hist.back().inHospital(day)) continue;
            const int32_t* locs = &patron[(size_t) pid * n_patron];
            int num_locs = 0; while (num_locs < (int) n_patron && locs[num_locs] >= 0) num_locs++;
            if (!num_locs) continue;
            for (auto& [loc, hours] : draw_visits(day, pid, num_locs, (double) loc_risk[home[pid]] < SD)) {
                const uint64_t c = (uint64_t) llrint((double) p.hist.back().rel_inf * (double) hours * 536870912.0);   // relInf*hours/8 :1033
                if (c > 0) flags_out.push_back(HotRec{ loc, (int32_t) pid, c, (uint8_t) p.hist.back().strain });
            }
        }
        work[CABM_W_FLAGS] += flags_out.size();
        return;
    }
    if (part == 1) {
        for (const HotRec& r : flags_out) flag(r);
        for (const HotRec& r : flags_in) flag(r);
        // public_activity (:1006-1039) for the owned agents: visitor lists stay local, the census is summed by the caller
        std::fill(census.begin(), census.end(), 0u);
        for (uint32_t pid = own_lo; pid < own_hi; ++pid) {
            const int32_t* locs = &patron[(size_t) pid * n_patron];
            int num_locs = 0; while (num_locs < (int) n_patron && locs[num_locs] >= 0) num_locs++;
            const Person& p = people[pid];
            const bool in_hosp = !p.hist.empty() && p.hist.back().inHospital(day);
            if (!num_locs || in_hosp) continue;
            work[CABM_W_PATRONS]++;
            for (auto& [loc, hours] : draw_visits(day, pid, num_locs, (double) loc_risk[home[pid]] < SD)) {
                (void) hours;
                visitors[loc].push_back((int32_t) pid);
                census[loc]++;
                work[CABM_W_VISITS]++;
            }
        }
        return;
    }
    if (part == 2) {
        std::map<int32_t, double> house_T, house_Ts;
        // within_household_transmission (:842-850)
        for (auto& [loc, H] : hotmap[CABM_HOUSE]) {
            const double T = hazard_to_T(par.haz_mult[CABM_HOUSE] * season * wsum(H));
            house_T[loc] = T;
            transmission(day, loc, &roster[roster_off[loc]], (size_t) (roster_off[loc + 1] - roster_off[loc]), H, T, CABM_P_EXPOSE_HH);
        }
        // between_household_transmission (:853-875)
        for (auto& [loc, H] : hotmap[CABM_HOUSE]) {
            const double w = wsum(H);
            double Ts = 0.0;
            if ((double) loc_risk[loc] > SD) {
                const int hh_size = (int) (roster_off[loc + 1] - roster_off[loc]);
                Ts = hazard_to_T(par.haz_mult[CABM_N_LOCTYPES] * season * w / hh_size);
                for (int64_t e = nbr_off[loc]; e < nbr_off[loc + 1]; ++e) {
                    const int32_t nb = nbr[e];
                    if (shard_rank == 0) work[CABM_W_NBR_EDGES]++;
                    if ((double) loc_risk[nb] > SD)
                        transmission(day, loc, &roster[roster_off[nb]], (size_t) (roster_off[nb + 1] - roster_off[nb]), H, Ts, CABM_P_EXPOSE_NB);
                }
            }
            house_Ts[loc] = Ts;
        }
        for (auto& [loc, H] : hotmap[CABM_HOUSE]) add_row(loc, H, (uint32_t) (roster_off[loc + 1] - roster_off[loc]), house_T[loc], house_Ts[loc]);
        // workplace_transmission (:878-911): size = staff + everybody's visitors
        for (auto& [loc, H] : hotmap[CABM_WORK]) {
            const std::vector<int32_t>& vis = visitors[loc];
            const int n_staff = (int) (roster_off[loc + 1] - roster_off[loc]);
            const int size = n_staff + (int) census[loc];
            double T = 0.0;
            work[CABM_W_HOT_VISITS] += vis.size();
            const bool closed = !loc_essential[loc] && ti[CABM_TI_BUSINESS_CLOSURE][day] != 0.0;
            if (!(size < 2 || closed)) {
                const double w = wsum(H);
                if (w > 0) {
                    T = hazard_to_T(par.haz_mult[CABM_WORK] * (loc_pubrisk[loc] == CABM_PT_HIGH ? 4.0 : 0.25) * season * w / (size - 1.0));
                    transmission(day, loc, vis.data(), vis.size(), H, T, CABM_P_EXPOSE_WORK_VIS);
                    transmission(day, loc, &roster[roster_off[loc]], (size_t) n_staff, H, T, CABM_P_EXPOSE_WORK);
                }
            }
            add_row(loc, H, (uint32_t) size, T, 0.0);
        }
        // school_transmission (:914-928) behind the weekday / closure gate (:1158)
        {
            const double closure = ti[CABM_TI_SCHOOL_CLOSURE][day];
            const bool open = is_weekday && closure < 1;
            const double coef = (1.0 - closure) * par.haz_mult[CABM_SCHOOL] * season;
            for (auto& [loc, H] : hotmap[CABM_SCHOOL]) {
                const int size = (int) (roster_off[loc + 1] - roster_off[loc]);
                double T = 0.0;
                if (open && coef != 0.0 && size >= 2) {
                    T = hazard_to_T(coef * wsum(H) / (size - 1.0));
                    transmission(day, loc, &roster[roster_off[loc]], (size_t) size, H, T, CABM_P_EXPOSE_SCHOOL);
                }
                add_row(loc, H, (uint32_t) size, T, 0.0);
            }
        }
        // nursinghome_transmission (:959-969)
        for (auto& [loc, H] : hotmap[CABM_NURSINGHOME]) {
            const int n = (int) (roster_off[loc + 1] - roster_off[loc]);
            double T = 0.0;
            if (n >= 2) {
                T = hazard_to_T(par.haz_mult[CABM_NURSINGHOME] * season * wsum(H) / (n - 1.0));
                transmission(day, loc, &roster[roster_off[loc]], (size_t) n, H, T, CABM_P_EXPOSE_NH);
            }
            add_row(loc, H, (uint32_t) n, T, 0.0);
        }
        std::fill(pat_delta.begin(), pat_delta.end(), 0);
        update_person_status(day);                                                                // :1165
        return;
    }
    // part 3: hospital_transmission (:931-941): census = staff + admitted patients of every instance
    for (size_t k = 0; k < hosp_ids.size(); ++k) pat_total[hosp_ids[k]] += pat_delta[k];
    for (auto& [loc, H] : hotmap[CABM_HOSPITAL]) {
        const std::vector<int32_t>& pat = patients[loc];
        const int n_staff = (int) (roster_off[loc + 1] - roster_off[loc]);
        const int n = n_staff + pat_total[loc];
        double T = 0.0;
        if (n >= 2) {
            T = hazard_to_T(par.haz_mult[CABM_HOSPITAL] * season * wsum(H) / (n - 1.0));
            transmission(day, loc, &roster[roster_off[loc]], (size_t) n_staff, H, T, CABM_P_EXPOSE_HOSP);
            transmission(day, loc, pat.data(), pat.size(), H, T, CABM_P_EXPOSE_HOSP_PAT);
        }
        add_row(loc, H, (uint32_t) n, T, 0.0);
    }
    if (shard_rank == 0) for (int t = 0; t < CABM_N_LOCTYPES; ++t) work[CABM_W_HOT_LOCS] += hotmap[t].size();
    std::vector<HotRec>().swap(hot[day]);                                                         // updateHotLocations :999-1003
    work[CABM_W_AGENTS] += own_hi - own_lo;
    for (uint32_t pid = own_lo; pid < own_hi; ++pid) {
        const Person& p = people[pid];
        if (p.hist.empty() || p.hist.back().t[T_INFECTED] > day) continue;
        const Infection& i = p.hist.back();
        const int until = std::max(std::max(i.t[T_INFECTIOUS_END], i.t[T_SYMPTOM_END]), std::max(i.t[T_SEVERE_END], i.t[T_CRITICAL_END]));
        if (day <= until) work[CABM_W_ACTIVE]++;
    }
}

// Poisson(mean) by products of uniforms over chunks of mean <= 16 (sum of independent Poissons);
// stands in for gsl_ran_poisson(RNG, expected_num_exposed) (simulator.h:209)
static int keyed_poisson(uint64_t seed, int day, double mean) {
    int total = 0; uint32_t chunk = 0;
    while (mean > 0.0) {
        const double m = mean > 16.0 ? 16.0 : mean;
        mean -= m;
        const double limit = cabm_exp(-m);
        Stream s; s.begin(seed, (uint32_t) day, CABM_P_IMPORT_COUNT, chunk++, 0);
        double prod = 1.0; int k = 0;
        for (;;) { prod *= philox_u01(s.next()); if (!(prod > limit)) break; ++k; }
        total += k;
    }
    return total;
}

// seed_epidemic(par, community, date, strain_weights) (simulator.h:201-223)
int Sim::import_day(int day, double expected, const double* strain_w) {
    int introduced = 0;
    if (expected > 0) {
        const int num_exposed = keyed_poisson(par.seed, day, expected);
        for (int i = 0; i < num_exposed; ++i) {
            Stream s; s.begin(par.seed, (uint32_t) day, CABM_P_IMPORT, (uint32_t) i, 0);
            const uint32_t pid = (uint32_t) (((uint64_t) s.next() * (uint64_t) n_people) >> 32);
            // covid::util::weighted_choice (Utility.h:93-105)
            const double sum = ((strain_w[0] + strain_w[1]) + strain_w[2]) + strain_w[3];
            double ran = sum * s.u01(); int strain = CABM_N_STRAINS - 1;
            for (int k = 0; k < CABM_N_STRAINS; ++k) { if (ran < strain_w[k]) { strain = k; break; } else ran -= strain_w[k]; }
            if (!owns(pid)) continue;
            if (infect((int) pid, day, -1, strain, false, s, CABM_P_IMPORT, (uint32_t) i, 0, 0)) introduced++;
        }
    }
    return introduced;
}

} // namespace cabm_oracle
