// oracle/cabm_oracle.h -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
//
// Sequential CPU restatement of the reference's daily transmission cycle with every random draw
// re-keyed on Philox(seed; day, purpose, person, location) -- the "Philox-instrumented" semantics
// the north star asks parity against. It keeps the reference's own shape on purpose: an event
// calendar of hot locations filled at infection time (Community.h:133, Community.cpp:788-791),
// location passes executed one after another in tick() order (Community.cpp:1147-1192), an
// at-risk loop that calls infect() directly so "first successful exposure wins" falls out of the
// control flow (Community.cpp:971-996). The CUDA path derives the same quantities in parallel
// (no calendar, atomicMin winner resolution); tests/ compare the two bit-for-bit.
//
// Pinned against the unmodified reference compiled and run in the build container (oracle/_ref): Person::infect bit-exact
// on 1500 scripted states (tests/test_reference_kat.py), every per-location exposure probability of Community::tick on a
// scripted 45-day epidemic (tests/test_foi_reference.py, 13 368 values, worst relative difference 3.4e-11), epidemic curves
// by KS test over 400 replicates (tools/ks_equivalence.py).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may link or load this.
//
// Exported C symbols mirror include/cabm.h with the prefix cabm_oracle_.
#ifndef CABM_ORACLE_H
#define CABM_ORACLE_H
#include <climits>
#include <cstdint>
#include <map>
#include <string>
#include <vector>
#include "../include/cabm.h"
#include "philox_ref.h"

namespace cabm_oracle {

enum { T_INFECTED = 0, T_INFECTIOUS_BEGIN, T_INFECTIOUS_END, T_SYMPTOM_BEGIN, T_SYMPTOM_END, T_SEVERE_BEGIN,
       T_SEVERE_END, T_HOSP_BEGIN, T_CRITICAL_BEGIN, T_CRITICAL_END, T_ICU_BEGIN, T_DEATH, T_N };

// Person.h:35-175
struct Infection {
    int t[T_N];
    int strain = 0;
    float rel_inf = 1.0f;
    int detected_state = CABM_N_OUTCOMES;
    int report_day = INT_MAX;
    int by = -1, place = -1;
    Infection() {
        for (int i = 0; i < T_N; ++i) t[i] = INT_MAX;
        t[T_INFECTIOUS_END] = t[T_SYMPTOM_END] = t[T_SEVERE_END] = t[T_CRITICAL_END] = INT_MIN;   // Person.h:50-53
    }
    bool isInfected(int now)   const { return t[T_INFECTED] <= now && now < t[T_INFECTIOUS_END]; }        // Person.h:149
    bool isInfectious(int now) const { return t[T_INFECTIOUS_BEGIN] <= now && now < t[T_INFECTIOUS_END]; } // :150
    bool isSevere(int now)     const { return t[T_SEVERE_BEGIN] <= now && now < t[T_SEVERE_END]; }        // :152
    bool isCritical(int now)   const { return t[T_CRITICAL_BEGIN] <= now && now < t[T_CRITICAL_END]; }    // :153
    bool inHospital(int now)   const { return t[T_HOSP_BEGIN] <= now && now < t[T_SEVERE_END]; }          // :154
    bool inIcu(int now)        const { return t[T_ICU_BEGIN] <= now && now < t[T_CRITICAL_END]; }         // :155
    bool isDead(int now)       const { return t[T_DEATH] <= now; }                                        // :156
};

// Person.h:353-381
struct Person {
    std::vector<Infection> hist;
    std::vector<int> vax_days;
    bool naive_vax = false;
    bool has_quar = false; int quar_from = 0, quar_until = 0;
    int hosp_count = 0;          // occurrences in the associated hospital's person list (Person.h:248-249)
};

// one entry of the hot-location calendar: Community::flagInfectedLocation (Community.cpp:788-791)
struct HotRec { int32_t loc; int32_t pid; uint64_t c_fx; uint8_t strain; };

struct LocHot { uint64_t w_fx[CABM_N_STRAINS] = {0, 0, 0, 0}; std::vector<HotRec> recs;
                uint64_t total() const { return w_fx[0] + w_fx[1] + w_fx[2] + w_fx[3]; } };

struct Stream {
    philox_stream_ref s;
    uint32_t sub;
    void begin(uint64_t seed, uint32_t day, uint32_t purpose, uint32_t a, uint32_t b, uint32_t sub_ = 0);
    uint32_t next();
    double u01() { return philox_u01(next()); }
};

struct Sim {
    cabm_params par{};
    bool have_par = false, have_pop = false;
    std::vector<double> pfd, imr, ti[3];
    // population
    uint32_t n_people = 0, n_loc = 0, n_patron = 0;
    std::vector<uint8_t> loc_type, loc_essential, loc_pubrisk; std::vector<float> loc_risk; std::vector<int32_t> loc_hospital;
    std::vector<int64_t> nbr_off, roster_off; std::vector<int32_t> nbr, roster;
    std::vector<int32_t> home, day_loc, patron; std::vector<uint8_t> age, sex, comorbid, ltc;
    std::vector<double> nat_eff; std::vector<float> xprot; std::vector<int32_t> nat_dur;
    // state
    std::vector<Person> people;
    std::vector<std::vector<HotRec>> hot;                 // [day] calendar
    std::map<int32_t, std::vector<int32_t>> patients;     // hospital -> admitted person list
    std::vector<uint64_t> ledger[CABM_N_COUNTERS];
    uint64_t cumul[CABM_N_OUTCOMES] = {0, 0, 0, 0, 0};
    uint64_t work[CABM_N_WORK] = {0};
    std::vector<cabm_foi_row> foi;
    std::vector<cabm_infection> log;                      // every infection, in creation order
    uint32_t seed_calls = 0;
    uint32_t last_rep_draws = 0;                          // REPORTING_RNG draws of the last infect() (known-answer tests)
    int pending_infector = -1;                            // infector traced for the exposure being applied (Community.cpp:795-825)
    std::string err;

    // one community over several instances: this one owns agents [own_lo, own_hi)
    uint32_t own_lo = 0, own_hi = 0; int shard_rank = 0, shard_world = 1;
    bool owns(uint32_t pid) const { return pid >= own_lo && pid < own_hi; }
    std::vector<HotRec> flags_out, flags_in;             // today's flag records: produced here / received from the other instances
    std::vector<uint32_t> census;                        // [n_loc] visitors of the day (summed over instances between part 1 and 2)
    std::vector<int32_t> hosp_ids, hosp_index, pat_delta, pat_total;   // hospitals; admissions - discharges of the day; patients of all instances
    std::map<int32_t, LocHot> hotmap[CABM_N_LOCTYPES];    // today's calendar slice by location type, then id (Community.h:133, Location.h:31)
    std::map<int32_t, std::vector<int32_t>> visitors;     // today's visitors among the owned agents
    void day_part(int part, int day, int julian_day, int is_weekday);
    std::map<int32_t, int> draw_visits(int day, uint32_t pid, int num_locs, bool avoiding) const;

    // the hot path
    void tick(int day, int julian_day, int is_weekday);
    int  import_day(int day, double expected, const double* strain_w);
    bool infect(int pid, int day, int source_loc, int strain, bool has_source, Stream& s, uint32_t purpose,
                uint32_t key_a, uint32_t key_b, uint32_t key_sub);
    void transmission(int day, int source, const int32_t* at_risk, size_t n, const LocHot& H, double T, uint32_t purpose);
    void update_person_status(int day);
    double remaining_efficacy(double start, double time_delta) const;
    bool is_cross_protected(const Person& p, int pid, int time, int strain, Stream& s) const;
    bool is_vaccine_protected(const Person& p, int time, int strain, Stream& s) const;
    bool quarantining(const Person& p, int time) const { return p.has_quar && time >= p.quar_from && time < p.quar_until; }
    static unsigned table_draw(const double* cdf, int n, double u) { int k = 0; while (k < n - 1 && !(u < cdf[k])) ++k; return (unsigned) k; }
};

} // namespace cabm_oracle
#endif
