// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
//
// AbcSmc-free driver for the UNMODIFIED reference core (Community.cpp, Person.cpp, Location.cpp,
// Parameters.cpp, Utility.cpp, Vac_Campaign.cpp + simulator.h), compiled from /root/reference by
// oracle/Makefile against oracle/gsl_shim and placed under oracle/_ref/ (git-ignored).
//
// This translation unit is compiled with -fno-access-control so that it can read and set the
// private state of the reference's classes (no reference source is edited or copied).
//
// Modes
//   run  <params.txt> <pop_dir> <tag> <out.csv>      whole epidemic with the reference's own RNG
//        streams; times ONLY community->tick() + daily seed_epidemic (simulator.h:338,374) and
//        writes one ledger row per day. This is the "reference" CPU baseline and the statistical
//        (KS) comparator.
//   kat  <params.txt> <cases.txt> <out.txt>           function-level known-answer generation in
//        TAPE MODE (see gsl_shim/gsl/gsl_rng.h): Person::infect / isCrossProtected /
//        isVaccineProtected evaluated on scripted person states with Philox-keyed draws.
//   foi  <params.txt> <pop_dir> <tag> <script.txt> <out.txt>   (binary ref_driver_probe only) per-location force of
//        infection of Community::tick on a scripted epidemic, read out by compile-time probes (foi_probe.h); see
//        ref_driver_kat.inc.
#include <chrono>
#include <map>
#include <string>
#include <vector>
#include <fstream>
#include <sstream>
#include <iostream>
#include "simulator.h"
#ifdef CABM_FOI_PROBE
#include "foi_probe_log.h"
#endif
#include "../include/cabm.h"

using namespace std;

namespace {

} // namespace
#include "../covid-abm_b200/facade/kv_parameters.h"
using namespace cabm_kv;
namespace {

// the strain "urn" update of simulate_epidemic (simulator.h:341-372)
void update_strain_mix(const Date* date, const vector<string>& mid, vector<double>& strains) {
    if (mid.size() == 0) return;
    if (*date >= mid[0] and *date < mid[1]) {
        if (strains[WILDTYPE] > 1) { strains[WILDTYPE]--; strains[ALPHA]++; }
    } else if (*date >= mid[1] and *date < mid[2]) {
        if (strains[WILDTYPE] > 1) { strains[WILDTYPE]--; strains[DELTA]++; }
        if (strains[ALPHA] > 1)    { strains[ALPHA]--;    strains[DELTA]++; }
    } else if (*date >= mid[2]) {
        for (int i = 0; i < 5; ++i) {
            if (strains[WILDTYPE] > 1) { strains[WILDTYPE]--; strains[OMICRON]++; }
            if (strains[ALPHA] > 1)    { strains[ALPHA]--;    strains[OMICRON]++; }
            if (strains[DELTA] > 1)    { strains[DELTA]--;    strains[OMICRON]++; }
        }
    }
}

int mode_run(int argc, char** argv) {
    if (argc < 6) { cerr << "run <params> <pop_dir> <tag> <out.csv> [max_days] [warmup_days]\n"; return 2; }
    const KV kv = read_kv(argv[2]);
    Parameters* par = make_parameters(kv);
    set_pop_files(par, argv[3], argv[4]);
    const long max_days = argc > 6 ? atol(argv[6]) : (long) par->runLength;
    const long warmup_days = argc > 7 ? atol(argv[7]) : 0;                     // excluded from the timed total
    gsl_rng_set(RNG, par->randomseed); gsl_rng_set(VAX_RNG, par->randomseed); gsl_rng_set(REPORTING_RNG, par->randomseed); // main.cpp:536-538

    auto t0 = chrono::steady_clock::now();
    Community* community = build_community(par);
    const double load_s = chrono::duration<double>(chrono::steady_clock::now() - t0).count();
    community->setSocialDistancingTimedIntervention(par->timedInterventions.at(SOCIAL_DISTANCING));

    vector<string> mid = kv.count("mutant_intro_dates") ? kv.at("mutant_intro_dates") : vector<string>();
    vector<double> strains = {50.0, 0.0, 0.0, 0.0};                                    // simulator.h:315
    Date* date = community->get_date();
    double hot_s = 0.0; long days = 0;
    for (; date->day() < (signed) par->runLength and days < max_days; date->increment(), ++days) {
        if (date->day() == 0) { seed_epidemic(par, community, WILDTYPE); }             // simulator.h:335
        auto a = chrono::steady_clock::now();
        community->tick();                                                             // simulator.h:338
        update_strain_mix(date, mid, strains);
        seed_epidemic(par, community, date, strains);                                  // simulator.h:374
        if (days >= warmup_days) hot_s += chrono::duration<double>(chrono::steady_clock::now() - a).count();
    }

    ofstream out(argv[5]);
    out << "day,inf,symp,severe,critical,dead,hosp_inc,hosp_prev,icu_inc,icu_prev,severe_prev,vax_cases,"
           "det_case_onset,det_case_report,det_hosp,det_death_onset,det_death_report,inf_wt,inf_alpha,inf_delta,inf_omicron\n";
    CommunityLedger* L = community->get_ledger();
    for (long d = 0; d < days; ++d) {
        out << d << ',' << L->_numNewlyInfected[d] << ',' << L->_numNewlySymptomatic[d] << ',' << L->_numNewlySevere[d] << ','
            << L->_numNewlyCritical[d] << ',' << L->_numNewlyDead[d] << ',' << L->_numHospInc[d] << ',' << L->_numHospPrev[d] << ','
            << L->_numIcuInc[d] << ',' << L->_numIcuPrev[d] << ',' << L->_numSeverePrev[d] << ',' << L->_numVaccinatedCases[d] << ','
            << L->_numDetectedCasesOnset[d] << ',' << L->_numDetectedCasesReport[d] << ',' << L->_numDetectedHospitalizations[d] << ','
            << L->_numDetectedDeathsOnset[d] << ',' << L->_numDetectedDeathsReport[d];
        for (int s = 0; s < NUM_OF_STRAIN_TYPES; ++s) out << ',' << L->_numNewInfectionsByStrain.at((StrainType) s)[d];
        out << '\n';
    }
    out.close();
    const double n = community->getNumPeople();
    printf("{\"impl\": \"reference\", \"agents\": %.0f, \"days\": %ld, \"load_s\": %.3f, \"tick_import_s\": %.6f, "
           "\"agent_days_per_s\": %.6g, \"cumul_asymptomatic\": %zu, \"cumul_death\": %zu}\n",
           n, days - warmup_days, load_s, hot_s, n * (days - warmup_days) / hot_s, community->getCumulIncidenceByOutcome(ASYMPTOMATIC),
           community->getCumulIncidenceByOutcome(DEATH));
    return 0;
}

#include "ref_driver_kat.inc"           // mode_kat, mode_foi (same translation unit: simulator.h defines globals)

} // namespace

int main(int argc, char** argv) {
    if (argc < 2) { cerr << "usage: ref_driver run|kat|foi ...\n"; return 2; }
    const string mode = argv[1];
    if (mode == "run") return mode_run(argc, argv);
    if (mode == "kat") return mode_kat(argc, argv);
    if (mode == "foi") return mode_foi(argc, argv);
    cerr << "unknown mode " << mode << endl; return 2;
}
