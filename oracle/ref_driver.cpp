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
//   foi  <params.txt> <pop_dir> <tag> <day> <out.txt> runs to <day>, then dumps the agent state
//        and the reference's own hot-location calendar slice _isHot[day] (Community.h:133) so the
//        oracle's derived per-location infectious pressure can be compared with it.
#include <chrono>
#include <map>
#include <string>
#include <vector>
#include <fstream>
#include <sstream>
#include <iostream>
#include "simulator.h"
#include "../include/cabm.h"

using namespace std;

namespace {

typedef map<string, vector<string>> KV;

KV read_kv(const string& path) {
    KV kv; ifstream f(path); string line;
    if (!f) { cerr << "cannot open " << path << endl; exit(2); }
    while (getline(f, line)) {
        istringstream ss(line); string k, v; if (!(ss >> k)) continue;
        vector<string> vals; while (ss >> v) vals.push_back(v);
        kv[k] = vals;
    }
    return kv;
}
double D(const KV& kv, const string& k, size_t i = 0) { return stod(kv.at(k).at(i)); }
long   I(const KV& kv, const string& k, size_t i = 0) { return stol(kv.at(k).at(i)); }
vector<double> DV(const KV& kv, const string& k) { vector<double> v; for (auto& s : kv.at(k)) v.push_back(stod(s)); return v; }

// Fill a reference Parameters object from the flat table file written by covid-abm_b200/params.py.
// Field-by-field this is what exp/active-vac/main.cpp:68-323 does in C++.
Parameters* make_parameters(const KV& kv) {
    Parameters* par = new Parameters();
    par->randomseed = I(kv, "seed");
    par->runLength = I(kv, "run_length");
    par->startJulianYear = I(kv, "start_julian_year");
    par->startDayOfYear = I(kv, "start_day_of_year");
    par->household_transmission_haz_mult   = D(kv, "haz_household");
    par->social_transmission_haz_mult      = D(kv, "haz_social");
    par->workplace_transmission_haz_mult   = D(kv, "haz_work");
    par->school_transmission_haz_mult      = D(kv, "haz_school");
    par->hospital_transmission_haz_mult    = D(kv, "haz_hospital");
    par->nursinghome_transmission_haz_mult = D(kv, "haz_nursinghome");
    par->seasonality = DV(kv, "seasonality");
    par->IEP = D(kv, "IEP"); par->IEH = D(kv, "IEH"); par->IEF = D(kv, "IEF"); par->IEI = D(kv, "IEI");
    par->vaccine_dose_to_protection_lag = I(kv, "vaccine_dose_to_protection_lag");
    par->immunityLeaky = I(kv, "immunity_leaky"); par->immunityWanes = I(kv, "immunity_wanes");
    par->retroactiveMatureVaccine = I(kv, "retroactive_mature_vaccine");
    par->seroconversionLag = D(kv, "seroconversion_lag"); par->seropositivityThreshold = D(kv, "seropositivity_threshold");
    par->symptomToTestLag = I(kv, "symptom_to_test_lag"); par->defaultReportingLag = I(kv, "default_reporting_lag");
    par->numSurveilledPeople = I(kv, "num_surveilled_people");
    par->numVaccineDoses = I(kv, "num_vaccine_doses");
    par->vaccineDoseInterval.clear(); for (double d : DV(kv, "vaccine_dose_interval")) par->vaccineDoseInterval.push_back((int) d);
    par->vaccineInfConstraint = (VaccineInfConstraint) I(kv, "vaccine_inf_constraint");
    par->vaccineSeroConstraint = (VaccineSeroConstraint) I(kv, "vaccine_sero_constraint");
    par->seroTestFalsePos = D(kv, "sero_false_pos"); par->seroTestFalseNeg = D(kv, "sero_false_neg");
    par->beginContactTracing = I(kv, "begin_contact_tracing");
    par->contactTracingCoverage = D(kv, "contact_tracing_coverage");
    par->contactTracingEV = DV(kv, "contact_tracing_ev");
    par->contactTracingDepth = I(kv, "contact_tracing_depth");
    par->quarantineProbability = DV(kv, "quarantine_probability");
    par->quarantineDuration = I(kv, "quarantine_duration");
    par->numInitialExposed = I(kv, "num_initial_exposed"); par->probInitialExposure = D(kv, "prob_initial_exposure");
    par->traceContacts = true;
    for (int s = 0; s < NUM_OF_STRAIN_TYPES; ++s) {
        StrainPars& sp = par->strainPars[s];
        sp.relInfectiousness = D(kv, "strain_rel_infectiousness", s); sp.relPathogenicity = D(kv, "strain_rel_pathogenicity", s);
        sp.relSeverity = D(kv, "strain_rel_severity", s); sp.relMortality = D(kv, "strain_rel_mortality", s);
        sp.relIcuMortality = D(kv, "strain_rel_icu_mortality", s); sp.immuneEscapeProb = D(kv, "strain_immune_escape", s);
        sp.symptomaticInfectiousPeriod = I(kv, "strain_infectious_period", s); sp.relSymptomOnset = D(kv, "strain_rel_symptom_onset", s);
    }
    par->crossProtectionMatrix.assign(NUM_OF_STRAIN_TYPES, vector<bool>(NUM_OF_STRAIN_TYPES, false));
    for (int a = 0; a < NUM_OF_STRAIN_TYPES; ++a) for (int b = 0; b < NUM_OF_STRAIN_TYPES; ++b)
        par->crossProtectionMatrix[a][b] = I(kv, "cross_protection", a * NUM_OF_STRAIN_TYPES + b) != 0;
    auto ve = [&](const string& key) {
        map<StrainType, vector<double>> m; const vector<double> v = DV(kv, key); const size_t nd = v.size() / NUM_OF_STRAIN_TYPES;
        for (int s = 0; s < NUM_OF_STRAIN_TYPES; ++s) m[(StrainType) s] = vector<double>(v.begin() + s * nd, v.begin() + (s + 1) * nd);
        return m;
    };
    par->VES = ve("VES"); par->VES_NAIVE = ve("VES_NAIVE"); par->VEP = ve("VEP"); par->VEH = ve("VEH"); par->VEI = ve("VEI"); par->VEF = ve("VEF");
    { const vector<double> v = DV(kv, "prob_first_detection");
      par->probFirstDetection.assign(par->runLength, vector<double>(NUM_OF_OUTCOME_TYPES));
      for (size_t d = 0; d < par->runLength; ++d) for (int o = 0; o < NUM_OF_OUTCOME_TYPES; ++o) par->probFirstDetection[d][o] = v[d * NUM_OF_OUTCOME_TYPES + o]; }
    par->timedInterventions[SCHOOL_CLOSURE] = DV(kv, "school_closure");
    par->timedInterventions[NONESSENTIAL_BUSINESS_CLOSURE] = DV(kv, "business_closure");
    par->timedInterventions[SOCIAL_DISTANCING] = DV(kv, "social_distancing");
    par->icuMortalityReduction = DV(kv, "icu_mortality_reduction");
    par->icuMortalityFraction = D(kv, "icu_mortality_fraction");
    par->define_susceptibility_and_pathogenicity();       // the reference's own tables (Parameters.cpp:128-195)
    par->probDailyExposure = DV(kv, "prob_daily_exposure");
    par->behaviorInputFilename = ""; par->behaviorOutputFilename = "";
    return par;
}

void set_pop_files(Parameters* par, const string& dir, const string& tag) {
    par->populationFilename     = dir + "/population-" + tag + ".txt";
    par->comorbidityFilename    = dir + "/comorbidity-" + tag + ".txt";
    par->locationFilename       = dir + "/locations-" + tag + ".txt";
    par->networkFilename        = dir + "/network-" + tag + ".txt";
    par->publicActivityFilename = dir + "/public-activity-" + tag + ".txt";
}

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
