// covid-abm_b200/facade/kv_parameters.h -- fill a reference `Parameters` object from the flat "key value..." file that
// covid-abm_b200/params.py write_text() produces. Field by field this is what exp/active-vac/main.cpp:68-323 does in
// C++ for the active-vac parameter set; drivers that already build their own Parameters do not need it.
// Include after simulator.h (it uses the reference's Parameters / enums). Shared by the facade driver and by the
// test-only oracle/ref_driver.cpp.
#ifndef CABM_KV_PARAMETERS_H
#define CABM_KV_PARAMETERS_H
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

namespace cabm_kv {
using namespace std;

typedef map<string, vector<string>> KV;

inline KV read_kv(const string& path) {
    KV kv; ifstream f(path); string line;
    if (!f) { cerr << "cannot open " << path << endl; exit(2); }
    while (getline(f, line)) {
        istringstream ss(line); string k, v; if (!(ss >> k)) continue;
        vector<string> vals; while (ss >> v) vals.push_back(v);
        kv[k] = vals;
    }
    return kv;
}
inline double D(const KV& kv, const string& k, size_t i = 0) { return stod(kv.at(k).at(i)); }
inline long   I(const KV& kv, const string& k, size_t i = 0) { return stol(kv.at(k).at(i)); }
inline vector<double> DV(const KV& kv, const string& k) { vector<double> v; for (auto& s : kv.at(k)) v.push_back(stod(s)); return v; }

// Fill a reference Parameters object from the flat table file written by covid-abm_b200/params.py.
// Field-by-field this is what exp/active-vac/main.cpp:68-323 does in C++.
inline Parameters* make_parameters(const KV& kv) {
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
    // optional: the behaviour autotuner of simulate_epidemic (exp/active-vac/main.cpp:93-97,257-262)
    if (kv.count("behavioral_autotuning") and I(kv, "behavioral_autotuning")) {
        par->behavioral_autotuning = true;
        par->tuning_window = I(kv, "tuning_window"); par->num_preview_windows = I(kv, "num_preview_windows");
        par->rCaseDeathFilename = kv.at("rcasedeath_file").at(0);
        par->behaviorOutputFilename = kv.at("behavior_output_file").at(0);
        par->behavior_fitting_data_target = CASES;
        par->death_tuning_offset = 18;
    }
    return par;
}

inline void set_pop_files(Parameters* par, const string& dir, const string& tag) {
    par->populationFilename     = dir + "/population-" + tag + ".txt";
    par->comorbidityFilename    = dir + "/comorbidity-" + tag + ".txt";
    par->locationFilename       = dir + "/locations-" + tag + ".txt";
    par->networkFilename        = dir + "/network-" + tag + ".txt";
    par->publicActivityFilename = dir + "/public-activity-" + tag + ".txt";
}

} // namespace cabm_kv
#endif
