// covid-abm_b200/facade/facade_driver.cpp -- runs the reference's UNMODIFIED simulator.h on top of the facade.
//
// This translation unit includes /root/reference/simulator.h exactly as the reference's experiment drivers do
// (exp/active-vac/main.cpp:4); `Community`, `Person` and `Location` resolve to the drop-in proxies of this directory,
// `Parameters`, `Date`, `Utility`, `Vac_Campaign` and `behavior_tuner.h` are the reference's own files. The day loop
// is the reference's `simulate_epidemic` (simulator.h:265-533): community->tick(), seed_epidemic(), the ~20 by-value
// ledger getters, calculate_vax_stats, two serosurveys and the quarantine census every day.
//
//   facade_driver sim <params.txt> <pop_dir> <tag> <out_prefix> [campaign]
//       writes <out_prefix>.plot.csv (simulate_epidemic's plot log) and <out_prefix>.ledger.csv (same columns as
//       oracle/ref_driver run) and prints one JSON line with wall-clock timings.
//   `campaign` = 1 attaches a general Vac_Campaign (the reference's class) so that Community::vaccinate and the dose
//       bookkeeping run; 2 adds ring vaccination of traced contacts (needs begin_contact_tracing < run_length in the file).
#include <chrono>
#include "simulator.h"
#include "kv_parameters.h"

using namespace std;
using namespace cabm_kv;

static Vac_Campaign* simple_campaign(Parameters* par, Community* community, int first_day, int doses_per_day, bool ring) {
    // the call sequence of generateVac_Campaign / parseVaccineFile (exp/active-vac/main.cpp:327-460,655-706) with a synthetic,
    // un-pooled standard allocation: `doses_per_day` first doses per day for adults from `first_day` on, boosters as they fall due
    Vac_Campaign* vc = new Vac_Campaign(par);
    vc->set_rng(VAX_RNG);
    vc->set_start_of_campaign(GENERAL_CAMPAIGN, first_day);
    vc->generate_age_bins(community, std::set<int>{ 18 }, std::set<int>{ NUM_AGE_CLASSES - 1 });
    Dose_Vals std_doses(par->runLength, vector<map<int, int>>(par->numVaccineDoses)), urg_doses(par->runLength, vector<map<int, int>>(par->numVaccineDoses));
    for (int d = 0; d < (int) par->runLength; ++d) for (int dose = 0; dose < par->numVaccineDoses; ++dose) for (int bin : vc->get_unique_age_bins()) {
        std_doses[d][dose][bin] = (d >= first_day and bin >= 18) ? doses_per_day : 0;
        urg_doses[d][dose][bin] = (ring and d >= par->beginContactTracing and dose == 0) ? doses_per_day : 0;
    }
    vc->set_pool_urg_doses(false); vc->set_pool_std_doses(false); vc->set_pool_all_doses(false);
    vc->init_doses_available(urg_doses, std_doses);
    vc->init_eligibility_queue(community);
    vc->set_prioritize_first_doses(false);
    vc->set_flexible_queue_allocation(false);
    if (ring) {                                                                           // exp/active-vac/main.cpp:609,658-674
        par->urgent_vax_dose_threshold = 1;
        vc->set_reactive_vac_strategy(RING_VACCINATION);
        vc->set_start_of_campaign(RING_VACCINATION, par->beginContactTracing);
        vc->set_end_of_campaign(RING_VACCINATION, par->runLength);
    } else {
        vc->set_reactive_vac_strategy(NO_CAMPAIGN);
    }
    vc->set_min_age(vector<int>(par->runLength, 5));
    vc->set_end_of_campaign(GENERAL_CAMPAIGN, par->runLength);
    return vc;
}

#include "../../include/cabm.h"
void cabm_facade_fill_params(const Parameters* par, cabm_params* out, std::vector<double>& pfd, std::vector<double>& imr);

// tables <params.txt> <out.txt>: the integer-support CDF tables params_bridge.cpp derives from `Parameters`, for comparison with
// covid-abm_b200/params.py build_tables() (tests/test_facade.py)
static int dump_tables(const char* params, const char* path) {
    Parameters* par = make_parameters(read_kv(params));
    cabm_params q; vector<double> pfd, imr;
    cabm_facade_fill_params(par, &q, pfd, imr);
    ofstream o(path); o.precision(17);
    auto row = [&](const char* name, const double* v, size_t n) { o << name; for (size_t i = 0; i < n; ++i) o << ' ' << v[i]; o << '\n'; };
    row("incub_cdf", &q.incub_cdf[0][0], CABM_N_STRAINS * CABM_INCUB_MAX); row("onset_cdf", &q.onset_cdf[0][0], (CABM_INCUB_MAX + 1) * (CABM_INCUB_MAX + 1));
    row("presev_cdf", q.presev_cdf, CABM_T_GAMMA); row("mild_cdf", q.mild_cdf, CABM_T_GAMMA); row("severe_only_cdf", q.severe_only_cdf, CABM_N_SEVERE_CDF);
    row("critical_cdf", q.critical_cdf, CABM_N_CRITICAL_CDF); row("icu_death_cdf", q.icu_death_cdf, CABM_T_NEGBIN);
    row("community_death_cdf", q.community_death_cdf, CABM_T_GEOM); row("death_lag_cdf", q.death_lag_cdf, CABM_T_DEATHLAG);
    row("visit_cdf", q.visit_cdf, CABM_T_POISSON); row("trace_ev_cdf", &q.trace_ev_cdf[0][0], CABM_N_LOCTYPES * CABM_T_POISSON);
    const double ph[2] = { q.p_high_asym, q.p_high_sym }; row("p_high", ph, 2);
    row("haz_mult", q.haz_mult, CABM_N_LOCTYPES + 1); row("VES", &q.VES[0][0], CABM_N_STRAINS * CABM_MAX_DOSES); row("VEH", &q.VEH[0][0], CABM_N_STRAINS * CABM_MAX_DOSES);
    vector<double> f(q.prob_icu_death[1], q.prob_icu_death[1] + CABM_N_AGES); row("prob_icu_death_comorbid", f.data(), f.size());
    row("prob_first_detection", pfd.data(), pfd.size()); row("icu_mortality_reduction", imr.data(), imr.size());
    return 0;
}

int main(int argc, char** argv) {
    if (argc == 4 and string(argv[1]) == "tables") return dump_tables(argv[2], argv[3]);
    if (argc < 6 or string(argv[1]) != "sim") { cerr << "usage: facade_driver sim <params> <pop_dir> <tag> <out_prefix> [campaign]\n"; return 2; }
    const KV kv = read_kv(argv[2]);
    Parameters* par = make_parameters(kv);
    set_pop_files(par, argv[3], argv[4]);
    const string prefix = argv[5];
    const int campaign = argc > 6 ? atoi(argv[6]) : 0;                                  // 0 none, 1 general campaign, 2 + ring vaccination of traced contacts
    gsl_rng_set(RNG, par->randomseed); gsl_rng_set(VAX_RNG, par->randomseed); gsl_rng_set(REPORTING_RNG, par->randomseed);   // main.cpp:536-538

    auto t0 = chrono::steady_clock::now();
    Community* community = build_community(par);                                       // simulator.h:63-91
    const double load_s = chrono::duration<double>(chrono::steady_clock::now() - t0).count();
    if (campaign) community->setVac_Campaign(simple_campaign(par, community, 20, max<int>(1, (int) community->getNumPeople() / 200), campaign == 2));
    const size_t n = community->getNumPeople();
    const vector<string> mid = kv.count("mutant_intro_dates") ? kv.at("mutant_intro_dates") : vector<string>();

    auto t1 = chrono::steady_clock::now();
    vector<string> plot_log = simulate_epidemic(par, community, "facade", mid);        // simulator.h:265-533
    const double sim_s = chrono::duration<double>(chrono::steady_clock::now() - t1).count();

    ofstream plot(prefix + ".plot.csv");
    for (const string& line : plot_log) plot << line << '\n';
    plot.close();
    ofstream out(prefix + ".ledger.csv");
    out << "day,inf,symp,severe,critical,dead,hosp_inc,hosp_prev,icu_inc,icu_prev,severe_prev,vax_cases,"
           "det_case_onset,det_case_report,det_hosp,det_death_onset,det_death_report,inf_wt,inf_alpha,inf_delta,inf_omicron\n";
    const vector<vector<size_t>> cols = { community->getNumNewlyInfected(), community->getNumNewlySymptomatic(), community->getNumNewlySevere(),
        community->getNumNewlyCritical(), community->getNumNewlyDead(), community->getNumHospInc(), community->getNumHospPrev(), community->getNumIcuInc(),
        community->getNumIcuPrev(), community->getNumSeverePrev(), community->getNumVaccinatedCases(), community->getNumDetectedCasesOnset(),
        community->getNumDetectedCasesReport(), community->getNumDetectedHospitalizations(), community->getNumDetectedDeathsOnset(),
        community->getNumDetectedDeathsReport(), community->getNumNewInfections(WILDTYPE), community->getNumNewInfections(ALPHA),
        community->getNumNewInfections(DELTA), community->getNumNewInfections(OMICRON) };
    for (size_t d = 0; d < par->runLength; ++d) { out << d; for (const auto& c : cols) out << ',' << c[d]; out << '\n'; }
    out.close();

    size_t n_hist = 0, n_vax = 0, n_secondary = 0, urg = 0;
    if (community->getVac_Campaign()) for (size_t d = 0; d < par->runLength; ++d) urg += community->getVac_Campaign()->get_doses_used(d, URGENT_ALLOCATION);
    double quar_days = 0.0;
    for (size_t d = 0; d < par->runLength; d += 5) quar_days += community->getNumPeopleQuarantining(d);
    for (Person* p : community->getPeople()) {
        n_hist += p->getNumNaturalInfections(); n_vax += p->getNumVaccinations();
        for (Infection* inf : p->getInfectionHistory()) n_secondary += inf->secondary_infection_tally();
    }
    printf("{\"impl\": \"facade\", \"agents\": %zu, \"days\": %zu, \"load_s\": %.3f, \"simulate_epidemic_s\": %.4f, \"agent_days_per_s\": %.6g, "
           "\"infections\": %zu, \"secondary_links\": %zu, \"doses\": %zu, \"urgent_doses\": %zu, \"quarantine_index\": %.6f, \"cumul_death\": %zu}\n",
           n, (size_t) par->runLength, load_s, sim_s, (double) n * par->runLength / sim_s, n_hist, n_secondary, n_vax, urg, quar_days,
           community->getCumulIncidenceByOutcome(DEATH));
    delete community;
    return 0;
}
