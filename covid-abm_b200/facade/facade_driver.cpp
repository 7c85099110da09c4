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
//   facade_driver cache_restore <params.txt> <pop_dir> <tag> <out_prefix> <cache_day> [campaign]
//       the reference's cache-and-restore check (tests/cache_and_restore/cache_and_restore_check.sh: equal checksums of the
//       plot logs written before and after a restore) on the facade: days [0, cache_day) run, the reference's own
//       SimulationCache (behavior_tuner.h:39-87) is taken, the tail runs to the end, restore_from_cache
//       (behavior_tuner.h:490-508) rewinds, the tail runs again -- twice. Writes <out_prefix>.pass{1,2,3}.csv; they must be equal.
//   `campaign` = 1 attaches a general Vac_Campaign (the reference's class) so that Community::vaccinate and the dose
//       bookkeeping run; 2 adds ring vaccination of traced contacts (needs begin_contact_tracing < run_length in the file).
#include <chrono>
#include <unistd.h>
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

// the strain urn of simulate_epidemic (simulator.h:341-372), which the function keeps inline
static void advance_strain_urn(const Date* date, const vector<string>& mid, vector<double>& strains) {
    if (mid.size() < 3) return;
    if (*date >= mid[0] and *date < mid[1]) {
        if (strains[WILDTYPE] > 1) { strains[WILDTYPE]--; strains[ALPHA]++; }
    } else if (*date >= mid[1] and *date < mid[2]) {
        if (strains[WILDTYPE] > 1) { strains[WILDTYPE]--; strains[DELTA]++; }
        if (strains[ALPHA] > 1)    { strains[ALPHA]--;    strains[DELTA]++; }
    } else if (*date >= mid[2]) {
        for (int i = 0; i < 5; ++i) for (StrainType st : { WILDTYPE, ALPHA, DELTA }) if (strains[st] > 1) { strains[st]--; strains[OMICRON]++; }
    }
}

static int cache_restore(int argc, char** argv) {
    if (argc < 7) { cerr << "usage: facade_driver cache_restore <params> <pop_dir> <tag> <out_prefix> <cache_day> [campaign]\n"; return 2; }
    const KV kv = read_kv(argv[2]);
    Parameters* par = make_parameters(kv);
    set_pop_files(par, argv[3], argv[4]);
    const string prefix = argv[5];
    const size_t cache_day = (size_t) atol(argv[6]);
    const int campaign = argc > 7 ? atoi(argv[7]) : 1;
    if (campaign < 1) { cerr << "SimulationCache needs a Vac_Campaign (behavior_tuner.h:56)\n"; return 2; }
    gsl_rng_set(RNG, par->randomseed); gsl_rng_set(VAX_RNG, par->randomseed); gsl_rng_set(REPORTING_RNG, par->randomseed);
    Community* community = build_community(par);
    community->setVac_Campaign(simple_campaign(par, community, 20, max<int>(1, (int) community->getNumPeople() / 200), campaign == 2));
    const vector<string> mid = kv.count("mutant_intro_dates") ? kv.at("mutant_intro_dates") : vector<string>();
    vector<TimeSeriesAnchorPoint> sd_anchors = { { Date::to_ymd(0, par), 0.0 }, { Date::to_ymd(par->runLength / 2, par), 0.4 }, { Date::to_ymd(par->runLength - 1, par), 0.2 } };
    community->setSocialDistancingTimedIntervention(sd_anchors);
    SimulationLedger* ledger = new SimulationLedger();
    ledger->strains = { 50.0, 0.0, 0.0, 0.0 };
    Date* date = community->get_date();

    auto run_to = [&](size_t end_day, ostream* out) {                                      // the hot-path part of simulate_epidemic's loop body
        for (; date->day() < (signed) end_day; date->increment()) {
            const size_t d = date->day();
            if (d == 0) seed_epidemic(par, community, WILDTYPE);
            community->tick();
            advance_strain_urn(date, mid, ledger->strains);
            seed_epidemic(par, community, date, ledger->strains);
            if (not out) continue;
            Vac_Campaign* vc = community->getVac_Campaign();
            *out << d << ',' << community->getNumNewlyInfected()[d] << ',' << community->getNumNewlySymptomatic()[d] << ',' << community->getNumNewlyDead()[d]
                 << ',' << community->getNumHospPrev()[d] << ',' << community->getNumDetectedCasesReport()[d] << ',' << community->getNumVaccinatedCases()[d]
                 << ',' << vc->get_doses_used(d, STANDARD_ALLOCATION) << ',' << vc->get_doses_used(d, URGENT_ALLOCATION)
                 << ',' << community->getNumPeopleQuarantining(d) << ',' << community->getNumInfected(d) << '\n';
        }
    };
    auto tail = [&](int pass) {
        ofstream out(prefix + ".pass" + to_string(pass) + ".csv");
        out << "day,inf,symp,dead,hosp_prev,det_case_report,vax_cases,std_doses,urg_doses,quarantining,infected\n";
        run_to(par->runLength, &out);
        unsigned long long h = 1469598103934665603ull; size_t n_inf = 0, n_vax = 0;       // every infection and dose of every person
        for (Person* p : community->getPeople()) {
            for (Infection* inf : p->getInfectionHistory()) {
                ++n_inf;
                for (long v : { (long) p->getID(), (long) inf->getInfectedTime(), (long) inf->getStrain(), (long) inf->getSymptomTime(), (long) inf->getDeathTime(),
                                (long) (inf->getInfectedBy() ? inf->getInfectedBy()->getID() : -1), (long) inf->secondary_infection_tally() })
                    h = (h ^ (unsigned long long) v) * 1099511628211ull;
            }
            n_vax += p->getNumVaccinations();
            h = (h ^ (unsigned long long) p->getNumVaccinations()) * 1099511628211ull;
        }
        out << "people," << n_inf << ',' << n_vax << ',' << h << ',' << community->getCumulIncidenceByOutcome(DEATH) << '\n';
        return n_inf;
    };

    run_to(cache_day, nullptr);
    SimulationCache* cache = new SimulationCache(community, ledger, RNG, REPORTING_RNG, VAX_RNG);      // behavior_tuner.h:52-64
    size_t n_inf[3];
    n_inf[0] = tail(1);
    for (int pass = 2; pass <= 3; ++pass) {
        restore_from_cache(community, date, cache, ledger, sd_anchors);                                 // behavior_tuner.h:490-508
        if ((size_t) date->day() != cache_day) { cerr << "restore did not rewind the date\n"; return 3; }
        n_inf[pass - 1] = tail(pass);
    }
    printf("{\"impl\": \"facade\", \"agents\": %zu, \"days\": %zu, \"cache_day\": %zu, \"infections\": [%zu, %zu, %zu]}\n",
           (size_t) community->getNumPeople(), (size_t) par->runLength, cache_day, n_inf[0], n_inf[1], n_inf[2]);
    delete cache; delete community;
    return 0;
}

int main(int argc, char** argv) {
    if (argc == 4 and string(argv[1]) == "tables") return dump_tables(argv[2], argv[3]);
    if (argc >= 2 and string(argv[1]) == "cache_restore") return cache_restore(argc, argv);
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

    // The facade uploads parameters and population to the device on first use (drivers still edit Parameters after
    // build_community). Trigger it here, outside the timed day loop, with the first question simulate_epidemic asks anyway
    // (simulator.h:122): a one-off cost of the run, reported on its own.
    auto tu = chrono::steady_clock::now();
    (void) community->getNumInfected(0);
    const double upload_s = chrono::duration<double>(chrono::steady_clock::now() - tu).count();

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

    // The reference's sim-data dump (generate_sim_data_db, simulator.h:642-761; exp/active-vac/main.cpp:814-824), unmodified: it
    // walks getPeople() -> getInfectionHistory() / get_infections_caused() / getDetection() / getVaccinationHistory(), i.e. the
    // proxies the facade fills from the device's infection log (cabm_sync_histories_from), writes seven CSV tables into the
    // working directory and hands them to the sqlite3 command. CABM_SIM_DB=<dir>: run it there (needs a campaign: it reads the
    // dose ledgers of the Vac_Campaign).
    if (const char* dbdir = getenv("CABM_SIM_DB")) {
        if (!community->getVac_Campaign()) { cerr << "CABM_SIM_DB needs campaign >= 1" << endl; return 2; }
        if (chdir(dbdir) != 0) { cerr << "cannot chdir to " << dbdir << endl; return 2; }
        generate_sim_data_db(par, community, (unsigned long) par->randomseed, { "infection_history", "secondary_infections", "infection_detection",
                                                                                 "vaccination_history", "age_bins", "doses_available", "doses_used" });
    }

    size_t n_hist = 0, n_vax = 0, n_secondary = 0, urg = 0;
    if (community->getVac_Campaign()) for (size_t d = 0; d < par->runLength; ++d) urg += community->getVac_Campaign()->get_doses_used(d, URGENT_ALLOCATION);
    double quar_days = 0.0;
    for (size_t d = 0; d < par->runLength; d += 5) quar_days += community->getNumPeopleQuarantining(d);
    for (Person* p : community->getPeople()) {
        n_hist += p->getNumNaturalInfections(); n_vax += p->getNumVaccinations();
        for (Infection* inf : p->getInfectionHistory()) n_secondary += inf->secondary_infection_tally();
    }
    // Rt by day of infection and the offspring distribution, as every exp/ driver reads them after the run (main.cpp:482,749)
    const vector<pair<size_t, double>> Rt = community->getMeanNumSecondaryInfections();
    const vector<size_t> offspring = community->generateOffspringDistribution();
    ofstream rt(prefix + ".rt.csv");
    rt << "day,infections,mean_secondary\n";
    for (size_t d = 0; d < Rt.size(); ++d) rt << d << ',' << Rt[d].first << ',' << Rt[d].second << '\n';
    rt << "offspring"; for (size_t k : offspring) rt << ',' << k; rt << '\n';
    vector<Person*> everybody;
    rt << "hsar," << community->getHouseholdSecondaryAttackRate(everybody) << '\n';
    rt.close();
    printf("{\"impl\": \"facade\", \"agents\": %zu, \"days\": %zu, \"load_s\": %.3f, \"upload_s\": %.3f, \"simulate_epidemic_s\": %.4f, \"agent_days_per_s\": %.6g, "
           "\"infections\": %zu, \"secondary_links\": %zu, \"doses\": %zu, \"urgent_doses\": %zu, \"quarantine_index\": %.6f, \"cumul_death\": %zu}\n",
           n, (size_t) par->runLength, load_s, upload_s, sim_s, (double) n * par->runLength / sim_s, n_hist, n_secondary, n_vax, urg, quar_days,
           community->getCumulIncidenceByOutcome(DEATH));
    delete community;
    return 0;
}
