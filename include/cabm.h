/* include/cabm.h -- C ABI of the B200-native covid-abm daily transmission cycle.
 *
 * The reference has no FFI for this path: its boundary is the C++ class surface of `Community`
 * (Community.h:138-441) consumed by simulator.h. A drop-in `Community` facade forwards each of the
 * calls below to this library (see INTEGRATION.md for the binding a maintainer adds). Every entry
 * point cites the reference interface it replaces.
 *
 * Conventions: plain pointers and sizes only; the library owns all device memory behind the
 * handle; caller owns every host buffer and may free it when the call returns; return 0 = ok,
 * <0 = error with text in cabm_last_error(). One host thread per handle; a handle is bound to the
 * CUDA device that was current at cabm_create(). There is NO CPU fallback: cabm_create fails with
 * CABM_E_NODEVICE when no CUDA device is usable.
 */
#ifndef CABM_H
#define CABM_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CABM_N_STRAINS   4      /* Parameters.h:168-174 */
#define CABM_N_AGES      121    /* Parameters.h:315 */
#define CABM_MAX_DOSES   4
#define CABM_N_OUTCOMES  5      /* Parameters.h:127-134 */
#define CABM_N_LOCTYPES  5      /* Parameters.h:74-81 */
#define CABM_INCUB_MAX   64
#define CABM_T_GAMMA     128
#define CABM_T_NEGBIN    128
#define CABM_T_GEOM      64
#define CABM_T_DEATHLAG  256
#define CABM_T_POISSON   32
#define CABM_N_SEVERE_CDF   71  /* Parameters.h:291-299 */
#define CABM_N_CRITICAL_CDF 60  /* Parameters.h:303-309 */

/* error codes */
#define CABM_OK          0
#define CABM_E_ARG      -1
#define CABM_E_NODEVICE -2
#define CABM_E_CUDA     -3
#define CABM_E_STATE    -4
#define CABM_E_NOMEM    -5

/* location types / public-transmission risk (Parameters.h:74-81, 98-103) */
enum { CABM_HOUSE = 0, CABM_WORK = 1, CABM_SCHOOL = 2, CABM_HOSPITAL = 3, CABM_NURSINGHOME = 4 };
enum { CABM_PT_NONE = 0, CABM_PT_LOW = 1, CABM_PT_HIGH = 2 };
/* timed interventions (Parameters.h:206-211) */
enum { CABM_TI_SCHOOL_CLOSURE = 0, CABM_TI_BUSINESS_CLOSURE = 1, CABM_TI_SOCIAL_DISTANCING = 2 };

/* Philox purpose slots: counter = (sub<<16 | block, a, b, day<<8 | purpose), key = seed */
enum {
    CABM_P_INIT = 1,          /* a = person                      : load-time immunity attributes            */
    CABM_P_VISIT = 2,         /* a = person                      : Community.cpp:1013-1017                  */
    CABM_P_EXPOSE_HH = 3,     /* a = person, b = source location : Community.cpp:842-850 -> :971-996        */
    CABM_P_EXPOSE_NB = 4,     /*                                   Community.cpp:853-875                    */
    CABM_P_EXPOSE_WORK_VIS = 5, /*                                 Community.cpp:904 (visitors)             */
    CABM_P_EXPOSE_WORK = 6,   /*                                   Community.cpp:905 (staff)                */
    CABM_P_EXPOSE_SCHOOL = 7, /*                                   Community.cpp:914-928                    */
    CABM_P_EXPOSE_NH = 8,     /*                                   Community.cpp:959-969                    */
    CABM_P_EXPOSE_HOSP = 9,   /* staff                             Community.cpp:931-941                    */
    CABM_P_EXPOSE_HOSP_PAT = 10, /* admitted patients (sub = occurrence in the hospital list)               */
    CABM_P_IMPORT = 11,       /* a = import index of the day     : simulator.h:201-223                      */
    CABM_P_SEED = 12,         /* a = person, b = call sequence   : Community::infect, simulator.h:94-131    */
    CABM_P_IMPORT_COUNT = 13, /* a = chunk                       : simulator.h:209 Poisson count            */
    CABM_P_TRACE = 14,        /* a = person                      : Community.cpp:1064                       */
    CABM_P_QUARANTINE = 15,   /* a = person                      : Community.cpp:1135                       */
    CABM_P_REPORTING = 0x80   /* OR-ed onto an event's purpose   : REPORTING_RNG draws (Person.cpp:424,452,459) */
};

/* Daily ledger series (CommunityLedger, Community.h:112-130). */
enum {
    CABM_C_NEW_INFECTED = 0, CABM_C_NEW_SYMPTOMATIC, CABM_C_NEW_SEVERE, CABM_C_NEW_CRITICAL, CABM_C_NEW_DEAD,
    CABM_C_VACCINATED_CASES, CABM_C_SEVERE_PREV, CABM_C_HOSP_INC, CABM_C_HOSP_PREV, CABM_C_ICU_INC, CABM_C_ICU_PREV,
    CABM_C_DET_CASES_ONSET, CABM_C_DET_CASES_REPORT, CABM_C_DET_HOSP, CABM_C_DET_DEATHS_ONSET, CABM_C_DET_DEATHS_REPORT,
    CABM_C_NEW_INF_STRAIN0, CABM_C_NEW_INF_STRAIN1, CABM_C_NEW_INF_STRAIN2, CABM_C_NEW_INF_STRAIN3,
    CABM_N_COUNTERS
};

/* Work counters of the byte model in SURVEY.md 8(d), accumulated on device every day. */
enum {
    CABM_W_AGENTS = 0,      /* N  : agents scanned                                   */
    CABM_W_ACTIVE,          /* A  : agents with an unresolved infection              */
    CABM_W_FLAGS,           /* F  : (infectious agent, location) pressure updates    */
    CABM_W_PATRONS,         /* P  : agents drawing visits                            */
    CABM_W_VISITS,          /* V0 : visits drawn                                     */
    CABM_W_HOT_VISITS,      /* V  : visits landing on hot locations                  */
    CABM_W_HOT_LOCS,        /* H  : hot locations                                    */
    CABM_W_NBR_EDGES,       /* E  : neighbour edges of hot risky households          */
    CABM_W_TRIALS,          /* X  : exposure trials                                  */
    CABM_W_INFECTIONS,      /* I  : new infections                                   */
    CABM_N_WORK
};

typedef struct cabm cabm_t;
typedef struct cabm_snap cabm_snap_t;

/* Flat model tables: the members of `Parameters` the hot path reads. */
typedef struct cabm_params {
    uint64_t seed;                        /* Parameters.h:465 randomseed -> Philox key                    */
    int32_t  run_length;                  /* Parameters.h:466                                              */
    int32_t  num_surveilled_people;       /* Parameters.h:661                                              */
    double   haz_mult[CABM_N_LOCTYPES + 1]; /* household, work, school, hospital, nursinghome, social (Parameters.h:467-472) */
    double   seasonality[366];            /* Parameters.h:473, indexed by julian day - 1 (Parameters.cpp:252) */
    float    susceptibility_by_age[CABM_N_AGES];      /* Parameters.h:476 */
    float    pathogenicity_by_age[CABM_N_AGES];       /* Parameters.h:477 */
    float    prob_severe[2][CABM_N_AGES];             /* probSeriousOutcome[SEVERE][comorbid][age]   Parameters.h:479 */
    float    prob_critical[2][CABM_N_AGES];           /* probSeriousOutcome[CRITICAL]                                 */
    float    prob_icu_death[2][CABM_N_AGES];          /* probSeriousOutcome[DEATH]                                    */
    double   VES[CABM_N_STRAINS][CABM_MAX_DOSES];     /* Parameters.h:481-486, stretchy_vector already applied        */
    double   VES_NAIVE[CABM_N_STRAINS][CABM_MAX_DOSES];
    double   VEP[CABM_N_STRAINS][CABM_MAX_DOSES];
    double   VEH[CABM_N_STRAINS][CABM_MAX_DOSES];
    double   VEF[CABM_N_STRAINS][CABM_MAX_DOSES];
    double   VEI[CABM_N_STRAINS][CABM_MAX_DOSES];
    double   IEP, IEH, IEF, IEI;                      /* Parameters.h:491-494 */
    int32_t  vaccine_dose_to_protection_lag;          /* Parameters.h:488 */
    int32_t  immunity_leaky, immunity_wanes, retroactive_mature_vaccine;  /* Parameters.h:572-573,649 */
    double   strain_rel_infectiousness[CABM_N_STRAINS];   /* StrainPars, Parameters.h:327-351 */
    double   strain_rel_pathogenicity[CABM_N_STRAINS];
    double   strain_rel_severity[CABM_N_STRAINS];
    double   strain_rel_mortality[CABM_N_STRAINS];
    double   strain_rel_icu_mortality[CABM_N_STRAINS];
    double   strain_immune_escape[CABM_N_STRAINS];
    int32_t  strain_infectious_period[CABM_N_STRAINS];
    int32_t  cross_protection[CABM_N_STRAINS][CABM_N_STRAINS]; /* Parameters.h:659 */
    int32_t  symptom_to_test_lag, default_reporting_lag;       /* Parameters.h:534,536 (no ReportingLagModel) */
    int32_t  begin_contact_tracing, contact_tracing_depth, quarantine_duration; /* Parameters.h:509-512,696 */
    double   contact_tracing_coverage;
    double   quarantine_probability[4];                       /* Parameters.h:695 */
    double   seroconversion_lag, seropositivity_threshold;    /* Parameters.h:650-651 */
    /* integer-support CDF tables for the discretised duration draws (see covid-abm_b200/params.py) */
    double   incub_cdf[CABM_N_STRAINS][CABM_INCUB_MAX];        /* 1+floor(Gamma)      Parameters.h:515-520 */
    double   onset_cdf[CABM_INCUB_MAX + 1][CABM_INCUB_MAX + 1];/* Binomial(p, n)      Parameters.h:521-524 */
    double   presev_cdf[CABM_T_GAMMA];                         /* Parameters.h:525-527 */
    double   mild_cdf[CABM_T_GAMMA];                           /* Parameters.h:528-530 */
    double   severe_only_cdf[CABM_N_SEVERE_CDF];               /* Parameters.h:531 */
    double   critical_cdf[CABM_N_CRITICAL_CDF];                /* Parameters.h:532 */
    double   p_high_asym, p_high_sym;                          /* Person.cpp:340-355 Weibull > 8.04 */
    double   icu_death_cdf[CABM_T_NEGBIN];                     /* Parameters.h:450-456 */
    double   community_death_cdf[CABM_T_GEOM];                 /* Parameters.h:458-461 */
    double   death_lag_cdf[CABM_T_DEATHLAG];                   /* Parameters.h:547-553 */
    double   visit_cdf[CABM_T_POISSON];                        /* Community.cpp:1012-1013 Poisson(1) */
    double   trace_ev_cdf[CABM_N_LOCTYPES][CABM_T_POISSON];    /* Community.cpp:1089,1105,1120 */
    /* per-day tables, length run_length; copied during the call */
    const double* prob_first_detection;   /* [run_length][5]  Parameters.h:566 */
    const double* icu_mortality_reduction;/* [run_length] or NULL  Parameters.h:671 */
} cabm_params;

/* Population as structure-of-arrays + CSR (what loadLocations/loadPopulation build, Community.cpp:174-524). */
typedef struct cabm_population {
    uint32_t n_people, n_loc;
    /* per location */
    const uint8_t*  loc_type;        /* CABM_HOUSE..            Location.h:97-109 */
    const uint8_t*  loc_essential;   /* 1 = essential            */
    const uint8_t*  loc_pubrisk;     /* CABM_PT_*                */
    const float*    loc_risk;        /* riskiness (float, Location.h:85) */
    const int32_t*  loc_hospital;    /* associated hospital id or -1 (Location.h:105) */
    const int64_t*  nbr_off;         /* [n_loc+1] neighbour CSR, ascending ids (std::set order, Location.h:60) */
    const int32_t*  nbr;
    const int64_t*  roster_off;      /* [n_loc+1] residents+workers in load order (Community.cpp:219-224) */
    const int32_t*  roster;
    /* per person */
    const int32_t*  home;            /* Person.h:355 */
    const int32_t*  day;             /* day location or -1 (Person.h:356) */
    const uint8_t*  age;
    const uint8_t*  sex;
    const uint8_t*  comorbid;
    const uint8_t*  ltc;             /* long-term-care resident */
    const int32_t*  patron;          /* [n_people][n_patron] patronised locations, -1 padded (Person.h:357) */
    uint32_t        n_patron;
    /* load-time immunity attributes (Community.cpp:212-215); NULL => drawn from Philox(CABM_P_INIT) */
    const double*   starting_natural_efficacy;
    const float*    cross_protection_probability;
    const int32_t*  natural_immunity_duration;
} cabm_population;

/* Daily diagnostics the caller computes over all people (simulator.h:402-410,438), as one device pass. */
enum cabm_diag {
    CABM_D_DOSE1 = 0,        /* people with >= 1, >= 2, >= 3 doses (calculate_vax_stats, Community.cpp:1351-1357) */
    CABM_D_DOSE2,
    CABM_D_DOSE3,
    CABM_D_FULLY_VAXD,       /* > 1 dose and days since last dose >= vaccine_dose_to_protection_lag (:1359) */
    CABM_D_BT_INF,           /* among those: latest infection began / symptoms began / hospitalised / died / reported today (:1361-1385) */
    CABM_D_BT_DIS,
    CABM_D_BT_HOSP,
    CABM_D_BT_DEATH,
    CABM_D_BT_REPORTED,
    CABM_D_UV_INF,           /* the same events among everybody else */
    CABM_D_UV_DIS,
    CABM_D_UV_HOSP,
    CABM_D_UV_DEATH,
    CABM_D_UV_REPORTED,
    CABM_D_INFECTED,         /* getNumInfected(day), getNumInfectious(day), getNumSymptomatic(day), getNumNaive() (:1196-1222) */
    CABM_D_INFECTIOUS,
    CABM_D_SYMPTOMATIC,
    CABM_D_NAIVE,
    CABM_D_QUARANTINING,     /* numerator of getNumPeopleQuarantining(day) (Community.h:437) */
    CABM_D_SEROPOS_ALL,      /* doSerosurvey(NATURAL, everybody, day) numerator (:1224-1274) */
    CABM_D_SEROPOS_PED,      /* the same over ages 0-17 (simulator.h:405-410) ... */
    CABM_D_N_PED,            /* ... and that cohort's size */
    CABM_N_DIAG = 24
};

/* One record of Person::infectionHistory, as returned by cabm_sync_histories (Person.h:35-175). */
typedef struct cabm_infection {
    uint32_t person;
    int32_t  infected_by;        /* -1: imported / unknown */
    int32_t  infected_place;     /* -1: none */
    int16_t  t[12];              /* infectedBegin, infectiousBegin, infectiousEnd, symptomBegin, symptomEnd, severeBegin,
                                    severeEnd, hospitalizedBegin, criticalBegin, criticalEnd, icuBegin, deathTime;
                                    32767 / -32768 stand for INT_MAX / INT_MIN */
    int16_t  report_day;         /* Detection::reported_time or 32767 */
    uint8_t  strain, detected_state; /* detected_state = CABM_N_OUTCOMES when undetected */
    float    rel_infectiousness;
} cabm_infection;

/* Parity hook: one row per hot location of the day (what _tally_infectiousness and the six
 * *_transmission functions compute, Community.cpp:833-969). */
typedef struct cabm_foi_row {
    uint32_t loc; uint8_t type; uint8_t pad[3];
    double   w[CABM_N_STRAINS];  /* infectious weight by strain */
    uint32_t size;               /* denominator census used (roster, + visitors for WORK) */
    double   T;                  /* exposure probability of the location's own pass (0 if gated off) */
    double   T_social;           /* HOUSE only: between-household T (0 if gated off) */
} cabm_foi_row;

/* lifecycle: replaces `new Community(par,date)` / `delete` (Community.cpp:32-73,126-152) */
int  cabm_create(cabm_t** out);
int  cabm_destroy(cabm_t* h);
const char* cabm_last_error(const cabm_t* h);

/* replaces loadLocations + loadPopulation (Community.cpp:174-524) */
int  cabm_load_population(cabm_t* h, const cabm_population* pop);
/* The built population (the arrays of cabm_population: per-location and per-person SoA, roster and neighbour CSRs) as ONE binary
 * file, so that a later run loads it with a read and the host-to-device copies instead of parsing the five text files and
 * rebuilding the CSRs (Community.cpp:174-524). cabm_population_save writes `pop` without its load-time
 * immunity attributes -- they depend on the run's seed (Community.cpp:212-215); cabm_load_population_file maps the file, checks
 * its header and sizes, and loads it with the attributes the caller passes (same requirements as cabm_load_population). The file
 * is little-endian, self-describing (magic "CABMPOP1", counts, then the arrays 8-byte aligned in the order of the struct). */
int  cabm_population_save(const cabm_population* pop, const char* path);
int  cabm_load_population_file(cabm_t* h, const char* path, const double* starting_natural_efficacy,
                               const float* cross_protection_probability, const int32_t* natural_immunity_duration);
/* replaces the `static const Parameters* _par` reads (Community.cpp:28, Person.cpp:24); re-callable */
int  cabm_set_params(cabm_t* h, const cabm_params* par);
/* replaces setSocialDistancingTimedIntervention / updateTimedIntervention (Community.h:396-425) */
int  cabm_set_timed_intervention(cabm_t* h, int which, const double* v, size_t n);

/* replaces Community::tick() (Community.cpp:1147-1192); julian_day/is_weekday come from Date (Date.h).
 * The arguments are checked at once; the day itself is staged and runs in the NEXT call into the handle: together with the
 * day's importation when that call is cabm_import for the same day (the simulate_epidemic loop, simulator.h:338-374: one
 * launch sequence and one host synchronisation per day), on its own before anything else (every getter and every mutator
 * sees the completed tick). A device-side error of the staged day (list overflow) is therefore returned by that next call.
 * Environment CABM_EAGER_TICK=1 makes cabm_tick run the day before it returns. */
int  cabm_tick(cabm_t* h, int sim_day, int julian_day, int is_weekday);
/* replaces seed_epidemic(par, community, date, strain_weights) (simulator.h:201-223); *out_infected = people whose
 * infection today came from an importation draw */
int  cabm_import(cabm_t* h, int sim_day, double expected, const double strain_w[CABM_N_STRAINS], int32_t* out_infected);
/* replaces Community::infect(id, strain) (Community.cpp:550-553) */
int  cabm_infect(cabm_t* h, uint32_t pid, int strain, int sim_day, int32_t* out_infected);
/* n consecutive Community::infect(id, strain) calls (Community.cpp:550-553) in one launch: the importation loop of
   seed_epidemic (simulator.h:209-219) and the initial seeding (simulator.h:101-103,113) call Person::infect once per case;
   a facade queues those calls and resolves them here with one round trip. Same result as calling cabm_infect n times in
   this order (the k-th call gets the k-th sequence number of the keyed stream; a person listed twice is tried again after
   its first try has been applied). out_infected[k] = 1 if the k-th call took (may be NULL). */
int  cabm_infect_batch(cabm_t* h, const uint32_t* pids, const uint8_t* strains, size_t n, int sim_day, uint8_t* out_infected);
/* run `n_days` of tick+import back to back without returning to the host (CUDA-graph replay) */
int  cabm_run_days(cabm_t* h, int first_day, int n_days, const int32_t* julian_day, const int32_t* is_weekday,
                   const double* expected_imports, const double* strain_w /* [n_days][4] */);

/* MANY COMMUNITIES, ONE LAUNCH PER KERNEL. Calibration runs thousands of replicates of one small population with different
   parameters and seeds (exp/abc-*, SLURM arrays; BASELINE configs[4]); a 100k-agent day is nine launches of a few
   microseconds each, so replicates on separate streams are bound by the device's launch rate. cabm_batch_run_days advances
   `n` handles of the same shape (same N, locations and rosters; own parameters, seed, state and ledger) together: each of
   the day's kernels is ONE launch whose blockIdx.y selects the community. Same arguments as cabm_run_days, shared by all
   handles; results are bit-identical to calling cabm_run_days on each handle. Handles must live on one device and must not
   be sharded. */
int  cabm_batch_run_days(cabm_t* const* handles, size_t n, int first_day, int n_days, const int32_t* julian_day,
                         const int32_t* is_weekday, const double* expected_imports, const double* strain_w);

/* replaces Person::vaccinate via Vac_Campaign::vaccinate (Person.cpp:611-647, Vac_Campaign.h:550) */
int  cabm_vaccinate(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint8_t* out_ok);
/* replaces Person::scheduleQuarantine (Person.cpp:649-651) */
int  cabm_quarantine(cabm_t* h, const uint32_t* pids, size_t n, int from_day, int until_day);

/* replaces the ledger getters (Community.h:353-375) */
int  cabm_ledger_series(cabm_t* h, int counter, uint64_t* out, size_t n);
/* the row of the day that was closed last is read from pinned host memory (left there by the kernel that closed the day) */
int  cabm_ledger_row(cabm_t* h, int sim_day, uint64_t out[CABM_N_COUNTERS]);
/* all series at once, counter-major: out[counter * run_length + day] (the 15+4 by-value getters of simulator.h:376-390,450-452) */
int  cabm_ledger_table(cabm_t* h, uint64_t* out, size_t n);
int  cabm_cumulative_outcomes(cabm_t* h, uint64_t out[CABM_N_OUTCOMES]);
int  cabm_work_counters(cabm_t* h, uint64_t out[CABM_N_WORK]);

/* lazy D2H of every infection record (Person::getInfectionHistory, Person.h:281) */
int  cabm_sync_histories(cabm_t* h, cabm_infection* out, size_t cap, size_t* out_n);
/* the same, starting at record `first` of the append-only infection log (incremental refresh of host-side proxies) */
int  cabm_sync_histories_from(cabm_t* h, size_t first, cabm_infection* out, size_t cap, size_t* out_n);
/* current (latest) infection of selected people; pids NULL => all */
int  cabm_get_infections(cabm_t* h, const uint32_t* pids, size_t n, cabm_infection* out);

/* checkpoint / restore (SimulationCache + Community::load_from_cache, behavior_tuner.h:39-87, Community.cpp:154-171) */
int  cabm_snapshot(cabm_t* h, cabm_snap_t** out);
int  cabm_restore(cabm_t* h, const cabm_snap_t* s);
int  cabm_snap_free(cabm_snap_t* s);

/* parity hook: hazards of the most recent tick */
int  cabm_dump_foi(cabm_t* h, cabm_foi_row* out, size_t cap, size_t* out_n);
/* the primary cases of traceForwardContacts (Community.cpp:1062-1070): living people whose latest infection is reported on
   sim_day, ascending ids (the reference walks _people in id order). The depth-limited contact search itself is sequential
   host logic over the static household / neighbour / day-location graph and stays with the caller. */
int  cabm_detected_today(cabm_t* h, int sim_day, uint32_t* out_pids, size_t cap, size_t* out_n);

/* caller-side daily diagnostics as device reductions: calculate_vax_stats (Community.cpp:1329-1423), doSerosurvey (:1224-1274),
   getNumPeopleQuarantining (Community.h:437), getNumInfected/Infectious/Symptomatic/Naive (Community.cpp:1196-1222) */
int  cabm_diagnostics(cabm_t* h, int sim_day, uint64_t out[CABM_N_DIAG]);
/* doSerosurvey(NATURAL, pop, day) for an arbitrary list of people (pids == NULL: everybody); returns the seropositive count */
int  cabm_serosurvey(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint64_t* out_seropos);

/* What Vac_Campaign asks of a person before scheduling a dose (Person::isAlive, isSeroEligible Person.cpp:562-583, isInfEligible
   :586-608, getNumVaccinations, Vac_Campaign.h:505-548), one byte per listed person (pids == NULL: everybody, n = n_people):
   CABM_EL_* bits. PRIOR_CASE looks at the person's latest infection (the device keeps one record per person; earlier ones are
   in the infection log, cabm_sync_histories). */
enum { CABM_EL_ALIVE = 1, CABM_EL_EVER_INFECTED = 2, CABM_EL_VACCINATED = 4, CABM_EL_PRIOR_CASE = 8,
       CABM_EL_DOSES_SHIFT = 4 /* bits 4-6: doses received, saturating at 7 */, CABM_EL_INFECTED_NOW = 128 };
int  cabm_eligibility(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint8_t* out_flags);
/* Counts over a list of people (pids == NULL: everybody) of one per-person predicate on sim_day: the population reductions of
   Community.cpp:1196-1274,1329-1423 and Community.h:437 for an arbitrary cohort (cabm_diagnostics gives them for everybody). */
enum cabm_reduce_what { CABM_R_SEROPOSITIVE = 0, CABM_R_QUARANTINING, CABM_R_INFECTED, CABM_R_INFECTIOUS, CABM_R_SYMPTOMATIC, CABM_R_NAIVE,
                        CABM_R_FULLY_VAXD, CABM_R_ALIVE, CABM_N_REDUCE };
int  cabm_reduce(cabm_t* h, int what, int sim_day, const uint32_t* pids, size_t n, double* out_count);

/* cabm_serosurvey for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that
   forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE. Same arguments and
   result as cabm_serosurvey, but the id list and the accumulator live in a file-scope `static __device__` array of the
   library (4 MiB + 256 B per GPU that loads the library, held until the process exits), so the call makes no cudaMalloc,
   cudaFree or allocating library call. Limit: n <= CABM_STATIC_MAX_IDS when pids != NULL (more: CABM_E_ARG with a message;
   pids == NULL, everybody, has no limit). Every call on a device shares the one array: it still holds the previous call's
   ids (the accumulator is cleared by the call), and two calls that overlap on the same GPU -- from different handles,
   streams or threads -- race; serialise them. */
#define CABM_STATIC_MAX_IDS (1u << 20)
int  cabm_serosurvey_static(cabm_t* h, int sim_day, const uint32_t* pids, size_t n, uint64_t* out_seropos);

/* ---- one community advanced by several handles, one per GPU -------------------------------------
   Every handle loads the whole population (static arrays and location state are replicated); handle `rank` OWNS the
   agents [pid_lo, pid_hi) -- it alone scans, exposes, infects and tallies them. Household members must share an owner.
   The day is advanced in four parts with three exchanges between them, done by the caller (NCCL / torch.distributed /
   plain copies between handles of one process):
     part 0  owned infectious agents -> flag records {place, agent, pressure, strain}
             exchange A: every handle receives every other handle's records (all-gather; cabm_shard_flags_out / _in)
     part 1  owned agents' visits (Community::public_activity) -> partial visitor census
             exchange B: the census buffer is summed over handles in place (all-reduce; cabm_shard_census)
     part 2  exposure passes HOUSE..NURSINGHOME, winners, status pass
             exchange C: the admissions-minus-discharges buffer is summed over handles in place (cabm_shard_patients)
     part 3  hospital pass, the day's importation (strain_w == NULL: none), end of day
   The exposure draw is keyed (seed, day, person, place), so the owner of an agent draws for it wherever the place's
   infectious people live: no exposure events are exchanged. Results are bit-identical to a single handle.
   Buffers are in the backend's memory (device memory for the CUDA library). cabm_infect must be called on every handle
   (each counts the call, the owner applies it); cabm_vaccinate / cabm_quarantine go to the owner of each agent;
   ledgers, outcome counts, histories and work counters are per handle and add up. */
int  cabm_shard_config(cabm_t* h, int rank, int world, uint32_t pid_lo, uint32_t pid_hi);
int  cabm_shard_day_part(cabm_t* h, int part, int sim_day, int julian_day, int is_weekday, double expected_imports,
                         const double* strain_w, int32_t* out_imported);
int  cabm_shard_flags_out(cabm_t* h, void** records, size_t* n);          /* CABM_SHARD_FLAG_BYTES each; valid until the next part 0 */
int  cabm_shard_flags_in(cabm_t* h, const void* records, size_t n);       /* between part 0 and part 1; may be called once per peer */
int  cabm_shard_census(cabm_t* h, void** buf, size_t* n_u32);
int  cabm_shard_patients(cabm_t* h, void** buf, size_t* n_i32);
#define CABM_SHARD_FLAG_BYTES 24

/* The same protocol driven inside the library over NCCL (one handle per process and GPU): no host round trip inside a batch
   of days. Rank 0 creates an id (cabm_shard_nccl_id), the launcher hands it to every rank by any transport, every rank
   calls cabm_shard_nccl_init (collective) after cabm_shard_config, then cabm_shard_run_days replaces cabm_run_days.
   Exchange A is an all-gather of the counts plus an all-gather of fixed-capacity record blocks; days are enqueued four at
   a time and between groups the host sizes the blocks at 8x the largest count it last saw (floor 8192 records, or the floor
   set with cabm_shard_exchange_capacity). A day that overflows the block fails the call (CABM_E_NOMEM): raise the floor and
   re-run from a snapshot. B and C are in-place ncclAllReduce.
   Where the GPUs of the job can map each other's memory (cudaIpc*, same node, all devices visible to every process: do not
   narrow CUDA_VISIBLE_DEVICES per rank) exchange A is instead one kernel that reads every peer's flag list in place over
   NVLink, exactly as many records as that peer counted: no capacity, no day groups. CABM_SHARD_PEER=0 forces the gather. NCCL is dlopen()ed (libnccl.so.2) at init; libraries without it return CABM_E_STATE. */
#define CABM_NCCL_ID_BYTES 128
#define CABM_MAX_SHARDS 16
int  cabm_shard_nccl_id(void* out_id);
int  cabm_shard_nccl_init(cabm_t* h, const void* id, int rank, int world);
int  cabm_shard_exchange_capacity(cabm_t* h, size_t records_per_rank);
int  cabm_shard_exchange_mode(cabm_t* h, int* out_peer_memory);     /* 1: exchange A reads peer memory, 0: NCCL gather */
int  cabm_shard_run_days(cabm_t* h, int first_day, int n_days, const int32_t* julian_day, const int32_t* is_weekday,
                         const double* expected_imports, const double* strain_w);

/* device time (ms, CUDA events) spent in the day pipeline since the last call; resets the accumulator */
int  cabm_device_time_ms(cabm_t* h, double* out_ms, uint64_t* out_launches);

/* per-kernel CUDA-event timing of the day pipeline (measurement only; no reference counterpart) */
int  cabm_set_profile(cabm_t* h, int on);
int  cabm_kernel_times(cabm_t* h, double* out_ms, uint64_t* out_count, const char** out_names, size_t cap, size_t* out_n);

#ifdef __cplusplus
}
#endif
#endif
