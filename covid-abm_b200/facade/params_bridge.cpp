// covid-abm_b200/facade/params_bridge.cpp -- reference `Parameters` -> flat `cabm_params` (include/cabm.h).
//
// Everything the device path reads from `Parameters` is copied field by field; the discretised
// duration draws of Person::infect (Parameters.h:515-532, 450-461, 547-553; Person.cpp:340-355; Community.cpp:1013)
// become integer-support CDF tables, the same construction as covid-abm_b200/params.py build_tables(): a draw is one
// table lookup on one 32-bit Philox word. The CDFs are evaluated here with textbook series (regularised incomplete
// gamma by series / continued fraction, binomial / negative binomial / Poisson by summing the pmf via lgamma).
#include <cmath>
#include <cstring>
#include <vector>
#include "Parameters.h"
#include "../../include/cabm.h"

namespace {

// regularised lower incomplete gamma P(a, x)
double gamma_p(double a, double x) {
    if (x <= 0.0) return 0.0;
    const double lg = std::lgamma(a);
    if (x < a + 1.0) {                                   // series
        double ap = a, sum = 1.0 / a, del = sum;
        for (int n = 0; n < 2000; ++n) { ap += 1.0; del *= x / ap; sum += del; if (std::fabs(del) < std::fabs(sum) * 1e-17) break; }
        return sum * std::exp(-x + a * std::log(x) - lg);
    }
    double b = x + 1.0 - a, c = 1.0 / 1e-300, d = 1.0 / b, h = d;   // Lentz continued fraction for Q(a, x)
    for (int i = 1; i < 2000; ++i) {
        const double an = -i * (i - a);
        b += 2.0;
        d = an * d + b; if (std::fabs(d) < 1e-300) d = 1e-300;
        c = b + an / c; if (std::fabs(c) < 1e-300) c = 1e-300;
        d = 1.0 / d;
        const double del = d * c; h *= del;
        if (std::fabs(del - 1.0) < 1e-17) break;
    }
    return 1.0 - std::exp(-x + a * std::log(x) - lg) * h;
}

// P(floor(Gamma(shape a, scale b)) <= k) = F(k + 1), k = 0..n-1, last entry forced to 1
void gamma_floor_cdf(double a, double b, double* out, int n) {
    for (int k = 0; k < n; ++k) out[k] = gamma_p(a, (k + 1.0) / b);
    out[n - 1] = 1.0;
}

void cumulate(const std::vector<double>& pmf, double* out, int n) {
    double s = 0.0;
    for (int k = 0; k < n; ++k) { s += k < (int) pmf.size() ? pmf[k] : 0.0; out[k] = s > 1.0 ? 1.0 : s; }
    out[n - 1] = 1.0;
}

std::vector<double> negbin_pmf(double p, double n, int len) {           // gsl_ran_negative_binomial(p, n): failures before n successes
    std::vector<double> v(len);
    for (int k = 0; k < len; ++k) v[k] = std::exp(std::lgamma(k + n) - std::lgamma(k + 1.0) - std::lgamma(n) + n * std::log(p) + k * std::log1p(-p));
    return v;
}
std::vector<double> poisson_pmf(double mu, int len) {
    std::vector<double> v(len);
    for (int k = 0; k < len; ++k) v[k] = mu > 0 ? std::exp(-mu + k * std::log(mu) - std::lgamma(k + 1.0)) : (k == 0 ? 1.0 : 0.0);
    return v;
}
std::vector<double> binomial_pmf(double p, int n) {
    std::vector<double> v(n + 1);
    for (int k = 0; k <= n; ++k)
        v[k] = std::exp(std::lgamma(n + 1.0) - std::lgamma(k + 1.0) - std::lgamma(n - k + 1.0) + k * std::log(p) + (n - k) * std::log1p(-p));
    return v;
}

template <typename T, size_t N>
void copy_vec(T (&dst)[N], const std::vector<T>& src, T fill) { for (size_t i = 0; i < N; ++i) dst[i] = i < src.size() ? src[i] : (src.empty() ? fill : src.back()); }

void copy_ve(double (&dst)[CABM_N_STRAINS][CABM_MAX_DOSES], const std::map<StrainType, std::vector<double>>& m, const Parameters* par) {
    for (int s = 0; s < CABM_N_STRAINS; ++s) for (int d = 0; d < CABM_MAX_DOSES; ++d) {
        auto it = m.find((StrainType) s);
        dst[s][d] = (it == m.end() || it->second.empty()) ? 0.0 : par->stretchy_vector(it->second, d);     // Parameters.h:496-500
    }
}

} // namespace

// pfd / imr receive the per-day tables the struct points to; they must outlive the cabm_set_params call
void cabm_facade_fill_params(const Parameters* par, cabm_params* out, std::vector<double>& pfd, std::vector<double>& imr) {
    std::memset(out, 0, sizeof(*out));
    cabm_params& q = *out;
    q.seed = (uint64_t) par->randomseed;
    q.run_length = (int32_t) par->runLength;
    q.num_surveilled_people = (int32_t) std::min<long long>((long long) par->numSurveilledPeople, 2147483647LL);
    q.haz_mult[CABM_HOUSE] = par->household_transmission_haz_mult; q.haz_mult[CABM_WORK] = par->workplace_transmission_haz_mult;
    q.haz_mult[CABM_SCHOOL] = par->school_transmission_haz_mult;   q.haz_mult[CABM_HOSPITAL] = par->hospital_transmission_haz_mult;
    q.haz_mult[CABM_NURSINGHOME] = par->nursinghome_transmission_haz_mult; q.haz_mult[CABM_N_LOCTYPES] = par->social_transmission_haz_mult;
    for (int d = 0; d < 366; ++d) q.seasonality[d] = d < (int) par->seasonality.size() ? par->seasonality[d] : 1.0;
    copy_vec(q.susceptibility_by_age, par->susceptibilityByAge, 1.0f);
    copy_vec(q.pathogenicity_by_age, par->pathogenicityByAge, 1.0f);
    for (int c = 0; c < 2; ++c) {
        copy_vec(q.prob_severe[c], par->probSeriousOutcome.at(SEVERE)[c], 0.0f);
        copy_vec(q.prob_critical[c], par->probSeriousOutcome.at(CRITICAL)[c], 0.0f);
        copy_vec(q.prob_icu_death[c], par->probSeriousOutcome.at(DEATH)[c], 0.0f);
    }
    copy_ve(q.VES, par->VES, par); copy_ve(q.VES_NAIVE, par->VES_NAIVE, par); copy_ve(q.VEP, par->VEP, par);
    copy_ve(q.VEH, par->VEH, par); copy_ve(q.VEF, par->VEF, par); copy_ve(q.VEI, par->VEI, par);
    q.IEP = par->IEP; q.IEH = par->IEH; q.IEF = par->IEF; q.IEI = par->IEI;
    q.vaccine_dose_to_protection_lag = par->vaccine_dose_to_protection_lag;
    q.immunity_leaky = par->immunityLeaky; q.immunity_wanes = par->immunityWanes; q.retroactive_mature_vaccine = par->retroactiveMatureVaccine;
    for (int s = 0; s < CABM_N_STRAINS; ++s) {
        const StrainPars& sp = par->strainPars[s];
        q.strain_rel_infectiousness[s] = sp.relInfectiousness; q.strain_rel_pathogenicity[s] = sp.relPathogenicity;
        q.strain_rel_severity[s] = sp.relSeverity; q.strain_rel_mortality[s] = sp.relMortality; q.strain_rel_icu_mortality[s] = sp.relIcuMortality;
        q.strain_immune_escape[s] = sp.immuneEscapeProb; q.strain_infectious_period[s] = (int32_t) sp.symptomaticInfectiousPeriod;
        for (int t = 0; t < CABM_N_STRAINS; ++t)
            q.cross_protection[s][t] = (s < (int) par->crossProtectionMatrix.size() && t < (int) par->crossProtectionMatrix[s].size() && par->crossProtectionMatrix[s][t]) ? 1 : 0;
    }
    q.symptom_to_test_lag = (int32_t) par->symptomToTestLag; q.default_reporting_lag = (int32_t) par->defaultReportingLag;
    q.begin_contact_tracing = par->beginContactTracing; q.contact_tracing_depth = (int32_t) par->contactTracingDepth;
    q.quarantine_duration = par->quarantineDuration; q.contact_tracing_coverage = par->contactTracingCoverage;
    for (int k = 0; k < 4; ++k) q.quarantine_probability[k] = k < (int) par->quarantineProbability.size() ? par->quarantineProbability[k] : 0.0;
    q.seroconversion_lag = par->seroconversionLag; q.seropositivity_threshold = par->seropositivityThreshold;

    // ---- integer-support tables ----
    for (int s = 0; s < CABM_N_STRAINS; ++s)                                                            // Parameters.h:515-520
        gamma_floor_cdf(WILDTYPE_SYMPTOM_ONSET_GAMMA_SHAPE, par->strainPars[s].relSymptomOnset * WILDTYPE_SYMPTOM_ONSET_GAMMA_SCALE, q.incub_cdf[s], CABM_INCUB_MAX);
    for (int n = 0; n <= CABM_INCUB_MAX; ++n) {                                                         // Parameters.h:521-524, n = incubation - 1
        cumulate(binomial_pmf(INFECTIOUSNESS_ONSET_FRACTION, n), q.onset_cdf[n], CABM_INCUB_MAX + 1);
        for (int k = n; k <= CABM_INCUB_MAX; ++k) q.onset_cdf[n][k] = 1.0;
    }
    gamma_floor_cdf(PRE_SEVERE_SYMPTOMATIC_GAMMA_SHAPE, PRE_SEVERE_SYMPTOMATIC_GAMMA_SCALE, q.presev_cdf, CABM_T_GAMMA);
    gamma_floor_cdf(WILDTYPE_SYMPTOM_DURATION_MILD_GAMMA_SHAPE, WILDTYPE_SYMPTOM_DURATION_MILD_GAMMA_SCALE, q.mild_cdf, CABM_T_GAMMA);
    for (int k = 0; k < CABM_N_SEVERE_CDF; ++k) q.severe_only_cdf[k] = k < (int) SEVERE_ONLY_DURATION_CDF.size() ? SEVERE_ONLY_DURATION_CDF[k] : 1.0;
    for (int k = 0; k < CABM_N_CRITICAL_CDF; ++k) q.critical_cdf[k] = k < (int) CRITICAL_DURATION_CDF.size() ? CRITICAL_DURATION_CDF[k] : 1.0;
    q.p_high_asym = std::exp(-std::pow(8.04 / 6.72, 3.33));                                              // Person.cpp:340-355
    q.p_high_sym  = std::exp(-std::pow(8.04 / 7.40, 3.81));
    cumulate(negbin_pmf(0.2, 2.0, CABM_T_NEGBIN), q.icu_death_cdf, CABM_T_NEGBIN);                       // Parameters.h:450-456
    for (int k = 0; k < CABM_T_GEOM; ++k) q.community_death_cdf[k] = 1.0 - std::pow(0.5, k + 1.0);       // Parameters.h:458-461, support 1..
    q.community_death_cdf[CABM_T_GEOM - 1] = 1.0;
    cumulate(negbin_pmf(0.1578952, 2.2786932, CABM_T_DEATHLAG), q.death_lag_cdf, CABM_T_DEATHLAG);       // Parameters.h:547-553
    cumulate(poisson_pmf(1.0, CABM_T_POISSON), q.visit_cdf, CABM_T_POISSON);                             // Community.cpp:1013
    for (int t = 0; t < CABM_N_LOCTYPES; ++t)
        cumulate(poisson_pmf(t < (int) par->contactTracingEV.size() ? par->contactTracingEV[t] : 0.0, CABM_T_POISSON), q.trace_ev_cdf[t], CABM_T_POISSON);

    pfd.assign((size_t) q.run_length * CABM_N_OUTCOMES, 0.0);
    for (int d = 0; d < q.run_length; ++d) for (int o = 0; o < CABM_N_OUTCOMES; ++o) {
        const std::vector<std::vector<double>>& m = par->probFirstDetection;
        if (!m.empty()) pfd[(size_t) d * CABM_N_OUTCOMES + o] = m[std::min<size_t>(d, m.size() - 1)][o];
    }
    imr.assign((size_t) q.run_length, 0.0);
    for (int d = 0; d < q.run_length; ++d) if (!par->icuMortalityReduction.empty())
        imr[d] = (size_t) d < par->icuMortalityReduction.size() ? par->icuMortalityReduction[d] : par->icuMortalityReduction.back();
    q.prob_first_detection = pfd.data();
    q.icu_mortality_reduction = imr.data();
}
