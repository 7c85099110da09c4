"""Model parameters for the daily transmission cycle, as flat tables.

This is the host-side mirror of the reference's `Parameters` members that the hot path reads
(Parameters.h:465-698). `default_params()` restates the parameter set the
active-vac driver builds (exp/active-vac/main.cpp:68-323) for a synthetic run; `build_tables()`
turns every continuous-then-discretised duration draw of Person::infect (Parameters.h:515-532,
450-461,547-553; Person.cpp:340-355) into an integer-support CDF table, so that CPU oracle, GSL-shim
tape mode and the CUDA kernels all resolve a draw with the same table lookup on one 32-bit value.

The same dict is (1) packed into the C struct `cabm_params` for the C-ABI, (2) written as a text
file for the reference driver under oracle/, (3) handed to the oracle.
"""
from __future__ import annotations

import math
import numpy as np

N_STRAINS = 4          # Parameters.h:168-174
N_AGES = 121           # Parameters.h:315
MAX_DOSES = 4
N_OUTCOMES = 5         # Parameters.h:127-134
INCUB_MAX = 64         # support cap of 1+floor(gamma) tables
T_GAMMA = 128
T_NEGBIN_ICU = 128
T_GEOM = 64
T_DEATHLAG = 256
T_POISSON = 32

WILDTYPE, ALPHA, DELTA, OMICRON = range(4)

# Parameters.h:291-299
SEVERE_ONLY_DURATION_CDF = [
    0.037435, 0.139665, 0.256730, 0.374955, 0.491195, 0.594245, 0.683795, 0.757310, 0.812225, 0.855110,
    0.886960, 0.909490, 0.925635, 0.935870, 0.943265, 0.949685, 0.954360, 0.958955, 0.963360, 0.967855,
    0.971540, 0.975125, 0.978495, 0.981425, 0.984275, 0.986235, 0.988115, 0.989515, 0.991115, 0.992335,
    0.993585, 0.994790, 0.995595, 0.996245, 0.996965, 0.997490, 0.997980, 0.998385, 0.998670, 0.998810,
    0.999055, 0.999170, 0.999255, 0.999305, 0.999335, 0.999420, 0.999545, 0.999585, 0.999585, 0.999645,
    0.999660, 0.999740, 0.999830, 0.999905, 0.999920, 0.999935, 0.999960, 0.999970, 0.999980, 0.999985,
    0.999985, 0.999985, 0.999985, 0.999985, 0.999985, 0.999985, 0.999985, 0.999985, 0.999985, 0.999985,
    1.000000]
# Parameters.h:303-309
CRITICAL_DURATION_CDF = [
    0.02395, 0.08514, 0.15554, 0.23370, 0.31259, 0.39308, 0.46988, 0.54266, 0.60962, 0.67088,
    0.72574, 0.77359, 0.81612, 0.85151, 0.88142, 0.90602, 0.92574, 0.94139, 0.95484, 0.96481,
    0.97280, 0.97898, 0.98391, 0.98735, 0.99029, 0.99231, 0.99374, 0.99496, 0.99584, 0.99670,
    0.99723, 0.99767, 0.99810, 0.99848, 0.99869, 0.99893, 0.99915, 0.99924, 0.99939, 0.99947,
    0.99955, 0.99962, 0.99969, 0.99977, 0.99980, 0.99984, 0.99986, 0.99987, 0.99990, 0.99990,
    0.99993, 0.99995, 0.99995, 0.99995, 0.99995, 0.99995, 0.99996, 0.99997, 0.99998, 1.00000]

f32 = np.float32


def _days_in_year(y):
    return 366 if (y % 4 == 0 and (y % 100 != 0 or y % 400 == 0)) else 365


_COMMON_END = [31, 59, 90, 120, 151, 181, 212, 243, 273, 304, 334, 365]
_LEAP_END = [31, 60, 91, 121, 152, 182, 213, 244, 274, 305, 335, 366]


def to_julian_day(ymd: str) -> int:
    """Date.h:260-274"""
    y, m, d = (int(x) for x in ymd.split("-"))
    ends = _LEAP_END if _days_in_year(y) == 366 else _COMMON_END
    return d if m == 1 else ends[m - 2] + d


def to_sim_day(start_year: int, start_day: int, ymd: str) -> int:
    """Date.h:303-315"""
    y = int(ymd.split("-")[0])
    tally = sum(_days_in_year(yy) for yy in range(start_year, y + 1))
    tally -= start_day
    tally -= _days_in_year(y) - to_julian_day(ymd)
    return tally


def calendar(start_year: int, start_day: int, run_length: int):
    """(julian_day[run_length], is_weekday[run_length]) following Date.h:30-180 (Conway doomsday, Monday=0)."""
    jd, jy = start_day, start_year
    out_j, out_w = [], []
    t = [6, 2, 1, 4, 6, 2, 4, 0, 3, 5, 1, 3]
    for _ in range(run_length):
        ends = _LEAP_END if _days_in_year(jy) == 366 else _COMMON_END
        m = next(i for i, e in enumerate(ends) if e >= jd) + 1
        dom = jd if m == 1 else jd - ends[m - 2]
        y = jy - (1 if m < 3 else 0)
        dow = (y + y // 4 - y // 100 + y // 400 + t[m - 1] + dom) % 7
        out_j.append(jd); out_w.append(0 if dow >= 5 else 1)
        jd += 1
        if jd > _days_in_year(jy):
            jd = 1; jy += 1
    return np.array(out_j, dtype=np.int32), np.array(out_w, dtype=np.int32)


def lin_interpolate_series(anchors, start_year, start_day):
    """Date::linInterpolateTimeSeries (Date.h:329-360)."""
    ap = sorted(anchors, key=lambda a: a[0])
    v = []
    for (da, va), (db, vb) in zip(ap[:-1], ap[1:]):
        n = to_sim_day(start_year, start_day, db) - to_sim_day(start_year, start_day, da)
        step = (vb - va) / n
        v.extend([va] + [va + i * step for i in range(1, n)])
    v.append(ap[-1][1])
    s0 = to_sim_day(start_year, start_day, ap[0][0])
    if s0 < 0:
        v = v[-s0:]
    return v


def _logistic(x):
    return 1.0 / (1.0 + math.exp(-x))


def _resize(vec, n, val):
    """std::vector::resize(n, val) semantics."""
    return vec[:n] + [val] * max(0, n - len(vec))


def age_tables(icu_mortality_fraction: float):
    """Parameters::define_susceptibility_and_pathogenicity (Parameters.cpp:128-195); float arithmetic as there."""
    def expand(bins, vals):
        out = []
        for up, v in zip(bins, vals):
            out = _resize(out, up + 1, v)
        return np.array(out, dtype=np.float32)
    b8 = [9, 19, 29, 39, 49, 59, 69, N_AGES - 1]
    susc = expand(b8, [f32(x) for x in [0.40, 0.38, 0.79, 0.86, 0.80, 0.82, 0.88, 0.74]])
    path = expand(b8, [f32(x) for x in [0.29, 0.21, 0.27, 0.33, 0.40, 0.49, 0.63, 0.69]])
    b9 = [9, 19, 29, 39, 49, 59, 69, 79, N_AGES - 1]
    sev_neg = [f32(x) for x in [0.03689, 0.02279, 0.02688, 0.04445, 0.06441, 0.09570, 0.15355, 0.27867, 0.30095]]
    sev_pos = [f32(x) for x in [0.22294, 0.14884, 0.17505, 0.24209, 0.29597, 0.36328, 0.49908, 0.64716, 0.62338]]
    cri_neg = [f32(x) for x in [0.00703, 0.00337, 0.00302, 0.00725, 0.01267, 0.02053, 0.03675, 0.07110, 0.05189]]
    cri_pos = [f32(x) for x in [0.05008, 0.03468, 0.03369, 0.05298, 0.06374, 0.08276, 0.10870, 0.11933, 0.07465]]
    fat_neg = [f32(x) for x in [0.00088, 0.00079, 0.00130, 0.00113, 0.00353, 0.00908, 0.02361, 0.10218, 0.29805]]
    fat_pos = [f32(x) for x in [0.00646, 0.00771, 0.01370, 0.02767, 0.04458, 0.07837, 0.16704, 0.31670, 0.49668]]
    severe = np.stack([expand(b9, sev_neg), expand(b9, sev_pos)])
    critical = np.stack([expand(b9, [f32(c / s) for c, s in zip(cri_neg, sev_neg)]),
                         expand(b9, [f32(c / s) for c, s in zip(cri_pos, sev_pos)])])
    b6 = [9, 19, 29, 39, 49, N_AGES - 1]
    imf = float(icu_mortality_fraction)
    # icuMortalityFraction is a double: double*float/float evaluated in double, then stored as float
    death = np.stack([expand(b6, [f32(imf * float(f) / float(c)) for f, c in zip(fat_neg, cri_neg)]),
                      expand(b6, [f32(imf * float(f) / float(c)) for f, c in zip(fat_pos, cri_pos)])])
    return susc, path, severe, critical, death


def detection_model(vals, inflection_days, slopes, run_length):
    """Parameters::createDetectionModel (Parameters.cpp:217-247)."""
    out = np.zeros((run_length, N_OUTCOMES))
    for d in range(run_length):
        for o in range(N_OUTCOMES):
            p = vals[0][o]
            for i in range(1, len(vals)):
                p += (vals[i][o] - vals[i - 1][o]) * _logistic(slopes[i - 1][o] * (float(d) - inflection_days[i - 1][o]))
            out[d, o] = p
    return out


def default_params(run_length: int = 200, seed: int = 1, T: float = 0.025, n_people: int | None = None,
                   mutant_intro_dates=(), prob_daily_exposure: float | None = None,
                   seasonality_amplitude: float = 0.15) -> dict:
    """Parameter set of exp/active-vac/main.cpp:68-323 (define_simulator_parameters + define_strain_parameters
    + the settings applied in simulator() at :598-605), without a reporting-lag csv (not shipped) so
    defaultReportingLag applies."""
    sy, sd = 2020, to_julian_day("2020-02-10")                      # main.cpp:89-90
    Tf = float(f32(T))                                              # `const float T` main.cpp:73
    P = dict(
        seed=int(seed), run_length=int(run_length), start_julian_year=sy, start_day_of_year=sd,
        haz_household=Tf, haz_social=Tf, haz_work=Tf / 2.0, haz_school=Tf, haz_hospital=Tf / 10.0,
        haz_nursinghome=Tf * 2.0,                                   # main.cpp:75-80
        seasonality=[1 - seasonality_amplitude * math.sin((4 * math.pi * (day - 60)) / 366) for day in range(366)],  # :113-117
        IEP=0.5, IEH=0.5, IEF=0.0, IEI=0.0,                         # main.cpp:290-291, Parameters.cpp:37-40
        vaccine_dose_to_protection_lag=10, immunity_leaky=1, immunity_wanes=1,   # main.cpp:598-604
        retroactive_mature_vaccine=0, seroconversion_lag=20.0, seropositivity_threshold=0.0,
        symptom_to_test_lag=2, default_reporting_lag=10, num_surveilled_people=2**31 - 1,
        num_vaccine_doses=3, vaccine_dose_interval=[21, 240], vaccine_inf_constraint=4, vaccine_sero_constraint=2,
        sero_false_pos=0.0, sero_false_neg=0.0,
        begin_contact_tracing=2**31 - 1, contact_tracing_coverage=0.7, contact_tracing_ev=[5.0, 3.0, 3.0, 0.0, 5.0],
        contact_tracing_depth=2, quarantine_probability=[0.0, 0.0, 0.0], quarantine_duration=10,  # main.cpp:101-111
        num_initial_exposed=0, prob_initial_exposure=0.0,
        mutant_intro_dates=list(mutant_intro_dates),
    )
    # strains (main.cpp:293-322)
    rel_inf = [1.0, 1.5, 1.5 * 1.5, 1.5 * 1.5 * 3.0]
    rel_path = [1.0, 1.2, 1.2 * 2.83, 1.2 * 2.83 * 0.5]
    rel_sev = [1.0, 1.0, 1.5, 1.5 * 0.5]
    P.update(
        strain_rel_infectiousness=rel_inf, strain_rel_pathogenicity=rel_path, strain_rel_severity=rel_sev,
        strain_rel_mortality=[1.0] * 4, strain_rel_icu_mortality=[1.0, 1.0, 3.25, 1.0],
        strain_immune_escape=[0.0, 0.15, 0.2, 0.5], strain_infectious_period=[7] * 4,
        strain_rel_symptom_onset=[1.0, 1.0, 1.0, 0.5],
        cross_protection=np.eye(4, dtype=np.int32),
        VES=[[0.40, 0.80]] * 4, VES_NAIVE=[[0.40, 0.80]] * 4,
        VEP=[[0.67, 0.75], [0.67, 0.75], [0.6, 0.7], [0.6, 0.7]],
        VEH=[[0.9, 1.0], [0.9, 1.0], [0.9, 0.93], [0.48, 0.96]],
        VEI=[[0.4, 0.8], [0.4, 0.8], [0.4, 0.8], [0.2, 0.4]],
        VEF=[[0.0, 0.0]] * 4,
    )
    # detection model (main.cpp:121-147, exp/lib/exp_util.h:4-27)
    fdp = [[0.0, 0.25, 0.7, 0.1], [0.1, 0.5, 0.5, 0.1], [0.3, 0.9, 0.5, 0.1], [0.1, 0.5, 0.9, 0.1],
           [0.1, 0.5, 0.9, 0.1], [0.03, 0.2, 0.95, 0.1]]
    for v in fdp:
        v.append(1.0 - (1.0 - 1.0) / ((1.0 - v[0]) * (1.0 - v[1]) * (1.0 - v[2]) * (1.0 - v[3])))
    infl = ["2020-06-01", "2020-10-01", "2021-02-15", "2021-10-01", "2022-01-19"]
    infl_m = [[to_sim_day(sy, sd, d)] * N_OUTCOMES for d in infl]
    slopes = [[0.1] * N_OUTCOMES for _ in range(len(infl) - 1)] + [[0.5] * N_OUTCOMES]
    P["prob_first_detection"] = detection_model(fdp, infl_m, slopes, run_length)
    # timed interventions (main.cpp:162-181) -- resize() semantics
    sc = _resize([], to_sim_day(sy, sd, "2020-03-15"), 0.0)
    sc = _resize(sc, min(to_sim_day(sy, sd, "2020-08-31"), run_length), 1.0)
    sc = _resize(sc, min(to_sim_day(sy, sd, "2021-06-16"), run_length), 0.5)
    sc = _resize(sc, min(to_sim_day(sy, sd, "2021-08-09"), run_length), 1.0)
    sc = _resize(sc, run_length, 0.0)
    nb = _resize([], to_sim_day(sy, sd, "2020-04-03"), 0.0)
    nb = _resize(nb, to_sim_day(sy, sd, "2020-05-04"), 1.0)
    nb = _resize(nb, run_length, 0.0)
    # social distancing anchors of simulate_epidemic (simulator.h:284-308)
    anchors = [("2020-01-01", 0.0), ("2020-03-10", 0.20), ("2020-03-15", 0.8), ("2020-04-01", 0.7), ("2020-05-01", 0.6),
               ("2020-06-01", 0.05), ("2020-07-01", 0.05), ("2020-08-01", 0.6), ("2020-09-01", 0.5), ("2020-10-01", 0.1),
               ("2020-11-01", 0.0), ("2020-12-01", 0.1), ("2021-01-01", 0.1), ("2021-02-01", 0.3), ("2021-03-01", 0.4),
               ("2021-04-01", 0.0), ("2021-05-01", 0.3), ("2021-06-01", 0.2), ("2021-07-01", 0.0), ("2021-08-01", 0.1),
               ("2021-09-01", 0.25), ("2021-10-01", 0.0)]
    sdv = lin_interpolate_series(anchors, sy, sd)
    sdv = _resize(sdv, run_length, sdv[-1])                          # Community.h:407-412
    P["school_closure"] = sc[:run_length]; P["business_closure"] = nb[:run_length]; P["social_distancing"] = sdv[:run_length]
    # ICU mortality reduction (main.cpp:226-229, Parameters.cpp:198-204)
    infl_day = to_sim_day(sy, sd, "2020-06-15")
    P["icu_mortality_reduction"] = [0.35 * _logistic(1.0 * (float(d) - infl_day)) for d in range(run_length)]
    P["icu_mortality_fraction"] = 0.43
    susc, path, sev, crit, death = age_tables(0.43)
    P.update(susceptibility_by_age=susc, pathogenicity_by_age=path, prob_severe=sev, prob_critical=crit, prob_icu_death=death)
    # importation (main.cpp:247-249)
    if prob_daily_exposure is None:
        j15 = to_sim_day(sy, sd, "2021-06-15")
        pde = _resize([1.0e-4] * min(j15, run_length), run_length, 2.0e-4)
    else:
        pde = [float(prob_daily_exposure)] * run_length
    P["prob_daily_exposure"] = pde
    return P


# ------------------------------------------------------------------------------------------------
# integer-support CDF tables
# ------------------------------------------------------------------------------------------------
def _gamma_floor_cdf(a, b, n):
    from scipy.stats import gamma
    c = gamma.cdf(np.arange(1, n + 1, dtype=np.float64), a, scale=b)   # P(floor(G) <= k) = F(k+1)
    c[-1] = 1.0
    return c


def _discrete_cdf(pmf_cdf, n):
    c = np.asarray(pmf_cdf(np.arange(n)), dtype=np.float64)
    c[-1] = 1.0
    return c


def build_tables(P: dict) -> dict:
    from scipy.stats import binom, nbinom, geom, poisson
    # constants are `static const float` in Parameters.h:263-285: float arithmetic, then promoted
    mean = f32(4.91); shape = f32(3.05); scale = f32(mean / shape)
    onset_frac = float(f32(f32(1.0) - f32(f32(2.3) / mean)))                 # Parameters.h:275-276
    T = {}
    T["incub_a"] = float(shape)
    T["incub_b"] = [float(r) * float(scale) for r in P["strain_rel_symptom_onset"]]   # Parameters.h:517
    T["incub_cdf"] = np.stack([_gamma_floor_cdf(float(shape), b, INCUB_MAX) for b in T["incub_b"]])
    T["onset_p"] = onset_frac
    onset = np.ones((INCUB_MAX + 1, INCUB_MAX + 1))
    for n in range(INCUB_MAX + 1):                                         # n = incubation_pd - 1
        onset[n, : n + 1] = binom.cdf(np.arange(n + 1), n, onset_frac); onset[n, n:] = 1.0
    T["onset_cdf"] = onset
    T["presev_a"], T["presev_b"] = float(f32(0.812130)), float(f32(7.199686))           # Parameters.h:284-285
    T["presev_cdf"] = _gamma_floor_cdf(T["presev_a"], T["presev_b"], T_GAMMA)
    T["mild_a"], T["mild_b"] = float(f32(4.070483)), float(f32(2.825217))               # Parameters.h:269-270
    T["mild_cdf"] = _gamma_floor_cdf(T["mild_a"], T["mild_b"], T_GAMMA)
    # Person.cpp:340-355: Weibull(scale, shape) > 8.04
    T["weibull_asym"] = (6.72, 3.33); T["weibull_sym"] = (7.40, 3.81)
    T["p_high_asym"] = math.exp(-((8.04 / 6.72) ** 3.33)); T["p_high_sym"] = math.exp(-((8.04 / 7.40) ** 3.81))
    T["icu_death_cdf"] = _discrete_cdf(lambda k: nbinom.cdf(k, 2, 0.2), T_NEGBIN_ICU)   # Parameters.h:450-456
    T["community_death_cdf"] = _discrete_cdf(lambda k: geom.cdf(k + 1, 0.5), T_GEOM)    # Parameters.h:458-461 (support 1..)
    T["death_lag_n"], T["death_lag_p"] = 2.2786932, 0.1578952                            # Parameters.h:550-552
    T["death_lag_cdf"] = _discrete_cdf(lambda k: nbinom.cdf(k, T["death_lag_n"], T["death_lag_p"]), T_DEATHLAG)
    T["visit_cdf"] = _discrete_cdf(lambda k: poisson.cdf(k, 1.0), T_POISSON)            # Community.cpp:1012-1013
    T["trace_ev_cdf"] = np.stack([
        _discrete_cdf(lambda k, m=m: poisson.cdf(k, m) if m > 0 else np.ones_like(k, dtype=np.float64), T_POISSON)
        for m in P["contact_tracing_ev"]])                                             # Community.cpp:1089,1105,1120
    T["severe_only_cdf"] = np.array(SEVERE_ONLY_DURATION_CDF); T["critical_cdf"] = np.array(CRITICAL_DURATION_CDF)
    return T


def _fmt(v):
    if isinstance(v, (list, tuple, np.ndarray)):
        return " ".join(_fmt(x) for x in np.asarray(v).ravel().tolist())
    if isinstance(v, (float, np.floating)):
        return repr(float(v))
    if isinstance(v, str):
        return v
    return str(int(v))


def write_text(P: dict, T: dict, path: str):
    """key value... lines, read by oracle/ref_driver.cpp (and only by it)."""
    with open(path, "w") as f:
        for k, v in list(P.items()) + [("T_" + k, v) for k, v in T.items()]:
            if isinstance(v, (list, tuple)) and len(v) == 0:
                f.write(f"{k}\n"); continue
            f.write(f"{k} {_fmt(v)}\n")
