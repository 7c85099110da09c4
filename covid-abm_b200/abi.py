"""ctypes mirror of include/cabm.h (struct layouts, prototypes, packing helpers).

Pure host-side plumbing: nothing here computes any part of the transmission cycle.
`bind(lib, prefix)` attaches prototypes to a loaded shared library; the product uses prefix
"cabm_", the test-only oracle library uses "cabm_oracle_".
"""
from __future__ import annotations

import ctypes as C
import numpy as np

N_STRAINS, N_AGES, MAX_DOSES, N_OUTCOMES, N_LOCTYPES = 4, 121, 4, 5, 5
INCUB_MAX, T_GAMMA, T_NEGBIN, T_GEOM, T_DEATHLAG, T_POISSON = 64, 128, 128, 64, 256, 32
N_SEVERE_CDF, N_CRITICAL_CDF = 71, 60

COUNTERS = ["new_infected", "new_symptomatic", "new_severe", "new_critical", "new_dead", "vaccinated_cases",
            "severe_prev", "hosp_inc", "hosp_prev", "icu_inc", "icu_prev", "det_cases_onset", "det_cases_report",
            "det_hosp", "det_deaths_onset", "det_deaths_report", "new_inf_strain0", "new_inf_strain1",
            "new_inf_strain2", "new_inf_strain3"]
N_COUNTERS = len(COUNTERS)
DIAG = ["dose_1", "dose_2", "dose_3", "fully_vaxd", "bt_inf", "bt_dis", "bt_hosp", "bt_death", "bt_reported", "uv_inf", "uv_dis", "uv_hosp",
        "uv_death", "uv_reported", "infected", "infectious", "symptomatic", "naive", "quarantining", "seropos_all", "seropos_ped", "n_ped"]
N_DIAG = 24
EL_ALIVE, EL_EVER_INFECTED, EL_VACCINATED, EL_PRIOR_CASE, EL_DOSES_SHIFT, EL_INFECTED_NOW = 1, 2, 4, 8, 4, 128
REDUCE = ["seropositive", "quarantining", "infected", "infectious", "symptomatic", "naive", "fully_vaxd", "alive"]
STATIC_MAX_IDS = 1 << 20
SHARD_FLAG_BYTES = 24
WORK = ["agents", "active", "flags", "patrons", "visits", "hot_visits", "hot_locs", "nbr_edges", "trials", "infections"]
N_WORK = len(WORK)

d, f, i32, u8 = C.c_double, C.c_float, C.c_int32, C.c_uint8


class Params(C.Structure):
    _fields_ = [
        ("seed", C.c_uint64), ("run_length", i32), ("num_surveilled_people", i32),
        ("haz_mult", d * (N_LOCTYPES + 1)), ("seasonality", d * 366),
        ("susceptibility_by_age", f * N_AGES), ("pathogenicity_by_age", f * N_AGES),
        ("prob_severe", (f * N_AGES) * 2), ("prob_critical", (f * N_AGES) * 2), ("prob_icu_death", (f * N_AGES) * 2),
        ("VES", (d * MAX_DOSES) * N_STRAINS), ("VES_NAIVE", (d * MAX_DOSES) * N_STRAINS), ("VEP", (d * MAX_DOSES) * N_STRAINS),
        ("VEH", (d * MAX_DOSES) * N_STRAINS), ("VEF", (d * MAX_DOSES) * N_STRAINS), ("VEI", (d * MAX_DOSES) * N_STRAINS),
        ("IEP", d), ("IEH", d), ("IEF", d), ("IEI", d),
        ("vaccine_dose_to_protection_lag", i32), ("immunity_leaky", i32), ("immunity_wanes", i32),
        ("retroactive_mature_vaccine", i32),
        ("strain_rel_infectiousness", d * N_STRAINS), ("strain_rel_pathogenicity", d * N_STRAINS),
        ("strain_rel_severity", d * N_STRAINS), ("strain_rel_mortality", d * N_STRAINS),
        ("strain_rel_icu_mortality", d * N_STRAINS), ("strain_immune_escape", d * N_STRAINS),
        ("strain_infectious_period", i32 * N_STRAINS), ("cross_protection", (i32 * N_STRAINS) * N_STRAINS),
        ("symptom_to_test_lag", i32), ("default_reporting_lag", i32),
        ("begin_contact_tracing", i32), ("contact_tracing_depth", i32), ("quarantine_duration", i32),
        ("contact_tracing_coverage", d), ("quarantine_probability", d * 4),
        ("seroconversion_lag", d), ("seropositivity_threshold", d),
        ("incub_cdf", (d * INCUB_MAX) * N_STRAINS), ("onset_cdf", (d * (INCUB_MAX + 1)) * (INCUB_MAX + 1)),
        ("presev_cdf", d * T_GAMMA), ("mild_cdf", d * T_GAMMA),
        ("severe_only_cdf", d * N_SEVERE_CDF), ("critical_cdf", d * N_CRITICAL_CDF),
        ("p_high_asym", d), ("p_high_sym", d),
        ("icu_death_cdf", d * T_NEGBIN), ("community_death_cdf", d * T_GEOM), ("death_lag_cdf", d * T_DEATHLAG),
        ("visit_cdf", d * T_POISSON), ("trace_ev_cdf", (d * T_POISSON) * N_LOCTYPES),
        ("prob_first_detection", C.POINTER(d)), ("icu_mortality_reduction", C.POINTER(d)),
    ]


class Population(C.Structure):
    _fields_ = [
        ("n_people", C.c_uint32), ("n_loc", C.c_uint32),
        ("loc_type", C.POINTER(u8)), ("loc_essential", C.POINTER(u8)), ("loc_pubrisk", C.POINTER(u8)),
        ("loc_risk", C.POINTER(f)), ("loc_hospital", C.POINTER(i32)),
        ("nbr_off", C.POINTER(C.c_int64)), ("nbr", C.POINTER(i32)),
        ("roster_off", C.POINTER(C.c_int64)), ("roster", C.POINTER(i32)),
        ("home", C.POINTER(i32)), ("day", C.POINTER(i32)), ("age", C.POINTER(u8)), ("sex", C.POINTER(u8)),
        ("comorbid", C.POINTER(u8)), ("ltc", C.POINTER(u8)), ("patron", C.POINTER(i32)), ("n_patron", C.c_uint32),
        ("starting_natural_efficacy", C.POINTER(d)), ("cross_protection_probability", C.POINTER(f)),
        ("natural_immunity_duration", C.POINTER(i32)),
    ]


INFECTION_DTYPE = np.dtype([("person", "<u4"), ("infected_by", "<i4"), ("infected_place", "<i4"), ("t", "<i2", (12,)),
                            ("report_day", "<i2"), ("strain", "u1"), ("detected_state", "u1"),
                            ("rel_infectiousness", "<f4")], align=True)
FOI_DTYPE = np.dtype([("loc", "<u4"), ("type", "u1"), ("pad", "u1", (3,)), ("w", "<f8", (4,)), ("size", "<u4"),
                      ("T", "<f8"), ("T_social", "<f8")], align=True)
assert INFECTION_DTYPE.itemsize == 44 and FOI_DTYPE.itemsize == 64

_vp = C.c_void_p
_PROTOS = {
    "create": [C.POINTER(_vp)], "destroy": [_vp], "load_population": [_vp, C.POINTER(Population)],
    "population_save": [C.POINTER(Population), C.c_char_p],
    "load_population_file": [_vp, C.c_char_p, C.POINTER(d), C.POINTER(f), C.POINTER(i32)],
    "set_params": [_vp, C.POINTER(Params)], "set_timed_intervention": [_vp, C.c_int, C.POINTER(d), C.c_size_t],
    "tick": [_vp, C.c_int, C.c_int, C.c_int], "import": [_vp, C.c_int, d, C.POINTER(d), C.POINTER(i32)],
    "infect": [_vp, C.c_uint32, C.c_int, C.c_int, C.POINTER(i32)],
    "infect_batch": [_vp, C.POINTER(C.c_uint32), C.POINTER(u8), C.c_size_t, C.c_int, C.POINTER(u8)],
    "run_days": [_vp, C.c_int, C.c_int, C.POINTER(i32), C.POINTER(i32), C.POINTER(d), C.POINTER(d)],
    "batch_run_days": [C.POINTER(_vp), C.c_size_t, C.c_int, C.c_int, C.POINTER(i32), C.POINTER(i32), C.POINTER(d), C.POINTER(d)],
    "vaccinate": [_vp, C.c_int, C.POINTER(C.c_uint32), C.c_size_t, C.POINTER(u8)],
    "quarantine": [_vp, C.POINTER(C.c_uint32), C.c_size_t, C.c_int, C.c_int],
    "ledger_series": [_vp, C.c_int, C.POINTER(C.c_uint64), C.c_size_t], "ledger_row": [_vp, C.c_int, C.POINTER(C.c_uint64)],
    "cumulative_outcomes": [_vp, C.POINTER(C.c_uint64)], "work_counters": [_vp, C.POINTER(C.c_uint64)],
    "sync_histories": [_vp, _vp, C.c_size_t, C.POINTER(C.c_size_t)], "get_infections": [_vp, C.POINTER(C.c_uint32), C.c_size_t, _vp],
    "snapshot": [_vp, C.POINTER(_vp)], "restore": [_vp, _vp], "snap_free": [_vp],
    "dump_foi": [_vp, _vp, C.c_size_t, C.POINTER(C.c_size_t)], "device_time_ms": [_vp, C.POINTER(d), C.POINTER(C.c_uint64)],
    "sync_histories_from": [_vp, C.c_size_t, _vp, C.c_size_t, C.POINTER(C.c_size_t)],
    "ledger_table": [_vp, C.POINTER(C.c_uint64), C.c_size_t],
    "detected_today": [_vp, C.c_int, C.POINTER(C.c_uint32), C.c_size_t, C.POINTER(C.c_size_t)],
    "shard_config": [_vp, C.c_int, C.c_int, C.c_uint32, C.c_uint32],
    "shard_day_part": [_vp, C.c_int, C.c_int, C.c_int, C.c_int, d, C.POINTER(d), C.POINTER(i32)],
    "shard_flags_out": [_vp, C.POINTER(_vp), C.POINTER(C.c_size_t)], "shard_flags_in": [_vp, _vp, C.c_size_t],
    "shard_census": [_vp, C.POINTER(_vp), C.POINTER(C.c_size_t)], "shard_patients": [_vp, C.POINTER(_vp), C.POINTER(C.c_size_t)],
    "eligibility": [_vp, C.c_int, C.POINTER(C.c_uint32), C.c_size_t, C.POINTER(u8)],
    "reduce": [_vp, C.c_int, C.c_int, C.POINTER(C.c_uint32), C.c_size_t, C.POINTER(d)],
    "diagnostics": [_vp, C.c_int, C.POINTER(C.c_uint64)], "serosurvey": [_vp, C.c_int, C.POINTER(C.c_uint32), C.c_size_t, C.POINTER(C.c_uint64)],
}
_OPTIONAL = {
    "shard_nccl_init": [_vp, _vp, C.c_int, C.c_int], "shard_exchange_capacity": [_vp, C.c_size_t], "shard_exchange_mode": [_vp, C.POINTER(C.c_int)],
    "shard_run_days": [_vp, C.c_int, C.c_int, C.POINTER(i32), C.POINTER(i32), C.POINTER(d), C.POINTER(d)],
    "set_profile": [_vp, C.c_int],
    "serosurvey_static": [_vp, C.c_int, C.POINTER(C.c_uint32), C.c_size_t, C.POINTER(C.c_uint64)],   # no-allocation variant (device scratch is a static array)
    "kernel_times": [_vp, C.POINTER(d), C.POINTER(C.c_uint64), C.POINTER(C.c_char_p), C.c_size_t, C.POINTER(C.c_size_t)],
}
SYMBOLS = sorted(list(_OPTIONAL) + list(_PROTOS) + ["last_error"])


class Bound:
    """Functions of one loaded library, addressed without the prefix."""
    def __init__(self, lib, prefix):
        for name, args in _PROTOS.items():
            fn = getattr(lib, prefix + name); fn.argtypes = args; fn.restype = C.c_int
            setattr(self, "import_" if name == "import" else name, fn)
        self.last_error = getattr(lib, prefix + "last_error"); self.last_error.argtypes = [_vp]; self.last_error.restype = C.c_char_p
        fn = getattr(lib, prefix + "shard_nccl_id", None)
        if fn is not None:
            fn.argtypes = [_vp]; fn.restype = C.c_int
        self.shard_nccl_id = fn
        for name, args in _OPTIONAL.items():            # measurement / NCCL entry points (absent from the oracle)
            fn = getattr(lib, prefix + name, None)
            if fn is not None:
                fn.argtypes = args; fn.restype = C.c_int
            setattr(self, name, fn)


def bind(lib, prefix="cabm_") -> Bound:
    return Bound(lib, prefix)


def _fill(dst, src):
    a = np.ascontiguousarray(np.asarray(src), dtype=np.ctypeslib.as_array(dst).dtype)
    np.ctypeslib.as_array(dst)[...] = a.reshape(np.ctypeslib.as_array(dst).shape)


def _stretch(rows, n=MAX_DOSES):
    """Parameters::stretchy_vector (Parameters.h:496-500) applied ahead of time."""
    return [list(r) + [r[-1]] * (n - len(r)) for r in rows]


def pack_params(P: dict, T: dict):
    """Return (Params, keepalive) for a parameter dict from params.default_params / build_tables."""
    s = Params()
    s.seed = int(P["seed"]); s.run_length = int(P["run_length"]); s.num_surveilled_people = int(min(P["num_surveilled_people"], 2**31 - 1))
    _fill(s.haz_mult, [P["haz_household"], P["haz_work"], P["haz_school"], P["haz_hospital"], P["haz_nursinghome"], P["haz_social"]])
    _fill(s.seasonality, P["seasonality"])
    _fill(s.susceptibility_by_age, P["susceptibility_by_age"]); _fill(s.pathogenicity_by_age, P["pathogenicity_by_age"])
    _fill(s.prob_severe, P["prob_severe"]); _fill(s.prob_critical, P["prob_critical"]); _fill(s.prob_icu_death, P["prob_icu_death"])
    for k in ("VES", "VES_NAIVE", "VEP", "VEH", "VEF", "VEI"):
        _fill(getattr(s, k), _stretch(P[k]))
    s.IEP, s.IEH, s.IEF, s.IEI = P["IEP"], P["IEH"], P["IEF"], P["IEI"]
    s.vaccine_dose_to_protection_lag = P["vaccine_dose_to_protection_lag"]
    s.immunity_leaky, s.immunity_wanes, s.retroactive_mature_vaccine = P["immunity_leaky"], P["immunity_wanes"], P["retroactive_mature_vaccine"]
    for k in ("rel_infectiousness", "rel_pathogenicity", "rel_severity", "rel_mortality", "rel_icu_mortality", "immune_escape", "infectious_period"):
        _fill(getattr(s, "strain_" + k), P["strain_" + k])
    _fill(s.cross_protection, P["cross_protection"])
    s.symptom_to_test_lag, s.default_reporting_lag = P["symptom_to_test_lag"], P["default_reporting_lag"]
    s.begin_contact_tracing = int(min(P["begin_contact_tracing"], 2**31 - 1)); s.contact_tracing_depth = P["contact_tracing_depth"]
    s.quarantine_duration = P["quarantine_duration"]; s.contact_tracing_coverage = P["contact_tracing_coverage"]
    _fill(s.quarantine_probability, list(P["quarantine_probability"]) + [0.0] * (4 - len(P["quarantine_probability"])))
    s.seroconversion_lag, s.seropositivity_threshold = P["seroconversion_lag"], P["seropositivity_threshold"]
    for k in ("incub_cdf", "onset_cdf", "presev_cdf", "mild_cdf", "severe_only_cdf", "critical_cdf", "icu_death_cdf",
              "community_death_cdf", "death_lag_cdf", "visit_cdf", "trace_ev_cdf"):
        _fill(getattr(s, k), T[k])
    s.p_high_asym, s.p_high_sym = T["p_high_asym"], T["p_high_sym"]
    pfd = np.ascontiguousarray(P["prob_first_detection"], dtype=np.float64)
    imr = np.ascontiguousarray(P["icu_mortality_reduction"], dtype=np.float64)
    s.prob_first_detection = pfd.ctypes.data_as(C.POINTER(d)); s.icu_mortality_reduction = imr.ctypes.data_as(C.POINTER(d))
    return s, (pfd, imr)


def immunity_attributes(n_people: int, seed: int):
    """Load-time per-person immunity attributes (Community.cpp:212-215; Parameters.h:598-635)."""
    rng = np.random.default_rng([int(seed), 0xC0F1D])
    neut = np.power(2.0, rng.normal(0.0, 1.0, n_people))
    aesp = 1.0 / (1.0 + np.exp(-(3.097703 * np.log10(neut / 0.2010885))))
    eff = np.maximum(0.0, 1.0 - (1.0 - aesp) / (1.0 - 0.75))
    q = rng.random(n_people)
    with np.errstate(divide="ignore", invalid="ignore"):
        est = lambda x: (np.log(1.0 / x - 1.0) + 2.4603816) / 0.2
        dur = (-1.0 / 0.2) * np.log(q / (1.0 - q)) + (est(0.5) - est(eff))
    dur = np.where(np.isfinite(dur) & (dur > 0), dur, 0.0)
    nat_dur = np.minimum(30 * dur, 2**31 - 1).astype(np.int32)
    xprot = rng.random(n_people).astype(np.float32)
    return eff.astype(np.float64), xprot, nat_dur


def pack_population(pop: dict, attrs):
    """Return (Population, keepalive) from a tools/synthpop.py dict and immunity_attributes()."""
    keep = {}
    def arr(key, dtype, src=None):
        a = np.ascontiguousarray(pop[key] if src is None else src, dtype=dtype); keep[key] = a; return a
    s = Population()
    s.n_people = int(pop["n_people"]); s.n_loc = int(pop["n_loc"])
    def ptr(a, ct): return a.ctypes.data_as(C.POINTER(ct))
    s.loc_type = ptr(arr("loc_type", np.uint8), u8); s.loc_essential = ptr(arr("loc_essential", np.uint8), u8)
    s.loc_pubrisk = ptr(arr("loc_pubrisk", np.uint8), u8); s.loc_risk = ptr(arr("loc_risk", np.float32), f)
    s.loc_hospital = ptr(arr("loc_hospital", np.int32), i32)
    s.nbr_off = ptr(arr("nbr_off", np.int64), C.c_int64); s.nbr = ptr(arr("nbr", np.int32), i32)
    s.roster_off = ptr(arr("roster_off", np.int64), C.c_int64); s.roster = ptr(arr("roster", np.int32), i32)
    s.home = ptr(arr("home", np.int32), i32); s.day = ptr(arr("day", np.int32), i32)
    s.age = ptr(arr("age", np.uint8), u8); s.sex = ptr(arr("sex", np.uint8), u8)
    s.comorbid = ptr(arr("comorbid", np.uint8), u8); s.ltc = ptr(arr("ltc", np.uint8), u8)
    pat = arr("patron", np.int32); s.patron = ptr(pat, i32); s.n_patron = int(pat.shape[1])
    eff, xprot, nat_dur = attrs
    s.starting_natural_efficacy = ptr(arr("_eff", np.float64, eff), d)
    s.cross_protection_probability = ptr(arr("_xprot", np.float32, xprot), f)
    s.natural_immunity_duration = ptr(arr("_dur", np.int32, nat_dur), i32)
    return s, keep
