"""Tick-level semantics of the CPU oracle, checked against the reference's formulas restated by hand (no GPU needed).

The oracle is the checker of the CUDA path, so its own behaviour is pinned three ways: Person::infect bit-for-bit against
the unmodified reference (test_reference_kat.py), epidemic curves by KS test against the unmodified reference
(test_ks_equivalence.py), and here the per-location exposure probability, the pass gating and the bookkeeping of
Community::tick (Community.cpp:842-1192) against closed-form expectations on populations small enough to reason about.
"""
import math

import numpy as np
import pytest

import covid_abm_b200 as cabm
import synthpop
import oracle_lib

HOUSE, WORK, SCHOOL, HOSPITAL, NURSINGHOME = 0, 1, 2, 3, 4
T_INFECTED, T_INFECTIOUS_BEGIN, T_INFECTIOUS_END, T_SYMPTOM_BEGIN, T_SYMPTOM_END, T_SEVERE_BEGIN, T_SEVERE_END, T_HOSP_BEGIN, \
    T_CRITICAL_BEGIN, T_CRITICAL_END, T_ICU_BEGIN, T_DEATH = range(12)
NEVER = 32767


def community(n=6000, days=60, seed=3, T=0.12, pde=0.0, pop_seed=5, **kw):
    pop = synthpop.generate(n, seed=pop_seed)
    P = cabm.params.default_params(run_length=days, seed=seed, T=T, prob_daily_exposure=pde, **kw)
    tab = cabm.params.build_tables(P)
    return pop, P, oracle_lib.oracle_community(P, pop, tab)


def test_no_infection_no_events():
    pop, P, c = community(pde=0.0, days=30)
    c.run_days(30)
    assert c.ledger_table().sum() == 0 and c.infection_histories().size == 0 and c.getCumulIncidenceByOutcome().sum() == 0
    assert c.foi().size == 0
    w = c.work_counters()
    assert w["trials"] == 0 and w["infections"] == 0 and w["flags"] == 0 and w["agents"] == 30 * int(pop["n_people"])


def test_household_force_of_infection_formula():
    """T_house = 1 - exp(-haz_mult[HOUSE] * seasonality * sum relInfectiousness) (Community.cpp:842-850);
    T_social = 1 - exp(-haz_mult[social] * seasonality * w / household size) where both homes are risk-tolerant (:853-875)."""
    pop, P, c = community(pde=0.0, days=40)
    pid = 17
    assert c.infect(pid, 0, 0)
    inf = c.current_infections([pid])[0]
    b, e = int(inf["t"][T_INFECTIOUS_BEGIN]), int(inf["t"][T_INFECTIOUS_END])
    assert 0 < b < e
    home = int(pop["home"][pid])
    hh = int(pop["roster_off"][home + 1] - pop["roster_off"][home])
    seen = 0
    for day in range(e + 2):
        c.tick(day); c.day = day + 1
        rows = c.foi()
        mine = rows[(rows["loc"] == home)]
        # only the index case's household is guaranteed: secondary cases may add their own rows
        if not (b <= day < e) or int(c.current_infections([pid])[0]["t"][T_HOSP_BEGIN]) <= day:
            continue
        assert mine.size == 1, f"day {day}: the index household must be hot"
        r = mine[0]
        assert r["type"] == HOUSE and r["size"] == hh
        w = float(r["w"].sum())
        assert w >= float(inf["rel_infectiousness"]) - 1e-9           # at least the index case's weight
        season = P["seasonality"][int(c.julian_day[day]) - 1]
        T_want = 1.0 - math.exp(-P["haz_household"] * season * w)
        assert abs(r["T"] - T_want) <= 1e-12 + 1e-9 * T_want
        sd = P["social_distancing"][day]
        if float(pop["loc_risk"][home]) > sd:
            Ts_want = 1.0 - math.exp(-P["haz_social"] * season * w / hh)
            assert abs(r["T_social"] - Ts_want) <= 1e-12 + 1e-9 * Ts_want
        else:
            assert r["T_social"] == 0.0
        seen += 1
    assert seen >= 2


def test_workplace_rows_and_gating():
    """WORK: T = 1 - exp(-haz * (4 | 0.25 by public-transmission risk) * season * w / (staff + visitors - 1)), zero when the
    business is closed and non-essential (Community.cpp:878-911); SCHOOL rows only on open weekdays (:1158, :914-928)."""
    pop, P, c = community(n=8000, pde=0.0, days=30, T=0.05)
    # infect many people so that plenty of workplaces and schools are hot on every day
    for pid in range(0, int(pop["n_people"]), 3):
        c.infect(pid, 0, 0)
    saw_school_weekday = saw_weekend = False
    for day in range(12):
        c.tick(day); c.day = day + 1
        rows = c.foi()
        season = P["seasonality"][int(c.julian_day[day]) - 1]
        wk = rows[rows["type"] == WORK]
        for r in wk[:200]:
            loc = int(r["loc"])
            staff = int(pop["roster_off"][loc + 1] - pop["roster_off"][loc])
            assert r["size"] >= staff
            closed = (not pop["loc_essential"][loc]) and P["business_closure"][day] != 0.0
            w = float(r["w"].sum())
            if r["size"] < 2 or closed or w == 0:
                assert r["T"] == 0.0
            else:
                mult = 4.0 if pop["loc_pubrisk"][loc] == 2 else 0.25
                want = 1.0 - math.exp(-P["haz_work"] * mult * season * w / (r["size"] - 1.0))
                assert abs(r["T"] - want) <= 1e-12 + 1e-9 * want
        school = rows[rows["type"] == SCHOOL]
        open_today = bool(c.is_weekday[day]) and P["school_closure"][day] < 1
        if open_today:
            saw_school_weekday |= bool(school.size and (school["T"] > 0).any())
        else:
            saw_weekend = True
            assert not (school["T"] > 0).any(), f"day {day}: schools must not transmit when closed"
    assert saw_school_weekday and saw_weekend


def test_infection_records_are_well_formed():
    pop, P, c = community(n=8000, days=80, T=0.15, pde=5e-4)
    c.run_days(80)
    h = c.infection_histories()
    assert h.size > 1000
    t = h["t"].astype(np.int64)
    assert np.all(t[:, T_INFECTIOUS_BEGIN] > t[:, T_INFECTED]) and np.all(t[:, T_INFECTIOUS_END] >= t[:, T_INFECTIOUS_BEGIN])
    sym = t[:, T_SYMPTOM_BEGIN] != NEVER
    assert np.all(t[sym, T_SYMPTOM_BEGIN] > t[sym, T_INFECTED]) and np.all(t[sym, T_SYMPTOM_END] >= t[sym, T_SYMPTOM_BEGIN])
    sev = t[:, T_SEVERE_BEGIN] != NEVER
    assert np.all(sym[sev]) and np.all(t[sev, T_SEVERE_BEGIN] >= t[sev, T_SYMPTOM_BEGIN])
    crit = t[:, T_CRITICAL_BEGIN] != NEVER
    assert np.all(sev[crit]) and np.all(t[crit, T_CRITICAL_BEGIN] >= t[crit, T_SEVERE_BEGIN])
    icu = t[:, T_ICU_BEGIN] != NEVER
    assert np.all(crit[icu])
    hosp = t[:, T_HOSP_BEGIN] != NEVER
    assert np.all(sev[hosp]) and np.all(t[hosp, T_HOSP_BEGIN] >= t[hosp, T_SEVERE_BEGIN])
    # transmission bookkeeping: imported cases have no infector, local ones were infected somewhere by someone infectious
    local = h["infected_by"] >= 0
    assert local.any() and (~local).any()
    assert np.all(h["infected_place"][local] >= 0)
    # ledger vs histories: updatePersonStatus tallies `infectedBegin == today` in the middle of the tick (Community.cpp:1170),
    # so infections made after it on the same day -- the hospital pass (:1172) and the day's importation (simulator.h:374) --
    # never enter _numNewlyInfected. That is the reference's behaviour and the restatement keeps it.
    led = c.ledger("new_infected")
    place = h["infected_place"]
    late = (place < 0) | (pop["loc_type"][np.maximum(place, 0)] == HOSPITAL)
    on_time = t[:, T_INFECTED][~late]
    assert np.array_equal(led.astype(np.int64), np.bincount(on_time[on_time < 80], minlength=80))
    assert late.sum() > 0
    dead = t[:, T_DEATH]
    assert c.ledger("new_dead").sum() == np.count_nonzero(dead < 80)
    by_strain = np.stack([c.ledger(f"new_inf_strain{s}") for s in range(4)]).sum(axis=0)
    assert np.array_equal(by_strain, led)


def test_infector_was_infectious_and_shared_the_place():
    pop, P, c = community(n=6000, days=50, T=0.2, pde=5e-4)
    c.run_days(50)
    h = c.infection_histories()
    by_person = {}
    for r in h:
        by_person.setdefault(int(r["person"]), []).append(r)
    checked = 0
    for r in h[h["infected_by"] >= 0][:3000]:
        day = int(r["t"][T_INFECTED])
        src = by_person[int(r["infected_by"])]
        assert any(int(s["t"][T_INFECTIOUS_BEGIN]) <= day < int(s["t"][T_INFECTIOUS_END]) and int(s["strain"]) == int(r["strain"]) for s in src), \
            "the recorded infector was not infectious with that strain on the day of transmission"
        checked += 1
    assert checked > 500


def test_run_days_equals_day_by_day_stepping_and_is_deterministic():
    pop, P, a = community(days=40, pde=1e-3, seed=11)
    _, _, b = community(days=40, pde=1e-3, seed=11)
    _, _, d = community(days=40, pde=1e-3, seed=12)
    a.run_days(40)
    for _ in range(40):
        b.step()
    d.run_days(40)
    assert np.array_equal(a.ledger_table(), b.ledger_table()) and np.array_equal(a.current_infections(), b.current_infections())
    assert not np.array_equal(a.ledger_table(), d.ledger_table())


def test_snapshot_restore_replays_identically():
    pop, P, c = community(days=50, pde=1e-3)
    c.run_days(20)
    snap = c.snapshot()
    c.run_days(30)
    first = c.ledger_table().copy(); state = c.current_infections().copy()
    c.restore(snap)
    assert c.day == 20 and c.ledger("new_infected")[20:].sum() == 0      # detection series are written ahead of time, not checked here
    c.run_days(30)
    assert np.array_equal(first, c.ledger_table()) and np.array_equal(state, c.current_infections())
    c.snap_free(snap)


def test_vaccination_and_quarantine_reduce_transmission():
    def total(vax=False, quar=False):
        pop, P, c = community(n=8000, days=70, T=0.25, pde=1e-3, seed=21)
        n = int(pop["n_people"])
        if vax:
            for dose_day in (0, 21):
                c.run_days(dose_day - c.day) if dose_day > c.day else None
                ok = c.vaccinate(np.arange(n, dtype=np.uint32), dose_day)
                assert ok.mean() > 0.9
        if quar:
            c.quarantine(np.arange(0, n, 2, dtype=np.uint32), 0, 70)
        c.run_days(70 - c.day)
        return int(c.ledger("new_infected").sum())
    base, v, q = total(), total(vax=True), total(quar=True)
    assert base > 2000, (base, v, q)
    assert v < 0.8 * base and q < 0.9 * base, (base, v, q)


def test_importation_only_hits_expected_count():
    """seed_epidemic: Poisson(prob_daily_exposure * N) people per day (simulator.h:201-223); T=0 leaves only imports."""
    pop, P, c = community(n=10000, days=60, T=0.0, pde=1e-3)
    c.run_days(60)
    n = int(pop["n_people"])
    assert c.ledger("new_infected").sum() == 0              # imports are made after the day's tally (see above)
    got = int(c.infection_histories().size)
    # each exposed person is infected with probability susceptibility_by_age (Person.cpp:333-336); nobody is immune yet
    m = float(np.asarray(P["susceptibility_by_age"])[pop["age"]].mean())
    assert abs(got - 60 * 1e-3 * n * m) < 5 * math.sqrt(60 * 1e-3 * n * m), (got, 60 * 1e-3 * n * m)
    assert np.all(c.infection_histories()["infected_by"] < 0)


def test_ragged_population_edges():
    """No patronised places, an empty workplace roster, single-person households and isolated houses are all legal."""
    pop = synthpop.generate(3000, seed=9)
    pop = dict(pop)
    pop["patron"] = np.full_like(pop["patron"], -1)                      # nobody visits anything
    pop["nbr_off"] = np.zeros_like(pop["nbr_off"]); pop["nbr"] = np.zeros(0, dtype=np.int32)
    P = cabm.params.default_params(run_length=30, seed=2, T=0.2, prob_daily_exposure=2e-3)
    c = oracle_lib.oracle_community(P, pop, cabm.params.build_tables(P))
    c.run_days(30)
    w = c.work_counters()
    assert w["patrons"] == 0 and w["visits"] == 0 and w["nbr_edges"] == 0 and w["infections"] > 0


def test_daily_diagnostics_against_brute_force():
    """cabm_diagnostics / cabm_serosurvey (calculate_vax_stats Community.cpp:1329-1423, doSerosurvey :1224-1274,
    getNumInfected... :1196-1222, getNumPeopleQuarantining Community.h:437) against a numpy restatement on the records."""
    pop, P, c = community(n=6000, days=90, T=0.2, pde=1e-3, seed=8)
    n = int(pop["n_people"])
    doses = np.zeros(n, dtype=np.int64); last = np.full(n, -10**6, dtype=np.int64)
    rng = np.random.default_rng(1)
    for dose_day, frac in ((5, 0.7), (26, 0.5), (50, 0.2)):
        c.run_days(dose_day - c.day)
        who = np.flatnonzero(rng.random(n) < frac).astype(np.uint32)
        ok = c.vaccinate(who, dose_day)
        doses[who[ok]] += 1; last[who[ok]] = dose_day
    c.quarantine(np.arange(0, n, 5, dtype=np.uint32), 60, 75)
    c.run_days(70 - c.day)
    day = 69
    d = c.diagnostics(day)
    rec = c.current_infections(); t = rec["t"].astype(np.int64)
    had = t[:, T_INFECTED] != NEVER
    assert d["dose_1"] == np.count_nonzero((doses >= 1) & (doses <= 3)) and d["dose_2"] == np.count_nonzero((doses >= 2) & (doses <= 3))
    assert d["dose_3"] == np.count_nonzero(doses == 3) and d["dose_1"] > d["dose_2"] > d["dose_3"] > 0
    full = (doses > 1) & (day - last >= P["vaccine_dose_to_protection_lag"])
    assert d["fully_vaxd"] == full.sum() > 0
    for key, col in (("inf", T_INFECTED), ("dis", T_SYMPTOM_BEGIN), ("hosp", T_HOSP_BEGIN), ("death", T_DEATH)):
        ev = had & (t[:, col] == day)
        assert d["bt_" + key] == np.count_nonzero(ev & full) and d["uv_" + key] == np.count_nonzero(ev & ~full), key
    rep = had & (rec["report_day"] == day) & (rec["detected_state"] < 5)
    assert d["bt_reported"] == np.count_nonzero(rep & full) and d["uv_reported"] == np.count_nonzero(rep & ~full)
    assert d["uv_inf"] + d["bt_inf"] > 0
    assert d["infected"] == np.count_nonzero(had & (t[:, T_INFECTED] <= day) & (day < t[:, T_INFECTIOUS_END]))
    assert d["infectious"] == np.count_nonzero(had & (t[:, T_INFECTIOUS_BEGIN] <= day) & (day < t[:, T_INFECTIOUS_END]))
    assert d["symptomatic"] == np.count_nonzero(had & (t[:, T_SYMPTOM_BEGIN] <= day) & (day < t[:, T_SYMPTOM_END]))
    assert d["naive"] == np.count_nonzero(~had & (doses == 0))
    assert d["quarantining"] == len(range(0, n, 5)) and c.diagnostics(59)["quarantining"] == 0 and c.diagnostics(75)["quarantining"] == 0
    ped = pop["age"] <= 17
    assert d["n_ped"] == ped.sum()
    # serosurvey: seroconverted (infected on or after day seroconversion_lag -- the reference's test) and remaining efficacy above threshold
    last_end = np.maximum(t[:, T_INFECTIOUS_END], np.where(t[:, T_SYMPTOM_END] == -32768, -10**9, t[:, T_SYMPTOM_END]))
    # with threshold 0 every converted person is positive unless their starting natural efficacy was clamped to 0
    # (Parameters.h:605: remainingEfficacy(0, .) = logistic(-inf) = 0, and 0 > 0 is false)
    nat_eff = cabm.abi.immunity_attributes(n, P["seed"])[0]
    conv = had & (t[:, T_INFECTED] >= P["seroconversion_lag"]) & (nat_eff > 0)
    assert P["seropositivity_threshold"] == 0.0 and (nat_eff == 0).sum() > 0
    assert d["seropos_all"] == conv.sum() > 0 and d["seropos_ped"] == (conv & ped).sum()
    assert abs(c.doSerosurvey(None, day) - conv.mean()) < 1e-15
    some = np.arange(3, n, 7, dtype=np.uint32)
    assert abs(c.doSerosurvey(some, day) - conv[some].mean()) < 1e-15
    vs = c.calculate_vax_stats(day)
    assert abs(vs["coverage"] - full.mean()) < 1e-15 and abs(vs["dose_2"] - d["dose_2"] / n) < 1e-15
    # incremental history sync returns the tail of the log
    allh = c.infection_histories(); tail = c.infection_histories_from(allh.size - 10)
    assert tail.size == 10 and np.array_equal(tail, allh[-10:]) and c.infection_histories_from(allh.size).size == 0


def test_infect_batch_is_the_calls_one_after_the_other():
    """Host-side contract of cabm_infect_batch on the oracle: same outcomes, sequence numbers and state as single calls."""
    import numpy as np
    import synthpop
    import covid_abm_b200 as cabm
    import oracle_lib
    pop = synthpop.generate(5000, seed=3)
    P = cabm.params.default_params(run_length=10, seed=2, T=0.1)
    a, b = oracle_lib.oracle_community(P, pop), oracle_lib.oracle_community(P, pop)
    rng = np.random.default_rng(1)
    pids = rng.integers(0, 5000, size=600).astype(np.uint32); pids[10:20] = pids[0:10]
    strains = rng.integers(0, 4, size=600).astype(np.uint8)
    ta = a.infect_many(pids, strains, 0)
    tb = np.array([b.infect(int(p), int(s), 0) for p, s in zip(pids, strains)])
    assert np.array_equal(ta, tb) and ta.sum() > 300 and not ta[10:20].any()
    a.step(); b.step()
    assert np.array_equal(a.current_infections(), b.current_infections()) and np.array_equal(a.ledger_table(), b.ledger_table())


def test_eligibility_and_reduce_follow_the_person_records():
    """cabm_eligibility / cabm_reduce on the oracle against the same predicates evaluated with numpy on the records the ABI returns."""
    import numpy as np
    import synthpop
    import covid_abm_b200 as cabm
    import oracle_lib
    pop = synthpop.generate(6000, seed=8)
    P = cabm.params.default_params(run_length=40, seed=4, T=0.25, prob_daily_exposure=2e-3)
    c = oracle_lib.oracle_community(P, pop)
    n = int(pop["n_people"]); rng = np.random.default_rng(2)
    for pid in range(0, 900, 3):
        c.infect(pid, 0, 0)
    dosed = rng.choice(n, size=n // 2, replace=False)
    for d in range(35):
        if d in (5, 30):
            c.vaccinate(dosed, d)
        if d == 20:
            c.quarantine(rng.choice(n, size=n // 5, replace=False), d, d + 30)
        c.step()
    day = 34
    el = c.eligibility(None, day)
    rec = c.current_infections()
    ever = rec["t"][:, 0] != 32767
    alive = ~(ever & (rec["t"][:, 11] <= day))
    assert np.array_equal((el & cabm.abi.EL_EVER_INFECTED) != 0, ever) and np.array_equal((el & cabm.abi.EL_ALIVE) != 0, alive)
    is_dosed = np.zeros(n, bool); is_dosed[dosed] = True
    doses = (el >> cabm.abi.EL_DOSES_SHIFT) & 7
    assert np.array_equal((el & cabm.abi.EL_VACCINATED) != 0, doses > 0) and set(np.unique(doses[is_dosed])) <= {1, 2} and (doses[~is_dosed] == 0).all()
    prior = ever & (rec["detected_state"] < 5) & (rec["report_day"] <= day)
    assert np.array_equal((el & cabm.abi.EL_PRIOR_CASE) != 0, prior) and prior.sum() > 10
    inf_now = ever & (rec["t"][:, 0] <= day) & (day < rec["t"][:, 2])
    assert np.array_equal((el & cabm.abi.EL_INFECTED_NOW) != 0, inf_now)
    diag = c.diagnostics(day)
    assert c.reduce("infected", None, day) == inf_now.sum() == diag["infected"]
    assert c.reduce("alive", None, day) == alive.sum() and c.reduce("quarantining", None, day) == diag["quarantining"] > 0
    assert c.reduce("fully_vaxd", None, day) == diag["fully_vaxd"] and c.reduce("naive", None, day) == diag["naive"]
    cohort = rng.choice(n, size=700, replace=False)
    assert c.reduce("infectious", cohort, day) == (ever & (rec["t"][:, 1] <= day) & (day < rec["t"][:, 2]))[cohort].sum()
    assert np.array_equal(c.eligibility(cohort, day), el[cohort])


def test_population_file_round_trip_and_rejections(tmp_path):
    """cabm_population_save / cabm_load_population_file (cabm.h, include/cabm_popfile.h): a community loaded
    from the binary file runs exactly like one loaded from the arrays; the file written by the CUDA library's host code equals
    the oracle's byte for byte; truncated files, foreign files and a missing file are refused with a message."""
    import ctypes as C
    pop = synthpop.generate(6000, seed=7)
    P = cabm.params.default_params(run_length=25, seed=3, T=0.15, prob_daily_exposure=3e-3)
    tab = cabm.params.build_tables(P)
    attrs = cabm.abi.immunity_attributes(int(pop["n_people"]), P["seed"])
    f_cuda, f_or = str(tmp_path / "pop_cuda.bin"), str(tmp_path / "pop_oracle.bin")
    cabm.save_population(pop, f_cuda)                                    # host code of libcabm.so: needs no GPU
    cabm.save_population(pop, f_or, backend=oracle_lib.oracle_backend())
    data = open(f_cuda, "rb").read()
    assert data == open(f_or, "rb").read() and data[:8] == b"CABMPOP1" and len(data) % 8 == 0
    a = oracle_lib.oracle_community(P, pop, tab, attrs=attrs)
    b = cabm.Community(P, pop, tab, backend=oracle_lib.oracle_backend(), attrs=attrs, population_file=f_cuda)
    for c in (a, b):
        for pid in range(0, 1500, 3):
            c.infect(pid, 0, 0)
        for _ in range(25):
            c.step()
    assert np.array_equal(a.ledger_table(), b.ledger_table()) and a.ledger_table()[:, 0].sum() > 100
    assert np.array_equal(a.current_infections(), b.current_infections())
    # refusals
    def load(path):
        h = cabm.Community.__new__(cabm.Community)
        be = oracle_lib.oracle_backend()
        hh = C.c_void_p(); assert be.fn.create(C.byref(hh)) == 0
        ne, xp, nd = (np.ascontiguousarray(attrs[0], dtype=np.float64), np.ascontiguousarray(attrs[1], dtype=np.float32), np.ascontiguousarray(attrs[2], dtype=np.int32))
        rc = be.fn.load_population_file(hh, path.encode(), ne.ctypes.data_as(C.POINTER(C.c_double)), xp.ctypes.data_as(C.POINTER(C.c_float)),
                                        nd.ctypes.data_as(C.POINTER(C.c_int32)))
        msg = (be.fn.last_error(hh) or b"").decode()
        be.fn.destroy(hh)
        return rc, msg
    rc, msg = load(str(tmp_path / "nothing.bin")); assert rc != 0 and "not found" in msg
    (tmp_path / "short.bin").write_bytes(data[: len(data) // 2]); rc, msg = load(str(tmp_path / "short.bin")); assert rc != 0 and "size does not match" in msg
    (tmp_path / "foreign.bin").write_bytes(b"NOTAPOPF" + data[8:]); rc, msg = load(str(tmp_path / "foreign.bin")); assert rc != 0 and "bad magic" in msg
    (tmp_path / "tiny.bin").write_bytes(b"CABM"); rc, msg = load(str(tmp_path / "tiny.bin")); assert rc != 0 and "too short" in msg
