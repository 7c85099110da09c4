"""Parity of the CUDA day pipeline against the CPU oracle, through the C ABI (run with -m gpu).

Bar (BASELINE.json north_star): integer state transitions bit-exact under the shared Philox stream;
per-location force of infection within rtol 1e-6 (here it is bit-equal as well, because both sides
accumulate pressure in 2^-32 fixed point and share cabm_exp).
"""
import numpy as np
import pytest

import covid_abm_b200 as cabm
import synthpop
import oracle_lib

pytestmark = pytest.mark.gpu

FOI_RTOL = 1e-6


def make_pair(n, days, seed=5, T=0.10, pde=1e-3, mutants=(), pop_seed=11):
    pop = synthpop.generate(n, seed=pop_seed)
    P = cabm.params.default_params(run_length=days, seed=seed, T=T, prob_daily_exposure=pde, mutant_intro_dates=mutants)
    tab = cabm.params.build_tables(P)
    attrs = cabm.abi.immunity_attributes(int(pop["n_people"]), P["seed"])
    g = cabm.Community(P, pop, tab, attrs=attrs)
    o = oracle_lib.oracle_community(P, pop, tab, attrs=attrs)
    return pop, P, g, o


def assert_same_state(g, o, day_label=""):
    lg, lo = g.ledger_table(), o.ledger_table()
    bad = np.argwhere(lg != lo)
    assert bad.size == 0, f"{day_label} ledger differs first at (day, counter) {bad[:5].tolist()}: gpu {lg[tuple(bad[0])]} oracle {lo[tuple(bad[0])]}"
    assert np.array_equal(g.getCumulIncidenceByOutcome(), o.getCumulIncidenceByOutcome())
    cg, co = g.current_infections(), o.current_infections()
    diff = np.flatnonzero(cg != co)
    assert diff.size == 0, f"{day_label} {diff.size} agents differ, first {diff[:3]}: gpu {cg[diff[:1]]} oracle {co[diff[:1]]}"


def sorted_hist(c):
    h = c.infection_histories()
    return h[np.lexsort((h["t"][:, 0], h["person"]))]


def test_small_community_daily_state_bit_exact():
    pop, P, g, o = make_pair(20000, 60, T=0.12, pde=2e-3)
    g.foi(); o.foi()
    for c in (g, o):
        for pid in range(0, 3000, 7):
            c.infect(pid, 0, 0)
    for d in range(60):
        assert g.step() == o.step(), f"day {d}: imported cases"       # tick + import of one day run as one day graph (cabm.cu, tick_pending)
        if d % 10 == 9 or d < 3:
            assert_same_state(g, o, f"day {d}")
    assert g.getNumNewlyInfected().sum() > 500          # the epidemic actually ran
    hg, ho = sorted_hist(g), sorted_hist(o)
    assert hg.size == ho.size and np.array_equal(hg, ho)


def test_force_of_infection_rows():
    pop, P, g, o = make_pair(30000, 30, T=0.15, pde=3e-3, seed=9)
    g.foi(); o.foi()
    for c in (g, o):
        for pid in range(0, 5000, 5):
            c.infect(pid, 0, 0)
    seen = set()
    for d in range(30):
        g.tick(d); o.tick(d)
        fg, fo = g.foi(), o.foi()
        assert fg.size == fo.size and np.array_equal(fg["loc"], fo["loc"]), f"day {d}: hot sets differ"
        assert np.array_equal(fg["size"], fo["size"])
        np.testing.assert_allclose(fg["w"], fo["w"], rtol=FOI_RTOL, atol=0)
        np.testing.assert_allclose(fg["T"], fo["T"], rtol=FOI_RTOL, atol=0)
        np.testing.assert_allclose(fg["T_social"], fo["T_social"], rtol=FOI_RTOL, atol=0)
        seen |= set(np.unique(fg["type"][fg["T"] > 0]).tolist())
        g.update_strain_mix(d); o.update_strain_mix(d)
        assert g.seed_epidemic(d) == o.seed_epidemic(d)
        g.day = o.day = d + 1
    assert seen == {0, 1, 2, 3, 4}, f"location types with positive hazard: {seen}"


def test_multi_variant_vaccination_quarantine():
    mut = ("2020-02-20", "2020-03-05", "2020-03-20")
    pop, P, g, o = make_pair(25000, 75, T=0.14, pde=3e-3, mutants=mut, seed=21)
    rng = np.random.default_rng(0)
    n = int(pop["n_people"])
    for c in (g, o):
        for pid in range(0, 4000, 9):
            c.infect(pid, 0, 0)
    for d in range(75):
        if d in (10, 31, 52):                      # three dose rounds (Person::vaccinate)
            pids = rng.choice(n, size=n // 3, replace=False)
            ok_g, ok_o = g.vaccinate(pids, d), o.vaccinate(pids, d)
            assert np.array_equal(ok_g, ok_o)
        if d in (20, 40):                          # self-quarantine windows (Person::scheduleQuarantine)
            pids = rng.choice(n, size=n // 10, replace=False)
            g.quarantine(pids, d, d + 10); o.quarantine(pids, d, d + 10)
        g.step(); o.step()
        if d % 15 == 14:
            assert_same_state(g, o, f"day {d}")
        if d % 7 == 3 or d in (10, 31, 52, 74):   # caller-side daily diagnostics as device reductions (cabm_diagnostics / cabm_serosurvey)
            dg, do = g.diagnostics(d), o.diagnostics(d)
            assert dg == do, f"day {d}: {[(k, dg[k], do[k]) for k in dg if dg[k] != do[k]]}"
            assert np.array_equal(g.detected_today(d), o.detected_today(d))
            some = rng.choice(n, size=n // 4, replace=False)
            assert g.doSerosurvey(some, d) == o.doSerosurvey(some, d) and g.doSerosurvey(None, d) == o.doSerosurvey(None, d)
            # cabm_eligibility / cabm_reduce: per-person predicates for a cohort and for everybody
            assert np.array_equal(g.eligibility(some, d), o.eligibility(some, d)) and np.array_equal(g.eligibility(None, d), o.eligibility(None, d))
            for what in cabm.abi.REDUCE:
                assert g.reduce(what, some, d) == o.reduce(what, some, d), what
                assert g.reduce(what, None, d) == o.reduce(what, None, d), what
            for what, key in (("infected", "infected"), ("infectious", "infectious"), ("symptomatic", "symptomatic"), ("naive", "naive"),
                              ("quarantining", "quarantining"), ("fully_vaxd", "fully_vaxd"), ("seropositive", "seropos_all")):
                assert g.reduce(what, None, d) == dg[key], (what, g.reduce(what, None, d), dg[key])
    assert_same_state(g, o, "end")
    assert sum(g.detected_today(d).size for d in range(40, 75)) > 20
    dg = g.diagnostics(74)
    assert dg["fully_vaxd"] > 0 and dg["seropos_all"] > 0 and dg["quarantining"] == 0 and dg["bt_inf"] + dg["uv_inf"] > 0
    hg, ho = g.infection_histories(), o.infection_histories()
    assert np.array_equal(g.infection_histories_from(hg.size - 50), hg[-50:]) and hg.size == ho.size
    strains = g.ledger_table()[:, 16:20].sum(axis=0)
    assert (strains > 0).sum() >= 3, f"variants seen: {strains}"
    assert g.ledger("vaccinated_cases").sum() > 0


def test_run_days_graph_equals_stepwise():
    pop, P, g, o = make_pair(20000, 50, T=0.12, pde=2e-3, seed=33)
    for c in (g, o):
        for pid in range(0, 2000, 4):
            c.infect(pid, 0, 0)
    g.run_days(25); g.run_days(25)
    for _ in range(50):
        o.step()
    assert_same_state(g, o, "run_days")
    w = g.work_counters()
    assert w["agents"] == 50 * int(pop["n_people"]) and w["infections"] == o.work_counters()["infections"]


def test_snapshot_restore_replays_identically():
    pop, P, g, o = make_pair(15000, 40, T=0.12, pde=2e-3, seed=41)
    for pid in range(0, 1500, 3):
        g.infect(pid, 0, 0)
    g.run_days(15)
    snap = g.snapshot()
    g.run_days(25)
    first = g.ledger_table().copy(); cur = g.current_infections().copy()
    g.restore(snap)
    g.run_days(25)
    assert np.array_equal(first, g.ledger_table()) and np.array_equal(cur, g.current_infections())
    g.snap_free(snap)


def test_edge_cases_empty_and_tiny():
    # no infections at all: nothing is hot, nothing happens
    pop, P, g, o = make_pair(3000, 10, T=0.1, pde=0.0, seed=2)
    for d in range(10):
        g.step(); o.step()
    assert g.ledger_table().sum() == 0 and o.ledger_table().sum() == 0
    assert g.work_counters()["hot_locs"] == 0
    # a tiny community below the nursing-home threshold, everyone seeded
    pop, P, g, o = make_pair(300, 30, T=0.3, pde=0.0, seed=4)
    for c in (g, o):
        for pid in range(int(pop["n_people"])):
            c.infect(pid, 1, 0)
    for d in range(30):
        g.step(); o.step()
    assert_same_state(g, o, "tiny")


def test_many_tiles_per_warp_state_and_work_counters():
    """1.5 M agents: every warp of k_visit walks several tiles, stages and flushes records repeatedly, and every block queues
    k >= 4 agents; a strong epidemic makes a large share of the visits land on hot locations. State, ledger and every device
    work counter (patrons, visits, hot visits, flags, trials, hot locations ...) must equal the oracle's."""
    pop, P, g, o = make_pair(1_500_000, 16, T=0.15, pde=2e-3, seed=71, pop_seed=3)
    n = int(pop["n_people"])
    for c in (g, o):
        for pid in range(0, n, 500):
            c.infect(pid, 0, 0)
    g.run_days(16)
    for _ in range(16):
        o.step()
    assert_same_state(g, o, "1.5M agents")
    wg, wo = g.work_counters(), o.work_counters()
    wg.pop("active"); wo.pop("active")      # the oracle counts agents whose window has not ended, the device agents inside their window
    assert wg == wo, {k: (wg[k], wo[k]) for k in wg if wg[k] != wo[k]}
    assert wg["hot_visits"] > 500_000 and wg["infections"] > 10_000, wg


def test_several_winners_per_warp(monkeypatch):
    """k_apply_status packs W = ceil(winners / warps) winners into the lanes of a warp once the winners outnumber the warps.
    A small grid (one block per SM for both apply launches) and a fast epidemic put that path, with thread- and warp-level
    infector searches, under the same bit-exact comparison as the rest."""
    monkeypatch.setenv("CABM_GRID_MULT", "8,8,8,4,1,1,2,1,4")
    pop, P, g, o = make_pair(80000, 40, T=0.35, pde=5e-3, seed=91, pop_seed=17)
    for c in (g, o):
        for pid in range(0, 8000, 4):
            c.infect(pid, 0, 0)
    for d in range(40):
        g.step(); o.step()
        if d % 8 == 7:
            assert_same_state(g, o, f"day {d}")
    assert g.getNumNewlyInfected().max() > 1500, g.getNumNewlyInfected().max()      # > 592 warps' worth of winners in a day
    hg, ho = sorted_hist(g), sorted_hist(o)
    assert hg.size == ho.size and np.array_equal(hg, ho)


def test_infection_log_overflow_is_an_error(monkeypatch):
    """The infection log is the only record of past infections: when it is full the history getters and the snapshot fail
    (CABM_E_NOMEM) instead of returning a truncated history (the capacity is 3 N + 1024 records; CABM_LOG_CAP forces it)."""
    monkeypatch.setenv("CABM_LOG_CAP", "200")
    pop, P, g, o = make_pair(8000, 30, T=0.3, pde=2e-3, seed=6)
    for pid in range(0, 3000, 5):
        g.infect(pid, 0, 0)
    with pytest.raises(cabm.community.CabmError, match="infection log capacity"):
        g.infection_histories()
    with pytest.raises(cabm.community.CabmError, match="infection log capacity"):
        g.snapshot()
    monkeypatch.delenv("CABM_LOG_CAP")
    g.close()


def test_infect_batch_equals_consecutive_calls():
    """cabm_infect_batch = the same calls one after the other: sequence numbers, people listed twice or three times (later tries
    see the earlier ones applied), mixed strains; on the CUDA library against the oracle's batch, against single CUDA calls."""
    pop, P, g, o = make_pair(30000, 30, T=0.12, pde=1e-3, seed=15)
    g1 = cabm.Community(P, pop, g.T, attrs=cabm.abi.immunity_attributes(int(pop["n_people"]), P["seed"]))
    rng = np.random.default_rng(8)
    for day in (0, 1, 4):                                                  # the day each community stands at (1 + 3 steps below)
        assert g.day == o.day == g1.day == day
        pids = rng.integers(0, 30000, size=4000).astype(np.uint32)
        pids[100:140] = pids[50:90]; pids[300:310] = pids[50:60]           # repeated people inside one batch
        strains = rng.integers(0, 4, size=pids.size).astype(np.uint8)
        tg, to = g.infect_many(pids, strains, day), o.infect_many(pids, strains, day)
        t1 = np.array([g1.infect(int(p), int(s), day) for p, s in zip(pids, strains)])
        assert np.array_equal(tg, to) and np.array_equal(tg, t1), f"day {day}: {np.flatnonzero(tg != to)[:5]} {np.flatnonzero(tg != t1)[:5]}"
        assert tg.sum() > 1000 and not tg[100:140].all()
        for c in (g, o, g1):
            for _ in range(3 if day else 1):
                c.step()
        assert_same_state(g, o, f"after day {day}")
    assert np.array_equal(g.ledger_table(), g1.ledger_table()) and np.array_equal(g.current_infections(), g1.current_infections())
    g1.close()


def test_ledger_row_of_the_closed_day_is_the_devices_row():
    """cabm_ledger_row of the day that was just closed is served from the row close_day left in pinned host memory (no device
    round trip); it must equal the device's ledger at that moment, fall back to the device read once another call has touched
    the ledger (cabm_infect on that day), and equal the oracle's row."""
    import ctypes as C
    pop, P, g, o = make_pair(20000, 40, T=0.15, pde=3e-3, seed=17)
    for c in (g, o):
        for pid in range(0, 3000, 5):
            c.infect(pid, 0, 0)
    row = np.zeros(cabm.abi.N_COUNTERS, dtype=np.uint64)

    def gpu_row(day):
        g._check(g.fn.ledger_row(g._h, int(day), row.ctypes.data_as(C.POINTER(C.c_uint64))))
        return row.copy()
    for d in range(30):
        g.tick(d)
        after_tick = gpu_row(d)
        assert np.array_equal(after_tick, g.ledger_table()[d]), f"day {d}: row after cabm_tick"
        g.update_strain_mix(d); g.seed_epidemic(d); g.day += 1
        o.step()
        r = gpu_row(d)
        assert np.array_equal(r, g.ledger_table()[d]) and np.array_equal(r, o.ledger_table()[d]), f"day {d}: row after cabm_import"
        if d >= 1:
            assert np.array_equal(gpu_row(d - 1), o.ledger_table()[d - 1])         # an older day: device read
        if d == 12:                                                                # a seed on the closed day: the pinned row is stale, the device read is not
            g.infect(19999, 1, d); o.infect(19999, 1, d)
            assert np.array_equal(gpu_row(d), g.ledger_table()[d]) and np.array_equal(gpu_row(d), o.ledger_table()[d])
    g.run_days(10)
    for _ in range(10):
        o.step()
    assert np.array_equal(gpu_row(39), o.ledger_table()[39]) and np.array_equal(gpu_row(38), o.ledger_table()[38])
    assert g.ledger_table()[:, 0].sum() > 300


def test_staged_tick_equals_eager_tick(monkeypatch):
    """cabm_tick stages its day and the next call launches it (fused with cabm_import of the same day, on its own before anything
    else). The same call sequence with CABM_EAGER_TICK=1 (every tick runs before it returns) must leave the same state, whatever
    comes between tick and import: a getter, a mutator, another tick, nothing."""
    pop = synthpop.generate(20000, seed=11)
    P = cabm.params.default_params(run_length=40, seed=23, T=0.15, prob_daily_exposure=3e-3)
    tab = cabm.params.build_tables(P)
    attrs = cabm.abi.immunity_attributes(int(pop["n_people"]), P["seed"])
    rng = np.random.default_rng(3)
    vax_ids = rng.choice(int(pop["n_people"]), size=2000, replace=False)
    quar_ids = rng.choice(int(pop["n_people"]), size=1500, replace=False)

    def run(eager):
        if eager:
            monkeypatch.setenv("CABM_EAGER_TICK", "1")
        else:
            monkeypatch.delenv("CABM_EAGER_TICK", raising=False)
        c = cabm.Community(P, pop, tab, attrs=attrs)
        for pid in range(0, 3000, 5):
            c.infect(pid, 0, 0)
        imported = []
        for d in range(30):
            c.tick(d)
            if d % 7 == 1:
                c.vaccinate(vax_ids[(d // 7) * 400:(d // 7 + 1) * 400], d)        # a mutator between tick and import
            elif d % 7 == 2:
                c.quarantine(quar_ids[(d // 7) * 300:(d // 7 + 1) * 300], d, d + 4)
            elif d % 7 == 3:
                c.diagnostics(d)                                                     # a getter
            elif d % 7 == 4:
                c.detected_today(d)
            c.update_strain_mix(d)
            imported.append(c.seed_epidemic(d))
            c.day = d + 1
        c.tick(30)                                                                   # a tick that no import follows: the ledger read launches it
        led = c.ledger_table().copy()
        state = c.current_infections().copy()
        hist = sorted_hist(c)
        c.close()
        return imported, led, state, hist
    a = run(True); b = run(False)
    assert a[0] == b[0], "imported cases per day"
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3])
    assert a[1][:, 0].sum() > 300 and a[1][30].sum() > 0


def test_community_loaded_from_the_population_file_equals_one_loaded_from_arrays(tmp_path):
    """cabm_load_population_file on the CUDA library: same state, ledger and histories as cabm_load_population of the arrays."""
    pop = synthpop.generate(20000, seed=11)
    P = cabm.params.default_params(run_length=30, seed=29, T=0.15, prob_daily_exposure=3e-3)
    tab = cabm.params.build_tables(P)
    attrs = cabm.abi.immunity_attributes(int(pop["n_people"]), P["seed"])
    path = str(tmp_path / "pop.bin")
    cabm.save_population(pop, path)
    a = cabm.Community(P, pop, tab, attrs=attrs)
    b = cabm.Community(P, pop, tab, attrs=attrs, population_file=path)
    for c in (a, b):
        for pid in range(0, 3000, 5):
            c.infect(pid, 0, 0)
        c.run_days(30)
    assert np.array_equal(a.ledger_table(), b.ledger_table()) and a.ledger_table()[:, 0].sum() > 300
    assert np.array_equal(a.current_infections(), b.current_infections())
    assert np.array_equal(sorted_hist(a), sorted_hist(b))
    with pytest.raises(cabm.CabmError, match="not found"):
        cabm.Community(P, pop, tab, attrs=attrs, population_file=str(tmp_path / "absent.bin"))
