"""CUDA path against the oracle AT THE SHAPES BASELINE.json names (run with -m gpu).

Every performance number of this repository is quoted on one of these configurations, so each of them is compared with
the CPU oracle here: state of every agent, the whole ledger table, cumulative outcomes and the device work counters.

  configs[0]  100 k agents, single variant, the full 200-day run
  configs[1]  2.7 M agents, wild type + alpha + delta: the bench workload itself (bench.make_workload), 60 days from a
              seeded epidemic
  configs[2]  21.5 M agents (Florida scale), multi-variant, 10 days from a seeded epidemic with three vaccination rounds
  configs[3]  ring vaccination + contact tracing through the facade drivers: tests/test_facade.py
              (test_ring_vaccination_at_scale_cuda_equals_oracle)
  configs[4]  256 communities of 100 k agents advanced by cabm_batch_run_days, eight of them checked against their own
              oracle runs, all of them against the structure of the batch (distinct seeds => distinct epidemics)

The oracle is sequential C++ (about 0.07 s per simulated day at 100 k agents, 2 s at 2.7 M, 15 s at 21.5 M), so the
small-population oracle runs are spread over the host cores with a process pool.
"""
import concurrent.futures as cf
import multiprocessing as mp
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)

pytestmark = pytest.mark.gpu

MUT3 = ("2020-03-01", "2020-03-21", "2020-04-10")          # alpha, delta, omicron inside a 200-day window from 2020-02-10


def _params_100k(seed, T, days, mutants):
    import covid_abm_b200 as cabm
    P = cabm.params.default_params(run_length=days, seed=seed, T=T, mutant_intro_dates=mutants)
    return P, cabm.params.build_tables(P)


def _oracle_100k(job):
    """One oracle run of the 100 k-agent population (worker process: no CUDA in here)."""
    seed, T, days, mutants, n_seed = job
    import covid_abm_b200 as cabm
    import synthpop
    import oracle_lib
    pop = synthpop.generate(100_000)
    P, tab = _params_100k(seed, T, days, mutants)
    o = oracle_lib.oracle_community(P, pop, tab)
    for pid in range(0, n_seed * 50, 50):
        o.infect(pid, 0, 0)
    for _ in range(days):
        o.step()
    out = (o.ledger_table(), o.current_infections(), o.getCumulIncidenceByOutcome(), o.work_counters())
    o.close()
    return out


def _oracle_pool(jobs):
    workers = max(1, min(len(jobs), (os.cpu_count() or 2) - 1))
    with cf.ProcessPoolExecutor(max_workers=workers, mp_context=mp.get_context("spawn")) as ex:
        return list(ex.map(_oracle_100k, jobs))


def _assert_equal(g, ref, label):
    lo, co, cumo, wo = ref
    lg = g.ledger_table()
    bad = np.argwhere(lg != lo)
    assert bad.size == 0, f"{label}: ledger differs first at (day, counter) {bad[:5].tolist()}"
    assert np.array_equal(g.getCumulIncidenceByOutcome(), cumo), label
    cg = g.current_infections()
    diff = np.flatnonzero(cg != co)
    assert diff.size == 0, f"{label}: {diff.size} agents differ, first {diff[:3]}"
    wg = dict(g.work_counters()); wo = dict(wo)
    wg.pop("active"); wo.pop("active")          # the oracle counts agents whose window has not ended, the device agents inside their window
    assert wg == wo, f"{label}: work counters {[(k, wg[k], wo[k]) for k in wg if wg[k] != wo[k]]}"


def test_config0_100k_agents_200_days_single_variant():
    import covid_abm_b200 as cabm
    import synthpop
    days = 200
    job = (1, 0.06, days, (), 40)
    ref = _oracle_pool([job])[0]
    pop = synthpop.generate(100_000)
    P, tab = _params_100k(job[0], job[1], days, job[3])
    g = cabm.Community(P, pop, tab)
    for pid in range(0, job[4] * 50, 50):
        g.infect(pid, 0, 0)
    g.run_days(120); g.run_days(80)
    _assert_equal(g, ref, "configs[0]")
    assert ref[0][:, 0].sum() > 20_000, "the 200-day epidemic did not take off"
    assert ref[0][:, 17:20].sum() == 0                    # single variant
    g.close()


def test_config4_256_batched_communities_of_100k():
    """256 communities, one launch per kernel for the whole batch; T ~ U(0.03, 0.07) and seed per replicate as in the ABC sweeps."""
    import covid_abm_b200 as cabm
    import synthpop
    days, R, checked = 100, 256, (0, 1, 37, 100, 128, 200, 254, 255)
    rng = np.random.default_rng(2024)
    Ts = rng.uniform(0.03, 0.07, size=R)
    jobs = [(1 + r, float(Ts[r]), days, MUT3, 20) for r in checked]
    refs = _oracle_pool(jobs)
    pop = synthpop.generate(100_000)
    comms = []
    for r in range(R):
        P, tab = _params_100k(1 + r, float(Ts[r]), days, MUT3)
        c = cabm.Community(P, pop, tab)
        for pid in range(0, 20 * 50, 50):
            c.infect(pid, 0, 0)
        comms.append(c)
    cabm.replicates.batch_run_days(comms, 60)
    cabm.replicates.batch_run_days(comms, days - 60)
    for r, ref in zip(checked, refs):
        _assert_equal(comms[r], ref, f"configs[4] replicate {r}")
    totals = [int(c.ledger_table()[:, 0].sum()) for c in comms]
    assert len(set(totals)) > R * 0.9 and min(totals) > 50, (len(set(totals)), min(totals))
    for c in comms:
        c.close()


def _run_pair(pop, P, tab, days, seed_stride, vaccinate_on=()):
    import covid_abm_b200 as cabm
    import oracle_lib
    n = int(pop["n_people"])
    attrs = cabm.abi.immunity_attributes(n, P["seed"])
    g = cabm.Community(P, pop, tab, attrs=attrs)
    o = oracle_lib.oracle_community(P, pop, tab, attrs=attrs)
    seeds = np.arange(0, n, seed_stride, dtype=np.int64)
    for c in (g, o):
        for k, pid in enumerate(seeds):
            c.infect(int(pid), k % 3, 0)
    rng = np.random.default_rng(7)
    d = 0
    stops = sorted(set(vaccinate_on) | {days})
    for stop in stops:
        if stop > d:
            g.run_days(stop - d)
            for _ in range(stop - d):
                o.step()
            d = stop
        if stop in vaccinate_on:
            pids = rng.choice(n, size=n // 8, replace=False).astype(np.uint32)
            ok_g, ok_o = g.vaccinate(pids, d), o.vaccinate(pids, d)
            assert np.array_equal(ok_g, ok_o)
    ref = (o.ledger_table(), o.current_infections(), o.getCumulIncidenceByOutcome(), o.work_counters())
    o.close()
    return g, ref


def test_config1_bench_workload_2p7M_agents_three_variants_60_days():
    """The population, parameters and variant calendar bench.py times (bench.make_workload), from a seeded epidemic."""
    import bench
    days = 60
    cabm, pop, P, tab = bench.make_workload(days, seed=1)
    assert int(pop["n_people"]) >= 2_600_000
    # alpha enters on day 40 of this calendar; seed all three strains so that every variant path runs inside the window
    g, ref = _run_pair(pop, P, tab, days, seed_stride=900)
    _assert_equal(g, ref, "configs[1]")
    led = ref[0]
    assert led[:, 0].sum() > 20_000 and (led[:, 16:19].sum(axis=0) > 100).all(), led[:, 16:20].sum(axis=0)
    g.close()


def test_config2_florida_scale_21p5M_agents_10_days_with_vaccination():
    import covid_abm_b200 as cabm
    import synthpop
    days = 10
    pop = synthpop.generate(21_500_000)
    assert int(pop["n_people"]) >= 21_000_000
    P = cabm.params.default_params(run_length=days, seed=3, mutant_intro_dates=("2020-02-12", "2020-02-14", "2020-02-16"),
                                   prob_daily_exposure=2e-4)
    tab = cabm.params.build_tables(P)
    g, ref = _run_pair(pop, P, tab, days, seed_stride=400, vaccinate_on=(2, 5, 8))
    _assert_equal(g, ref, "configs[2]")
    led = ref[0]
    assert led[:, 0].sum() > 50_000 and ref[3]["infections"] > 60_000, (led[:, 0].sum(), ref[3])
    g.close()
