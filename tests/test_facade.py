"""The C++ drop-in facade (covid-abm_b200/facade/): the reference's UNMODIFIED simulator.h running on the C ABI.

facade_driver_{cuda,oracle} are built by `make -C covid-abm_b200/facade` from the reference's own simulator.h /
Parameters / Utility / Vac_Campaign / behavior_tuner sources plus the proxies Community / Person / Location of this
repository (needs /root/reference at BUILD time only; the binaries travel to the GPU box). The same host code links
against the CUDA library and, prefix cabm_oracle_, against the CPU oracle:

  CPU : host logic of the facade on the oracle backend (loaders, tables, lazy history mirror, campaign, diagnostics) and
        coarse agreement of simulate_epidemic-through-the-facade with the unmodified reference;
  GPU : simulate_epidemic through the facade gives byte-identical output on the CUDA library and on the oracle.
"""
import json
import os
import subprocess

import numpy as np
import pytest

import covid_abm_b200 as cabm
import synthpop
import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "covid-abm_b200", "facade", "_build")
DRV_ORACLE = os.path.join(BUILD, "facade_driver_oracle")
DRV_CUDA = os.path.join(BUILD, "facade_driver_cuda")
MUT = ("2020-03-11", "2020-04-20", "2030-01-01")

needs_oracle_driver = pytest.mark.skipif(not os.path.exists(DRV_ORACLE), reason="facade driver not built (make -C covid-abm_b200/facade needs /root/reference)")


def _case(tmp, n=12000, days=90, seed=5, T=0.2, pde=3e-4, tracing_from=None):
    pop = synthpop.generate(n, seed=4242)
    P = cabm.params.default_params(run_length=days, seed=seed, T=T, prob_daily_exposure=pde, mutant_intro_dates=MUT)
    if tracing_from is not None:                         # exp/active-vac/main.cpp:101-111,711: tracing + self-quarantine of traced contacts
        P["begin_contact_tracing"] = tracing_from; P["quarantine_probability"] = [0.9, 0.75, 0.5]
    tab = cabm.params.build_tables(P)
    synthpop.write_reference_files(pop, str(tmp), "f")
    return pop, P, tab


def _run(driver, tmp, P, tab, seed, campaign, tag):
    Pr = dict(P); Pr["seed"] = seed
    pf = os.path.join(str(tmp), f"params_{tag}.txt"); cabm.params.write_text(Pr, tab, pf)
    prefix = os.path.join(str(tmp), f"out_{tag}")
    r = subprocess.run([driver, "sim", pf, str(tmp), "f", prefix, str(int(campaign))], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    info = json.loads(r.stdout.strip().splitlines()[-1])
    led = np.genfromtxt(prefix + ".ledger.csv", delimiter=",", names=True)
    return info, led, open(prefix + ".plot.csv").read(), open(prefix + ".ledger.csv").read(), r.stderr


@needs_oracle_driver
def test_bridge_tables_match_python_tables(tmp_path):
    """params_bridge.cpp (C++ series for the gamma / binomial / negative-binomial / Poisson CDFs) vs params.py (scipy)."""
    pop, P, tab = _case(tmp_path, n=2000, days=40)
    pf = tmp_path / "p.txt"; cabm.params.write_text(P, tab, str(pf))
    out = tmp_path / "tables.txt"
    subprocess.check_call([DRV_ORACLE, "tables", str(pf), str(out)])
    got = {l.split()[0]: np.array(l.split()[1:], dtype=np.float64) for l in open(out)}
    for key in ("incub_cdf", "onset_cdf", "presev_cdf", "mild_cdf", "severe_only_cdf", "critical_cdf", "icu_death_cdf", "community_death_cdf",
                "death_lag_cdf", "visit_cdf", "trace_ev_cdf"):
        want = np.asarray(tab[key], dtype=np.float64).ravel()
        assert got[key].shape == want.shape, key
        np.testing.assert_allclose(got[key], want, rtol=0, atol=2e-13, err_msg=key)
    np.testing.assert_allclose(got["p_high"], [tab["p_high_asym"], tab["p_high_sym"]], rtol=1e-15)
    s, _keep = cabm.abi.pack_params(P, tab)
    np.testing.assert_array_equal(got["haz_mult"], np.ctypeslib.as_array(s.haz_mult))
    np.testing.assert_array_equal(got["VES"], np.ctypeslib.as_array(s.VES).ravel())
    np.testing.assert_array_equal(got["VEH"], np.ctypeslib.as_array(s.VEH).ravel())
    np.testing.assert_array_equal(got["prob_icu_death_comorbid"], np.ctypeslib.as_array(s.prob_icu_death)[1].astype(np.float64))
    np.testing.assert_allclose(got["prob_first_detection"], np.asarray(P["prob_first_detection"]).ravel(), rtol=0, atol=1e-15)
    np.testing.assert_allclose(got["icu_mortality_reduction"], np.asarray(P["icu_mortality_reduction"]), rtol=0, atol=1e-15)


@needs_oracle_driver
def test_simulate_epidemic_runs_through_the_facade(tmp_path):
    pop, P, tab = _case(tmp_path)
    info, led, plot, _, err = _run(DRV_ORACLE, tmp_path, P, tab, 5, campaign=True, tag="a")
    n, days = int(pop["n_people"]), P["run_length"]
    assert info["agents"] == n and info["days"] == days
    assert len(plot.strip().splitlines()) == days + 1                      # header + one plot-log row per simulated day
    assert "symptomatic infections, total infections" in err                # simulate_epidemic's own end-of-run summary ran
    assert info["infections"] > 1000 and info["doses"] > 1000 and info["cumul_death"] > 0
    # host mirror of the device's infection log: one Infection per ledger entry plus the ones made after the daily tally
    assert info["infections"] >= led["inf"].sum() > 0.8 * info["infections"]
    assert 0.8 * info["infections"] < info["secondary_links"] < info["infections"]      # every local infection is linked to its infector
    assert np.array_equal(led["inf"], led["inf_wt"] + led["inf_alpha"] + led["inf_delta"] + led["inf_omicron"])
    # the plot log carries coverage computed by the device reduction (calculate_vax_stats): monotone, reaching > 10 %
    rows = [l.split(",") for l in plot.strip().splitlines()[1:]]
    cov1 = np.array([float(r[25]) for r in rows])
    assert np.all(np.diff(cov1) >= 0) and cov1[-1] > 0.1 and cov1[15] == 0.0
    seroprev = np.array([float(r[28]) for r in rows])
    assert seroprev[-1] > 0.02 and np.all((seroprev >= 0) & (seroprev <= 1))
    # post-run analytics every exp/ driver reads (getMeanNumSecondaryInfections, generateOffspringDistribution, hSAR)
    rt = open(os.path.join(str(tmp_path), "out_a.rt.csv")).read().strip().splitlines()
    body = np.array([[float(x) for x in r.split(",")] for r in rt[1:1 + days]])
    assert body[:, 1].sum() == info["infections"] and round((body[:, 1] * body[:, 2]).sum()) == info["secondary_links"]
    early = body[5:40]; early = early[early[:, 1] > 0]
    assert 0.8 < np.average(early[:, 2], weights=early[:, 1]) < 6.0                    # a growing epidemic: Rt above ~1 early on
    off = np.array([int(x) for x in rt[1 + days].split(",")[1:]])
    assert rt[1 + days].startswith("offspring") and len(off) >= 10
    assert off.sum() == info["infections"] and (off * np.arange(len(off))).sum() == info["secondary_links"]
    hsar = float(rt[2 + days].split(",")[1])
    assert 0.0 < hsar <= 1.0
    # same seed, same output; another seed, another epidemic
    info2, _, plot2, _, _ = _run(DRV_ORACLE, tmp_path, P, tab, 5, campaign=True, tag="b")
    assert plot2 == plot
    _, _, plot3, _, _ = _run(DRV_ORACLE, tmp_path, P, tab, 6, campaign=True, tag="c")
    assert plot3 != plot


@needs_oracle_driver
def test_contact_tracing_and_ring_vaccination_through_the_facade(tmp_path):
    """traceForwardContacts (Community.cpp:1054-1145) on the proxies with device-supplied primary cases, quarantines written
    through, and the reference's own Vac_Campaign::ring_scheduling (Vac_Campaign.cpp:24-48) turning traced sets into urgent doses."""
    pop, P, tab = _case(tmp_path, tracing_from=30)
    info, led, plot, _, _ = _run(DRV_ORACLE, tmp_path, P, tab, 5, campaign=2, tag="ring")
    base, _, plot0, _, _ = _run(DRV_ORACLE, tmp_path, P, tab, 5, campaign=1, tag="noring")
    assert info["urgent_doses"] > 100 and base["urgent_doses"] == 0
    assert info["quarantine_index"] > 0 and base["quarantine_index"] > 0          # tracing quarantines in both (it is a Parameters setting)
    rows = [l.split(",") for l in plot.strip().splitlines()[1:]]
    urg = np.array([float(r[24]) for r in rows])                                    # urg_doses per 10k in simulate_epidemic's plot log
    assert urg[:30].sum() == 0 and urg[30:].sum() > 0
    # without tracing nobody quarantines
    pop2, P2, tab2 = _case(tmp_path)
    none, _, _, _, _ = _run(DRV_ORACLE, tmp_path, P2, tab2, 5, campaign=1, tag="notrace")
    assert none["quarantine_index"] == 0.0


DRV_REF = os.path.join(BUILD, "facade_driver_ref")


def _ks_campaign():
    import importlib.util
    spec = importlib.util.spec_from_file_location("ks_campaign", os.path.join(ROOT, "tools", "ks_campaign.py"))
    m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
    return m


@needs_oracle_driver
@pytest.mark.skipif(not os.path.exists(DRV_REF), reason="facade_driver_ref not built (needs /root/reference at build time)")
@pytest.mark.parametrize("campaign", [0, 1, 2])
def test_facade_statistically_equivalent_to_the_unmodified_reference(campaign):
    """The SAME driver source (unmodified simulate_epidemic, Vac_Campaign, reactive_strategy) linked against the reference's own
    Community / Person / Location (arm A) and against the facade over the C ABI (arm B, oracle backend): no campaign, general
    campaign, general campaign + contact tracing + self-quarantine + ring vaccination. 48 replicates
    per arm; two-sample KS per statistic, Holm-Bonferroni at family-wise 0.05 over the pre-declared family (tools/ks_campaign.py):
    15 headline curve statistics + doses / urgent doses / quarantine index where the mode produces them."""
    rep = _ks_campaign().run(backend="oracle", campaign=campaign, agents=12000, days=90, replicates=48)
    t = rep["tests"]
    assert rep["family"]["tests"] == 15 + (0, 1, 3)[campaign]
    assert t["incidence.total"]["ref_mean"] > 1500 and t["incidence.total"]["ours_mean"] > 1500       # an epidemic happened in both arms
    if campaign:
        assert t["campaign.doses"]["ref_mean"] > 5000
    if campaign == 2:
        assert t["campaign.urgent_doses"]["ref_mean"] > 1000 and t["campaign.quarantine_index"]["ref_mean"] > 0.1
    assert rep["pass"], rep["family"]


@pytest.mark.gpu
@pytest.mark.skipif(not (os.path.exists(DRV_CUDA) and os.path.exists(DRV_REF)), reason="facade drivers not built")
def test_cuda_facade_statistically_equivalent_to_the_unmodified_reference_with_ring_vaccination():
    """The same comparison with arm B on the CUDA library (32 replicates per arm; the 100-replicate record is under profiles/)."""
    rep = _ks_campaign().run(backend="cuda", campaign=2, agents=12000, days=90, replicates=32)
    assert rep["family"]["tests"] == 18 and rep["tests"]["campaign.urgent_doses"]["ours_mean"] > 1000
    assert rep["pass"], rep["family"]


def _cache_restore(driver, tmp, P, tab, cache_day, campaign, tag):
    pf = os.path.join(str(tmp), f"params_{tag}.txt"); cabm.params.write_text(P, tab, pf)
    prefix = os.path.join(str(tmp), f"cr_{tag}")
    r = subprocess.run([driver, "cache_restore", pf, str(tmp), "f", prefix, str(cache_day), str(campaign)], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    info = json.loads(r.stdout.strip().splitlines()[-1])
    return info, [open(f"{prefix}.pass{k}.csv").read() for k in (1, 2, 3)]


@needs_oracle_driver
def test_cache_and_restore_check_through_the_facade(tmp_path):
    """The reference's own regression check (tests/cache_and_restore/: equal checksums of the plot logs before and after a
    restore) with its own SimulationCache / restore_from_cache (behavior_tuner.h:39-87,490-508): CommunityLedger copy = device
    snapshot, load_from_cache = device restore + roll-back of the host mirrors, Vac_Campaign::quick_cache, the three generators."""
    pop, P, tab = _case(tmp_path, n=12000, days=100, tracing_from=35)
    info, passes = _cache_restore(DRV_ORACLE, tmp_path, P, tab, cache_day=45, campaign=2, tag="a")
    assert passes[0] == passes[1] == passes[2], "a restored run differs from the run the cache was taken from"
    rows = passes[0].strip().splitlines()
    assert len(rows) == 1 + (100 - 45) + 1 and rows[1].startswith("45,")
    body = np.array([[float(x) for x in r.split(",")] for r in rows[1:-1]])
    assert body[:, 1].sum() > 300 and body[:, 7].sum() > 1000 and body[:, 8].sum() > 0 and body[:, 9].max() > 0   # infections, doses, ring doses, quarantine
    assert info["infections"][0] == info["infections"][1] == info["infections"][2] > 1000
    # the cache matters: with another cache day the tail rows of the shared days are still the same epidemic
    _, other = _cache_restore(DRV_ORACLE, tmp_path, P, tab, cache_day=70, campaign=2, tag="b")
    assert other[0] == other[1] == other[2]
    assert other[0].strip().splitlines()[1:-1] == rows[1 + 25:-1]


def _autotuned(driver, tmp, P, tab, cases, tag, n_agents):
    """simulate_epidemic with par->behavioral_autotuning (exp/active-vac/main.cpp:93-97): the reference's behaviour autotuner
    (behavior_tuner.h) bisects the social-distancing anchor of every 14-day window against reported cases, re-running each window
    from its SimulationCache until the fit is good. The 'empirical' series is half of what the untuned run reported, scaled to
    Florida's population as score_fit expects (behavior_tuner.h:261-265)."""
    import datetime
    d0 = datetime.date(2020, 2, 5)
    csv = os.path.join(str(tmp), f"rcasedeath_{tag}.csv")
    with open(csv, "w") as f:
        f.write("Date,rcase,x,y,rdeath\n")
        for d, c in enumerate(cases):
            f.write(f"{d0 + datetime.timedelta(days=d)},{int(round(c * 0.5 * 21538187 / n_agents))},0,0,0\n")
    Pa = dict(P); Pa.update(seed=5, behavioral_autotuning=1, tuning_window=14, num_preview_windows=3, rcasedeath_file=csv,
                            behavior_output_file=os.path.join(str(tmp), f"anchors_{tag}.csv"))
    pf = os.path.join(str(tmp), f"params_{tag}.txt"); cabm.params.write_text(Pa, tab, pf)
    prefix = os.path.join(str(tmp), f"auto_{tag}")
    r = subprocess.run([driver, "sim", pf, str(tmp), "f", prefix, "1"], capture_output=True, text=True, timeout=1500, cwd=str(tmp))   # gen_simvis writes ./plot_log.csv
    assert r.returncode == 0, r.stderr[-2000:]
    info = json.loads(r.stdout.strip().splitlines()[-1])
    return info, open(prefix + ".plot.csv").read(), open(prefix + ".ledger.csv").read(), open(Pa["behavior_output_file"]).read(), r.stderr


@needs_oracle_driver
def test_behaviour_autotuner_runs_on_the_facade(tmp_path):
    pop, P, tab = _case(tmp_path, n=12000, days=120)
    base, led, _, _, _ = _run(DRV_ORACLE, tmp_path, P, tab, 5, campaign=1, tag="untuned")
    info, plot, ledger, anchors, err = _autotuned(DRV_ORACLE, tmp_path, P, tab, led["det_case_report"], "a", 12000)
    rows = anchors.strip().splitlines()
    assert rows[0] == "date,anchor_val" and len(rows) >= 6
    vals = [float(r.split(",")[1]) for r in rows[1:]]
    assert max(vals) > 0.3 and all(0.0 <= v <= 1.0 for v in vals)            # the tuner had to raise distancing to halve the cases
    assert err.count("Decision made.") >= 4                                   # windows tuned, each after several restores
    assert info["infections"] < 0.7 * base["infections"]
    info2, plot2, ledger2, anchors2, _ = _autotuned(DRV_ORACLE, tmp_path, P, tab, led["det_case_report"], "b", 12000)
    assert (plot2, ledger2, anchors2) == (plot, ledger, anchors)              # restores are exact, so the search is reproducible


@pytest.mark.gpu
@pytest.mark.skipif(not (os.path.exists(DRV_CUDA) and os.path.exists(DRV_ORACLE)), reason="facade drivers not built")
def test_behaviour_autotuner_cuda_equals_oracle(tmp_path):
    pop, P, tab = _case(tmp_path, n=20000, days=120)
    _, led, _, _, _ = _run(DRV_ORACLE, tmp_path, P, tab, 5, campaign=1, tag="untuned")
    ic, plot_c, led_c, anch_c, _ = _autotuned(DRV_CUDA, tmp_path, P, tab, led["det_case_report"], "cuda", 20000)
    io, plot_o, led_o, anch_o, _ = _autotuned(DRV_ORACLE, tmp_path, P, tab, led["det_case_report"], "oracle", 20000)
    assert anch_c == anch_o and led_c == led_o and plot_c == plot_o
    assert ic["infections"] == io["infections"] > 500


@pytest.mark.gpu
@pytest.mark.skipif(not (os.path.exists(DRV_CUDA) and os.path.exists(DRV_ORACLE)), reason="facade drivers not built")
def test_cache_and_restore_check_on_the_cuda_library(tmp_path):
    pop, P, tab = _case(tmp_path, n=30000, days=110, tracing_from=40)
    info, passes = _cache_restore(DRV_CUDA, tmp_path, P, tab, cache_day=50, campaign=2, tag="cuda")
    assert passes[0] == passes[1] == passes[2]
    _, ref = _cache_restore(DRV_ORACLE, tmp_path, P, tab, cache_day=50, campaign=2, tag="oracle")
    assert passes[0] == ref[0], "CUDA library and oracle disagree behind the same facade"
    assert info["infections"][0] > 3000


@pytest.mark.gpu
@pytest.mark.skipif(not (os.path.exists(DRV_CUDA) and os.path.exists(DRV_ORACLE)), reason="facade drivers not built")
@pytest.mark.parametrize("campaign", [0, 1, 2])
def test_facade_cuda_equals_facade_oracle(tmp_path, campaign):
    pop, P, tab = _case(tmp_path, n=30000, days=120, T=0.2, tracing_from=40 if campaign == 2 else None)
    ic, _, plot_c, led_c, _ = _run(DRV_CUDA, tmp_path, P, tab, 9, campaign, tag="cuda")
    io, _, plot_o, led_o, _ = _run(DRV_ORACLE, tmp_path, P, tab, 9, campaign, tag="oracle")
    assert led_c == led_o, "ledger differs between the CUDA library and the oracle behind the same facade"
    assert open(os.path.join(str(tmp_path), "out_cuda.rt.csv")).read() == open(os.path.join(str(tmp_path), "out_oracle.rt.csv")).read(), "Rt / offspring / hSAR differ"
    assert plot_c == plot_o, "simulate_epidemic's plot log differs"
    for k in ("infections", "secondary_links", "doses", "urgent_doses", "quarantine_index", "cumul_death"):
        assert ic[k] == io[k], k
    assert ic["infections"] > 3000 and (ic["doses"] > 0) == (campaign > 0) and (ic["urgent_doses"] > 0) == (campaign == 2)


@pytest.mark.gpu
@pytest.mark.skipif(not (os.path.exists(DRV_CUDA) and os.path.exists(DRV_ORACLE)), reason="facade drivers not built")
def test_ring_vaccination_at_scale_cuda_equals_oracle(tmp_path):
    """BASELINE configs[3] at a size the oracle arm finishes in about two minutes: 300 k agents, 100 days, general campaign +
    contact tracing from day 30 with self-quarantine and ring vaccination of the traced contacts (exp/active-vac/main.cpp:
    101-111,711), the reference's unmodified simulate_epidemic and Vac_Campaign on both backends. Everything the run writes
    must be byte-identical: ledger, plot log, Rt / offspring / household attack rate, dose and quarantine totals."""
    pop, P, tab = _case(tmp_path, n=300_000, days=100, T=0.12, pde=2e-4, tracing_from=30)
    ic, _, plot_c, led_c, _ = _run(DRV_CUDA, tmp_path, P, tab, 3, 2, tag="cuda")
    io, _, plot_o, led_o, _ = _run(DRV_ORACLE, tmp_path, P, tab, 3, 2, tag="oracle")
    assert led_c == led_o and plot_c == plot_o
    assert open(os.path.join(str(tmp_path), "out_cuda.rt.csv")).read() == open(os.path.join(str(tmp_path), "out_oracle.rt.csv")).read()
    for k in ("infections", "secondary_links", "doses", "urgent_doses", "quarantine_index", "cumul_death"):
        assert ic[k] == io[k], k
    assert ic["infections"] > 20_000 and ic["urgent_doses"] > 500 and ic["quarantine_index"] > 0, ic


def _sim_db(driver, tmp, P, tab, tag):
    """Run the driver with the reference's generate_sim_data_db on; a stub `sqlite3` on PATH keeps the CSV tables it is handed."""
    keep = os.path.join(str(tmp), f"db_{tag}"); os.makedirs(keep, exist_ok=True)
    stub = os.path.join(keep, "sqlite3")
    with open(stub, "w") as f:
        f.write("#!/bin/sh\n# stands in for the sqlite3 shell: keep what would have been imported\n"
                "for a in \"$@\"; do case \"$a\" in .import*) set -- $a; cp \"$2\" \"$(dirname \"$0\")/$3.table.csv\";; esac; done\nexit 0\n")
    os.chmod(stub, 0o755)
    pf = os.path.join(str(tmp), f"params_db_{tag}.txt"); cabm.params.write_text(P, tab, pf)
    env = dict(os.environ, PATH=keep + os.pathsep + os.environ["PATH"], CABM_SIM_DB=keep)
    r = subprocess.run([driver, "sim", pf, str(tmp), "f", os.path.join(str(tmp), f"outdb_{tag}"), "1"], capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    tables = {}
    for name in ("infection_history", "secondary_infections", "infection_detection", "vaccination_history", "age_bins", "doses_available", "doses_used"):
        tables[name] = open(os.path.join(keep, name + ".table.csv")).read().splitlines()
    return json.loads(r.stdout.strip().splitlines()[-1]), tables


def _canonical(tables):
    """Object addresses (Infection*, Detection*) printed by the dump differ from run to run: number them in order of first appearance."""
    ids = {}
    def canon(tok):
        if tok.startswith("0x"):
            return "#%d" % ids.setdefault(tok, len(ids))
        return tok
    # ids are handed out while walking infection_history (person order, then history order: the same on every backend); the
    # lines of a table are then sorted, because the order of an infection's secondary cases follows the device log, whose order
    # within a day is not defined
    return {k: sorted(",".join(canon(t) for t in line.split(",")) for line in v) for k, v in tables.items()}


@needs_oracle_driver
def test_sim_data_dump_runs_unmodified_on_the_facade(tmp_path):
    """generate_sim_data_db (simulator.h:642-761) through the facade: every infection of the device log appears once with its
    place / infector / owner ids and timeline, secondary-infection links match the infector column, detections and doses are there."""
    pop, P, tab = _case(tmp_path, n=8000, days=70, T=0.25, pde=5e-4)
    info, t = _sim_db(DRV_ORACLE, tmp_path, P, tab, "oracle")
    hist = [l.split(",") for l in t["infection_history"]]
    assert len(hist) == info["infections"] > 1500
    by_ptr = {h[0]: h for h in hist}
    assert len(by_ptr) == len(hist)
    owners = np.array([int(h[3]) for h in hist]); assert owners.min() >= 0 and owners.max() < int(pop["n_people"])
    n_links = 0
    for line in t["secondary_infections"]:
        src, dst = line.split(",")
        assert by_ptr[dst][2] == by_ptr[src][3]                      # the secondary case's infector id is the source infection's owner
        n_links += 1
    assert n_links == info["secondary_links"] == sum(int(h[19]) for h in hist) > 500
    det = [l.split(",") for l in t["infection_detection"]]
    assert 0 < len(det) <= len(hist) and all(d[1] in by_ptr for d in det)
    assert len(t["vaccination_history"]) == info["doses"] > 0 and len(t["doses_used"]) == len(t["doses_available"]) > 0


@pytest.mark.gpu
@pytest.mark.skipif(not (os.path.exists(DRV_CUDA) and os.path.exists(DRV_ORACLE)), reason="facade drivers not built")
def test_sim_data_dump_cuda_equals_oracle(tmp_path):
    pop, P, tab = _case(tmp_path, n=20000, days=80, T=0.25, pde=5e-4)
    ic, tc = _sim_db(DRV_CUDA, tmp_path, P, tab, "cuda")
    io, to = _sim_db(DRV_ORACLE, tmp_path, P, tab, "oracle")
    assert ic["infections"] == io["infections"] > 3000
    assert _canonical(tc) == _canonical(to)


# ---- binary cache of the population files (covid-abm_b200/facade/pop_cache.h) ----
def _run_env(driver, tmp, P, tab, tag, env):
    pf = os.path.join(str(tmp), f"params_{tag}.txt"); cabm.params.write_text(P, tab, pf)
    prefix = os.path.join(str(tmp), f"out_{tag}")
    e = dict(os.environ); e.update(env); e["CABM_FACADE_TIMING"] = "1"
    r = subprocess.run([driver, "sim", pf, str(tmp), "f", prefix, "1"], capture_output=True, text=True, timeout=900, env=e)
    assert r.returncode == 0, r.stderr[-2000:]
    hits = [json.loads(l.split(" ", 1)[1]) for l in r.stderr.splitlines() if l.startswith("facade_pop_cache ")]
    outs = tuple(open(prefix + ext).read() for ext in (".ledger.csv", ".plot.csv", ".rt.csv"))
    return hits[-1], outs


@needs_oracle_driver
def test_population_cache_gives_the_same_run_and_notices_edited_files(tmp_path):
    """First run with CABM_POP_CACHE parses the text and writes the cache, the second reads the cache: ledger, plot log and Rt
    file are byte-identical to a run without a cache (same RNG consumption at load time: the cache holds what the files say,
    not the drawn attributes). Editing a file (size / mtime) makes the next run parse again; a truncated cache file is ignored."""
    pop, P, tab = _case(tmp_path, n=8000, days=40)
    cache = tmp_path / "cache"; cache.mkdir()
    st0, plain = _run_env(DRV_ORACLE, tmp_path, P, tab, "plain", {})
    assert st0 == {"locations": 0, "population": 0} and not list(cache.iterdir())
    st1, cold = _run_env(DRV_ORACLE, tmp_path, P, tab, "cold", {"CABM_POP_CACHE": str(cache)})
    files = sorted(f.name for f in cache.iterdir())
    assert st1 == {"locations": 0, "population": 0} and len(files) == 2 and files[0].startswith("cabm_loc_") and files[1].startswith("cabm_pop_")
    st2, warm = _run_env(DRV_ORACLE, tmp_path, P, tab, "warm", {"CABM_POP_CACHE": str(cache)})
    assert st2 == {"locations": 1, "population": 1}
    assert plain == cold == warm
    assert float(plain[0].count("\n")) > 30
    # an edited population file: new key, parsed again (the locations cache still serves), results follow the edit
    popfile = tmp_path / "population-f.txt"
    lines = popfile.read_text().splitlines()
    pid, hid, sex, age, did = lines[-1].split()[:5]
    lines[-1] = " ".join([pid, hid, sex, str((int(age) + 30) % 90), did])
    popfile.write_text("\n".join(lines) + "\n")
    st3, edited = _run_env(DRV_ORACLE, tmp_path, P, tab, "edited", {"CABM_POP_CACHE": str(cache)})
    assert st3 == {"locations": 1, "population": 0} and len(list(cache.iterdir())) == 3
    st3b, edited_plain = _run_env(DRV_ORACLE, tmp_path, P, tab, "edited_plain", {})
    assert edited == edited_plain
    # a damaged cache file is not trusted
    for f in cache.iterdir():
        data = f.read_bytes(); f.write_bytes(data[: len(data) // 2])
    st4, again = _run_env(DRV_ORACLE, tmp_path, P, tab, "damaged", {"CABM_POP_CACHE": str(cache)})
    assert st4 == {"locations": 0, "population": 0} and again == edited
