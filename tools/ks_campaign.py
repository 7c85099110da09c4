#!/usr/bin/env python3
"""Statistical equivalence under vaccination campaign, contact tracing and ring vaccination: facade vs the UNMODIFIED reference.

    python tools/ks_campaign.py [--backend oracle|cuda] [--campaign 0|1|2] [--agents N] [--days D] [--replicates R] [--out file.json]

Both arms run the SAME driver source (covid-abm_b200/facade/facade_driver.cpp, i.e. the reference's unmodified
simulate_epidemic, Vac_Campaign and reactive_strategy) on the same synthetic population and parameter file:
  arm A  facade_driver_ref   : linked against the reference's own Community.cpp / Person.cpp / Location.cpp (seeds 1..R)
  arm B  facade_driver_cuda / facade_driver_oracle : linked against the drop-in facade over the C ABI (seeds 1001..1000+R)
campaign 1 attaches a general Vac_Campaign (Community::vaccinate, dose bookkeeping), 2 adds contact tracing from day 30 with
self-quarantine and ring vaccination of the traced contacts (exp/active-vac/main.cpp:101-111,711) -- SURVEY.md 8a rows a18-a20.

Decision rule (fixed before the runs, same as tools/ks_equivalence.py): two-sample KS per statistic, Holm-Bonferroni at
family-wise 0.05 over the pre-declared family = 15 headline curve statistics ({incidence, hospitalisation, death} x {total,
cum_at_1/3, cum_at_2/3, peak_7d_mean, peak_day}) + the campaign statistics the mode produces (doses, urgent doses,
quarantine index). `pass` is that decision.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
BUILD = os.path.join(ROOT, "covid-abm_b200", "facade", "_build")
CURVES = {"incidence": "inf", "hospitalisation": "hosp_inc", "death": "dead"}
CAMPAIGN_STATS = {0: (), 1: ("doses",), 2: ("doses", "urgent_doses", "quarantine_index")}


def run_one(driver, tmp, P, tab, seed, campaign, tag, cabm, ks):
    Pr = dict(P); Pr["seed"] = seed
    pf = os.path.join(tmp, f"params_{tag}_{seed}.txt"); cabm.params.write_text(Pr, tab, pf)
    prefix = os.path.join(tmp, f"out_{tag}_{seed}")
    r = subprocess.run([driver, "sim", pf, tmp, "f", prefix, str(campaign)], capture_output=True, text=True, timeout=1800)
    if r.returncode != 0:
        raise RuntimeError(f"{driver} seed {seed}: {r.stderr[-500:]}")
    info = json.loads(r.stdout.strip().splitlines()[-1])
    led = np.genfromtxt(prefix + ".ledger.csv", delimiter=",", names=True)
    days = int(P["run_length"])
    out = {}
    for name, col in CURVES.items():
        s = ks.summaries(np.asarray(led[col], dtype=np.float64), days)
        out.update({f"{name}.{k}": v for k, v in s.items() if k in ks.HEADLINE_STATS})
    for k in CAMPAIGN_STATS[campaign]:
        out[f"campaign.{k}"] = float(info[k])
    for f in (pf, prefix + ".ledger.csv", prefix + ".plot.csv", prefix + ".rt.csv"):
        if os.path.exists(f):
            os.remove(f)
    return out


def compare(A, B, ks):
    from scipy.stats import ks_2samp
    tests, ps = {}, {}
    for key in A[0]:
        a = np.array([x[key] for x in A]); b = np.array([x[key] for x in B])
        r = ks_2samp(a, b)
        tests[key] = {"p": float(r.pvalue), "D": float(r.statistic), "ref_mean": float(a.mean()), "ours_mean": float(b.mean())}
        ps[key] = float(r.pvalue)
    adj = ks.holm(ps)
    return tests, {"tests": len(ps), "min_raw_p": min(ps.values()), "min_holm_p": min(adj.values()), "pass": bool(min(adj.values()) >= 0.05),
                   "smallest": sorted(adj.items(), key=lambda kv: kv[1])[:3]}


def run(backend="oracle", campaign=2, agents=12000, days=90, replicates=48, threads=None, T=0.2, pde=3e-4, tracing_from=30):
    import importlib.util
    import covid_abm_b200 as cabm
    import synthpop
    spec = importlib.util.spec_from_file_location("ks_equivalence", os.path.join(ROOT, "tools", "ks_equivalence.py"))
    ks = importlib.util.module_from_spec(spec); spec.loader.exec_module(ks)
    ref, ours = os.path.join(BUILD, "facade_driver_ref"), os.path.join(BUILD, f"facade_driver_{backend}")
    for d in (ref, ours):
        if not os.path.exists(d):
            raise FileNotFoundError(f"{d}: make -C covid-abm_b200/facade (needs /root/reference at build time)")
    pop = synthpop.generate(agents, seed=4242)
    P = cabm.params.default_params(run_length=days, seed=1, T=T, prob_daily_exposure=pde, mutant_intro_dates=("2020-03-11", "2020-04-20", "2030-01-01"))
    if campaign == 2:
        P["begin_contact_tracing"] = tracing_from; P["quarantine_probability"] = [0.9, 0.75, 0.5]
    tab = cabm.params.build_tables(P)
    tmp = tempfile.mkdtemp(prefix="cabm_ksc_")
    synthpop.write_reference_files(pop, tmp, "f")
    threads = threads or (os.cpu_count() or 1)
    with ThreadPoolExecutor(max_workers=threads) as ex:
        A = list(ex.map(lambda s: run_one(ref, tmp, P, tab, s, campaign, "ref", cabm, ks), range(1, replicates + 1)))
    with ThreadPoolExecutor(max_workers=threads if backend == "oracle" else min(threads, 8)) as ex:
        B = list(ex.map(lambda s: run_one(ours, tmp, P, tab, s, campaign, backend, cabm, ks), range(1001, 1001 + replicates)))
    subprocess.run(["rm", "-rf", tmp])
    tests, family = compare(A, B, ks)
    return {"what": "facade (arm B) vs unmodified reference (arm A) through the same facade_driver.cpp", "backend": backend, "campaign": campaign,
            "agents": int(pop["n_people"]), "days": days, "replicates": replicates, "T": T, "pde": pde,
            "procedure": "Holm-Bonferroni at family-wise 0.05 over the pre-declared family (module docstring)", "family": family,
            "pass": family["pass"], "tests": tests}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="oracle", choices=["oracle", "cuda"])
    ap.add_argument("--campaign", type=int, default=2, choices=[0, 1, 2])
    ap.add_argument("--agents", type=int, default=12000)
    ap.add_argument("--days", type=int, default=90)
    ap.add_argument("--replicates", type=int, default=48)
    ap.add_argument("--threads", type=int, default=None)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    rep = run(a.backend, a.campaign, a.agents, a.days, a.replicates, a.threads)
    txt = json.dumps(rep, indent=1)
    if a.out:
        open(a.out, "w").write(txt + "\n")
    print(txt)


if __name__ == "__main__":
    main()
