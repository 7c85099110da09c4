"""Per-location force of infection against the reference's own Community::tick (north_star: rtol 1e-6).

tests/golden/foi_reference.npz was written by tests/golden/make_foi.py from an instrumented build of the UNMODIFIED
reference (Community.cpp compiled with oracle/foi_probe.h: each `T = 1.0 - exp(-hazard)` of Community.cpp:846,868,903,923,
937,965 logs location, -hazard and infectious weight) running the scripted epidemic of tests/golden/foi_case.py: 1200
scripted infections in three waves and four strains, 45 days, nobody infected except by script, so the state of every day is
the same function of the script on both sides. Here the same script runs through the C ABI (cabm_infect / cabm_tick /
cabm_dump_foi) and every one of the reference's 13 368 exposure probabilities must be reproduced, and nothing else reported:

  CPU : the oracle restatement;   GPU : the CUDA library.

What this pins beyond the formulas: the hot-location calendar (who is infectious where on which day, hospital stays, severe
cases staying home), the public-activity visitor census in the workplace denominators (the reference's visit draws are served
from the per-person Philox streams), hospital census after the status pass, gates (school closure, weekday, social
distancing vs household riskiness), seasonality, and Person::infect for every scripted case (final records compared).
The reference's own ledger of the run (20 daily series + cumulative outcome tallies, dumped by the same probe run) must be
reproduced bit for bit as well: that pins updatePersonStatus and the ledger writers.
"""
import os

import numpy as np
import pytest

import covid_abm_b200 as cabm
import oracle_lib

HERE = os.path.dirname(os.path.abspath(__file__))
import sys
sys.path.insert(0, os.path.join(HERE, "golden"))
import foi_case                                                   # noqa: E402

GOLD = os.path.join(HERE, "golden", "foi_reference.npz")
RTOL = 1e-6                     # north_star's tolerance; the observed agreement is 3.4e-11 (asserted at 1e-9 below)
# source line of Community.cpp -> (location type, field of cabm_foi_row)
LINE = {846: (0, "T"), 868: (0, "T_social"), 903: (1, "T"), 923: (2, "T"), 937: (3, "T"), 965: (4, "T")}


def _run(make):
    pop, P, tab, seeds = foi_case.build()
    P0 = foi_case.no_susceptibility(P)
    c = make(P, pop, tab)
    by_day = {}
    for d, pid, st in seeds:
        by_day.setdefault(d, []).append((pid, st))
    took, rows = [], []
    for d in range(foi_case.DAYS):
        c.set_params(P, tab)
        for pid, st in by_day.get(d, []):
            took.append((d, pid, int(c.infect(pid, st, d))))
        c.set_params(P0, tab)                                      # Person.cpp:280-282: every exposure of the day fails
        c.tick(d)
        rows.append(c.foi().copy())
    hist = c.infection_histories()
    cur = c.current_infections()
    ledger = c.ledger_table(); cumul = c.getCumulIncidenceByOutcome()
    c.close()
    return took, rows, hist, cur, pop, P, ledger, cumul


def _compare(took, rows, hist, cur, pop, P, ledger, cumul):
    G = np.load(GOLD)
    # The reference's own CommunityLedger of the same run (Community.h:112-130; the 13 tallies of updatePersonStatus,
    # Community.cpp:682-741, and the future-dated reportCase / reportDeath writes, :657-679), number for number, every day.
    bad = np.argwhere(ledger != G["ledger"])
    assert bad.size == 0, f"ledger differs from the reference's at (day, counter) {bad[:6].tolist()}"
    assert np.array_equal(cumul, G["cumul"]), (cumul, G["cumul"])
    assert G["ledger"][:, 0].sum() == G["seeds"][:, 2].sum() and G["ledger"][:, 8].sum() > 100 and G["ledger"][:, 12].sum() > 100
    risk, nbr_off, nbr = np.asarray(pop["loc_risk"]), np.asarray(pop["nbr_off"]), np.asarray(pop["nbr"])
    assert np.array_equal(np.array(took, dtype=np.int32), G["seeds"]), "Person::infect outcome of a scripted case differs"
    assert len(hist) == int(G["seeds"][:, 2].sum()), "somebody was infected outside the script"
    # final infection record of every infected person (timeline, strain, relative infectiousness)
    for k, pid in enumerate(G["hist"][:, 0]):
        rec = cur[pid]
        want_t = np.clip(G["hist"][k, 1:13].astype(np.int64), -32768, 32767)
        assert np.array_equal(rec["t"].astype(np.int64), want_t), pid
        assert rec["strain"] == G["hist"][k, 13] and np.float32(rec["rel_infectiousness"]) == G["hist_rel"][k], pid
    worst, n_checked = 0.0, 0
    for d in range(foi_case.DAYS):
        r = rows[d]
        sel = G["foi_day"] == d
        seen = set()
        for line, loc, x, w in zip(G["foi_line"][sel], G["foi_loc"][sel], G["foi_x"][sel], G["foi_w"][sel]):
            typ, field = LINE[int(line)]
            i = np.searchsorted(r["loc"], loc)
            assert i < len(r) and r["loc"][i] == loc and r["type"][i] == typ, f"day {d}: reference computed a T for location {loc} (line {line}), not reported here"
            t_ref = 1.0 - np.exp(x)                                 # the reference's own expression on its own -hazard
            t_our = float(r[field][i])
            assert t_our > 0 or t_ref == 0, f"day {d} loc {loc} line {line}: gated off here, T={t_ref} in the reference"
            err = abs(t_our - t_ref) / max(t_ref, 1e-300)
            assert err <= RTOL, f"day {d} loc {loc} line {line}: T {t_our} vs reference {t_ref}"
            worst = max(worst, err)
            assert abs(float(r["w"][i].sum()) - w) <= 1e-9 * max(w, 1.0), f"day {d} loc {loc}: infectious weight {r['w'][i].sum()} vs {w}"
            seen.add((int(loc), field)); n_checked += 1
        # nothing the reference did not compute: every positive T reported here has its record
        for field in ("T", "T_social"):
            for loc in r["loc"][r[field] > 0]:
                if field == "T_social" and (int(loc), field) not in seen:
                    # the row carries the household's between-household T as soon as its own gate is open; the reference
                    # evaluates it per neighbour that passes the same gate (Community.cpp:863-869): none here
                    sd = P["social_distancing"][d]
                    assert not np.any(risk[nbr[nbr_off[loc]:nbr_off[loc + 1]]].astype(np.float64) > sd), \
                        f"day {d}: household {loc} has an eligible neighbour, the reference computed no between-household T"
                    continue
                assert (int(loc), field) in seen, f"day {d}: {field} > 0 at location {loc}, the reference computed none"
    assert n_checked == len(G["foi_x"]) > 13000
    # Observed 3.4e-11: infectious pressure is summed in 2^-32 fixed-point units (order-independent atomics), i.e. up to 2^-33
    # absolute on a weight of >= 0.25, and 1 - exp(-h) loses relative precision for small h on both sides.
    assert worst < 1e-9, worst
    return worst


@pytest.mark.skipif(not os.path.exists(GOLD), reason="golden file missing")
def test_oracle_force_of_infection_equals_the_reference():
    _compare(*_run(lambda P, pop, tab: oracle_lib.oracle_community(P, pop, tab)))


@pytest.mark.gpu
def test_cuda_force_of_infection_equals_the_reference():
    _compare(*_run(lambda P, pop, tab: cabm.Community(P, pop, tab)))
