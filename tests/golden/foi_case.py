"""The scripted epidemic behind tests/golden/foi_reference.npz: shared by make_foi.py (reference arm, run once where
/root/reference exists) and tests/test_foi_reference.py (oracle and CUDA arms, run anywhere).

A synthetic community, three waves of scripted infections (days 0, 6 and 15; four strains), no importation, and
susceptibility switched off whenever the day ticks, so that nobody is infected except by script: the state of every day is a
deterministic function of the script on both sides and the per-location exposure probabilities can be compared number for
number. Social distancing ramps up over the run so that the between-household gate opens and closes; schools close on day 30.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))

N_AGENTS, DAYS, SEED = 6000, 45, 77
MUT = ("2020-02-25", "2020-03-10", "2020-03-25")


def build():
    import covid_abm_b200 as cabm
    import synthpop
    pop = synthpop.generate(N_AGENTS, seed=909)
    P = cabm.params.default_params(run_length=DAYS, seed=SEED, T=0.2, prob_daily_exposure=0.0, mutant_intro_dates=MUT)
    sd = np.clip(np.linspace(-0.1, 0.9, DAYS), 0.0, 1.0)
    P["social_distancing"] = [float(x) for x in sd]
    sc = np.zeros(DAYS); sc[30:] = 1.0
    P["school_closure"] = [float(x) for x in sc]
    tab = cabm.params.build_tables(P)
    n = int(pop["n_people"])
    rng = np.random.default_rng(5)
    seeds = []
    for day, k in ((0, 500), (6, 350), (15, 350)):
        for pid in rng.choice(n, size=k, replace=False):
            seeds.append((day, int(pid), int(rng.integers(0, 4))))
    return pop, P, tab, seeds


def no_susceptibility(P):
    Q = dict(P)
    Q["susceptibility_by_age"] = [0.0] * len(P["susceptibility_by_age"])
    return Q
