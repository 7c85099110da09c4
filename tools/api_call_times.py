#!/usr/bin/env python3
"""Wall time per call of the C-ABI entry points a facade day uses besides cabm_tick (2.7 M agents by default)."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--agents", type=int, default=2_700_000)
    ap.add_argument("--day", type=int, default=60)
    a = ap.parse_args()
    import numpy as np
    import covid_abm_b200 as cabm
    import synthpop
    import bench
    pop = synthpop.generate(a.agents)
    P = cabm.params.default_params(run_length=a.day + 40, seed=1, mutant_intro_dates=bench.MUTANTS)
    T = cabm.params.build_tables(P)
    n = int(pop["n_people"])
    c = cabm.Community(P, pop, T, attrs=cabm.abi.immunity_attributes(n, P["seed"]))
    c.run_days(a.day)
    rng = np.random.default_rng(0)
    out = {}

    def timeit(name, fn, reps=20):
        fn()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        out[name] = round((time.perf_counter() - t0) / reps * 1e6, 1)
    d = a.day - 1
    timeit("diagnostics_us", lambda: c.diagnostics(d))
    timeit("ledger_table_us", lambda: c.ledger_table())
    timeit("detected_today_us", lambda: c.detected_today(d))
    ids10 = rng.choice(n, size=10, replace=False).astype(np.uint32)
    day = [a.day]

    def inf():
        c.infect_many(ids10, 0, c.day)
    timeit("infect_batch_10_us", inf)
    timeit("histories_new_us", lambda: c.infection_histories(), reps=3)
    ids = rng.choice(n, size=n // 100, replace=False).astype(np.uint32)
    timeit("vaccinate_1pct_us", lambda: c.vaccinate(ids, c.day), reps=5)
    timeit("quarantine_1pct_us", lambda: c.quarantine(ids, c.day, c.day + 3), reps=5)
    timeit("serosurvey_1pct_us", lambda: c.doSerosurvey(ids, d), reps=5)
    timeit("step_us", lambda: c.step(), reps=20)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
