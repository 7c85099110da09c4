#!/usr/bin/env python3
"""Host wall time of the per-day C-ABI calls (cabm_tick, cabm_import, cabm_ledger_row) against the device time of the same days.

    python tools/e2e_breakdown.py [--agents N] [--start-day D] [--days K]
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--agents", type=int, default=2_700_000)
    ap.add_argument("--start-day", type=int, default=30)
    ap.add_argument("--days", type=int, default=200)
    a = ap.parse_args()
    import numpy as np
    import covid_abm_b200 as cabm
    import synthpop
    import bench
    pop = synthpop.generate(a.agents)
    P = cabm.params.default_params(run_length=a.start_day + a.days + 1, seed=1, mutant_intro_dates=bench.MUTANTS)
    T = cabm.params.build_tables(P)
    n = int(pop["n_people"])
    c = cabm.Community(P, pop, T, attrs=cabm.abi.immunity_attributes(n, P["seed"]))
    c.run_days(a.start_day)
    c.device_time_ms()
    row = np.zeros(cabm.abi.N_COUNTERS, dtype=np.uint64)
    prow = row.ctypes.data_as(C.POINTER(C.c_uint64))
    t_tick = t_mix = t_imp = t_row = 0.0
    t0 = time.perf_counter()
    for _ in range(a.days):
        d = c.day
        s = time.perf_counter(); c.tick(d); e = time.perf_counter(); t_tick += e - s
        s = e; c.update_strain_mix(d); e = time.perf_counter(); t_mix += e - s
        s = e; c.seed_epidemic(d); e = time.perf_counter(); t_imp += e - s
        c.day += 1
        s = e; c._check(c.fn.ledger_row(c._h, d, prow)); e = time.perf_counter(); t_row += e - s
    wall = time.perf_counter() - t0
    ms, launches = c.device_time_ms()
    k = 1e6 / a.days
    print(json.dumps({"agents": n, "days": a.days, "wall_us_per_day": wall * k, "device_us_per_day": ms * 1e3 / a.days, "tick_us": t_tick * k, "strain_mix_us": t_mix * k,
                      "import_us": t_imp * k, "ledger_row_us": t_row * k}))


if __name__ == "__main__":
    main()
