#!/usr/bin/env python3
"""Per-kernel timing of the day pipeline at a chosen epidemic phase (development harness, not the benchmark).

    python tools/profile_day.py [--agents N] [--start-day D] [--days K] [--graph]

Runs D days through the CUDA-graph path, then K days either with one CUDA event in front of every kernel
(default; prints microseconds per kernel per day and the work counters of those days) or, with --graph,
through graph replay (prints device microseconds per day). Under `ncu`, kernel launches of day d start at
launch index 1 + 9 d (one k_init_state launch, then 9 kernels per day).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--agents", type=int, default=2_700_000)
    ap.add_argument("--start-day", type=int, default=150)
    ap.add_argument("--days", type=int, default=10)
    ap.add_argument("--graph", action="store_true")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--chunk", type=int, default=0, help="print one line per CHUNK days instead of one for the whole window")
    args = ap.parse_args()
    import covid_abm_b200 as cabm
    import synthpop
    import bench
    pop = synthpop.generate(args.agents)
    P = cabm.params.default_params(run_length=args.start_day + args.days + 1, seed=args.seed, mutant_intro_dates=bench.MUTANTS)
    T = cabm.params.build_tables(P)
    n = int(pop["n_people"])
    A = cabm.Community(P, pop, T, attrs=cabm.abi.immunity_attributes(n, P["seed"]))
    if args.start_day:
        A.run_days(args.start_day)
    A.device_time_ms()
    if not args.graph:
        A.set_profile(True); A.kernel_times()
    chunk = args.chunk if args.chunk > 0 else args.days
    for d0 in range(0, args.days, chunk):
        nd = min(chunk, args.days - d0)
        w0 = A.work_counters()
        A.run_days(nd)
        ms, launches = A.device_time_ms()
        w1 = A.work_counters()
        out = {"agents": n, "start_day": args.start_day + d0, "days": nd, "device_us_per_day": round(ms * 1e3 / nd, 2), "launches": launches,
               "agent_days_per_s": n * nd / (ms * 1e-3), "work_per_day": {k: (w1[k] - w0[k]) // nd for k in w1}}
        if not args.graph:
            out["kernels_us_per_day"] = {k: round(v[0] * 1e3 / nd, 2) for k, v in A.kernel_times().items() if v[1]}
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
