#!/usr/bin/env python3
"""End-to-end timing of the reference's unmodified simulate_epidemic (simulator.h:265-533) through the C++ facade.

    python tools/facade_e2e.py [--agents N] [--days D] [--backend cuda|oracle] [--campaign 0|1]

Writes the synthetic population in the reference's five text formats, runs covid-abm_b200/facade/_build/facade_driver_*
and prints its JSON line (wall-clock seconds of simulate_epidemic: tick + importation + every per-day getter and
diagnostic the reference's loop calls, with the host-side proxies kept in sync).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--agents", type=int, default=2_700_000)
    ap.add_argument("--days", type=int, default=120)
    ap.add_argument("--backend", default="cuda", choices=["cuda", "oracle"])
    ap.add_argument("--campaign", type=int, default=0)
    ap.add_argument("--seed", type=int, default=1)
    args = ap.parse_args()
    import covid_abm_b200 as cabm
    import synthpop
    import bench
    drv = os.path.join(ROOT, "covid-abm_b200", "facade", "_build", f"facade_driver_{args.backend}")
    if not os.path.exists(drv):
        sys.exit(f"{drv} missing: make -C covid-abm_b200/facade (needs /root/reference at build time)")
    pop = synthpop.generate(args.agents)
    P = cabm.params.default_params(run_length=args.days, seed=args.seed, mutant_intro_dates=bench.MUTANTS)
    T = cabm.params.build_tables(P)
    tmp = tempfile.mkdtemp(prefix="cabm_facade_")
    t0 = time.time()
    synthpop.write_reference_files(pop, tmp, "e2e")
    cabm.params.write_text(P, T, os.path.join(tmp, "params.txt"))
    write_s = time.time() - t0
    r = subprocess.run([drv, "sim", os.path.join(tmp, "params.txt"), tmp, "e2e", os.path.join(tmp, "out"), str(args.campaign)], capture_output=True, text=True,
                       env=dict(os.environ, CABM_FACADE_TIMING="1"))
    if r.returncode != 0:
        sys.exit(r.stderr[-3000:])
    out = json.loads(r.stdout.strip().splitlines()[-1])
    for line in r.stderr.splitlines():
        if line.startswith("facade_timing "):                  # seconds and calls per facade entry point (CABM_FACADE_TIMING)
            out["facade_timing_s_calls"] = json.loads(line[len("facade_timing "):])
    out.update(backend=args.backend, campaign=args.campaign, write_population_files_s=round(write_s, 2),
               ms_per_day=out["simulate_epidemic_s"] * 1e3 / out["days"])
    subprocess.run(["rm", "-rf", tmp])
    print(json.dumps(out))


if __name__ == "__main__":
    main()
