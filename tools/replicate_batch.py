#!/usr/bin/env python3
"""configs[4]-style calibration batch: R independent replicates of a small community advancing concurrently on ONE GPU.

    python tools/replicate_batch.py [--agents 100000] [--days 200] [--replicates 1,4,16,32]

Every replicate is its own handle (own stream, own dynamic state; include/cabm.h: distinct handles are independent), driven
by its own host thread; the day pipelines of different replicates overlap on the device. Prints one JSON line per R with
the aggregate agent-days/s by wall clock around the whole batch (graph replay, state resident) -- the single-community
pipeline is latency-bound at this size, so concurrent replicates fill the idle SMs.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--agents", type=int, default=100_000)
    ap.add_argument("--days", type=int, default=200)
    ap.add_argument("--replicates", default="1,4,16,32")
    ap.add_argument("--mode", default="streams", choices=["streams", "batched"],
                    help="streams: one handle, stream and host thread per replicate; batched: cabm_batch_run_days, one launch per kernel for all replicates")
    args = ap.parse_args()
    import numpy as np
    import covid_abm_b200 as cabm
    import synthpop
    pop = synthpop.generate(args.agents)
    n = int(pop["n_people"])
    for R in (int(x) for x in args.replicates.split(",")):
        comms = []
        for r in range(R):
            P = cabm.params.default_params(run_length=args.days + 20, seed=cabm.replicates.replicate_seed(1, 0, r, 1), T=0.035, prob_daily_exposure=1e-4)
            comms.append(cabm.Community(P, pop, cabm.params.build_tables(P) if r == 0 else comms[0].T))
        for c in comms:
            c.run_days(20)                                   # warm-up: graph capture, first touch
        if args.mode == "batched":
            cabm.replicates.batch_run_days(comms, 5)          # first touch of the batch path
            for c in comms:
                c.device_time_ms()
            t0 = time.perf_counter()
            cabm.replicates.batch_run_days(comms, args.days - 5)
            wall = (time.perf_counter() - t0) * args.days / (args.days - 5)
        else:
            barrier = threading.Barrier(R + 1)

            def work(c):
                barrier.wait()
                c.run_days(args.days)
            th = [threading.Thread(target=work, args=(c,)) for c in comms]
            for t in th:
                t.start()
            barrier.wait(); t0 = time.perf_counter()
            for t in th:
                t.join()
            wall = time.perf_counter() - t0
        dev = [c.device_time_ms()[0] for c in comms]
        inf = [int(c.ledger("new_infected").sum()) for c in comms]
        print(json.dumps({"mode": args.mode, "replicates": R, "agents": n, "days": args.days, "wall_s": round(wall, 4), "agent_days_per_s": R * n * args.days / wall,
                          "per_replicate_device_ms_max": round(max(dev), 2), "us_per_replicate_day_wall": round(wall * 1e6 / args.days, 1),
                          "infections_min_max": [min(inf), max(inf)]}), flush=True)
        for c in comms:
            c.close()


if __name__ == "__main__":
    main()
