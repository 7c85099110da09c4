#!/usr/bin/env python3
"""A small run that touches every kernel and every exchange path, for compute-sanitizer (SURVEY.md 5 "race detection").

    compute-sanitizer --tool memcheck|racecheck|initcheck|synccheck python tools/sanitize_case.py

One tool per gpurun call (B200_PROFILING.md). The case: a 12 000-agent community seeded hard enough that every location type
is hot within days; cabm_infect / cabm_infect_batch, per-day cabm_tick + cabm_import, cabm_run_days (graph replay),
vaccination, quarantine, diagnostics, both serosurveys, detected_today, snapshot/restore, the force-of-infection dump; a batch of
four communities through cabm_batch_run_days; and one community advanced by two handles on this device with the exchange
buffers copied between them (the sharded day parts, cabm_shard_flags_in). Prints one JSON line with checksums so that a
sanitizer run can be compared with a plain run.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    import numpy as np
    import covid_abm_b200 as cabm
    import synthpop
    days = int(os.environ.get("CABM_SANITIZE_DAYS", 14))
    pop = synthpop.generate(12000, seed=5)
    n = int(pop["n_people"])
    mut = ("2020-02-12", "2020-02-15", "2020-02-18")
    P = cabm.params.default_params(run_length=days + 4, seed=7, T=0.3, prob_daily_exposure=3e-3, mutant_intro_dates=mut)
    tab = cabm.params.build_tables(P)
    out = {}
    # ---- one handle: every entry point of the day path ----
    c = cabm.Community(P, pop, tab)
    c.foi()
    for pid in range(0, 1200, 3):
        c.infect(pid, 0, 0)
    more = np.arange(1, 3000, 7, dtype=np.uint32)
    c.infect_many(more, (np.arange(more.size) % 4).astype(np.uint8), 0)
    rng = np.random.default_rng(0)
    for d in range(days // 2):
        c.step()
        if d == 2:
            c.vaccinate(rng.choice(n, size=n // 3, replace=False), d)
            c.quarantine(rng.choice(n, size=n // 10, replace=False), d, d + 5)
        if d == 3:
            c.foi(); c.diagnostics(d); c.doSerosurvey(rng.choice(n, size=500, replace=False), d); c.detected_today(d)
    snap = c.snapshot()
    c.run_days(days - days // 2)
    first = c.ledger_table().copy()
    c.restore(snap)
    c.run_days(days - days // 2)
    assert np.array_equal(first, c.ledger_table())
    c.snap_free(snap)
    import ctypes as C
    pos = C.c_uint64(0)
    ids = np.arange(0, n, 3, dtype=np.uint32)
    c._check(c.fn.serosurvey_static(c._h, days - 1, ids.ctypes.data_as(C.POINTER(C.c_uint32)), ids.size, C.byref(pos)))
    out["single"] = {"infections": int(first[:, 0].sum()), "deaths": int(first[:, 4].sum()), "histories": int(c.infection_histories().size),
                     "seropos_static": int(pos.value), "work": c.work_counters()}
    c.close()
    # ---- a batch of four communities, one launch per kernel ----
    comms = []
    for r in range(4):
        Pr = cabm.params.default_params(run_length=days, seed=20 + r, T=0.25 + 0.02 * r, prob_daily_exposure=3e-3, mutant_intro_dates=mut)
        b = cabm.Community(Pr, pop, cabm.params.build_tables(Pr))
        for pid in range(0, 900, 3):
            b.infect(pid, 0, 0)
        comms.append(b)
    cabm.replicates.batch_run_days(comms, days)
    out["batch"] = [int(b.ledger_table()[:, 0].sum()) for b in comms]
    for b in comms:
        b.close()
    # ---- one community over two handles on this device ----
    import torch
    sh = cabm.sharded.local_sharded(P, pop, tab, 2, device=torch.device("cuda", 0))
    for pid in range(0, 1200, 3):
        sh.infect(pid, 0, 0)
    for d in range(days):
        if d == 3:
            sh.quarantine(rng.choice(n, size=n // 10, replace=False), d, d + 5)
        sh.step()
    out["sharded"] = {"infections": int(sh.ledger_table()[:, 0].sum())}
    sh.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
