#!/usr/bin/env python3
"""Generates tests/golden/foi_reference.npz from the UNMODIFIED reference (needs /root/reference; run once, commit the file).

    python tests/golden/make_foi.py

oracle/_ref/ref_driver_probe is the reference's own Community.cpp / Person.cpp / ... compiled where they lie, Community.cpp
with the compile-time probes of oracle/foi_probe.h: every `T = 1.0 - exp(-hazard)` of Community::tick logs (source line,
location, -hazard, infectious weight). Its mode `foi` (oracle/ref_driver_kat.inc) runs the scripted epidemic of foi_case.py
through Community::infect and Community::tick. Stored per record: day, source line, location, -hazard, infectious weight;
per scripted infection whether Person::infect succeeded; per person the final infection record; and the reference's own
ledger of the run (the 20 daily series of CommunityLedger in CABM_C_* order + the cumulative outcome tallies).
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import foi_case                                                    # noqa: E402  (also puts the repo root and tools/ on sys.path)

DRIVER = os.path.join(foi_case.ROOT, "oracle", "_ref", "ref_driver_probe")


def main():
    import covid_abm_b200 as cabm
    import synthpop
    if not os.path.exists(DRIVER):
        subprocess.check_call(["make", "-C", os.path.join(foi_case.ROOT, "oracle"), "ref"])
    pop, P, tab, seeds = foi_case.build()
    with tempfile.TemporaryDirectory() as tmp:
        synthpop.write_reference_files(pop, tmp, "f")
        pf = os.path.join(tmp, "params.txt"); cabm.params.write_text(P, tab, pf)
        sf = os.path.join(tmp, "script.txt")
        with open(sf, "w") as f:
            for d, pid, st in seeds:
                f.write(f"seed {d} {pid} {st}\n")
            f.write(f"days {foi_case.DAYS}\n")
        eff, xprot, dur = cabm.abi.immunity_attributes(int(pop["n_people"]), P["seed"])       # what Community() hands the C ABI by default
        af = os.path.join(tmp, "attrs.txt")
        with open(af, "w") as f:
            for pid in range(int(pop["n_people"])):
                f.write(f"{pid} {float(eff[pid]):.17g} {float(xprot[pid]):.9g} {int(dur[pid])}\n")
        out = os.path.join(tmp, "foi.txt")
        subprocess.check_call([DRIVER, "foi", pf, tmp, "f", sf, out, af])
        seed_rows, foi_rows, hist_rows, ledger_rows, cumul = [], set(), [], [], []
        for line in open(out):
            w = line.split()
            if w[0] == "seed":
                seed_rows.append((int(w[1]), int(w[2]), int(w[3])))
            elif w[0] == "foi":                                     # the between-household line repeats per neighbour: keep one
                foi_rows.add((int(w[1]), int(w[2]), int(w[3]), float(w[4]), float(w[5])))
            elif w[0] == "hist":
                hist_rows.append([int(x) for x in w[1:15]] + [float(w[15])])
            elif w[0] == "ledger":                                  # the reference's own CommunityLedger, day by day (CABM_C_* order)
                ledger_rows.append([int(x) for x in w[2:]])
            elif w[0] == "cumul":
                cumul = [int(x) for x in w[1:]]
    foi = np.array(sorted(foi_rows), dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "foi_reference.npz"),
                        seeds=np.array(seed_rows, dtype=np.int32),
                        foi_day=foi[:, 0].astype(np.int16), foi_line=foi[:, 1].astype(np.int16), foi_loc=foi[:, 2].astype(np.int32),
                        foi_x=foi[:, 3], foi_w=foi[:, 4],
                        ledger=np.array(ledger_rows, dtype=np.uint64), cumul=np.array(cumul, dtype=np.uint64),
                        hist=np.array([r[:14] for r in hist_rows], dtype=np.int32), hist_rel=np.array([r[14] for r in hist_rows], dtype=np.float32))
    lines, counts = np.unique(foi[:, 1].astype(int), return_counts=True)
    print(f"{len(seed_rows)} scripted infections ({sum(r[2] for r in seed_rows)} took), {len(foi)} force-of-infection records "
          f"by source line {dict(zip(lines.tolist(), counts.tolist()))}, {len(hist_rows)} infected people")


if __name__ == "__main__":
    main()
