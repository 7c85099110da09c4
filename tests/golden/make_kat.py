#!/usr/bin/env python3
"""Generate the Person::infect known-answer fixture from the UNMODIFIED reference (run in the build container).

    python tests/golden/make_kat.py          # needs oracle/_ref/ref_driver (built by `make -C oracle ref`)

Writes, next to this script:
    kat_params.txt    parameter/table file handed to the reference driver (covid-abm_b200/params.py write_text)
    kat_cases.txt     scripted person states and Philox keys (format: oracle/ref_driver_kat.inc)
    kat_expected.txt  what the reference's own Person::infect produced for them in tape mode

tests/test_reference_kat.py replays the cases through the CPU oracle and compares every field; the GPU parity tests
then compare the CUDA path with the oracle. /root/reference is needed only here, never at test time.
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
INT_MIN = -2147483648
N_CASES = 1500
RUN_LENGTH = 260
MUTANTS = ("2020-03-21", "2020-05-20", "2020-07-15")


def kat_params():
    import covid_abm_b200 as cabm
    P = cabm.params.default_params(run_length=RUN_LENGTH, seed=20240607, mutant_intro_dates=MUTANTS)
    T = cabm.params.build_tables(P)
    return cabm, P, T


def make_cases(rng):
    lines = []
    for i in range(N_CASES):
        age = int(rng.integers(0, 101)); comorbid = int(rng.random() < 0.35); ltc = int(age >= 65 and rng.random() < 0.1)
        time = int(rng.integers(0, 200)); strain = int(rng.integers(0, 4)); has_source = int(rng.random() < 0.5)
        nat_eff = float(np.round(rng.uniform(0.3, 0.999), 6)); xprot = float(np.float32(rng.random()))
        n_hist = int(rng.choice([0, 0, 1, 1, 2]))
        hist = []
        t_prev = -60
        for k in range(n_hist):
            last = k == n_hist - 1
            # most histories ended before `time`; some are still running (exercises the isInfected bail, Person.cpp:326)
            if last and rng.random() < 0.15:
                ib = time - int(rng.integers(0, 6)); ie = time + int(rng.integers(1, 8))
            else:
                ib = int(rng.integers(t_prev + 1, max(t_prev + 2, time - 12))); ie = ib + int(rng.integers(6, 12))
            se = INT_MIN if rng.random() < 0.4 else ie + int(rng.integers(-3, 12))
            hist += [ib, ie, se, int(rng.integers(0, 4))]
            t_prev = ie
        n_dose = int(rng.choice([0, 0, 1, 2, 3]))
        doses = sorted(int(x) for x in rng.integers(max(0, time - 150), time + 1, size=n_dose))
        naive = int(n_dose > 0 and rng.random() < 0.5)
        fields = [i, age, comorbid, ltc, time, strain, has_source, strain, repr(nat_eff), repr(xprot), n_hist] + hist + [n_dose] + doses + \
                 [naive, int(rng.integers(0, 2**31)), int(rng.integers(0, 2**31))]
        lines.append(" ".join(str(f) for f in fields))
    return lines


def main():
    cabm, P, T = kat_params()
    rng = np.random.default_rng(777)
    driver = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    if not os.path.exists(driver):
        sys.exit("oracle/_ref/ref_driver is missing: run `make -C oracle ref` where /root/reference exists")
    cabm.params.write_text(P, T, os.path.join(HERE, "kat_params.txt"))
    with open(os.path.join(HERE, "kat_cases.txt"), "w") as f:
        f.write("# id age comorbid ltc time strain has_source src_strain nat_eff xprot n_hist {ib ie se strain}* n_dose {day}* naive key_a key_b\n")
        f.write("\n".join(make_cases(rng)) + "\n")
    subprocess.check_call([driver, "kat", os.path.join(HERE, "kat_params.txt"), os.path.join(HERE, "kat_cases.txt"),
                           os.path.join(HERE, "kat_expected.txt")])
    ok = sum(1 for l in open(os.path.join(HERE, "kat_expected.txt")) if l.split()[1:2] == ["1"])
    print(f"{N_CASES} cases, {ok} infections took")


if __name__ == "__main__":
    main()
