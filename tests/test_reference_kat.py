"""Pins the CPU oracle against the UNMODIFIED reference's own Person::infect (no GPU needed).

tests/golden/kat_expected.txt was produced by the reference sources themselves (oracle/_ref/ref_driver kat, built from
/root/reference against oracle/gsl_shim in tape mode -- see tests/golden/make_kat.py): 1500 scripted person states
(age, comorbidity, long-term care, prior infections incl. ones still running, 0-3 vaccine doses, naive flag, all four
strains, with and without a source) pushed through Person::infect (Person.cpp:320-494) with every draw served from the
Philox stream keyed (seed; day, purpose, a, b). The oracle must reproduce every output bit for bit: success, the twelve
timeline days, strain, relInfectiousness (float), detection state and report day, the NUMBER of draws consumed from each
of the two streams (pins the draw order), the outcome tallies, the future hot-location calendar entries
(Person.cpp:475-486) and the detection ledger series (Community.cpp:657-679).
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import covid_abm_b200 as cabm
import oracle_lib

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
INT_MIN = -2147483648


def _load_golden_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_kat", os.path.join(GOLD, "make_kat.py"))
    m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
    return m


def _cases():
    out = []
    for line in open(os.path.join(GOLD, "kat_cases.txt")):
        if line.startswith("#") or not line.strip():
            continue
        f = line.split()
        c = dict(id=int(f[0]), age=int(f[1]), comorbid=int(f[2]), ltc=int(f[3]), time=int(f[4]), strain=int(f[5]), has_source=int(f[6]),
                 nat_eff=float(f[8]), xprot=float(f[9]))
        n_hist = int(f[10]); k = 11
        c["hist"] = [int(x) for x in f[k:k + 4 * n_hist]]; k += 4 * n_hist
        n_dose = int(f[k]); k += 1
        c["doses"] = [int(x) for x in f[k:k + n_dose]]; k += n_dose
        c["naive"], c["key_a"], c["key_b"] = int(f[k]), int(f[k + 1]), int(f[k + 2])
        out.append(c)
    return out


def _expected():
    rows, ledger = [], None
    for line in open(os.path.join(GOLD, "kat_expected.txt")):
        f = line.split()
        if f[0] == "ledger":
            ledger = [int(x) for x in f[1:]]
        else:
            rows.append(f)
    return rows, ledger


def _tiny_population(cases):
    n = len(cases)
    return {
        "n_people": n, "n_loc": 3,
        "loc_type": np.array([0, 1, 3], dtype=np.uint8), "loc_essential": np.ones(3, dtype=np.uint8), "loc_pubrisk": np.zeros(3, dtype=np.uint8),
        "loc_risk": np.zeros(3, dtype=np.float32), "loc_hospital": np.array([2, -1, -1], dtype=np.int32),
        "nbr_off": np.zeros(4, dtype=np.int64), "nbr": np.zeros(0, dtype=np.int32),
        "roster_off": np.array([0, n, 2 * n, 2 * n], dtype=np.int64), "roster": np.concatenate([np.arange(n), np.arange(n)]).astype(np.int32),
        "home": np.zeros(n, dtype=np.int32), "day": np.ones(n, dtype=np.int32),
        "age": np.array([c["age"] for c in cases], dtype=np.uint8), "sex": np.zeros(n, dtype=np.uint8),
        "comorbid": np.array([c["comorbid"] for c in cases], dtype=np.uint8), "ltc": np.array([c["ltc"] for c in cases], dtype=np.uint8),
        "patron": np.full((n, 5), -1, dtype=np.int32),
    }


def test_oracle_reproduces_reference_person_infect():
    mk = _load_golden_module()
    _, P, T = mk.kat_params()
    cases = _cases(); exp, exp_ledger = _expected()
    assert len(cases) == len(exp) == mk.N_CASES
    pop = _tiny_population(cases)
    attrs = (np.array([c["nat_eff"] for c in cases]), np.array([c["xprot"] for c in cases], dtype=np.float32), np.zeros(len(cases), dtype=np.int32))
    o = oracle_lib.oracle_community(P, pop, T, attrs=attrs)
    lib = o.backend.lib
    lib.cabm_oracle_kat_set_person.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.POINTER(C.c_int32), C.c_int, C.POINTER(C.c_int32), C.c_int]
    lib.cabm_oracle_kat_infect.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.POINTER(C.c_int64),
                                           C.POINTER(C.c_float), C.POINTER(C.c_int32), C.c_size_t]
    lib.cabm_oracle_kat_ledger_digest.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    out = (C.c_int64 * 32)(); rel = C.c_float(0); flags = (C.c_int32 * 4096)()
    took = 0; branches = set()
    for c, e in zip(cases, exp):
        hist = (C.c_int32 * max(1, len(c["hist"])))(*c["hist"]); doses = (C.c_int32 * max(1, len(c["doses"])))(*c["doses"])
        assert lib.cabm_oracle_kat_set_person(o._h, c["id"], len(c["hist"]) // 4, hist, len(c["doses"]), doses, c["naive"]) == 0
        assert lib.cabm_oracle_kat_infect(o._h, c["id"], c["strain"], c["time"], c["has_source"], c["key_a"], c["key_b"], out, C.byref(rel), flags, 2048) == 0
        got = list(out[:25])
        want_ok = int(e[1])
        label = f"case {c['id']} (age {c['age']}, t {c['time']}, strain {c['strain']}, src {c['has_source']}, hist {c['hist']}, doses {c['doses']})"
        assert got[0] == want_ok, f"{label}: success {got[0]} != reference {want_ok}"
        want_draws = [int(e[18]), int(e[19])]
        assert got[16:18] == want_draws, f"{label}: draws consumed {got[16:18]} != reference {want_draws}"
        if not want_ok:
            continue
        took += 1
        want_t = [int(x) for x in e[2:14]]
        assert got[1:13] == want_t, f"{label}: timeline {got[1:13]} != reference {want_t}"
        assert got[13] == int(e[14]), f"{label}: strain"
        assert np.float32(rel.value) == np.float32(float(e[15])), f"{label}: relInfectiousness {rel.value} != {e[15]}"
        assert got[14:16] == [int(e[16]), int(e[17])], f"{label}: detection {got[14:16]} != reference {e[16:18]}"
        assert got[18:23] == [int(x) for x in e[20:25]], f"{label}: outcome tallies"
        n_flags = int(e[25])
        assert got[23] == n_flags, f"{label}: {got[23]} calendar entries != reference {n_flags}"
        want_flags = sorted((int(e[26 + 2 * k]), int(e[27 + 2 * k])) for k in range(n_flags))
        assert sorted((flags[2 * k], flags[2 * k + 1]) for k in range(n_flags)) == want_flags, f"{label}: calendar entries"
        branches.add((want_t[3] != 2147483647, want_t[5] != 2147483647, want_t[8] != 2147483647, want_t[10] != 2147483647, want_t[11] != 2147483647,
                      int(e[16])))
    assert took > 500
    # every branch of the disease course and of the detection cascade was exercised
    seen = {b[5] for b in branches}
    assert seen >= {0, 1, 2, 5} and (seen & {3, 4}), sorted(seen)      # undetected, asymptomatic, mild, severe, critical-or-death
    assert any(b[4] for b in branches) and any(b[3] for b in branches) and any(b[2] and not b[3] for b in branches)
    # detection ledger series: sum and position-sensitive digest
    names = ["det_cases_onset", "det_cases_report", "det_hosp", "det_deaths_onset", "det_deaths_report"]
    for k, name in enumerate(names):
        s, hsh = C.c_uint64(0), C.c_uint64(0)
        assert lib.cabm_oracle_kat_ledger_digest(o._h, cabm.abi.COUNTERS.index(name), C.byref(s), C.byref(hsh)) == 0
        assert [s.value, hsh.value] == exp_ledger[2 * k:2 * k + 2], f"ledger series {name}"


@pytest.mark.skipif(not os.path.exists(oracle_lib.REF_DRIVER), reason="reference driver not built (no /root/reference on this machine)")
def test_golden_fixture_is_what_the_reference_produces(tmp_path):
    """Where the reference sources are available, re-run them and make sure the committed fixture is not stale."""
    out = tmp_path / "kat.txt"
    subprocess.check_call([oracle_lib.REF_DRIVER, "kat", os.path.join(GOLD, "kat_params.txt"), os.path.join(GOLD, "kat_cases.txt"), str(out)])
    assert open(out).read() == open(os.path.join(GOLD, "kat_expected.txt")).read()
