"""The C-ABI boundary and the host-side mirror, without a GPU.

* every function include/cabm.h declares is exported by covid-abm_b200/csrc/libcabm.so (and, under the prefix
  cabm_oracle_, by the test-only oracle library) -- load + symbol lookup only, no compute calls;
* the ctypes mirror of the ABI structs (covid-abm_b200/abi.py) has the layout gcc gives include/cabm.h;
* the product fails loudly where there is no CUDA device (no CPU fallback anywhere in the package);
* host logic of the mirror: calendar (Date.h:30-180,260-315), strain urn (simulator.h:341-372), importation
  rate (simulator.h:204-206), parameter tables.
"""
import ctypes as C
import datetime
import os
import re
import subprocess

import numpy as np
import pytest

import covid_abm_b200 as cabm
import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "cabm.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    names = re.findall(r"^\s*(?:int|void|const char\s*\*|double|uint64_t)\s+(cabm_[a-z_0-9]+)\s*\(", src, flags=re.M)
    assert len(names) >= 25, names
    return names


def test_product_library_exports_every_declared_symbol():
    assert os.path.exists(cabm.community.LIB_PATH), "libcabm.so not built: python -c 'import __graft_entry__ as g; g.build()'"
    lib = C.CDLL(cabm.community.LIB_PATH)
    missing = [n for n in _declared_functions() if not hasattr(lib, n)]
    assert not missing, f"include/cabm.h declares, libcabm.so lacks: {missing}"


def test_oracle_library_exports_the_same_abi_under_its_prefix():
    lib = oracle_lib.oracle_backend().lib
    skip = {"cabm_set_profile", "cabm_kernel_times",            # device profiling: nothing to mirror on the CPU
            "cabm_serosurvey_static",                           # where the device scratch lives: nothing to mirror on the CPU
            "cabm_shard_nccl_id", "cabm_shard_nccl_init", "cabm_shard_exchange_capacity", "cabm_shard_exchange_mode", "cabm_shard_run_days"}   # NCCL-driven batch loop
    missing = [n for n in _declared_functions() if n not in skip and not hasattr(lib, n.replace("cabm_", "cabm_oracle_", 1))]
    assert not missing, missing


def test_product_library_is_device_code_for_sm_100a():
    """The shared library carries sm_100a SASS for the day-pipeline kernels (cuobjdump is part of the CUDA toolkit)."""
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("no cuobjdump on this machine")
    out = subprocess.run([cuobjdump, "-lelf", cabm.community.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out, out
    res = subprocess.run([cuobjdump, "-res-usage", cabm.community.LIB_PATH], capture_output=True, text=True)
    text = res.stdout + res.stderr
    for k in ("k_scan", "k_active", "k_visit", "k_hazard", "k_trials", "k_gate", "k_apply_status", "k_hosp_import"):
        assert k in text, f"kernel {k} not found in libcabm.so"
    # k_visit stages its rows and the hot bitmap with bulk copies completing on mbarriers (TMA; SASS UBLKCP / SYNCS)
    sass = subprocess.run([cuobjdump, "-sass", "-fun", "_Z7k_visit3Devj", cabm.community.LIB_PATH], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass and "SYNCS" in sass, "k_visit carries no bulk-copy instructions"


def test_ctypes_structs_match_the_header_layout(tmp_path):
    prog = tmp_path / "layout.c"
    prog.write_text(r'''
#include <stdio.h>
#include <stddef.h>
#include "cabm.h"
int main(void) {
    printf("params %zu\n", sizeof(cabm_params));
    printf("population %zu\n", sizeof(cabm_population));
    printf("infection %zu\n", sizeof(cabm_infection));
    printf("foi_row %zu\n", sizeof(cabm_foi_row));
    printf("params.seasonality %zu\n", offsetof(cabm_params, seasonality));
    printf("params.VES %zu\n", offsetof(cabm_params, VES));
    printf("params.incub_cdf %zu\n", offsetof(cabm_params, incub_cdf));
    printf("params.visit_cdf %zu\n", offsetof(cabm_params, visit_cdf));
    printf("params.prob_first_detection %zu\n", offsetof(cabm_params, prob_first_detection));
    printf("n_counters %d\n", (int) CABM_N_COUNTERS);
    printf("n_work %d\n", (int) CABM_N_WORK);
    return 0;
}
''')
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])
    got = dict(line.split() for line in subprocess.check_output([str(exe)], text=True).splitlines())
    A = cabm.abi
    assert int(got["params"]) == C.sizeof(A.Params)
    assert int(got["population"]) == C.sizeof(A.Population)
    assert int(got["infection"]) == A.INFECTION_DTYPE.itemsize == 44
    assert int(got["foi_row"]) == A.FOI_DTYPE.itemsize == 64
    for field in ("seasonality", "VES", "incub_cdf", "visit_cdf", "prob_first_detection"):
        assert int(got[f"params.{field}"]) == getattr(A.Params, field).offset, field
    assert int(got["n_counters"]) == A.N_COUNTERS and int(got["n_work"]) == A.N_WORK


def test_header_compiles_as_plain_c_and_cxx(tmp_path):
    for comp, std, ext in (("gcc", "-std=c99", "c"), ("g++", "-std=c++11", "cpp")):
        f = tmp_path / f"t.{ext}"
        f.write_text('#include "cabm.h"\nint main(void) { return 0; }\n')
        subprocess.check_call([comp, std, "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", str(f), "-o", str(tmp_path / "t.o")])


def test_product_fails_loudly_without_a_cuda_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import synthpop
    pop = synthpop.generate(2000, seed=1)
    P = cabm.params.default_params(run_length=10, seed=1)
    with pytest.raises(cabm.community.CabmError):
        cabm.Community(P, pop)


def test_missing_library_is_an_error_not_a_fallback(tmp_path):
    with pytest.raises(cabm.community.CabmError):
        cabm.community.Backend(str(tmp_path / "nope.so"), "cabm_")


def test_package_has_no_cpu_implementation_of_the_path():
    """Nothing under the package imports or mentions the oracle."""
    pkg = os.path.join(ROOT, "covid-abm_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, fn)).read()
                assert "libcabm_oracle" not in text and "oracle_lib" not in text and "import oracle" not in text, fn


# ---- host logic ---------------------------------------------------------------------------------------------------
def test_calendar_matches_python_datetime():
    jd, wd = cabm.params.calendar(2020, 36, 800)            # the active-vac run starts 2020-02-05 (julian day 36)
    d0 = datetime.date(2020, 2, 5)
    for k in range(800):
        d = d0 + datetime.timedelta(days=k)
        assert jd[k] == d.timetuple().tm_yday
        assert wd[k] == (1 if d.weekday() < 5 else 0)


def test_to_sim_day_and_julian_day():
    assert cabm.params.to_julian_day("2020-03-01") == 61 and cabm.params.to_julian_day("2021-03-01") == 60
    d0 = datetime.date(2020, 2, 5)
    for ymd in ("2020-02-05", "2020-12-31", "2021-01-01", "2022-07-04", "2024-02-29"):
        want = (datetime.date(*map(int, ymd.split("-"))) - d0).days
        assert cabm.params.to_sim_day(2020, 36, ymd) == want, ymd


def test_lin_interpolate_series():
    v = cabm.params.lin_interpolate_series([("2020-02-05", 0.0), ("2020-02-15", 1.0), ("2020-02-20", 0.5)], 2020, 36)
    assert len(v) == 16 and v[0] == 0.0 and abs(v[5] - 0.5) < 1e-12 and v[10] == 1.0 and abs(v[15] - 0.5) < 1e-12
    assert all(abs((v[k + 1] - v[k]) - 0.1) < 1e-12 for k in range(10))


def test_poisson_visit_thresholds_are_a_cdf():
    P = cabm.params.default_params(run_length=30, seed=1)
    T = cabm.params.build_tables(P)
    cdf = np.asarray(T["visit_cdf"])
    assert np.all(np.diff(cdf) >= 0) and abs(cdf[0] - np.exp(-1.0)) < 1e-15 and cdf[-1] == 1.0
    for name in ("presev_cdf", "mild_cdf", "icu_death_cdf", "community_death_cdf", "death_lag_cdf"):
        c = np.asarray(T[name]); assert np.all(np.diff(c) >= 0) and c[-1] == 1.0, name


class _UrnOnly(cabm.Community):
    def __init__(self, mid):                                  # host logic only: no library handle is created
        self.strains = [50.0, 0.0, 0.0, 0.0]; self._mid = mid; self._h = None


def test_strain_urn_follows_simulator_h():
    u = _UrnOnly([10, 40, 90])
    for day in range(200):
        u.update_strain_mix(day)
        assert sum(u.strains) == 50.0 and min(u.strains) >= 0
        if day == 9:
            assert u.strains == [50.0, 0, 0, 0]
        if day == 39:
            assert u.strains == [20.0, 30.0, 0, 0]             # one wild-type ball per day moves to alpha
    assert u.strains[3] == 47.0 and u.strains[:3] == [1.0, 1.0, 1.0]
