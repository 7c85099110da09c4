"""cabm_serosurvey_static: the no-allocation variant of cabm_serosurvey (include/cabm.h), checked against it (-m gpu).

no counterpart in the reference. Its device scratch (id list + accumulator)
is a file-scope `static __device__` array of the library instead of memory from cudaMalloc. The counts are integers, so the
bar is equality. Two calls in a row share the array, which keeps the first call's ids: the second call uses a SHORTER,
different list, so stale ids or a stale accumulator would show. This file is named to be collected last (pytest -x
must not hide other tests behind a failure here) and builds its own community, so it passes alone in a fresh
process as well as after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

import covid_abm_b200 as cabm
import synthpop

pytestmark = pytest.mark.gpu


def _survey(c, fn, pids, day):
    out = C.c_uint64(0xDEADBEEF)
    if pids is None:
        ptr, n = None, 0
    else:
        pids = np.ascontiguousarray(pids, dtype=np.uint32)
        ptr, n = pids.ctypes.data_as(C.POINTER(C.c_uint32)), pids.size
    rc = fn(c._h, int(day), ptr, n, C.byref(out))
    return rc, out.value


def test_static_scratch_serosurvey_equals_the_allocating_one_twice_in_a_row():
    pop = synthpop.generate(20000, seed=11)
    P = cabm.params.default_params(run_length=50, seed=5, T=0.15, prob_daily_exposure=3e-3)
    c = cabm.Community(P, pop, cabm.params.build_tables(P))
    assert c.fn.serosurvey_static is not None, "libcabm.so does not export cabm_serosurvey_static"
    n = int(pop["n_people"])
    for pid in range(0, 3000, 5):
        c.infect(pid, 0, 0)
    c.run_days(45)
    day = 44
    rng = np.random.default_rng(1)
    first = rng.choice(n, size=n // 2, replace=False)
    second = rng.choice(n, size=n // 7, replace=False)             # shorter than the first: the array still holds the first list's tail
    results = []
    for pids in (first, second, None, second):                      # back to back on one handle
        rc_s, got = _survey(c, c.fn.serosurvey_static, pids, day)
        rc_a, want = _survey(c, c.fn.serosurvey, pids, day)
        assert rc_s == 0 and rc_a == 0, (rc_s, rc_a, c.fn.last_error(c._h))
        assert got == want, f"static {got} != allocating {want} (ids: {None if pids is None else len(pids)})"
        results.append(got)
    assert results[0] > 50 and results[2] >= results[0] > results[1] > 0 and results[1] == results[3], results
    # an empty list is zero, a list beyond the stated maximum is rejected with a clear error and runs nothing
    rc, got = _survey(c, c.fn.serosurvey_static, np.zeros(0, dtype=np.uint32), day)
    assert rc == 0 and got == 0
    too_many = np.zeros(cabm.abi.STATIC_MAX_IDS + 1, dtype=np.uint32)
    rc, got = _survey(c, c.fn.serosurvey_static, too_many, day)
    assert rc == -1 and got == 0 and b"CABM_STATIC_MAX_IDS" in c.fn.last_error(c._h)
    at_limit = rng.integers(0, n, size=cabm.abi.STATIC_MAX_IDS, dtype=np.uint32)      # exactly the maximum (ids may repeat: doSerosurvey takes any list)
    rc_s, got = _survey(c, c.fn.serosurvey_static, at_limit, day)
    rc_a, want = _survey(c, c.fn.serosurvey, at_limit, day)
    assert rc_s == 0 and rc_a == 0 and got == want > 0
    # and the next small call is not disturbed by the full array the previous one left behind
    rc_s, got = _survey(c, c.fn.serosurvey_static, second, day)
    assert rc_s == 0 and got == results[1]
    c.close()
