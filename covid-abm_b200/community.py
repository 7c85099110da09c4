"""Host-side mirror of the reference's `Community` for the daily transmission cycle.

Same names and argument meaning as the reference interface the path sits behind
(Community.h:316-375, simulator.h:201-223,329-374); every method forwards to the C ABI of
include/cabm.h. There is no CPU implementation here: constructing a `Community` without the CUDA
library (covid-abm_b200/csrc/libcabm.so) raises.

`Backend` lets the tests drive a second shared library exporting the same ABI under another prefix
(the CPU oracle); the product never does that.
"""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

from . import abi
from . import params as _params

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CABM_LIB", os.path.join(_HERE, "csrc", "libcabm.so"))    # CABM_LIB: development override (kernel experiments)


class CabmError(RuntimeError):
    pass


class Backend:
    def __init__(self, path: str, prefix: str):
        if not os.path.exists(path):
            raise CabmError(f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
        self.lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
        self.fn = abi.bind(self.lib, prefix)
        self.path = path


_cuda_backend = None


def cuda_backend() -> Backend:
    global _cuda_backend
    if _cuda_backend is None:
        _cuda_backend = Backend(LIB_PATH, "cabm_")
    return _cuda_backend


def save_population(pop: dict, path: str, backend: Backend | None = None):
    """Write the built population (SoA + CSRs of a tools/synthpop.py dict) as one binary file (cabm_population_save, cabm.h);
    host code only: works without a GPU."""
    be = backend or cuda_backend()
    n = int(pop["n_people"])
    dummy = (np.zeros(n), np.zeros(n, dtype=np.float32), np.zeros(n, dtype=np.int32))          # the attributes are not part of the file
    pstruct, keep = abi.pack_population(pop, dummy)
    rc = be.fn.population_save(C.byref(pstruct), os.fsencode(path))
    del keep
    if rc != 0:
        raise CabmError(f"cabm_population_save failed rc={rc} ({path})")


class Community:
    """One simulated community bound to one GPU (reference: `class Community`, Community.h:138)."""

    def __init__(self, P: dict, pop: dict, tables: dict | None = None, backend: Backend | None = None, attrs=None, population_file: str | None = None):
        """pop: a tools/synthpop.py dict; with population_file only its n_people / n_loc are read and the arrays come from the
        binary file written by save_population (cabm_load_population_file: one read + the device copies)."""
        self.backend = backend or cuda_backend()
        self.fn = self.backend.fn
        self.P = P
        self.T = tables if tables is not None else _params.build_tables(P)
        self._h = C.c_void_p()
        self._check(self.fn.create(C.byref(self._h)), create=True)
        self.n_people = int(pop["n_people"]); self.n_loc = int(pop["n_loc"])
        self.run_length = int(P["run_length"])
        self._par_struct, self._par_keep = abi.pack_params(P, self.T)
        self._check(self.fn.set_params(self._h, C.byref(self._par_struct)))
        attrs = attrs if attrs is not None else abi.immunity_attributes(self.n_people, P["seed"])
        if population_file is None:
            pstruct, keep = abi.pack_population(pop, attrs)
            self._check(self.fn.load_population(self._h, C.byref(pstruct)))
            del keep
        else:
            ne = np.ascontiguousarray(attrs[0], dtype=np.float64)
            xp = np.ascontiguousarray(attrs[1], dtype=np.float32)
            nd = np.ascontiguousarray(attrs[2], dtype=np.int32)
            assert ne.size == xp.size == nd.size == self.n_people
            self._check(self.fn.load_population_file(self._h, os.fsencode(population_file), ne.ctypes.data_as(C.POINTER(C.c_double)),
                                                     xp.ctypes.data_as(C.POINTER(C.c_float)), nd.ctypes.data_as(C.POINTER(C.c_int32))))
        for which, key in ((0, "school_closure"), (1, "business_closure"), (2, "social_distancing")):
            self.setTimedIntervention(which, P[key])
        self.julian_day, self.is_weekday = _params.calendar(P["start_julian_year"], P["start_day_of_year"], self.run_length)
        self.strains = [50.0, 0.0, 0.0, 0.0]                                   # simulator.h:315
        self._mid = [(_params.to_sim_day(P["start_julian_year"], P["start_day_of_year"], d)) for d in P.get("mutant_intro_dates", [])]
        self.day = 0

    # ---- plumbing ----
    def _check(self, rc, create=False):
        if rc != 0:
            msg = "" if create or not self._h else (self.fn.last_error(self._h) or b"").decode()
            raise CabmError(f"cabm call failed rc={rc} {msg}")

    def close(self):
        if getattr(self, "_h", None):
            self.fn.destroy(self._h); self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_params(self, P: dict, tables: dict | None = None):
        """Push an edited parameter set (drivers edit Parameters between calls, exp/active-vac/main.cpp:598-605); run_length is fixed."""
        self.P = P
        self.T = tables if tables is not None else _params.build_tables(P)
        self._par_struct, self._par_keep = abi.pack_params(P, self.T)
        self._check(self.fn.set_params(self._h, C.byref(self._par_struct)))

    # ---- reference-facing calls ----
    def setTimedIntervention(self, which: int, values):
        """Community::setSocialDistancingTimedIntervention / updateTimedIntervention (Community.h:396-412)."""
        v = np.ascontiguousarray(values, dtype=np.float64)
        self._check(self.fn.set_timed_intervention(self._h, which, v.ctypes.data_as(C.POINTER(C.c_double)), v.size))

    def infect(self, pid: int, strain: int = 0, day: int | None = None) -> bool:
        """Community::infect(id, strain) (Community.cpp:550-553)."""
        out = C.c_int32(0)
        self._check(self.fn.infect(self._h, int(pid), int(strain), self.day if day is None else int(day), C.byref(out)))
        return bool(out.value)

    def infect_many(self, pids, strains, day: int | None = None) -> np.ndarray:
        """len(pids) consecutive Community::infect(id, strain) calls in one launch (cabm_infect_batch)."""
        p = np.ascontiguousarray(pids, dtype=np.uint32)
        st = np.ascontiguousarray(np.broadcast_to(np.asarray(strains, dtype=np.uint8), p.shape))
        took = np.zeros(p.size, dtype=np.uint8)
        self._check(self.fn.infect_batch(self._h, p.ctypes.data_as(C.POINTER(C.c_uint32)), st.ctypes.data_as(C.POINTER(C.c_uint8)), p.size,
                                         self.day if day is None else int(day), took.ctypes.data_as(C.POINTER(C.c_uint8))))
        return took.astype(bool)

    def tick(self, day: int | None = None):
        """Community::tick() (Community.cpp:1147-1192)."""
        dday = self.day if day is None else int(day)
        self._check(self.fn.tick(self._h, dday, int(self.julian_day[dday]), int(self.is_weekday[dday])))

    def update_strain_mix(self, day: int):
        """The strain urn of simulate_epidemic (simulator.h:341-372), on sim-day numbers."""
        s, m = self.strains, self._mid
        if not m:
            return
        if m[0] <= day < m[1]:
            if s[0] > 1: s[0] -= 1; s[1] += 1
        elif m[1] <= day < m[2]:
            if s[0] > 1: s[0] -= 1; s[2] += 1
            if s[1] > 1: s[1] -= 1; s[2] += 1
        elif day >= m[2]:
            for _ in range(5):
                for k in (0, 1, 2):
                    if s[k] > 1: s[k] -= 1; s[3] += 1

    def expected_imports(self, day: int) -> float:
        pde = self.P["prob_daily_exposure"]
        return float(pde[day % len(pde)]) * self.n_people                       # simulator.h:204-206

    def seed_epidemic(self, day: int | None = None) -> int:
        """seed_epidemic(par, community, date, strain_weights) (simulator.h:201-223)."""
        dday = self.day if day is None else int(day)
        out = C.c_int32(0)
        w = (C.c_double * 4)(*self.strains)
        self._check(self.fn.import_(self._h, dday, self.expected_imports(dday), w, C.byref(out)))
        return out.value

    def step(self):
        """One iteration of the simulate_epidemic day loop body that belongs to the hot path (simulator.h:338-374)."""
        self.tick(self.day)
        self.update_strain_mix(self.day)
        n = self.seed_epidemic(self.day)
        self.day += 1
        return n

    def run_days(self, n_days: int):
        """n_days of step() without returning to Python between days."""
        first = self.day
        strain_w = np.zeros((n_days, 4)); exp = np.zeros(n_days)
        for i in range(n_days):
            self.update_strain_mix(first + i)
            strain_w[i] = self.strains; exp[i] = self.expected_imports(first + i)
        jd = np.ascontiguousarray(self.julian_day[first:first + n_days]); wd = np.ascontiguousarray(self.is_weekday[first:first + n_days])
        self._check(self.fn.run_days(self._h, first, n_days, jd.ctypes.data_as(C.POINTER(C.c_int32)), wd.ctypes.data_as(C.POINTER(C.c_int32)),
                                     exp.ctypes.data_as(C.POINTER(C.c_double)), strain_w.ctypes.data_as(C.POINTER(C.c_double))))
        self.day += n_days

    def vaccinate(self, pids, day: int | None = None):
        p = np.ascontiguousarray(pids, dtype=np.uint32); ok = np.zeros(p.size, dtype=np.uint8)
        self._check(self.fn.vaccinate(self._h, self.day if day is None else int(day), p.ctypes.data_as(C.POINTER(C.c_uint32)), p.size,
                                      ok.ctypes.data_as(C.POINTER(C.c_uint8))))
        return ok.astype(bool)

    def quarantine(self, pids, from_day: int, until_day: int):
        p = np.ascontiguousarray(pids, dtype=np.uint32)
        self._check(self.fn.quarantine(self._h, p.ctypes.data_as(C.POINTER(C.c_uint32)), p.size, int(from_day), int(until_day)))

    # ---- ledger getters (Community.h:353-375) ----
    def ledger(self, counter) -> np.ndarray:
        idx = abi.COUNTERS.index(counter) if isinstance(counter, str) else int(counter)
        out = np.zeros(self.run_length, dtype=np.uint64)
        self._check(self.fn.ledger_series(self._h, idx, out.ctypes.data_as(C.POINTER(C.c_uint64)), out.size))
        return out

    def ledger_table(self) -> np.ndarray:
        """[run_length][N_COUNTERS], one device read (cabm_ledger_table)."""
        out = np.zeros((abi.N_COUNTERS, self.run_length), dtype=np.uint64)
        self._check(self.fn.ledger_table(self._h, out.ctypes.data_as(C.POINTER(C.c_uint64)), out.size))
        return np.ascontiguousarray(out.T)

    def getNumNewlyInfected(self): return self.ledger("new_infected")
    def getNumNewlySymptomatic(self): return self.ledger("new_symptomatic")
    def getNumNewlySevere(self): return self.ledger("new_severe")
    def getNumNewlyCritical(self): return self.ledger("new_critical")
    def getNumNewlyDead(self): return self.ledger("new_dead")
    def getNumHospInc(self): return self.ledger("hosp_inc")
    def getNumHospPrev(self): return self.ledger("hosp_prev")
    def getNumDetectedCasesReport(self): return self.ledger("det_cases_report")

    def getCumulIncidenceByOutcome(self) -> np.ndarray:
        out = np.zeros(abi.N_OUTCOMES, dtype=np.uint64)
        self._check(self.fn.cumulative_outcomes(self._h, out.ctypes.data_as(C.POINTER(C.c_uint64))))
        return out

    def work_counters(self) -> dict:
        out = np.zeros(abi.N_WORK, dtype=np.uint64)
        self._check(self.fn.work_counters(self._h, out.ctypes.data_as(C.POINTER(C.c_uint64))))
        return dict(zip(abi.WORK, (int(x) for x in out)))

    def infection_histories(self) -> np.ndarray:
        """Every infection so far (Person::getInfectionHistory over all people, Person.h:281)."""
        n = C.c_size_t(0)
        self._check(self.fn.sync_histories(self._h, None, 0, C.byref(n)))
        out = np.zeros(n.value, dtype=abi.INFECTION_DTYPE)
        if n.value:
            self._check(self.fn.sync_histories(self._h, out.ctypes.data, n.value, C.byref(n)))
        return out

    def current_infections(self, pids=None) -> np.ndarray:
        if pids is None:
            n = self.n_people; ptr = None
        else:
            p = np.ascontiguousarray(pids, dtype=np.uint32); n = p.size; ptr = p.ctypes.data_as(C.POINTER(C.c_uint32))
        out = np.zeros(n, dtype=abi.INFECTION_DTYPE)
        self._check(self.fn.get_infections(self._h, ptr, n, out.ctypes.data))
        return out

    # ---- caller-side daily diagnostics (simulator.h:402-410,438), device reductions ----
    def diagnostics(self, day: int | None = None) -> dict:
        out = np.zeros(abi.N_DIAG, dtype=np.uint64)
        self._check(self.fn.diagnostics(self._h, max(self.day - 1, 0) if day is None else int(day), out.ctypes.data_as(C.POINTER(C.c_uint64))))
        return dict(zip(abi.DIAG, (int(x) for x in out)))

    def calculate_vax_stats(self, day: int) -> dict:
        """Community::calculate_vax_stats (Community.cpp:1329-1423): same keys, same arithmetic on the device counts."""
        c = self.diagnostics(day); n = float(self.n_people)
        full = float(c["fully_vaxd"]); unv = n - full

        def ve(bt, uv):
            with np.errstate(divide="ignore", invalid="ignore"):
                return float(1.0 - (np.float64(c[bt]) / np.float64(full)) / (np.float64(c[uv]) / np.float64(unv)))
        with np.errstate(divide="ignore", invalid="ignore"):
            brk = float(np.float64(c["bt_reported"]) / np.float64(c["bt_reported"] + c["uv_reported"]))
        return {"coverage": full / n, "dose_1": c["dose_1"] / n, "dose_2": c["dose_2"] / n, "dose_3": c["dose_3"] / n,
                "VES": ve("bt_inf", "uv_inf"), "VEP": ve("bt_dis", "uv_dis"), "VEH": ve("bt_hosp", "uv_hosp"), "VED": ve("bt_death", "uv_death"),
                "breakthruRatio": brk, "vaxInfs": float(c["bt_inf"]), "unvaxInfs": float(c["uv_inf"]), "vaxHosp": float(c["bt_hosp"]),
                "unvaxHosp": float(c["uv_hosp"])}

    def doSerosurvey(self, pids, day: int) -> float:
        """Community::doSerosurvey(NATURAL, pop, day) (Community.cpp:1224-1274); pids None = everybody."""
        out = C.c_uint64(0)
        if pids is None:
            n = self.n_people; ptr = None
        else:
            p = np.ascontiguousarray(pids, dtype=np.uint32); n = p.size; ptr = p.ctypes.data_as(C.POINTER(C.c_uint32))
        self._check(self.fn.serosurvey(self._h, int(day), ptr, n, C.byref(out)))
        return out.value / n if n else float("nan")

    def eligibility(self, pids, day: int) -> np.ndarray:
        """CABM_EL_* byte per person (Person::isAlive / isSeroEligible / isInfEligible inputs, Person.cpp:562-608); pids None = everybody."""
        if pids is None:
            n = self.n_people; ptr = None
        else:
            p = np.ascontiguousarray(pids, dtype=np.uint32); n = p.size; ptr = p.ctypes.data_as(C.POINTER(C.c_uint32))
        out = np.zeros(n, dtype=np.uint8)
        self._check(self.fn.eligibility(self._h, int(day), ptr, n, out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    def reduce(self, what, pids, day: int) -> int:
        """Count of one per-person predicate (abi.REDUCE) over a cohort; pids None = everybody (cabm_reduce)."""
        w = abi.REDUCE.index(what) if isinstance(what, str) else int(what)
        if pids is None:
            n = self.n_people; ptr = None
        else:
            p = np.ascontiguousarray(pids, dtype=np.uint32); n = p.size; ptr = p.ctypes.data_as(C.POINTER(C.c_uint32))
        out = C.c_double(0)
        self._check(self.fn.reduce(self._h, w, int(day), ptr, n, C.byref(out)))
        return int(out.value)

    def getNumPeopleQuarantining(self, day: int) -> float:
        return self.diagnostics(day)["quarantining"] / self.n_people

    def getNumInfected(self, day: int) -> int: return self.diagnostics(day)["infected"]
    def getNumInfectious(self, day: int) -> int: return self.diagnostics(day)["infectious"]
    def getNumSymptomatic(self, day: int) -> int: return self.diagnostics(day)["symptomatic"]
    def getNumNaive(self) -> int: return self.diagnostics(max(self.day - 1, 0))["naive"]

    def detected_today(self, day: int) -> np.ndarray:
        """Primary cases of traceForwardContacts (Community.cpp:1062-1070) before the coverage draw, ascending ids."""
        n = C.c_size_t(0)
        self._check(self.fn.detected_today(self._h, int(day), None, 0, C.byref(n)))
        out = np.zeros(n.value, dtype=np.uint32)
        if n.value:
            self._check(self.fn.detected_today(self._h, int(day), out.ctypes.data_as(C.POINTER(C.c_uint32)), n.value, C.byref(n)))
        return out

    def infection_histories_from(self, first: int) -> np.ndarray:
        n = C.c_size_t(0)
        self._check(self.fn.sync_histories_from(self._h, int(first), None, 0, C.byref(n)))
        out = np.zeros(n.value, dtype=abi.INFECTION_DTYPE)
        if n.value:
            self._check(self.fn.sync_histories_from(self._h, int(first), out.ctypes.data, n.value, C.byref(n)))
        return out

    def foi(self) -> np.ndarray:
        n = C.c_size_t(0)
        self._check(self.fn.dump_foi(self._h, None, 0, C.byref(n)))
        out = np.zeros(n.value, dtype=abi.FOI_DTYPE)
        if n.value:
            self._check(self.fn.dump_foi(self._h, out.ctypes.data, n.value, C.byref(n)))
        return out[np.argsort(out["loc"], kind="stable")]

    def device_time_ms(self):
        ms = C.c_double(0); k = C.c_uint64(0)
        self._check(self.fn.device_time_ms(self._h, C.byref(ms), C.byref(k)))
        return ms.value, k.value

    def set_profile(self, on: bool):
        self._check(self.fn.set_profile(self._h, 1 if on else 0))

    def kernel_times(self) -> dict:
        """{kernel name: (total ms, launches)} accumulated since the last call (profile mode only)."""
        ms = (C.c_double * 32)(); cnt = (C.c_uint64 * 32)(); names = (C.c_char_p * 32)(); n = C.c_size_t(0)
        self._check(self.fn.kernel_times(self._h, ms, cnt, names, 32, C.byref(n)))
        return {names[i].decode(): (ms[i], int(cnt[i])) for i in range(n.value)}

    def snapshot(self):
        s = C.c_void_p()
        self._check(self.fn.snapshot(self._h, C.byref(s)))
        return (s, self.day, list(self.strains))

    def restore(self, snap):
        s, day, strains = snap
        self._check(self.fn.restore(self._h, s)); self.day = day; self.strains = list(strains)

    def snap_free(self, snap):
        self.fn.snap_free(snap[0])
