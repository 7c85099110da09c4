"""Test-only loader of the CPU oracle (oracle/libcabm_oracle.so) behind the same host mirror."""
import os
import subprocess

import covid_abm_b200 as cabm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libcabm_oracle.so")
REF_DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
_backend = None


def oracle_backend():
    global _backend
    if _backend is None:
        if not os.path.exists(ORACLE_SO):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])
        _backend = cabm.Backend(ORACLE_SO, "cabm_oracle_")
    return _backend


def oracle_community(P, pop, tables=None, attrs=None):
    return cabm.Community(P, pop, tables, backend=oracle_backend(), attrs=attrs)
