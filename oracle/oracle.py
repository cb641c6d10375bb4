"""Loader for the CPU oracle (test infrastructure only — see crm_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
legs may import this module.
"""
from __future__ import annotations

import subprocess
from pathlib import Path

from cr_mech_coli_b200._capi import Library, Sim  # struct definitions only

_DIR = Path(__file__).resolve().parent
_LIBS: dict[bool, Library] = {}


def build():
    subprocess.run(["make", "-C", str(_DIR)], check=True, capture_output=True)


def library(openmp: bool = False) -> Library:
    if openmp not in _LIBS:
        name = "libcrm_oracle_omp.so" if openmp else "libcrm_oracle.so"
        path = _DIR / name
        if not path.exists():
            build()
        if not path.exists() and openmp:
            path = _DIR / "libcrm_oracle.so"
        _LIBS[openmp] = Library(path, "oracle_")
    return _LIBS[openmp]


def sim(cfg, openmp: bool = False) -> Sim:
    return Sim(library(openmp), cfg)
