"""Builds (if needed) and loads the CPU oracle.  TEST INFRASTRUCTURE ONLY: import this from
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, nowhere else."""
from __future__ import annotations

import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "libfimo_worlds_oracle.so")


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".cpp", ".h"))]
    srcs.append(os.path.join(_HERE, "..", "include", "fimo_worlds_b200.h"))
    stale = force or not os.path.exists(LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if stale:
        subprocess.run(["make", "-C", _HERE, "-s"], check=True)
    return LIB_PATH


def load():
    from fimo_b200.abi import WorldsLib

    return WorldsLib(build())
