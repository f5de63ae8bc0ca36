"""Builds (in-tree, for sm_100a) and loads libfimo_worlds_b200.so — the product.

The shared library is the drop-in boundary (include/fimo_worlds_b200.h); this module only
compiles and dlopen()s it.  There is no fallback: if the library cannot be built or loaded,
or a batch world is created on a machine without a CUDA device, the error propagates.
"""
from __future__ import annotations

import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "libfimo_worlds_b200.so")


def _sources() -> list[str]:
    out = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cpp", ".cu", ".h")) or f == "Makefile"]
    out.append(os.path.join(_HERE, "..", "include", "fimo_worlds_b200.h"))
    return out


def build(force: bool = False) -> str:
    """nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo (see csrc/Makefile).  Cross-compiles without a GPU."""
    stale = force or not os.path.exists(LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in _sources())
    if stale:
        if not os.path.exists("/usr/local/cuda/bin/nvcc"):  # never hand out a library older than its sources
            raise RuntimeError("libfimo_worlds_b200.so is missing or older than its sources and there is no nvcc to rebuild it")
        subprocess.run(["make", "-C", CSRC, "-s", "-j8"], check=True)
    return LIB_PATH


def build_examples() -> None:
    """examples/wind_system/libwind_system.so (nvcc, sm_100a): the user-supplied device system the tests register."""
    subprocess.run(["make", "-C", os.path.join(_HERE, "..", "examples", "wind_system"), "-s"], check=True)


def load():
    from .abi import WorldsLib

    lib = WorldsLib(build())
    if not lib.is_product:
        raise RuntimeError("libfimo_worlds_b200.so does not export the device entry points")
    return lib
