"""CPU-side checks of the product library (no GPU needed): it loads, exports every symbol
include/fimo_worlds_b200.h declares, and its host scheduler (host systems on the pool) passes the
reference's own behavioural tests — the same cases the CPU oracle is pinned with."""
import ctypes as C
import os
import re

import pytest

import reference_goldens as G
from fimo_b200 import abi, library

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    return library.load()


def test_every_declared_symbol_is_exported(lib):
    header = open(os.path.join(ROOT, "include", "fimo_worlds_b200.h")).read()
    declared = set(re.findall(r"FIMO_B200_EXPORT[^;(]*?\b(fimo_[a-z0-9_]+)\s*\(", header, flags=re.S))
    assert len(declared) >= 32 + 8 + 7
    missing = [name for name in sorted(declared) if not hasattr(lib.dll, name)]
    assert not missing, missing


def test_export_table_lists_the_32_fimo_worlds_symbols_in_order(lib):
    rows = lib.exports()
    assert [r[0] for r in rows] == [name for name, _, _ in abi.WORLDS_SYMBOLS]  # symbols.zig:22-58
    assert all(ns == "fimo-worlds" for _, ns, _ in rows)  # package.h:12
    for name, _, fn in rows:
        assert fn == C.cast(getattr(lib.dll, "fimo_worlds_" + name), C.c_void_p).value


def test_no_cpu_fallback_for_batch_worlds(lib):
    """Without a CUDA device a batch world must fail loudly, not degrade to a CPU path."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from fimo_b200.worlds import FimoError, World

    batch = abi.WorldBatchExt(abi.ExtHeader(abi.STYPE_WORLD_BATCH, 0, None), 4, 64, 0, 0)
    with pytest.raises(FimoError) as e:
        World.init(lib, "w", batch=batch)
    assert e.value.name == "NoDevice"


def test_product_does_not_link_the_oracle(lib):
    import subprocess

    out = subprocess.run(["ldd", lib.path], capture_output=True, text=True).stdout
    assert "oracle" not in out
    syms = subprocess.run(["nm", "-D", lib.path], capture_output=True, text=True).stdout
    assert "fimo_oracle" not in syms and "fimo_ref_" not in syms


@pytest.mark.parametrize("case", G.ALL, ids=[f.__name__ for f in G.ALL])
@pytest.mark.timeout(120)
def test_product_host_path_matches_reference_golden(lib, case):
    case(lib)
