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


class _Version(C.Structure):  # FimoVersion, version.h:14-22
    _fields_ = [("major", C.c_size_t), ("minor", C.c_size_t), ("patch", C.c_size_t), ("pre", C.c_char_p), ("pre_len", C.c_size_t),
                ("build", C.c_char_p), ("build_len", C.c_size_t)]


class _SymbolImport(C.Structure):  # FimoModulesSymbolImport
    _fields_ = [("version", _Version), ("name", C.c_char_p), ("ns", C.c_char_p)]


class _SymbolExport(C.Structure):  # FimoModulesSymbolExport
    _fields_ = [("symbol", C.c_void_p), ("linkage", C.c_int32), ("version", _Version), ("name", C.c_char_p), ("ns", C.c_char_p)]


class _NamespaceImport(C.Structure):
    _fields_ = [("name", C.c_char_p)]


class _ModuleExport(C.Structure):  # struct FimoModulesExport, modules.h:942-1002
    _fields_ = [("next", C.c_void_p), ("version", _Version), ("name", C.c_char_p), ("description", C.c_char_p), ("author", C.c_char_p),
                ("license", C.c_char_p), ("parameters", C.c_void_p), ("parameters_count", C.c_size_t), ("resources", C.c_void_p),
                ("resources_count", C.c_size_t), ("namespace_imports", C.POINTER(_NamespaceImport)), ("namespace_imports_count", C.c_size_t),
                ("symbol_imports", C.POINTER(_SymbolImport)), ("symbol_imports_count", C.c_size_t),
                ("symbol_exports", C.POINTER(_SymbolExport)), ("symbol_exports_count", C.c_size_t),
                ("dynamic_symbol_exports", C.c_void_p), ("dynamic_symbol_exports_count", C.c_size_t), ("modifiers", C.c_void_p),
                ("modifiers_count", C.c_size_t)]


def test_module_record_in_the_fimo_module_section(lib):
    """What the fimo loader does with `module.fimo_module` (impl/module.h:40,61-62; exports.zig:2016-2044): iterate the
    `fimo_module` section; exactly one record, exporting the 32 fimo-worlds symbols and importing the futex symbols."""
    import subprocess

    path = os.path.join(ROOT, "fimo_b200", "libfimo_worlds_b200.so")
    sections = subprocess.run(["readelf", "-S", "-W", path], capture_output=True, text=True).stdout
    assert re.search(r"\bfimo_module\s+PROGBITS", sections), "no fimo_module ELF section"
    inspector_t = C.CFUNCTYPE(C.c_bool, C.POINTER(_ModuleExport), C.c_void_p)
    seen = []

    def inspect(module, _data):
        m = module.contents
        seen.append({
            "name": m.name.decode(), "version": (m.version.major, m.version.minor, m.version.patch), "next": m.next,
            "namespaces": [m.namespace_imports[i].name.decode() for i in range(m.namespace_imports_count)],
            "imports": [(m.symbol_imports[i].name.decode(), m.symbol_imports[i].ns.decode()) for i in range(m.symbol_imports_count)],
            "exports": [(m.symbol_exports[i].name.decode(), m.symbol_exports[i].ns.decode(), m.symbol_exports[i].symbol,
                         m.symbol_exports[i].linkage) for i in range(m.symbol_exports_count)],
            "dynamic": m.dynamic_symbol_exports_count, "modifiers": m.modifiers_count,
        })
        return True

    fn = lib.dll.fimo_impl_module_export_iterator
    fn.restype, fn.argtypes = None, [inspector_t, C.c_void_p]
    fn(inspector_t(inspect), None)
    assert len(seen) == 1  # the NULL entry of the section is skipped
    m = seen[0]
    assert m["name"] == "fimo_worlds" and m["next"] is None and m["dynamic"] == 0 and m["modifiers"] == 0
    assert m["namespaces"] == ["fimo-tasks"]
    assert m["imports"] == [("futex_wait", "fimo-tasks"), ("futex_wake", "fimo-tasks")]  # fimo_tasks_meta symbols.zig:132-160
    assert [e[0] for e in m["exports"]] == [name for name, _, _ in abi.WORLDS_SYMBOLS]
    assert all(e[1] == "fimo-worlds" and e[3] == 0 for e in m["exports"])
    table = {name: fn for name, _, fn in lib.exports()}
    assert all(e[2] == table[e[0]] for e in m["exports"])
    stopped = []
    fn(inspector_t(lambda module, _d: bool(stopped.append(1))), None)  # returning false stops the iteration
    assert stopped == [1]


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
