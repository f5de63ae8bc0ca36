"""the module packaging a fimo maintainer drops into `modules/` (integration/fimo_worlds_b200/).

No Zig toolchain exists in this image, so the files cannot be compiled here; what can be checked is that the shim
stays in step with the library: it maps exactly the symbols the library's export table lists, in the reference's
order (pkgs/fimo_worlds_meta/src/symbols.zig:22-58), onto C functions the shared library really exports, and every
helper it calls exists."""
import ctypes as C
import os
import re

from fimo_b200 import library

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "integration", "fimo_worlds_b200", "src", "FimoWorldsB200.zig")


def test_shim_exports_match_the_library_table():
    lib = library.load()
    text = open(SHIM).read()
    pairs = re.findall(r"export_\(symbols\.(\w+), &c\.(\w+)\)", text)
    n = C.c_size_t()
    table = lib.fimo_worlds_b200_exports(C.byref(n))
    names = [table[i].name.decode() for i in range(n.value)]
    assert n.value == 32 and [p[0] for p in pairs] == names
    for sym, fn in pairs:
        assert fn == f"fimo_worlds_{sym}"
        assert getattr(lib.dll, fn)  # exported by the shared library
    for helper in set(re.findall(r"c\.(fimo_worlds_b200_\w+)", text)):
        assert getattr(lib.dll, helper), helper


def test_build_script_names_the_makefile_and_header():
    text = open(os.path.join(ROOT, "integration", "fimo_worlds_b200", "build.zig")).read()
    assert os.path.exists(os.path.join(ROOT, "fimo_b200", "csrc", "Makefile")) and '"fimo_b200", "csrc"' in text
    assert os.path.exists(os.path.join(ROOT, "include", "fimo_worlds_b200.h")) and 'linkSystemLibrary("fimo_worlds_b200"' in text
    assert "FIMO_B200_CONTEXT_VERSION_MAJOR" in text and "FIMO_B200_CONTEXT_VERSION_MAJOR" in open(
        os.path.join(ROOT, "fimo_b200", "csrc", "module.cpp")).read()
