"""The drop-in boundary is byte-identical to the reference's C headers.

tests/golden/abi_layout.json holds sizes, alignments, field offsets and enumerator values measured on the
reference's own headers (scripts/make_abi_golden.py, run where /root/reference exists).  Here the same probe
is compiled against include/fimo_worlds_b200.h and compared field by field.  When the reference is present
(this container, not the GPU box) the golden itself is re-derived and must not have drifted."""
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import make_abi_golden as G  # noqa: E402

GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "abi_layout.json")))


@pytest.fixture(scope="module")
def product_layout():
    return G.product_layout()


@pytest.mark.parametrize("struct", sorted(G.STRUCTS))
def test_struct_layout_matches_reference(product_layout, struct):
    want, got = GOLDEN["structs"][struct], product_layout["structs"][struct]
    assert got["size"] == want["size"], f"sizeof({struct})"
    assert got["align"] == want["align"], f"alignof({struct})"
    for field, (offset, size) in want["fields"].items():
        assert got["fields"][field] == [offset, size], f"{struct}.{field} (offset, size)"


def test_enumerator_values_match_reference(product_layout):
    assert product_layout["constants"] == GOLDEN["constants"]


@pytest.mark.skipif(not os.path.isdir(G.REF), reason="reference tree not present (GPU box)")
def test_golden_is_what_the_reference_headers_say():
    ref = G.reference_layout()
    assert ref["structs"] == GOLDEN["structs"] and ref["constants"] == GOLDEN["constants"]
