"""Pins the CPU oracle against the reference's own golden tests for the hot path
(SURVEY.md §8c): every case in tests/reference_goldens.py cites the Zig test it restates."""
import pytest

import reference_goldens as G


@pytest.mark.parametrize("case", G.ALL, ids=[f.__name__ for f in G.ALL])
@pytest.mark.timeout(120)
def test_oracle_matches_reference_golden(oracle, case):
    case(oracle)
