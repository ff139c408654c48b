"""GPU parity of the drop-in sr.recognition API: every case of tests/mirror_cases.py through libsrb200 on the
B200 against the golden vectors frozen from the unmodified reference."""
import pytest

import mirror_cases

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", mirror_cases.ALL, ids=lambda f: f.__name__)
def test_mirror_matches_reference(case, golden, capsys):
    import sr.recognition as rec
    case(rec, golden)
