"""Host logic of the sr.recognition mirror on the CPU: the device entry points of srb200.engine are replaced by the
numpy stand-in of tests/engine_double.py (test infrastructure, see its header).  The same cases run on the B200
through libsrb200 in tests/test_gpu_mirror.py."""
import pytest

import engine_double
import mirror_cases


@pytest.fixture
def rec(monkeypatch):
    engine_double.install(monkeypatch)
    import sr.recognition as rec
    return rec


@pytest.mark.parametrize("case", mirror_cases.ALL, ids=lambda f: f.__name__)
def test_mirror_host_logic(case, rec, golden, capsys):
    case(rec, golden)
