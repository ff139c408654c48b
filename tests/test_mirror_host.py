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
    if case is mirror_cases.case_fit_words:
        case(rec, golden, fused_align=False)       # the fused alignment kernel has no numpy stand-in: K1 + K2 path
    else:
        case(rec, golden)


def test_pickles_load_on_both_sides(rec, tmp_path):
    """A model pickled by the mirror loads in the reference AND evaluates there (the reference's pdf reads the dense
    inv_cov, hmm_state.py:42), and a reference pickle loads in the mirror.  Needs the reference tree (container only)."""
    import os
    import pickle
    import subprocess
    import sys
    import numpy as np
    if not os.path.isdir("/root/reference/sr"):
        pytest.skip("the reference tree is not present on this machine")
    rng = np.random.default_rng(5)
    g = rec.GMM(rng.normal(size=39), rng.uniform(0.5, 1.5, size=39), 2)
    g.update_models(rng.normal(size=(2, 39)), rng.uniform(0.5, 1.5, size=(2, 39)), np.array([0.4, 0.6]))
    x = rng.normal(size=39)
    mine = g.evaluate(x)
    pm, pr = tmp_path / "mirror.pkl", tmp_path / "ref.pkl"
    pickle.dump(g, open(pm, "wb"))
    code = (
        "import sys, pickle, numpy as np\n"
        "sys.path.insert(0, %r)\n"
        "from _ref_shim import load_reference\n"
        "rec = load_reference()\n"
        "g = pickle.load(open(%r, 'rb'))\n"
        "assert type(g).__module__ == 'sr.recognition.hmm_state' and g.dists[0].inv_cov.shape == (39, 39)\n"
        "x = np.array(%r)\n"
        "print(repr(float(g.evaluate(x))))\n"
        "pickle.dump(g, open(%r, 'wb'))\n"
    ) % (os.path.join(os.path.dirname(__file__), "golden"), str(pm), x.tolist(), str(pr))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    assert abs(float(r.stdout.strip().splitlines()[-1]) - mine) <= 1e-4 * abs(mine)
    back = pickle.load(open(pr, "rb"))                       # the reference's pickle, loaded by the mirror
    assert isinstance(back, rec.GMM) and abs(back.evaluate(x) - mine) <= 1e-4 * abs(mine)
