"""Host-only check of the mirror's pooling state machine (sr/recognition/continuous_speech.py:_pool_segments) against
the literal loop of the reference (continuous_speech.py:91-106, restated in oracle.pool_segments) on random forced-
alignment paths, including the cases where the last states of a word last ONE frame each (the same-column hop through
a NES then delays the flush by one point)."""
import numpy as np
import pytest

import engine_double
from oracle import oracle as orc


def simple_chain_path(rng, n_words, S, short_p):
    """Same shape of path, built directly: list of (row, frame)."""
    pts = []
    c = 0
    row0 = 1
    for w in range(n_words):
        s = 0
        states = []
        while s < S:
            states.append(s)
            s += 2 if (rng.random() < 0.2 and s + 2 < S) else 1
        for k, s in enumerate(states):
            dur = 1 if rng.random() < short_p else int(rng.integers(1, 6))
            for j in range(dur):
                if not (k == 0 and j == 0 and w > 0):       # a word's first state re-emits the boundary frame (A9)
                    c += 1 if pts else 0
                pts.append((row0 + s, c))
        pts.append((row0 + S, c))                            # NES at the same frame
        row0 += S + 1
    return np.array(pts)


@pytest.fixture
def rec(monkeypatch):
    engine_double.install(monkeypatch)
    import sr.recognition.continuous_speech as cs
    return cs


@pytest.mark.parametrize("short_p", [0.0, 0.3, 0.8, 1.0])
def test_pool_segments_matches_reference_loop(rec, short_p):
    rng = np.random.default_rng(int(short_p * 10) + 5)
    S = 5
    for trial in range(3000):
        n_words = int(rng.integers(1, 5))
        path = simple_chain_path(rng, n_words, S, short_p)
        R = 1 + n_words * (S + 1)
        is_nes = np.zeros(R, bool)
        is_nes[::S + 1] = True
        assert not is_nes[path[0, 0]]
        assert rec._pool_segments(path, is_nes) == orc.pool_segments(path, is_nes), (trial, path.tolist())


def test_pool_segments_word_loop_paths(rec):
    """Random left-to-right paths without NES and with arbitrary NES positions (any row may be non-emitting)."""
    rng = np.random.default_rng(99)
    for trial in range(3000):
        n = int(rng.integers(1, 40))
        rows = np.cumsum(rng.integers(0, 3, n))
        frames = np.cumsum(rng.integers(0, 2, n))
        is_nes = rng.random(int(rows.max()) + 1) < 0.25
        path = np.array([rows, frames]).T
        assert rec._pool_segments(path, is_nes) == orc.pool_segments(path, is_nes), (trial, path.tolist())
