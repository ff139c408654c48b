#!/usr/bin/env python
"""Golden vectors of the MFCC front-end by EXECUTING the unmodified reference (container only).

    python tests/golden/make_golden_feature.py        # writes tests/golden/ref_feature.npz

Synthetic 16 kHz int16 signals are written as wav files; the reference's own `mfcc_features(path)`
(/root/reference/sr/feature/feature.py:43-82) and `load_wav_as_mfcc1(path)` (/root/reference/sr/core.py:25-31, with
`delta_feature` core.py:13-22 and `standardize` feature.py:85-88) read them back.  Inputs and outputs are frozen
exactly as produced (int16 samples, float64 results)."""
import os
import sys
import tempfile

import numpy as np
from scipy.io import wavfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from _ref_shim import load_reference  # noqa: E402

load_reference()
import sr.core as ref_core  # noqa: E402  (the reference: /root/reference/sr/core.py)
import sr.feature.feature as ref_feat  # noqa: E402

assert ref_core.__file__.startswith("/root/reference") and ref_feat.__file__.startswith("/root/reference")


def synth(rng, n):
    t = np.arange(n) / 16000.0
    f0 = rng.uniform(90, 240)
    sig = sum(a * np.sin(2 * np.pi * f0 * h * t + rng.uniform(0, 6.28)) for h, a in
              zip(range(1, 9), rng.uniform(0.05, 1.0, size=8)))
    sig *= 0.5 + 0.5 * np.sin(2 * np.pi * rng.uniform(2, 6) * t) ** 2          # syllable-like envelope
    sig += 0.02 * rng.normal(size=n)
    return np.clip(sig / np.abs(sig).max() * rng.uniform(3000, 20000), -32768, 32767).astype(np.int16)


def main():
    rng = np.random.default_rng(2024)
    lengths = [16000, 8007, 3217, 400, 481, 23999, 161 * 3]
    out = {"n": np.array(len(lengths))}
    with tempfile.TemporaryDirectory() as tmp:
        for i, n in enumerate(lengths):
            sig = synth(rng, n)
            path = os.path.join(tmp, f"u{i}.wav")
            wavfile.write(path, 16000, sig)
            fb, mfcc = ref_feat.mfcc_features(path)
            feat = ref_core.load_wav_as_mfcc1(path)
            out[f"sig{i}"], out[f"fbank{i}"], out[f"mfcc{i}"], out[f"feat{i}"] = sig, fb, mfcc, feat
            print(i, n, fb.shape, mfcc.shape, feat.shape)
    np.savez_compressed(os.path.join(HERE, "ref_feature.npz"), **out)


if __name__ == "__main__":
    main()
