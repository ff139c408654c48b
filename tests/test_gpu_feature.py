"""GPU parity of the MFCC front-end: srb_mfcc + srb_delta_standardize through the C ABI against the
reference's frozen outputs (tests/golden/ref_feature.npz) and the float64 oracle; the sr.feature / sr.core drop-in."""
import os

import numpy as np
import pytest
from scipy.io import wavfile

from oracle import feature_oracle as fo

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_feature.npz")
# FP32 FFT + log10 + DCT against float64: tolerances relative to the scale of the quantity (log10 energies ~ 1..10,
# cepstra ~ 0.1..10, standardised features ~ 1)
TOL_FBANK, TOL_MFCC, TOL_FEAT = 1e-4, 1e-4, 1e-3


def _close(a, b, tol):
    scale = max(1.0, float(np.abs(b).max()))
    err = float(np.abs(a - b).max()) / scale
    assert err < tol, err
    return err


def test_front_end_matches_reference_golden():
    from srb200.frontend import MfccFrontEnd
    g = np.load(GOLD)
    n = int(g["n"])
    sigs = [g[f"sig{i}"] for i in range(n)]
    out = MfccFrontEnd().features(sigs, standardize=True, want_fbank=True)
    fo_ = out["frame_off"]
    fb, mf, ft = (out[k].double().cpu().numpy() for k in ("fbank", "mfcc", "features"))
    for i in range(n):
        sl = slice(fo_[i], fo_[i + 1])
        assert fb[sl].shape == g[f"fbank{i}"].shape
        _close(fb[sl], g[f"fbank{i}"], TOL_FBANK)
        _close(mf[sl], g[f"mfcc{i}"], TOL_MFCC)
        if g[f"feat{i}"].shape[0] >= 8:            # very short utterances have near-zero std in some dimensions
            _close(ft[sl], g[f"feat{i}"], TOL_FEAT)


def test_front_end_matches_oracle_on_ragged_batch():
    from srb200.frontend import MfccFrontEnd
    rng = np.random.default_rng(5)
    lens = [int(v) for v in rng.integers(700, 40000, size=37)] + [320, 321, 160 * 50, 160 * 50 + 1]
    sigs = [(rng.normal(size=n).cumsum() * 40 + 2000 * np.sin(np.arange(n) * rng.uniform(0.01, 0.3))).astype(np.int16)
            for n in lens]
    sigs.append(np.zeros(4000, np.int16))              # silence: every filter hits the eps guard (feature.py:77)
    out = MfccFrontEnd().features(sigs, standardize=False, want_fbank=True)
    fo_ = out["frame_off"]
    fb, mf, ft = (out[k].double().cpu().numpy() for k in ("fbank", "mfcc", "features"))
    for i, s in enumerate(sigs):
        sl = slice(fo_[i], fo_[i + 1])
        rfb, rmf = fo.mfcc_features(s)
        assert fb[sl].shape == rfb.shape
        _close(fb[sl], rfb, TOL_FBANK)
        _close(mf[sl], rmf, TOL_MFCC)
        d = fo.delta_feature(rmf)
        _close(ft[sl], np.concatenate([rmf, d, fo.delta_feature(d)], axis=1), TOL_MFCC)
    # standardised features of the non-degenerate utterances
    out = MfccFrontEnd().features(sigs[:-1], standardize=True)
    ft = out["features"].double().cpu().numpy()
    for i, s in enumerate(sigs[:-1]):
        if len(s) >= 160 * 20:
            _close(ft[out["frame_off"][i]:out["frame_off"][i + 1]], fo.load_features(s), TOL_FEAT)


def test_sr_feature_and_core_drop_in(tmp_path):
    """The reference-facing functions: mfcc_features(path) and load_wav_as_mfcc1(path) on a wav file; core.test()
    classifies a folder of files with models trained through make_HMM (core.py:48-94) on the device back-end."""
    import sr
    from sr.feature import mfcc_features
    from sr.core import load_wav_as_mfcc1, make_HMM, test as core_test
    g = np.load(GOLD)
    p = str(tmp_path / "u0.wav")
    wavfile.write(p, 16000, g["sig0"])
    fb, mf = mfcc_features(p)
    _close(fb, g["fbank0"], TOL_FBANK)
    _close(mf, g["mfcc0"], TOL_MFCC)
    _close(load_wav_as_mfcc1(p), g["feat0"], TOL_FEAT)
    assert sr.make_HMM is make_HMM
    # two synthetic "digits": different pitch contours; 6 training + 3 test files each
    rng = np.random.default_rng(3)

    def word(kind, n=9000):
        t = np.arange(n) / 16000.0
        f = (150 + 200 * t / t[-1]) if kind == 0 else (420 - 260 * t / t[-1])
        s = sum(np.sin(2 * np.pi * np.cumsum(f * h) / 16000.0) / h for h in (1, 2, 3, 5))
        return (6000 * s + 300 * rng.normal(size=n)).astype(np.int16)

    models = []
    for kind in (0, 1):
        files = []
        for k in range(6):
            fpath = str(tmp_path / f"train_{kind}_{k}.wav")
            wavfile.write(fpath, 16000, word(kind))
            files.append(fpath)
        models.append(make_HMM(files, 5, use_gmm=True, use_em=True))
    os.mkdir(tmp_path / "test")
    for kind in (0, 1):
        for k in range(3):
            wavfile.write(str(tmp_path / "test" / f"{kind}_{k}.wav"), 16000, word(kind))
    acc = core_test(models, str(tmp_path / "test"), [r"0_\d\.wav", r"1_\d\.wav"] + [r"^$"] * 9)
    assert acc == 1.0
