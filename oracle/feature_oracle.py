"""CPU restatement (float64 numpy) of the reference's MFCC front-end -- TEST INFRASTRUCTURE ONLY.

Follows /root/reference/sr/feature/feature.py:7-88 and /root/reference/sr/core.py:13-31 function by function, taking
the sample array instead of a wav path (`mfcc_features` reads the file with scipy.io.wavfile, feature.py:44).
Pinned by tests/test_oracle_golden.py against tests/golden/ref_feature.npz, which tests/golden/make_golden.py
produces by running the UNMODIFIED reference functions on wav files it writes.  Only tests/, smoke() and bench.py's CPU
legs may import this module."""
import math

import numpy as np


def segment(sig, sample_rate=16000, frame_len=0.025, frame_step=0.01):
    """feature.py:7-22.  NB `slen > frame_len` compares samples with seconds (always true for a non-empty signal)."""
    slen = sig.size
    fl, fs = int(frame_len * sample_rate), int(frame_step * sample_rate)
    nf = math.ceil(slen / fs) if slen > frame_len else 1
    pad = np.concatenate([sig, np.zeros(int((nf - 1) * fs + fl) - slen)])
    idx = np.arange(nf)[:, None] * fs + np.arange(fl)[None, :]
    return pad[idx]


def zero_padding(frames, frame_len=None):
    """feature.py:25-40: centre every frame in the next power of two."""
    h, w = frames.shape
    if frame_len is None:
        frame_len = 1 << (w - 1).bit_length()
    left = (frame_len - w) // 2
    f = np.zeros((h, frame_len))
    f[:, left:left + w] = frames
    return f


def mel_filterbank(sample_rate=16000, nfft=512, nfilt=40, low_freq=80, high_freq=None):
    """feature.py:57-75 -> fbank[nfilt, nfft/2+1] (float64) with the reference's floor()ed bin edges."""
    high_freq = high_freq or sample_rate / 2
    lo, hi = 2595 * np.log10(1 + low_freq / 700), 2595 * np.log10(1 + high_freq / 700)
    hz = 700 * (10 ** (np.linspace(lo, hi, nfilt + 2) / 2595) - 1)
    b = np.floor((nfft + 1) * hz / sample_rate)
    fbank = np.zeros((nfilt, nfft // 2 + 1))
    for m in range(1, nfilt + 1):
        l, c, r = int(b[m - 1]), int(b[m]), int(b[m + 1])
        for k in range(l, c):
            fbank[m - 1, k] = (k - b[m - 1]) / (b[m] - b[m - 1])
        for k in range(c, r):
            fbank[m - 1, k] = (b[m + 1] - k) / (b[m + 1] - b[m])
    return fbank, b


def mfcc_features(signal, sample_rate=16000):
    """feature.py:43-82 on a sample array.  -> (log10 filter-bank energies [T,40], mfcc [T,13])."""
    signal = np.asarray(signal, np.float64)
    emph = np.append(signal[0], signal[1:] - 0.97 * signal[:-1])
    frames = zero_padding(segment(emph, sample_rate))
    frames = frames * np.hamming(frames.shape[1])
    nfft = 512
    powf = (1.0 / nfft) * np.absolute(np.fft.rfft(frames, nfft)) ** 2
    fbank, _ = mel_filterbank(sample_rate, nfft)
    fb = powf @ fbank.T
    fb = np.log10(np.where(fb == 0, np.finfo(float).eps, fb))
    n = fb.shape[1]
    k = np.arange(1, 14)[:, None]
    basis = np.sqrt(2.0 / n) * np.cos(np.pi * (2 * np.arange(n)[None, :] + 1) * k / (2 * n))   # scipy dct II, ortho, k >= 1
    return fb, fb @ basis.T


def delta_feature(feat):
    """core.py:13-22."""
    d = np.zeros(feat.shape)
    d[0], d[-1] = feat[1] - feat[0], feat[-1] - feat[-2]
    d[1:-1] = feat[2:] - feat[:-2]
    return d


def standardize(data):
    """feature.py:85-88."""
    return (data - data.mean(axis=0)) / data.std(axis=0)


def load_features(signal, sample_rate=16000):
    """core.py:25-31 (load_wav_as_mfcc1) on a sample array -> standardised [T, 39]."""
    _, c = mfcc_features(signal, sample_rate)
    d = delta_feature(c)
    return standardize(np.concatenate([c, d, delta_feature(d)], axis=1))
