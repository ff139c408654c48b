"""Drop-in for the reference's sr/feature/feature.py on the device front-end (srb200.frontend).

Same names and call signatures: `segment`, `zero_padding` (host helpers, feature.py:7-40), `mfcc_features(path)`
(feature.py:43-82: wav file -> (log10 filter banks [T,40], mfcc [T,13])), `standardize` (feature.py:85-88).
`mfcc_batch` is the batched form the reference lacks."""
import math

import numpy as np
from scipy.io import wavfile

from srb200 import frontend as _fe

_front = {}


def _frontend(sample_rate):
    if sample_rate not in _front:
        _front[sample_rate] = _fe.MfccFrontEnd(sample_rate)
    return _front[sample_rate]


def segment(sig, sample_rate=16000, frame_len=0.025, frame_step=0.01):
    slen = sig.size
    fl, fs = int(frame_len * sample_rate), int(frame_step * sample_rate)
    num_frames = math.ceil(slen / fs) if slen > frame_len else 1
    pad_sig = np.concatenate([sig, np.zeros(int((num_frames - 1) * fs + fl) - slen)])
    return pad_sig[np.arange(num_frames)[:, None] * fs + np.arange(fl)[None, :]]


def zero_padding(frames, frame_len=None):
    height, width = frames.shape
    if frame_len is None:
        frame_len = 1 << (width - 1).bit_length()
    left = (frame_len - width) // 2
    f = np.zeros((height, frame_len))
    f[:, left:left + width] = frames
    return f


def mfcc_batch(signals, sample_rate=16000, standardized=False):
    """list of sample arrays -> list of (filter_banks [T,40], mfcc [T,13]) float64 pairs, or, with standardized=True,
    list of the [T,39] feature matrices of core.load_wav_as_mfcc1."""
    out = _frontend(sample_rate).features(signals, standardize=standardized, want_fbank=not standardized)
    fo = out["frame_off"]
    if standardized:
        feat = out["features"].double().cpu().numpy()
        return [feat[fo[u]:fo[u + 1]] for u in range(len(signals))]
    fb, mf = out["fbank"].double().cpu().numpy(), out["mfcc"].double().cpu().numpy()
    return [(fb[fo[u]:fo[u + 1]], mf[fo[u]:fo[u + 1]]) for u in range(len(signals))]


def mfcc_features(path_file, frame_size=0.025, frame_stride=0.01, low_freq=80, high_freq=None):
    if (frame_size, frame_stride, low_freq, high_freq) != (0.025, 0.01, 80, None):
        raise NotImplementedError("the device front-end implements the reference's default framing and filter bank")
    sample_rate, signal = wavfile.read(path_file)
    return mfcc_batch([signal], sample_rate)[0]


def standardize(data):
    data = data - np.mean(data, axis=0)
    return data / np.std(data, axis=0)


__all__ = ["segment", "zero_padding", "mfcc_features", "mfcc_batch", "standardize"]
