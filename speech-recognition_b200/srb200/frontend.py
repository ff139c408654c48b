"""Batched MFCC front-end on the device (srb_mfcc + srb_delta_standardize; sr/feature/feature.py:7-88, core.py:13-31)."""
import numpy as np
import torch

from . import lib as _l
from .engine import device

NFFT, NFILT, NCEP = 512, 40, 13


def mel_filterbank(sample_rate=16000, low_freq=80, high_freq=None):
    """feature.py:57-75 in float64 -> (dense weights float32 [40, 257], [first, last) non-zero bin per filter int32)."""
    high_freq = high_freq or sample_rate / 2
    lo, hi = 2595 * np.log10(1 + low_freq / 700), 2595 * np.log10(1 + high_freq / 700)
    hz = 700 * (10 ** (np.linspace(lo, hi, NFILT + 2) / 2595) - 1)
    b = np.floor((NFFT + 1) * hz / sample_rate)
    w = np.zeros((NFILT, NFFT // 2 + 1))
    rng = np.zeros((NFILT, 2), np.int32)
    for m in range(1, NFILT + 1):
        l, c, r = int(b[m - 1]), int(b[m]), int(b[m + 1])
        k = np.arange(l, c)
        w[m - 1, k] = (k - b[m - 1]) / (b[m] - b[m - 1])
        k = np.arange(c, r)
        w[m - 1, k] = (b[m + 1] - k) / (b[m + 1] - b[m])
        rng[m - 1] = (l, r)
    return w.astype(np.float32), rng


class MfccFrontEnd:
    """signals (list of 1-D sample arrays, int16 or float) -> standardised [T, 39] features per utterance, all
    utterances in two launches.  `features()` returns one device tensor + frame offsets ready for the scoring path."""

    def __init__(self, sample_rate=16000):
        if sample_rate != 16000:
            raise NotImplementedError("the device front-end is built for 16 kHz (feature.py defaults)")
        self.sample_rate = sample_rate
        w, r = mel_filterbank(sample_rate)
        self.fb_w = torch.from_numpy(w).to(device())
        self.fb_range = torch.from_numpy(r).to(device())

    def stage(self, signals):
        """Host sample arrays -> device buffers + offsets (one H2D copy of the concatenated float32 samples)."""
        L = _l.load()
        n = np.array([len(s) for s in signals], np.int64)
        nf = np.array([L.srb_mfcc_num_frames(int(k), self.sample_rate) for k in n], np.int64)
        if (nf < 2).any():
            raise IndexError("an utterance gives fewer than 2 frames (core.py:13-22 indexes feat[1])")
        sig_off = np.zeros(len(n) + 1, np.int64)
        np.cumsum(n, out=sig_off[1:])
        frame_off = np.zeros(len(n) + 1, np.int64)
        np.cumsum(nf, out=frame_off[1:])
        dev = device()
        if len(signals) and all(getattr(s, "dtype", None) == np.int16 for s in signals):
            # wav samples: copy the int16 bytes, widen on the device (half the PCIe traffic, no host conversion)
            sig = torch.from_numpy(np.concatenate(signals)).to(dev).float()
        else:
            flat = np.concatenate([np.asarray(s, np.float32) for s in signals]) if len(signals) else np.zeros(0, np.float32)
            sig = torch.from_numpy(flat).to(dev)
        return {"sig": sig, "so": torch.from_numpy(sig_off).to(dev),
                "fo": torch.from_numpy(frame_off).to(dev), "B": len(n), "F": int(frame_off[-1]), "frame_off": frame_off}

    def run(self, st, standardize=True, want_fbank=False, out=None):
        """srb_mfcc + srb_delta_standardize on staged buffers.  `out` (a previous result) is reused when given."""
        L, p, dev, F = _l.load(), _l.ptr, device(), st["F"]
        if out is None:
            out = {"mfcc": torch.empty((max(F, 1), NCEP), dtype=torch.float32, device=dev),
                   "features": torch.empty((max(F, 1), 3 * NCEP), dtype=torch.float32, device=dev),
                   "fbank": torch.empty((max(F, 1), NFILT), dtype=torch.float32, device=dev) if want_fbank else None}
        _l.check(L.srb_mfcc(p(st["sig"]), p(st["so"]), p(st["fo"]), st["B"], F, self.sample_rate, p(self.fb_w),
                            p(self.fb_range), p(out["fbank"]), p(out["mfcc"]), _l.stream_ptr()))
        _l.check(L.srb_delta_standardize(p(out["mfcc"]), p(st["fo"]), st["B"], p(out["features"]),
                                         int(bool(standardize)), _l.stream_ptr()))
        return out

    def features(self, signals, standardize=True, want_fbank=False):
        st = self.stage(signals)
        o = self.run(st, standardize, want_fbank)
        F = st["F"]
        out = {"features": o["features"][:F], "frame_off": st["frame_off"], "mfcc": o["mfcc"][:F]}
        if want_fbank:
            out["fbank"] = o["fbank"][:F]
        return out
