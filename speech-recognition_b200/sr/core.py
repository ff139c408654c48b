"""Drop-in for the reference's sr/core.py (workflow façade) on the device back-end.

Same functions and arguments: `delta_feature`, `load_wav_as_mfcc1`, `load_wav_as_mfcc`, `make_HMM`, `train`, `test`,
`aurora_continuous_train`.  The MFCCs come from the device front-end (srb200.frontend), whole file lists per call;
`test` scores all files against all models with the fused classifier instead of the per-file, per-model loop of
core.py:79-87.  Model pickles keep the reference's module paths (`sr.recognition.hmm.HMM`), so `*.pkl` files written
by either side load on the other.

`load_wav_as_mfcc` in the reference uses `python_speech_features` (core.py:34-46: preemph 0.95, no zero padding of the
window); that package is not available here, so it maps to `load_wav_as_mfcc1`, the reference's own implementation
(core.py:25-31), and says so once."""
import os
import pickle
import re
import warnings

import numpy as np
from scipy.io import wavfile

from .feature import mfcc_batch, standardize  # noqa: F401
from .recognition import HMM, continuous_train
from .recognition.hmm import classify

try:                                                   # core.py:9 `from config import *`
    from config import digit_names, filename_index_map
except ImportError:                                    # the reference's config.py values
    digit_names = ['1', '2', '3', '4', '5', '6', '7', '8', '9', 'O', 'Z']
    filename_index_map = {d: i for i, d in enumerate(digit_names)}


def delta_feature(feat):
    """core.py:13-22."""
    feat = np.asarray(feat)
    delta = np.zeros(feat.shape)
    delta[0], delta[-1] = feat[1] - feat[0], feat[-1] - feat[-2]
    delta[1:-1] = feat[2:] - feat[:-2]
    return delta


def load_wavs_as_mfcc(paths):
    """Batched core.py:25-31: wav files -> list of standardised [T, 39] float64 feature matrices."""
    signals, rate = [], None
    for p in paths:
        sr_, sig = wavfile.read(p)
        if rate is not None and sr_ != rate:
            raise ValueError("all files of one call must share the sample rate")
        rate = sr_
        signals.append(sig)
    return mfcc_batch(signals, rate, standardized=True) if signals else []


def load_wav_as_mfcc1(path):
    return load_wavs_as_mfcc([path])[0]


_warned = []


def load_wav_as_mfcc(path):
    if not _warned:
        warnings.warn("python_speech_features is not available: load_wav_as_mfcc uses the reference's own "
                      "implementation (load_wav_as_mfcc1)")
        _warned.append(True)
    return load_wav_as_mfcc1(path)


def make_HMM(filenames, n_segs, use_gmm, use_em):
    print('Loading wav files to mfcc features')
    ys = load_wavs_as_mfcc(list(filenames))
    m = HMM(n_segs)
    print('Fitting HMMs')
    return m.fit(ys, n_gaussians=4, use_gmm=use_gmm, use_em=use_em)


def train(filenames, output_path, model_name, n_segs, use_gmm, use_em):
    models = make_HMM(filenames, n_segs, use_gmm=use_gmm, use_em=use_em)
    with open(os.path.join(output_path, model_name + '.pkl'), 'wb') as file:
        pickle.dump(models, file)


def test(models, folder, file_patterns):
    """core.py:63-94: fraction of files whose lowest-cost model is their digit (strict '<', first model wins)."""
    files, truth = [], []
    listing = os.listdir(folder)
    for digit in range(len(digit_names)):
        names = [os.path.join(folder, f) for f in listing if re.match(file_patterns[digit], f)]
        files += names
        truth += [digit] * len(names)
    xs = load_wavs_as_mfcc(files)
    _, best = classify(models, xs)
    for b, d in zip(best, truth):
        if b != d:
            print("Digit:", digit_names[d], "is wrong")
    return float(np.sum(np.asarray(best) == np.asarray(truth))) / len(files)


def aurora_continuous_train(input_path, output_path):
    """core.py:97-141: isolated models `<input_path><digit>.pkl` -> embedded training on ./train with cache/data.pkl."""
    models = []
    for hmm_index in range(11):
        with open(input_path + str(hmm_index) + '.pkl', 'rb') as file:
            model = pickle.load(file)
        for s in model.gmm_states:
            s.parent = hmm_index
        models.append(model)
    sequence_regex = re.compile('(?<=_)[OZ0-9]+(?=[AB])')
    filenames = sorted(f for f in os.listdir('train') if os.path.isfile(os.path.join('train', f)))
    sequences = [re.search(sequence_regex, f).group(0) for f in filenames]
    labels = [list(map(lambda x: filename_index_map[x], s)) for s in sequences]
    if not os.path.isdir('cache'):
        os.mkdir('cache')
    if os.path.isfile('cache/data.pkl'):
        print('using cache/data.pkl')
        with open('cache/data.pkl', 'rb') as f:
            data = pickle.load(f)
    else:
        print('loading data')
        data = load_wavs_as_mfcc([os.path.join('train', f) for f in filenames])
        with open('cache/data.pkl', 'wb') as f:
            pickle.dump(data, f)
    continuous_train(data, models, labels, output_path)
