"""Seeded synthetic models and utterances of the shapes BASELINE.json names.

Pure numpy; shared by tests/ and bench.py so that the CUDA path and its CPU checker see identical inputs."""
import numpy as np

D = 39


def ltr_transitions(S, p_stay=0.9):
    """Left-to-right costs in the reference convention trans[i, j] = cost j -> i (kmeans.py:85-94 shape)."""
    tr = np.full((S, S), np.inf)
    for i in range(S):
        if i < S - 1:
            tr[i, i] = -np.log(p_stay)
            tr[i + 1, i] = -np.log(1 - p_stay)
        else:
            tr[i, i] = 0.0
    return tr


def make_models(seed, W=11, S=5, M=4, dim=D, spread=0.5, normalised=True, base_scale=1.0):
    """-> (word_models, base[W,S,dim]).  word model = (mu[S,M,dim], var[S,M,dim], w[S,M], trans[S,S]).
    base_scale < 1 moves the state means together (0.12: neighbouring states about one standard deviation apart in
    the 39-dimensional space), which makes alignments and posteriors genuinely soft -- the hard case for parity."""
    rng = np.random.default_rng(seed)
    base = base_scale * rng.normal(0.0, 1.0, size=(W, S, dim))
    models = []
    for wd in range(W):
        mu = base[wd][:, None, :] + spread * rng.normal(size=(S, M, dim))
        var = rng.uniform(0.5, 1.5, size=(S, M, dim))
        w = rng.dirichlet(np.ones(M), size=S)
        if not normalised:          # appendix A2: untrained components keep 1/M, sum is 1.5
            w[:, M // 2:] = 1.0 / M
        models.append((mu, var, w, ltr_transitions(S)))
    return models, base


def make_utterances(seed, base, n_utts, T=100, noise=0.7, ragged=None, dtype=np.float32):
    """Utterances walking left-to-right through a random word's state means.
    -> (X[N,dim] dtype, off[B+1] int32, words[B]).  ragged=(lo, hi) draws lengths uniformly."""
    rng = np.random.default_rng(seed)
    W, S, dim = base.shape
    words = rng.integers(0, W, size=n_utts)
    lens = np.full(n_utts, T) if ragged is None else rng.integers(ragged[0], ragged[1] + 1, size=n_utts)
    off = np.zeros(n_utts + 1, np.int64)
    np.cumsum(lens, out=off[1:])
    N = int(off[-1])
    # state of every frame: uniform segments
    u_of = np.repeat(np.arange(n_utts), lens)
    t_in = np.arange(N) - off[u_of]
    seg = np.minimum((t_in * S) // lens[u_of], S - 1)
    X = base[words[u_of], seg].astype(dtype)
    # noise in blocks to bound peak memory
    step = 1 << 20
    for i in range(0, N, step):
        X[i:i + step] += (noise * rng.standard_normal(size=(min(step, N - i), dim), dtype=np.float32)).astype(dtype)
    return X, off.astype(np.int32), words.astype(np.int64)


def split_utts(X, off):
    return [X[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def make_strings(seed, base, n_utts, n_words=7, T=500, noise=0.7):
    """Connected utterances (BASELINE config 4): n_words random words back to back, T frames in total, each word
    walking left-to-right through its state means.  -> (X[N,dim] float32, off[B+1] int32, strings[B,n_words])"""
    rng = np.random.default_rng(seed)
    W, S, dim = base.shape
    strings = rng.integers(0, W, size=(n_utts, n_words))
    t = np.arange(T)
    word_pos = np.minimum(t * n_words // T, n_words - 1)                 # word slot of every frame
    w_start = (np.arange(n_words) * T + n_words - 1) // n_words
    w_len = np.diff(np.append(w_start, T))
    seg = np.minimum((t - w_start[word_pos]) * S // w_len[word_pos], S - 1)
    X = base[strings[:, word_pos], seg[None, :]].reshape(n_utts * T, dim).astype(np.float32)
    step = 1 << 20
    for i in range(0, X.shape[0], step):
        X[i:i + step] += noise * rng.standard_normal(size=(min(step, X.shape[0] - i), dim), dtype=np.float32)
    off = (np.arange(n_utts + 1) * T).astype(np.int32)
    return X, off, strings
