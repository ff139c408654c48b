"""Model layer of the sr.recognition mirror (reference: sr/recognition/hmm.py:8-135).

`HMM` keeps the reference's attributes (`mu`, `sigma`, `transitions`, `segments`, `gmm_states`, `use_gmm`,
`use_em`) and methods; fitting and evaluation are batched device operations (skmeans / kmeans / GMM.em /
Viterbi re-alignment from .kmeans and .hmm_state, the fused score+Viterbi classifier for `evaluate`).
"""
import numpy as np
import torch

from . import _ops
from ._ops import engine
from .decode import decode_hmm_states, dtw
from .hmm_state import GMM, mahalanobis
from .kmeans import _kmeans_device, align_gmm_states, skmeans


class HMM:
    """Left-to-right word model of `n_segments` states (hmm.py:8-55)."""

    def __init__(self, n_segments):
        self.n_segments = n_segments
        self.mu = None
        self.sigma = None
        self.transitions = None
        self.segments = []
        self.gmm_states = None
        self.use_gmm = True
        self.use_em = True

    def __eq__(self, other):
        if not self.use_gmm:
            return np.allclose(self.mu, other.mu) and np.allclose(self.sigma, other.sigma)
        if not other.use_gmm or self.n_segments != other.n_segments:
            return False
        return all(self.gmm_states[i] == other.gmm_states[i] for i in range(self.n_segments))

    def reset(self):
        self.mu = self.sigma = self.transitions = self.gmm_states = None
        self.segments = []

    def __getitem__(self, item):
        assert self.use_gmm == True  # noqa: E712 (hmm.py:51)
        if type(item) is int or type(item) is slice:
            return self.gmm_states[item]
        raise TypeError('The type of index is not supported')

    def fit(self, ys, n_gaussians, use_gmm=True, use_em=True):
        """Fit from a list of [T, D] utterances (hmm.py:57-78).  use_gmm=False: one Gaussian per state from
        segmental K-means; otherwise binary-split K-means (+ EM when use_em) per state."""
        self.use_em = use_em
        self.use_gmm = use_gmm
        if use_gmm:
            self.fit_GMM(ys, n_gaussians)
        else:
            self.mu, self.sigma, self.transitions, self.segments = skmeans(ys, self.n_segments,
                                                                           return_segmented_data=True)
        return self

    def _init_gmm(self, n_gaussians):
        self.gmm_states = [GMM(self.mu[i, :], self.sigma[i, :], n_gaussians) for i in range(self.n_segments)]

    def fit_GMM(self, ys, n_gaussians):
        """skmeans -> per-state split / K-means / EM -> Viterbi re-alignment (hmm.py:83-95), all states of the word at
        once on the device (fit_words with one word: same draws of the global numpy RNG, same "EM converged" lines)."""
        m = fit_words([ys], self.n_segments, n_gaussians, use_em=self.use_em)[0]
        self.mu, self.sigma, self.transitions = m.mu, m.sigma, m.transitions
        self.gmm_states, self.segments = m.gmm_states, m.segments
        ft = fit_words.last
        for g in range(self.n_segments):                       # the reference prints this per state and split (hmm_state.py:153)
            for its, conv in zip(ft.iterations.get("em", []), ft.converged.get("em", [])):
                if conv[g]:
                    print("EM converged at iteration:", int(its[g]))

    def fit_GMM_per_state(self, ys, n_gaussians):
        """The same fit one state at a time (the reference's own loop order, one launch per K-means / EM iteration and
        state): kept for callers that hook `_fit_GMM`."""
        self.mu, self.sigma, self.transitions, self.segments = skmeans(ys, self.n_segments,
                                                                       return_segmented_data=True)
        self._init_gmm(n_gaussians)
        for i in range(len(self.segments)):
            self._fit_GMM(self.segments[i], n_gaussians, i)
        self.segments = align_gmm_states(ys, self.gmm_states, self.transitions, self.n_segments)

    def _fit_GMM(self, data, n_gaussians, seg_i):
        """One state (hmm.py:97-124): int(log(n)) binary splits (natural log, A1); after each split K-means on the
        x0.9 / x1.1 centroids, weights from cluster counts, then EM over the 2^(i+1) trained components."""
        n_splits = int(np.log(n_gaussians))
        assert n_splits > 0
        Xd = _ops.to_device(data)
        n = data.shape[0]
        state = self.gmm_states[seg_i]
        centroids = np.array([self.mu[seg_i, :]])
        weights = np.full(n_gaussians, 1 / n)
        for i in range(n_splits):
            k = 2 ** (i + 1)
            centroids = np.concatenate([centroids * 0.9, centroids * 1.1], axis=0)
            clusters, centroids, variance = _kmeans_device(Xd, data, k, centroids, mahalanobis, 1000)
            cs, c_counts = np.unique(clusters, return_counts=True)
            for c in cs:
                weights[c] = c_counts[c] / n              # indexed by label as in hmm.py:116-118 (A4)
            state.update_models(centroids, variance, weights[:k])
            if self.use_em:
                state._em_device(Xd, k, 10000)

    def evaluate(self, x):
        """Cost of the best path ending in the last state at the last frame, costs[-1, -1] (hmm.py:126-135)."""
        if self.use_gmm:
            return float(self.evaluate_batch([x])[0])
        costs, _ = dtw(x, self.mu, mahalanobis, self.transitions, self.sigma)
        return costs[-1, -1]

    def evaluate_batch(self, xs):
        """evaluate() of many utterances in one launch -> float64 [B]."""
        return classify([self], xs)[0][:, 0]


def classify(models, xs):
    """The isolated-word test loop (core.py:79-87) for many utterances: every x against every model.
    -> (costs [B, W] float64, labels [B] int64 = first argmin, strict '<').  Models of up to
    SRB_DP_MAX_STATES equal-sized GMM states go through the fused score+Viterbi kernel; others through K1 + K2."""
    S = models[0].n_segments
    fused = all(m.use_gmm and m.n_segments == S for m in models) and S <= 8 and all(len(x) >= 2 for x in xs)
    if fused:
        X, off = _ops.cat_frames(xs)
        pm = _ops.pack_gmms([s for m in models for s in m.gmm_states])
        trans = torch.as_tensor(np.stack([np.asarray(m.transitions, np.float64) for m in models])) \
            .to(pm.ab.device, torch.float32).contiguous()
        final, label = engine.classify(X, off, pm, trans, len(models), S)
        return final.double().cpu().numpy(), label.cpu().numpy().astype(np.int64)
    costs = np.empty((len(xs), len(models)))
    for w, m in enumerate(models):
        for u, x in enumerate(xs):
            if m.use_gmm:
                costs[u, w] = decode_hmm_states(x, m.gmm_states, m.transitions)[0][-1, -1]
            else:
                costs[u, w] = m.evaluate(x)
    return costs, np.argmin(costs, axis=1)


def fit_words(ys_per_word, n_segments, n_gaussians, use_em=True, seeds=None, group=None, fused_align=True,
              precision="f64"):
    """`HMM(n_segments).fit(ys, n_gaussians)` for EVERY word at once (hmm.py:57-124 per word): all words' utterances are
    one device-resident corpus and every phase -- skmeans, binary-split K-means, per-state EM, re-alignment -- runs for
    all states of all words together (srb200.fit.WordFitter; with torch.distributed initialised the utterances of this
    process are its shard and the statistics are summed over the ranks).
    ys_per_word: list (one entry per word) of lists of [T, D] utterances; seeds: optional per-word seeds of the global
    numpy RNG, applied where a caller of the reference would seed before each word's fit.  -> list of HMM"""
    from srb200.fit import WordFitter
    from .hmm_state import GMM
    from .kmeans import segment_data
    utts = [u for ys in ys_per_word for u in ys]
    words = np.repeat(np.arange(len(ys_per_word)), [len(ys) for ys in ys_per_word])
    X, off = _ops.cat_frames(utts)
    ft = WordFitter(X, off, words, len(ys_per_word), n_segments, n_gaussians, use_em=use_em, seeds=seeds, group=group,
                    fused_align=fused_align, precision=precision)
    models = ft.fit()
    out, S = [], n_segments
    for w, (mu, var, wt, trans) in enumerate(models):
        m = HMM(S)
        m.use_gmm, m.use_em = True, use_em
        m.mu, m.sigma, m.transitions = ft.hmm_mu[w].copy(), ft.hmm_sigma[w].copy(), trans.copy()
        m.gmm_states = []
        for s in range(S):
            g = GMM(m.mu[s], m.sigma[s], n_gaussians)
            g.update_models(mu[s].copy(), var[s].copy(), wt[s].copy())
            g.mu_old, g.sigma_old, g.w_old = ft.mu_old[w * S + s].copy(), ft.var_old[w * S + s].copy(), ft.w_old[w * S + s].copy()
            m.gmm_states.append(g)
        a, b = ft.word_utts[w]
        m.segments = segment_data(ys_per_word[w], b - a, S, ft.final_starts[a:b]) if b > a else []
        out.append(m)
    fit_words.last = ft
    return out


__all__ = ["HMM", "classify", "fit_words"]
