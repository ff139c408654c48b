"""GPU: one Baum-Welch iteration (srb200.trainer.BaumWelchTrainer: K1 + K4 + srb_gmm_stats_soft + M-step) against a
float64 numpy restatement built on the oracle's forward-backward.  The reference has no Baum-Welch, so
parity cannot be pinned to reference outputs; it is pinned as far as it can be: the oracle's forward-backward is
checked against explicit path enumeration and is bit-identical to decode_hmm_states in the (min,+) semiring
(tests/test_oracle_golden.py), the kernels are checked against both (tests/test_gpu_dp.py), and the soft statistics
reduce to the hard-count kernel (pinned to GMM.em) in the 0/1 limit (below).  Tolerance: BASELINE north_star's 1e-4
relative on re-estimated weights / means / variances / transition probabilities (means: 1e-4 of max(|mu|, sigma), the
scale a mean is known to)."""
import numpy as np
import pytest

import engine_double as ed
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def numpy_iteration(X, off, words, models, S, ncomp):
    W, M, D = len(models), models[0][0].shape[1], X.shape[1]
    mu = np.concatenate([m[0] for m in models]).copy()
    var = np.concatenate([m[1] for m in models]).copy()
    w = np.concatenate([m[2] for m in models]).copy()
    pm = ed.PackedModel(mu, var, w)
    l = ed._log_comps(X, pm)                                     # [N, R, M]
    E = ed._neg_lse(l, 2)
    S0 = np.zeros((W * S, ncomp)); S1 = np.zeros((W * S, ncomp, D)); S2 = np.zeros((W * S, ncomp, D))
    cnt = np.zeros(W * S)
    xi = [np.zeros((S, S)) for _ in range(W)]
    nlz_sum = 0.0
    for u in range(len(words)):
        wd, a, b = words[u], off[u], off[u + 1]
        tr = models[wd][3]
        Eu = E[a:b, wd * S:(wd + 1) * S]
        nlz, al, be, ga = orc.forward_backward(Eu, tr)
        nlz_sum += nlz
        for c in range(1, b - a):
            for r in range(S):
                for o in range(S):
                    if np.isfinite(tr[r, o]):
                        xi[wd][r, o] += np.exp(-(al[c - 1, o] + tr[r, o] + Eu[c, r] + be[c, r] - nlz))
        for s in range(S):
            g = wd * S + s
            lr = l[a:b, g, :ncomp]
            p = np.exp(lr - lr.max(1, keepdims=True))
            p = p / p.sum(1, keepdims=True) * ga[:, [s]]
            cnt[g] += ga[:, s].sum()
            for m in range(ncomp):
                e = X[a:b] - mu[g, m]
                S0[g, m] += p[:, m].sum()
                S1[g, m] += (p[:, [m]] * e).sum(0)
                S2[g, m] += (p[:, [m]] * e * e).sum(0)
    for g in range(W * S):
        for m in range(ncomp):
            m1 = S1[g, m] / S0[g, m]
            mu[g, m] += m1
            var[g, m] = S2[g, m] / S0[g, m] - m1 * m1
            w[g, m] = S0[g, m] / cnt[g]
    trans = []
    for wd in range(W):
        t = models[wd][3].copy()
        out = xi[wd].sum(0)                                      # mass leaving each origin
        for r in range(S):
            for o in range(S):
                if np.isfinite(t[r, o]) and xi[wd][r, o] > 0:
                    t[r, o] = -np.log(xi[wd][r, o] / out[o])
        trans.append(t)
    return mu, var, w, np.stack(trans), nlz_sum / len(words)


@pytest.mark.parametrize("base_scale", [1.0, 0.12])
@pytest.mark.parametrize("S,M,ncomp,tensor", [(4, 4, 2, False), (8, 8, 4, False), (5, 4, 3, False), (3, 8, 8, False),
                                               (6, 2, 1, False), (8, 16, 4, True), (5, 4, 2, True), (4, 8, 4, True)])
def test_baum_welch_iteration(S, M, ncomp, tensor, base_scale):
    """ncomp = 1, 2, 3 (padded lane group), 4, 8 cover every lane-group width of the soft-count kernel; S = 3 .. 8 the
    4- and 8-lane groups of the banded forward-backward kernel."""
    from srb200 import engine, synth
    from srb200.trainer import BaumWelchTrainer, sort_by_word
    W = 2
    # base_scale 0.12: neighbouring states overlap, the posteriors are genuinely soft (synth.make_models)
    models, base = synth.make_models(81, W=W, S=S, M=M, base_scale=base_scale, spread=0.5 * base_scale)
    X, off, words = synth.make_utterances(82, base, 24, ragged=(12, 40), noise=1.1)
    X, off, words, _ = sort_by_word(X, off, words)
    pm, _ = engine.PackedModel.from_words(models)
    tr = BaumWelchTrainer(X, off, words, pm, np.stack([m[3] for m in models]), S, ncomp, gmin=0.0, tensor=tensor)
    assert bool(tr.tensor) == tensor           # tensor: states scored as Mp/4 sub-states on the tcgen05 scorer
    tr.em_iteration()
    mu, var, w, trans, nlz = numpy_iteration(X.astype(np.float64), off.astype(np.int64), words, models, S, ncomp)
    np.testing.assert_allclose(tr.mean_neg_log_z(), nlz, rtol=1e-6)
    np.testing.assert_allclose(pm.w.cpu().numpy(), w, rtol=1e-4)
    scale = np.maximum(np.abs(mu), np.sqrt(var))
    assert np.max(np.abs(pm.mu.cpu().numpy() - mu) / scale) < 1e-4
    np.testing.assert_allclose(pm.var.cpu().numpy(), var, rtol=1e-4)
    fin = np.isfinite(trans)
    assert np.array_equal(np.isfinite(tr.trans), fin)
    np.testing.assert_allclose(np.exp(-tr.trans[fin]), np.exp(-trans[fin]), rtol=1e-4)      # transition probabilities
    # EM monotonicity: the data log-likelihood does not decrease
    c1 = tr.mean_neg_log_z()
    tr.em_iteration()
    assert tr.mean_neg_log_z() <= c1 + 1e-3


@pytest.mark.parametrize("M,ncomp", [(4, 2), (4, 4), (4, 1), (4, 3), (8, 8), (8, 5)])
def test_soft_stats_reduce_to_hard_counts(M, ncomp):
    """gamma = 0/1 indicator of an alignment => srb_gmm_stats_soft == srb_gmm_stats (which is pinned to GMM.em)."""
    import torch
    from srb200 import engine, lib, synth
    W, S = 1, 5
    models, base = synth.make_models(91, W=W, S=S, M=M)
    X, off, words = synth.make_utterances(92, base, 10, ragged=(20, 50))
    pm, _ = engine.PackedModel.from_words(models)
    N = X.shape[0]
    rng = np.random.default_rng(3)
    lab = rng.integers(0, S, size=N).astype(np.int32)
    hard = engine.gmm_stats(X, lab, pm, ncomp).cpu().numpy()
    gamma = torch.zeros((N, S), dtype=torch.float32, device="cuda")
    gamma[torch.arange(N), torch.from_numpy(lab).long()] = 1.0
    L = lib.load()
    row = L.srb_stats_row(pm.D)
    out = torch.zeros((S * M, row), dtype=torch.float64, device="cuda")
    partial = torch.empty(L.srb_stats_parts() * S * M * row, dtype=torch.float32, device="cuda")
    lc = engine.gauss_consts(pm.var, pm.w)
    Xd = engine.as_frames(X)
    lib.check(L.srb_gmm_stats_soft(lib.ptr(Xd), N, 0, N, pm.D, lib.ptr(gamma), S, M, ncomp, lib.ptr(pm.mu), lib.ptr(pm.var),
                                   lib.ptr(lc), 0.0, lib.ptr(partial), lib.ptr(out), 0, lib.stream_ptr()))
    soft = out.cpu().numpy()
    D = pm.D
    scale = np.abs(hard).max(axis=1, keepdims=True) + 1e-30
    assert np.max(np.abs(soft[:, :2 * D + 2] - hard[:, :2 * D + 2]) / scale) < 2e-5
