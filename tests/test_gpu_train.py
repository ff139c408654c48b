"""GPU parity: one batched segmental-K-means + EM iteration (srb200.trainer.IsolatedTrainer) against the oracle's
per-word restatement of align_gmm_states (kmeans.py:196-205) + GMM.em (hmm_state.py:122-159)."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu
RTOL = 1e-4


def oracle_iteration(X, off, words, models, S, ncomp):
    """-> new (mu, cov, w) per word and mean Viterbi cost; frames pooled per state by the Viterbi alignment."""
    new, total = [], 0.0
    for wd, m in enumerate(models):
        pools = [[] for _ in range(S)]
        for u in np.nonzero(words == wd)[0]:
            x = X[off[u]:off[u + 1]].astype(np.float64)
            E = orc.emissions(x, [m], np.arange(S), log_domain=True)
            costs, path, _ = orc.decode_hmm_states(E, m[3])
            total += costs[-1, -1]
            lab = np.empty(len(x), np.int64)
            lab[-1] = S - 1
            for r, c in path:
                lab[c] = r
            for s in range(S):
                pools[s].append(x[lab == s])
        mu, cov, w = m[0].copy(), m[1].copy(), m[2].copy()
        for s in range(S):
            st = orc.GMMState(mu[s, 0], cov[s, 0], mu.shape[1])
            st.mu, st.cov, st.w = mu[s], cov[s], w[s]
            data = np.vstack(pools[s])
            st.em(data, ncomp, 1)
        new.append((mu, cov, w))
    return new, total / len(words)


@pytest.mark.parametrize("shuffle", [False, True])
def test_em_iteration_matches_oracle(shuffle):
    from srb200 import engine, synth
    from srb200.trainer import IsolatedTrainer
    W, S, M, ncomp = 4, 5, 4, 2
    models, base = synth.make_models(61, W=W, S=S, M=M)
    X, off, words = synth.make_utterances(62, base, 120, ragged=(25, 70), noise=0.9)
    if not shuffle:
        from srb200.trainer import sort_by_word
        X, off, words, _ = sort_by_word(X, off, words)
    pm, _ = engine.PackedModel.from_words(models)
    trans = np.stack([m[3] for m in models])
    tr = IsolatedTrainer(X, off, words, pm, trans, S, ncomp)
    tr.em_iteration()
    ref, mean_cost = oracle_iteration(X, off, words, models, S, ncomp)
    np.testing.assert_allclose(tr.mean_cost(), mean_cost, rtol=RTOL)
    mu, var, w = pm.mu.cpu().numpy(), pm.var.cpu().numpy(), pm.w.cpu().numpy()
    for wd in range(W):
        np.testing.assert_allclose(mu[wd * S:(wd + 1) * S], ref[wd][0], rtol=RTOL, atol=2e-5)
        np.testing.assert_allclose(var[wd * S:(wd + 1) * S], ref[wd][1], rtol=RTOL)
        np.testing.assert_allclose(w[wd * S:(wd + 1) * S], ref[wd][2], rtol=RTOL)
    # a second iteration keeps improving the alignment cost (EM monotonicity on the hard-alignment objective)
    c1 = tr.mean_cost()
    tr.em_iteration()
    assert tr.mean_cost() <= c1 + 1e-3
