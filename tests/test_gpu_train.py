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


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("shuffle", [False, True])
def test_em_iteration_matches_oracle(shuffle, fused):
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
    tr = IsolatedTrainer(X, off, words, pm, trans, S, ncomp, fused=fused)
    assert tr.fused == fused
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


@pytest.mark.parametrize("W,S,M,ncomp,skip,lens", [(3, 5, 4, 2, False, (2, 120)), (2, 3, 4, 1, False, (2, 120)),
                                                   (2, 8, 8, 4, True, (2, 120)), (1, 5, 4, 4, True, (2, 120)),
                                                   (2, 6, 4, 2, False, (2, 120)), (3, 5, 4, 2, True, (3, 250))])
def test_fused_pass_equals_unfused_kernels(W, S, M, ncomp, skip, lens):
    """srb_em_isolated against the K1 + K2 + K5 kernels it fuses: identical alignments (labels, boundaries),
    statistics within FP32 accumulation noise; ragged lengths incl. utterances too short to reach the last state."""
    import torch
    from srb200 import engine, synth
    from srb200.trainer import IsolatedTrainer, sort_by_word
    models, base = synth.make_models(71 + S, W=W, S=S, M=M)
    if skip:                                   # skip arcs of calc_transition_costs (kmeans.py:85-94, jump 2)
        models = [(m[0], m[1], m[2], m[3].copy()) for m in models]
        for m in models:
            m[3][2, 0] = 1.3
            m[3][S - 1, S - 3] = 0.7
    X, off, words = synth.make_utterances(72, base, 90, ragged=lens, noise=0.9)
    X, off, words, _ = sort_by_word(X, off, words)
    trans = np.stack([m[3] for m in models])
    pm_a, _ = engine.PackedModel.from_words(models)
    pm_b, _ = engine.PackedModel.from_words(models)
    a = IsolatedTrainer(X, off, words, pm_a, trans, S, ncomp, fused=True)
    assert a.tile_frames == (128 if lens[1] <= 128 else 256)
    b = IsolatedTrainer(X, off, words, pm_b, trans, S, ncomp, fused=False)
    seg = torch.zeros((a.B, S), dtype=torch.int32, device="cuda")
    lab = torch.full((a.N,), -7, dtype=torch.int32, device="cuda")
    a.fused_pass(seg_start=seg, frame_label=lab)
    b.align()
    la, lb = lab.cpu().numpy(), b.plan.label.cpu().numpy()
    ca, cb = a.best_cost.cpu().numpy(), b.plan.best_cost.cpu().numpy()
    assert np.array_equal(np.isfinite(ca), np.isfinite(cb))
    fin = np.isfinite(cb)
    assert fin.sum() > 40 and ((~fin).sum() > 0 or lens[0] > 2)
    # the fused kernel chains per-warp (min,+) segment matrices, K2 runs frame by frame: the same minimum summed in a
    # different FP32 order
    np.testing.assert_allclose(ca[fin], cb[fin], rtol=2e-5)
    # labels: state index within the word in the fused kernel's global numbering w*S+s == graph offset + s.  Equal
    # except where two paths tie within the FP32 re-association noise; there the fused path's own cost (FP64, from the
    # K1 emissions) must equal the optimum within the same tolerance.
    offn = np.asarray(off, np.int64)
    En = b.E.cpu().numpy().astype(np.float64)

    def path_cost(lbl, u):
        l = lbl[offn[u]:offn[u + 1]] - words[u] * S
        e = En[offn[u]:offn[u + 1]]
        tr = trans[words[u]].astype(np.float64)
        return e[np.arange(len(l)), l].sum() + tr[l[1:], l[:-1]].sum()

    ties = [u for u in np.nonzero(fin)[0] if not np.array_equal(la[offn[u]:offn[u + 1]], lb[offn[u]:offn[u + 1]])]
    assert len(ties) <= max(1, fin.sum() // 20)
    for u in list(np.nonzero(fin)[0][:8]) + ties:
        assert abs(path_cost(lb, u) - cb[u]) <= 2e-5 * abs(cb[u])
        assert abs(path_cost(la, u) - cb[u]) <= 2e-5 * abs(cb[u])
        assert la[offn[u]] == words[u] * S and la[offn[u + 1] - 1] == words[u] * S + S - 1
    assert np.array_equal(la[np.repeat(~fin, np.diff(offn))], lb[np.repeat(~fin, np.diff(offn))])
    # boundaries == get_segments_from_path of the fused labels
    sg = seg.cpu().numpy()
    for u in np.nonzero(fin)[0]:
        l = la[offn[u]:offn[u + 1]] - words[u] * S
        assert np.all(np.diff(l) >= 0)
        cnt = np.bincount(l, minlength=S)
        assert np.array_equal(sg[u, 1:], np.cumsum(cnt)[:-1]) and sg[u, 0] == 0
    # statistics: K5 on the fused kernel's alignment
    b.plan.label.copy_(lab)
    b.plan.best_cost.copy_(a.best_cost)
    b.accumulate()
    G, row = a.G, a.row
    sa, sb = a.stats[:G * row].cpu().numpy().reshape(G, row), b.stats[:G * row].cpu().numpy().reshape(G, row)
    scale = np.abs(sb).max(axis=1, keepdims=True) + 1e-30
    assert np.max(np.abs(sa - sb) / scale) < 2e-5
    np.testing.assert_allclose(sa[:, 0], sb[:, 0], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(a.stats[-2:].cpu().numpy(), b.stats[-2:].cpu().numpy(), rtol=1e-6)


def test_pipelined_host_iteration_equals_resident():
    """em_iteration_from_host (chunked H2D overlapped with the kernel) == em_iteration on the resident shard."""
    import torch
    from srb200 import engine, synth
    from srb200.trainer import IsolatedTrainer, sort_by_word
    W, S, M, ncomp = 3, 5, 4, 2
    models, base = synth.make_models(61, W=W, S=S, M=M)
    X, off, words = synth.make_utterances(63, base, 300, ragged=(30, 110), noise=0.9)
    X, off, words, _ = sort_by_word(X, off, words)
    trans = np.stack([m[3] for m in models])
    pa, _ = engine.PackedModel.from_words(models)
    pb, _ = engine.PackedModel.from_words(models)
    a = IsolatedTrainer(X, off, words, pa, trans, S, ncomp, fused=True)
    b = IsolatedTrainer(X, off, words, pb, trans, S, ncomp, fused=True)
    Xh = torch.from_numpy(np.ascontiguousarray(X)).pin_memory()
    b.X.zero_()
    a.em_iteration()
    b.em_iteration_from_host(Xh, chunks=5)
    torch.cuda.synchronize()
    np.testing.assert_allclose(b.stats.cpu().numpy(), a.stats.cpu().numpy(), rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(pb.mu.cpu().numpy(), pa.mu.cpu().numpy(), rtol=1e-7, atol=1e-9)
