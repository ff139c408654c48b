"""GPU: size-independent properties at benchmark-like scale (2 M frames), where the oracle would take too long:
cross-kernel identities, conservation laws, shard additivity (what the multi-GPU allreduce relies on)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, S, M, NCOMP, B, T = 11, 5, 4, 2, 20000, 100


@pytest.fixture(scope="module")
def corpus():
    from srb200 import synth
    from srb200.trainer import sort_by_word
    models, base = synth.make_models(1, W=W, S=S, M=M)
    X, off, words = synth.make_utterances(777, base, B, T=T)
    X, off, words, _ = sort_by_word(X, off, words)
    return models, X, off, words


def test_fused_pass_invariants_and_cross_kernel_identity(corpus):
    import torch
    from srb200 import engine
    from srb200.trainer import IsolatedTrainer
    models, X, off, words = corpus
    pm, trans_d = engine.PackedModel.from_words(models)
    tr = IsolatedTrainer(X, off, words, pm, np.stack([m[3] for m in models]), S, NCOMP, fused=True)
    seg = torch.zeros((B, S), dtype=torch.int32, device="cuda")
    lab = torch.empty(B * T, dtype=torch.int32, device="cuda")
    tr.fused_pass(seg_start=seg, frame_label=lab)
    G, row, D = tr.G, tr.row, pm.D
    st = tr.stats[:G * row].cpu().numpy().reshape(W * S, M, row)
    # conservation: every frame is counted once; posteriors of the trained components sum to the state count
    np.testing.assert_allclose(st[:, 0, 2 * D + 1].sum(), B * T, rtol=1e-7)
    np.testing.assert_allclose(st[:, :NCOMP, 0].sum(axis=1), st[:, 0, 2 * D + 1], rtol=2e-5)
    assert np.all(st[:, NCOMP:, :] == 0)
    # alignments: labels are non-decreasing inside an utterance, start in state 0, end in state S-1
    l = lab.cpu().numpy().reshape(B, T) - (words * S)[:, None]
    assert l.min() == 0 and l.max() == S - 1
    assert np.all(np.diff(l, axis=1) >= 0) and np.all(l[:, 0] == 0) and np.all(l[:, -1] == S - 1)
    sg = seg.cpu().numpy()
    assert np.array_equal(sg[:, 1:], np.stack([(l < k).sum(axis=1) for k in range(1, S)], axis=1))
    # cross-kernel identity: costs[-1,-1] of the utterance's own word from the fused training kernel == the fused
    # classifier's cost for that word (two independent kernels, same recursion)
    final, label = engine.classify(X, off, pm, trans_d, W, S)
    own = final.cpu().numpy()[np.arange(B), words]
    np.testing.assert_allclose(tr.best_cost.cpu().numpy(), own, rtol=2e-6)
    assert (label.cpu().numpy() == words).mean() > 0.999


def test_statistics_are_additive_over_shards(corpus):
    """stats(shard 0) + stats(shard 1) == stats(whole corpus): the property the one-allreduce design rests on."""
    from srb200 import engine
    from srb200.trainer import IsolatedTrainer, shard_range
    models, X, off, words = corpus
    trans = np.stack([m[3] for m in models])

    def run(lo, hi):
        pm, _ = engine.PackedModel.from_words(models)
        o = off[lo:hi + 1] - off[lo]
        tr = IsolatedTrainer(X[off[lo]:off[hi]], o, words[lo:hi], pm, trans, S, NCOMP, fused=True)
        tr.fused_pass()
        return tr.stats.cpu().numpy()

    whole = run(0, B)
    parts = sum(run(*shard_range(B, r, 2)) for r in range(2))
    scale = np.abs(whole).max()
    assert np.max(np.abs(parts - whole)) / scale < 1e-6
    np.testing.assert_allclose(parts[-2:], whole[-2:], rtol=1e-9)
