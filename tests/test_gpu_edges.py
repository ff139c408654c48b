"""GPU: edge cases of the C ABI the reference's callers can produce -- empty batches, one-frame utterances, shapes the
kernels are not built for (error codes, never a crash), utterances that cannot reach the end state."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_empty_and_degenerate_batches():
    import torch
    from srb200 import engine, synth
    from srb200.graph import GraphBatch, StateGraph
    models, base = synth.make_models(1, W=2, S=5, M=4)
    pm, trans = engine.PackedModel.from_words(models)
    # N = 0
    E = engine.gmm_score(np.zeros((0, 39), np.float32), pm)["E"]
    assert tuple(E.shape) == (0, 10)
    final, label = engine.classify(np.zeros((0, 39), np.float32), np.zeros(1, np.int32), pm, trans, 2, 5)
    assert final.shape[0] == 0 and label.shape[0] == 0
    # utterances of 1 and 2 frames next to a normal one: no path for the short ones (S = 5 needs >= 3 frames with skips
    # absent), costs +inf, path length 0, labels -1
    X, off, words = synth.make_utterances(5, base, 3, T=30)
    off = np.array([0, 1, 3, 33], np.int32)
    X = X[:33]
    gb = GraphBatch([StateGraph(models[0][3], np.arange(5))], engine.device())
    E = engine.gmm_score(X, pm)["E"]
    fl = torch.full((33,), 7, dtype=torch.int32, device="cuda")
    r = engine.viterbi(E, off, gb, want_path=True, frame_label=fl)
    bc = r.best_cost.cpu().numpy()
    assert np.isinf(bc[0]) and np.isinf(bc[1]) and np.isfinite(bc[2])
    assert list(r.path_len.cpu().numpy()[:2]) == [0, 0] and r.path_len.cpu().numpy()[2] == 29
    assert np.all(fl.cpu().numpy()[:3] == -1) and np.all(fl.cpu().numpy()[3:] >= 0)
    # the same through the fused classifier: +inf costs, label 0 (core.py:82-87 keeps best_model = 0)
    final, label = engine.classify(X, off, pm, trans, 2, 5)
    f = final.cpu().numpy()
    assert np.isinf(f[1]).all() and np.isfinite(f[2]).all() and label.cpu().numpy()[1] == 0


def test_unsupported_shapes_return_error_codes():
    import torch
    from srb200 import engine, lib
    L = lib.load()
    X = torch.zeros((8, 40), dtype=torch.float32, device="cuda")
    ab = torch.zeros(4 * 40 * 2, dtype=torch.float32, device="cuda")
    c = torch.zeros(4, dtype=torch.float32, device="cuda")
    sh = torch.zeros(40, dtype=torch.float32, device="cuda")
    E = torch.zeros((8, 1), dtype=torch.float32, device="cuda")
    rc = L.srb_gmm_score(X.data_ptr(), 8, 40, ab.data_ptr(), c.data_ptr(), sh.data_ptr(), 1, 4, E.data_ptr(), 1, None,
                         None, None, lib.stream_ptr())
    assert rc == -2 and b"dimension" in L.srb_last_error_string()
    rc = L.srb_gmm_score(None, 8, 39, ab.data_ptr(), c.data_ptr(), sh.data_ptr(), 1, 4, E.data_ptr(), 1, None, None,
                         None, lib.stream_ptr())
    assert rc == -1
    with pytest.raises(lib.SrbError):
        lib.check(rc)
    torch.cuda.synchronize()                        # the context is still healthy
    assert float(torch.ones(4, device="cuda").sum()) == 4.0


def test_mirror_short_and_empty_inputs():
    import warnings
    import sr.recognition as rec
    from srb200 import synth
    models, base = synth.make_models(1, W=1, S=5, M=4)
    mu, cov, w, tr = models[0]
    m = rec.HMM(5)
    m.transitions = tr
    m.gmm_states = []
    for s in range(5):
        g = rec.GMM(mu[s, 0], cov[s, 0], 4)
        g.update_models(mu[s], cov[s], w[s])
        m.gmm_states.append(g)
    x = synth.make_utterances(3, base, 1, T=3)[0].astype(np.float64)
    with warnings.catch_warnings(record=True) as wr:
        warnings.simplefilter("always")
        costs, path = rec.decode_hmm_states(x, m.gmm_states, m.transitions)     # 3 frames cannot reach state 4
    assert any("Cannot find a path" in str(v.message) for v in wr)
    assert costs.shape == (5, 3) and np.isinf(costs[-1, -1]) and path.shape == (0, 2)
    assert np.isinf(m.evaluate(x))
    with pytest.raises(AssertionError):
        rec.dtw(x[:1], mu[:, 0], rec.mahalanobis, tr, cov[:, 0])               # decode.py:22
    with pytest.raises(TypeError):
        m["0"]
