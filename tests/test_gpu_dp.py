"""GPU parity: K2 graph Viterbi (decode_hmm_states / dtw) and K4 forward-backward against the oracle and the
frozen reference outputs, through the C ABI."""
import numpy as np
import pytest

from conftest import unragged
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
RTOL = 1e-4


@pytest.fixture(scope="module")
def eng():
    from srb200 import engine
    return engine


def models_from(mu, cov, w, tr):
    return [(mu[i], cov[i], w[i], tr[i]) for i in range(mu.shape[0])]


def dev_E(eng, X, models):
    pm, _ = eng.PackedModel.from_words(models)
    return eng.gmm_score(X.astype(np.float32), pm)["E"]


def check_costs(costs_flat, ws, off, R, ref_list):
    c = costs_flat.cpu().numpy()
    for u, ref in enumerate(ref_list):                     # ref: [T, R]
        got = c[ws[u]:ws[u + 1]].reshape(-1, R)
        f = np.isfinite(ref)
        assert np.array_equal(np.isfinite(got), f)
        np.testing.assert_allclose(got[f], ref[f], rtol=RTOL)


def test_isolated_decode_golden(eng, golden):
    """decode_hmm_states on 5 utterances x 3 isolated models: costs and paths frozen from the reference."""
    from srb200.graph import GraphBatch, StateGraph
    g = golden["decode"]
    models = models_from(g["iso_mu"], g["iso_cov"], g["iso_w"], g["iso_trans"])
    E = dev_E(eng, g["iso_X"], models)                     # [N, 15]
    rc = unragged(g["iso_costs"], g["iso_costs_off"])
    rp = unragged(g["iso_paths"], g["iso_paths_off"])
    B = len(g["iso_off"]) - 1
    for w, m in enumerate(models):
        gb = GraphBatch([StateGraph(m[3], np.arange(5) + 5 * w)], E.device)
        res = eng.viterbi(E, g["iso_off"].astype(np.int32), gb, want_costs=True)
        check_costs(res.costs, res.ws_off, g["iso_off"], 5, [rc[u * 3 + w] for u in range(B)])
        for u, p in enumerate(res.paths()):
            assert np.array_equal(p, rp[u * 3 + w])
        np.testing.assert_allclose(res.best_cost.cpu().numpy(), g["iso_final"][:, w], rtol=RTOL)


def test_graph_decode_golden(eng, golden):
    """Word-loop graph with NES (build_state_sequences) and per-utterance forced-alignment chains."""
    from srb200.graph import GraphBatch, StateGraph
    g = golden["decode"]
    models = models_from(g["iso_mu"], g["iso_cov"], g["iso_w"], g["iso_trans"])
    E = dev_E(eng, g["gr_X"], models)
    off = g["gr_off"].astype(np.int32)
    gb = GraphBatch([StateGraph(g["gr_trans"], g["gr_emit"], g["gr_ends"])], E.device)
    res = eng.viterbi(E, off, gb, want_costs=True)
    check_costs(res.costs, res.ws_off, off, 49, unragged(g["gr_costs"], g["gr_costs_off"]))
    rp = unragged(g["gr_paths"], g["gr_paths_off"])
    for u, p in enumerate(res.paths()):
        assert np.array_equal(p, rp[u])
        assert orc.path_to_words(p, g["gr_emit"], 5) == list(g["gr_strings"][u])
    # forced alignment: one chain graph per utterance (continuous_speech.py:80,89)
    graphs = [StateGraph(g["fa_trans"][i], g["fa_emit"][i], g["fa_ends"][i]) for i in range(2)]
    gb = GraphBatch(graphs, E.device)
    res = eng.viterbi(E, off, gb, utt_graph=[0, 1], want_costs=True)
    check_costs(res.costs, res.ws_off, off, 19, unragged(g["fa_costs"], g["fa_costs_off"]))
    fp = unragged(g["fa_paths"], g["fa_paths_off"])
    for u, p in enumerate(res.paths()):
        assert np.array_equal(p, fp[u])


def test_digit_loop_graph_vs_oracle(eng):
    """Config-4 shape: 7 layers x 11 words x 5 states = 393 states, ragged utterances."""
    from srb200 import synth
    from srb200.graph import GraphBatch, StateGraph
    models, base = synth.make_models(31, W=11, S=5, M=4)
    rng = np.random.default_rng(7)
    utts = []
    for _ in range(4):
        digits = rng.integers(0, 11, size=7)
        parts = [synth.make_utterances(int(rng.integers(1 << 30)), base[d:d + 1], 1, T=int(rng.integers(12, 25)))[0]
                 for d in digits]
        utts.append(np.concatenate(parts))
    X = np.concatenate(utts)
    off = np.concatenate([[0], np.cumsum([len(u) for u in utts])]).astype(np.int32)
    emit, trans, ends = orc.build_state_sequences(5, [m[3] for m in models], [list(range(11))] * 7)
    assert len(emit) == 393
    E = dev_E(eng, X, models)
    gb = GraphBatch([StateGraph(trans, emit, ends)], E.device)
    res = eng.viterbi(E, off, gb)
    paths = res.paths()
    bc = res.best_cost.cpu().numpy()
    for u, x in enumerate(utts):
        Eo = orc.emissions(x, models, emit, log_domain=True)
        costs, path, best = orc.decode_hmm_states(Eo, trans, emit < 0, [[e, -1] for e in ends])
        ref = costs[best[0], best[1]]
        np.testing.assert_allclose(bc[u], ref, rtol=RTOL)
        if not np.array_equal(paths[u], path):             # only inside the decision margin
            assert abs(bc[u] - ref) <= 1e-5 * abs(ref)
        assert orc.path_to_words(paths[u], emit, 5) == orc.path_to_words(path, emit, 5)


def test_dtw_golden(eng, golden):
    """dtw (decode.py:7-77) == the same kernel on local distances with inf transitions kept as candidates."""
    import torch
    from srb200.graph import GraphBatch, StateGraph
    g = golden["dtw_helpers"]
    xs = unragged(g["dtw_X"], g["dtw_off"])
    off = g["dtw_off"].astype(np.int32)
    for nm, var in (("dtw_e", None), ("dtw_m", g["dtw_var"])):
        Dl = np.concatenate([orc.local_dist(x, g["dtw_y"], var) for x in xs]).astype(np.float32)
        E = torch.from_numpy(Dl).cuda()
        gb = GraphBatch([StateGraph(g["dtw_trans"], np.arange(5), keep_inf=True)], E.device)
        res = eng.viterbi(E, off, gb, want_costs=True)
        check_costs(res.costs, res.ws_off, off, 5, unragged(g[nm + "_costs"], g[nm + "_costs_off"]))
        rp = unragged(g[nm + "_paths"], g[nm + "_paths_off"])
        for u, p in enumerate(res.paths()):
            assert np.array_equal(p, rp[u])


@pytest.mark.parametrize("banded", [False, True])
@pytest.mark.parametrize("base_scale", [1.0, 0.12])
@pytest.mark.parametrize("S,M", [(5, 4), (8, 16), (3, 2), (12, 2), (20, 2)])
def test_forward_backward_vs_oracle(eng, S, M, banded, base_scale):
    """Both forward-backward kernels (generic CSR: banded=False; one lane per state, 32 / pow2(S) utterances per warp:
    banded=True) against the float64 oracle (itself pinned by path enumeration and by the exact (min,+) identity with
    decode_hmm_states, tests/test_oracle_golden.py): -log Z, state posteriors, summed edge posteriors at 1e-4 relative
    (BASELINE north_star's FP32 rule).  base_scale 0.12 moves the state means together so that ~10 % of the posteriors
    are genuinely soft; the oracle runs on the SAME float32 emission costs the kernels read (K1's own 1e-4 parity is
    tests/test_gpu_score.py's business), so what is measured here is the recursion.
    Measured on the B200 (tools/measure_fb_error.py, profiles/r02_fb_error.json): gamma <= 2e-6 relative above the
    1e-4 floor, edge sums <= 1e-6, -log Z <= 3e-8."""
    from srb200 import synth
    from srb200.graph import GraphBatch, StateGraph
    models, base = synth.make_models(41, W=3, S=S, M=M, base_scale=base_scale, spread=0.5 * base_scale)
    X, off, words = synth.make_utterances(42, base, 12, ragged=(S + 3, 90), noise=1.0 if base_scale < 1 else 0.7)
    E = dev_E(eng, X, models)
    Eh = E.double().cpu().numpy()
    graphs = [StateGraph(m[3], np.arange(S) + S * w) for w, m in enumerate(models)]
    gb = GraphBatch(graphs, E.device)
    assert gb.banded
    gamma, ws, nlz, xi = eng.forward_backward(E, off, gb, utt_graph=words.astype(np.int32), want_xi=True, banded=banded)
    xi_again = eng.forward_backward(E, off, gb, utt_graph=words.astype(np.int32), want_xi=True, banded=banded)[3]
    assert np.array_equal(xi.cpu().numpy(), xi_again.cpu().numpy())       # no atomics: bit-reproducible run to run
    gamma, nlz, xi = gamma.cpu().numpy(), nlz.cpu().numpy(), xi.cpu().numpy()
    xi_ref = np.zeros_like(xi)
    for u in range(len(off) - 1):
        m = models[words[u]]
        Eo = Eh[off[u]:off[u + 1], S * words[u]:S * words[u] + S]
        z, a, b, go = orc.forward_backward(Eo, m[3])
        np.testing.assert_allclose(nlz[u], z, rtol=1e-6)
        got = gamma[ws[u]:ws[u + 1]].reshape(-1, S)
        np.testing.assert_allclose(got, go, rtol=RTOL, atol=1e-4 * RTOL)   # 1e-4 relative above a floor of 1e-4
        np.testing.assert_allclose(got.sum(axis=1), 1.0, rtol=1e-5)
        sg = graphs[words[u]]
        e0 = gb.edge_off[words[u]]
        for j in range(sg.nnz):
            o, r = sg.pred_idx[j], sg.pred_tgt[j]
            v = a[:-1, o] + m[3][r, o] + Eo[1:, r] + b[1:, r] - z
            xi_ref[e0 + j] += np.exp(-v).sum()
    np.testing.assert_allclose(xi, xi_ref, rtol=RTOL, atol=1e-3 * RTOL)


@pytest.mark.parametrize("banded", [False, True])
def test_forward_backward_vs_path_enumeration(eng, banded):
    """The kernels against EXPLICIT enumeration of every state sequence (oracle.forward_backward_bruteforce: no
    recursion shared with anything): T <= 9 frames, S <= 4 states, left-to-right with skip arcs."""
    import torch
    from srb200.graph import GraphBatch, StateGraph
    rng = np.random.default_rng(78)
    for T, R in ((2, 2), (3, 3), (5, 2), (6, 3), (7, 4), (8, 4), (9, 3)):
        E = rng.uniform(0.5, 6.0, size=(T, R)).astype(np.float32)
        tr = np.full((R, R), np.inf)
        for i in range(R):
            tr[i, i] = rng.uniform(0.05, 2.0)
            if i:
                tr[i, i - 1] = rng.uniform(0.05, 2.5)
            if i > 1:
                tr[i, i - 2] = rng.uniform(0.5, 3.0)
        tr = np.where(np.isinf(tr), np.inf, tr.astype(np.float32).astype(np.float64))
        sg = StateGraph(tr, np.arange(R))
        gb = GraphBatch([sg], torch.device("cuda"))
        gamma, ws, nlz, xi = eng.forward_backward(torch.from_numpy(E).cuda(), np.array([0, T], np.int32), gb,
                                                  want_xi=True, banded=banded)
        z, g, x = orc.forward_backward_bruteforce(E.astype(np.float64), tr)
        np.testing.assert_allclose(nlz.cpu().numpy()[0], z, rtol=1e-6)
        np.testing.assert_allclose(gamma.cpu().numpy().reshape(T, R), g, rtol=RTOL, atol=1e-8)
        got = np.zeros((R, R))
        got[sg.pred_tgt, sg.pred_idx] = xi.cpu().numpy()[:sg.nnz]
        np.testing.assert_allclose(got, x, rtol=RTOL, atol=1e-8)


@pytest.mark.parametrize("layers", [1, 3])
def test_banded_kernel_equals_generic_kernel(layers):
    """srb_viterbi_banded (three data-parallel sweeps per column) against srb_viterbi_graph (one sequential run per
    NES) on word-loop graphs with skip arcs: cost matrices, back-pointers, best ends and paths must be identical."""
    import torch
    from srb200 import engine, synth
    from srb200.graph import GraphBatch, StateGraph, word_loop_graph
    W, S = 4, 5
    models, base = synth.make_models(33, W=W, S=S, M=4)
    trans = np.stack([m[3] for m in models]).copy()
    rng = np.random.default_rng(5)
    for w in range(W):                                   # skip arcs and unequal costs (calc_transition_costs style)
        trans[w][2, 0] = 1.0 + rng.random()
        trans[w][S - 1, S - 3] = 0.5 + rng.random()
        trans[w][1, 1] = 0.2 + rng.random()
    g, _ = word_loop_graph(trans, layers)
    iso = StateGraph(trans[1], np.arange(S))             # a NES-free left-to-right graph in the same batch
    iso.emit = (np.arange(S) + S).astype(np.int32)
    gb = GraphBatch([g, iso], engine.device())
    assert gb.banded
    X, off, strings = synth.make_strings(34, base, 12, n_words=layers, T=40 * layers, noise=1.0)
    pm, _ = engine.PackedModel.from_words(models)
    E = engine.gmm_score(X, pm)["E"]
    ug = (np.arange(12) % 3 == 2).astype(np.int32)       # every third utterance on the isolated graph
    out = []
    for banded in (True, False):
        fl = torch.full((E.shape[0],), -9, dtype=torch.int32, device="cuda")
        r = engine.viterbi(E, off, gb, utt_graph=ug, want_costs=True, want_path=True, frame_label=fl, banded=banded)
        out.append((r.costs.cpu().numpy(), r.bp.cpu().numpy(), r.best_cost.cpu().numpy(), r.best_end.cpu().numpy(),
                    r.paths(), fl.cpu().numpy()))
    a, b = out
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3]) and np.array_equal(a[5], b[5])
    assert all(np.array_equal(x, y) for x, y in zip(a[4], b[4]))
    assert np.isfinite(a[2]).all()


@pytest.mark.parametrize("W,layers,why", [(4, 3, "register descriptors, staged emission rows"),
                                          (2, 40, "41 NES: more than one pass of the NES sweep"),
                                          (11, 10, "110 entry states: descriptors stay in shared memory"),
                                          (30, 2, "150 emission columns: rows gathered from global")])
def test_one_graph_banded_kernel_equals_generic_kernel(W, layers, why):
    """viterbi_banded1_kernel (one graph for the whole batch: tables in shared memory, lane-group NES reduction,
    windowed warp traceback) against srb_viterbi_graph: cost matrices, back-pointers, ends, paths and frame labels must
    be identical, including 1-frame utterances, utterances too short for any path, and an empty one."""
    import torch
    from srb200 import engine, synth
    from srb200.graph import GraphBatch, word_loop_graph
    S = 5
    models, base = synth.make_models(41, W=W, S=S, M=4)
    trans = np.stack([m[3] for m in models]).copy()
    rng = np.random.default_rng(6)
    for w in range(W):
        trans[w][2, 0] = 1.0 + rng.random()
        trans[w][S - 1, S - 3] = 0.5 + rng.random()
    g, _ = word_loop_graph(trans, layers)
    gb = GraphBatch([g], engine.device())
    assert gb.banded
    B = 40
    X, off, _ = synth.make_strings(42, base, B, n_words=layers, T=12 * layers + 30, noise=1.0)
    lens = np.diff(off).copy()
    lens[3], lens[7], lens[11] = 1, 2, 0                 # too short for a path / empty
    lens[5:] = np.maximum(1, lens[5:] - rng.integers(0, 25, size=B - 5)) * (lens[5:] > 0)
    lens[11] = 0
    off2 = np.zeros(B + 1, np.int32)
    np.cumsum(lens, out=off2[1:])
    X = np.ascontiguousarray(X[:off2[-1]])
    pm, _ = engine.PackedModel.from_words(models)
    E = engine.gmm_score(X, pm)["E"]
    out = []
    for banded in (True, False):
        fl = torch.full((E.shape[0],), -9, dtype=torch.int32, device="cuda")
        r = engine.viterbi(E, off2, gb, want_costs=True, want_path=True, frame_label=fl, banded=banded)
        out.append((r.costs.cpu().numpy(), r.bp.cpu().numpy(), r.best_cost.cpu().numpy(), r.best_end.cpu().numpy(),
                    r.paths(), fl.cpu().numpy(), r.path_len.cpu().numpy()))
    a, b = out
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3]) and np.array_equal(a[5], b[5])
    assert np.array_equal(a[6], b[6])
    assert all(np.array_equal(x, y) for x, y in zip(a[4], b[4]))
    assert np.isfinite(a[2]).sum() >= B - 8 and not np.isfinite(a[2][[3, 11]]).any()


def test_path_words_kernel_matches_main_py_postprocessing():
    """srb_path_words against the host form of main.py:59-67 (reverse, drop consecutive duplicates, split at the NES,
    first state of every piece -> word) on random paths: runs longer than a warp, NES runs, paths that start or end in
    a NES, empty paths, and the count of a string longer than max_words."""
    import ctypes
    import torch
    from srb200 import lib as _l
    rng = np.random.default_rng(17)
    R, B, max_words = 23, 300, 6
    state_word = np.where(rng.random(R) < 0.25, -1, rng.integers(0, 11, size=R)).astype(np.int32)
    state_word[0] = -1
    paths = []
    for u in range(B):
        n_runs = int(rng.integers(0, 9)) if u % 17 else 0
        fwd = []
        for _ in range(n_runs):
            fwd += [int(rng.integers(0, R))] * int(rng.integers(1, 70))
        if u == 1:                                                           # 10 words: more than max_words
            emitting = int(np.nonzero(state_word >= 0)[0][0])
            fwd = ([0] * 3 + [emitting] * 40) * 10
        paths.append(np.asarray(fwd[::-1], np.int32))                       # stored reversed, like decode.py:142-146
    ln = np.array([len(p) for p in paths], np.int32)
    off = np.zeros(B + 1, np.int64)
    np.cumsum(ln + 3, out=off[1:])                                          # slack between paths, like the decoder's layout
    flat = np.full(int(off[-1]), -7, np.int32)
    for u, p in enumerate(paths):
        flat[off[u]:off[u] + len(p)] = p
    want = []
    for p in paths:
        rows = p[::-1]
        if rows.size:
            rows = rows[np.concatenate([[True], rows[1:] != rows[:-1]])]
        ws, found = [], False
        for r in rows:                                                       # split_result of main.py:38-52
            if state_word[r] >= 0:
                if not found:
                    ws.append(int(state_word[r]))
                    found = True
            else:
                found = False
        want.append(ws)
    dev = "cuda"
    t = lambda a: torch.from_numpy(a).to(dev)
    words = torch.full((B, max_words), -5, dtype=torch.int32, device=dev)
    nw = torch.empty(B, dtype=torch.int32, device=dev)
    gso, d_flat, d_off, d_ln, d_sw = t(np.array([0, R], np.int32)), t(flat), t(off), t(ln), t(state_word)   # kept alive
    _l.check(_l.load().srb_path_words(_l.ptr(d_flat), _l.ptr(d_off), _l.ptr(d_ln), B, None, _l.ptr(gso),
                                      _l.ptr(d_sw), max_words, _l.ptr(words), _l.ptr(nw), _l.stream_ptr()))
    wn, nn = words.cpu().numpy(), nw.cpu().numpy()
    assert any(len(w) > max_words for w in want) and any(len(w) == 0 for w in want)
    for u in range(B):
        assert nn[u] == len(want[u])
        k = min(len(want[u]), max_words)
        assert wn[u, :k].tolist() == want[u][:k]
        assert (wn[u, k:] == -5).all()


def test_dtw_beam_batch_against_oracle(eng):
    """srb_dtw_beam on a ragged batch against the literal restatement of decode.py:19-77 (oracle.dtw_beam): random
    templates of 2..40 rows, dense transitions with +inf entries, beams from 1 to past the row count and none."""
    import torch
    rng = np.random.default_rng(5)
    for S, beam in ((2, 1), (5, 2), (7, 3), (16, 4), (40, 6), (40, None), (9, 100), (300, 8)):
        T = rng.integers(2, 40, size=12)
        off = np.zeros(len(T) + 1, np.int32)
        off[1:] = np.cumsum(T)
        Dl = rng.uniform(0.5, 3.0, size=(off[-1], S)).astype(np.float32)
        tr = np.full((S, S), np.inf)
        for i in range(S):
            tr[i, i] = rng.uniform(0.1, 1.0)
            if i:
                tr[i, i - 1] = rng.uniform(0.1, 2.0)
            if i > 1 and rng.random() < 0.4:
                tr[i, i - 2] = rng.uniform(1.0, 3.0)
        tr = tr.astype(np.float32).astype(np.float64)
        costs, ws, paths, status = eng.dtw_beam(torch.from_numpy(Dl).cuda(), off, tr, beam)
        c = costs.cpu().numpy()
        for u in range(len(T)):
            try:
                rc, rp = orc.dtw_beam(Dl[off[u]:off[u + 1]].astype(np.float64), tr, np.inf if beam is None else beam)
            except RuntimeError:
                assert status[u] == -3 and paths[u] is None     # the reference's walk leaves column 0 off row 0
                continue
            got = c[ws[u]:ws[u + 1]].reshape(-1, S).T
            f = np.isfinite(rc)
            assert np.array_equal(np.isfinite(got), f), (S, beam, u)
            np.testing.assert_allclose(got[f], rc[f], rtol=RTOL)
            assert np.array_equal(paths[u], rp), (S, beam, u)    # random uniform costs: no near-ties at these sizes


def test_end_frames_through_the_abi(eng, golden):
    """srb_viterbi_graph's end_frame: best end point over any [state, frame] list, path from that frame."""
    from srb200.graph import GraphBatch, StateGraph
    import torch
    g = golden["ends"]
    models = models_from(g["en_mu"], g["en_cov"], g["en_w"], g["en_wtrans"])
    E = dev_E(eng, g["en_X"], models)
    T = len(g["en_X"])
    off = np.array([0, T], np.int32)
    for k in range(int(g["en_n"])):
        el = g[f"en_list{k}"]
        gb = GraphBatch([StateGraph(g["en_trans"], g["en_emit"], end_states=el[:, 0], end_frames=el[:, 1])], E.device)
        assert "end_frame" in gb.dev
        fl = torch.full((T,), -7, dtype=torch.int32, device=E.device)
        res = eng.viterbi(E, off, gb, want_costs=True, frame_label=fl)
        ref = g[f"en_costs{k}"]
        got = res.costs.cpu().numpy().reshape(T, -1).T
        f = np.isfinite(ref)
        np.testing.assert_allclose(got[f], ref[f], rtol=RTOL)
        rp = g[f"en_path{k}"]
        if np.isfinite(res.best_cost.cpu().numpy()[0]):
            assert np.array_equal(res.paths()[0], rp)
            lab = fl.cpu().numpy()
            assert (lab[np.asarray(rp)[:, 1].max() + 2:] == -1).all()
        else:
            assert res.path_len.cpu().numpy()[0] == 0
    with pytest.raises(Exception):
        eng.viterbi(E, off, gb, want_costs=False)          # end frames need the cost matrix
