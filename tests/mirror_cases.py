"""Parity cases of the sr.recognition mirror against the golden vectors frozen from the UNMODIFIED reference
(tests/golden/make_golden.py).  Each case is a plain function so that the same body runs
  * through libsrb200 on the B200 (tests/test_gpu_mirror.py, -m gpu: the parity tests proper), and
  * through tests/engine_double.py on the CPU (tests/test_mirror_host.py: host logic only).
Tolerance: 1e-4 relative on costs and re-estimated parameters (BASELINE north_star, FP32); paths, segment
boundaries and cluster assignments exact.
"""
import os
import pickle
import tempfile

import numpy as np

RTOL = 1e-4


def unrag(cat, off):
    return [cat[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def close(a, b, rtol=RTOL, atol=0.0):
    np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=rtol, atol=atol)


def em_stops(fn, *a, **k):
    """Runs fn with stdout captured -> (result, the k of every "EM converged at iteration: k" it printed)."""
    import contextlib
    import io
    import re
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        out = fn(*a, **k)
    return out, np.array([int(x) for x in re.findall(r"EM converged at iteration: (\d+)", buf.getvalue())])


def check_stops(stops, ref):
    """The 2e-3 tolerance of converged points must not hide an early stop: every EM loop has to stop within a few
    iterations of the reference's (np.allclose on FP32-accumulated statistics may trip a little earlier or later,
    SURVEY A14)."""
    ref = np.asarray(ref)
    assert len(stops) == len(ref), (stops.tolist(), ref.tolist())
    assert np.all(np.abs(stops - ref) <= np.maximum(3, 0.1 * ref)), (stops.tolist(), ref.tolist())


def load_models(rec, mu, cov, w, trans):
    """Rebuild mirror HMM objects from dumped parameters [W,S,M,D]."""
    models = []
    for wd in range(mu.shape[0]):
        S, M = mu.shape[1], mu.shape[2]
        m = rec.HMM(S)
        m.transitions = trans[wd].copy()
        m.gmm_states = []
        for s in range(S):
            g = rec.GMM(mu[wd, s, 0].copy(), cov[wd, s, 0].copy(), M)
            g.update_models(mu[wd, s].copy(), cov[wd, s].copy(), w[wd, s].copy())
            m.gmm_states.append(g)
        models.append(m)
    return models


# ------------------------------------------------------------------------------------------------ emission layer
def case_emission(rec, G):
    z = G["emission"]
    g = rec.GMM(z["em_mu"][0], z["em_cov"][0], 4)
    g.update_models(z["em_mu"], z["em_cov"], z["em_w"])
    fin = np.isfinite(z["em_cost"])
    close(g.evaluate(z["em_X"])[fin], z["em_cost"][fin])
    close(g.evaluate(z["em_X"][3]), z["em_cost"][3])
    comps = g.evaluate(z["em_X"], return_neg_log_likelihood=False)
    big = z["em_comps"] > 1e-300
    close(np.log(comps[big]), np.log(z["em_comps"][big]), atol=1e-3)
    n = rec.MultivariateNormal(z["em_mu"][0], z["em_cov"][0])
    ok = z["em_pdf0"] > 1e-300
    close(np.log(n.pdf(z["em_X"])[ok]), np.log(z["em_pdf0"][ok]), atol=1e-3)
    assert n.inv_cov.shape == (39, 39)
    try:
        n.pdf(np.zeros(5))
        raise AssertionError("dimension mismatch must raise NameError (hmm_state.py:45)")
    except NameError:
        pass
    for i in range(len(z["mh_out"])):
        close(rec.mahalanobis(z["mh_v1"][i], z["mh_v2"][i], z["mh_var"][i]), z["mh_out"][i])
    assert rec.NES().evaluate(None) == 0 and type(rec.NES()) == rec.NES


# ------------------------------------------------------------------------------------------------ DP layer
def case_isolated_decode(rec, G):
    z = G["decode"]
    models = load_models(rec, z["iso_mu"], z["iso_cov"], z["iso_w"], z["iso_trans"])
    xs = unrag(z["iso_X"], z["iso_off"])
    k = 0
    for u, x in enumerate(xs):
        for w, m in enumerate(models):
            costs, path = rec.decode_hmm_states(x, m.gmm_states, m.transitions)
            ref = z["iso_costs"][z["iso_costs_off"][k]:z["iso_costs_off"][k + 1]].T       # [S, T]
            fin = np.isfinite(ref)
            assert np.array_equal(np.isfinite(costs), fin)
            close(costs[fin], ref[fin])
            assert np.array_equal(path, z["iso_paths"][z["iso_paths_off"][k]:z["iso_paths_off"][k + 1]])
            close(m.evaluate(x), z["iso_final"][u, w])
            k += 1
    costs, labels = rec.classify(models, xs)
    close(costs, z["iso_final"])
    assert np.array_equal(labels, np.argmin(z["iso_final"], axis=1))


def case_graph_decode(rec, G):
    z = G["decode"]
    models = load_models(rec, z["iso_mu"], z["iso_cov"], z["iso_w"], z["iso_trans"])
    W = len(models)
    seq, trans, ends = rec.build_state_sequences(models, [list(range(W))] * 3)
    fin = np.isfinite(z["gr_trans"])
    assert np.array_equal(np.isfinite(trans), fin) and np.allclose(trans[fin], z["gr_trans"][fin])
    assert np.array_equal([type(s) == rec.NES for s in seq], z["gr_is_nes"])
    assert list(ends) == list(z["gr_ends"])
    xs = unrag(z["gr_X"], z["gr_off"])
    for u, x in enumerate(xs):
        costs, path = rec.decode_hmm_states(x, seq, trans, end_points=[[e, -1] for e in ends])
        ref = z["gr_costs"][z["gr_costs_off"][u]:z["gr_costs_off"][u + 1]].T
        f = np.isfinite(ref)
        assert np.array_equal(np.isfinite(costs), f)
        close(costs[f], ref[f])
        assert np.array_equal(path, z["gr_paths"][z["gr_paths_off"][u]:z["gr_paths_off"][u + 1]])
    # main.py:61-67 post-processing: the decoded word strings (the synthetic strings are recognised correctly)
    for i, m in enumerate(models):
        for s in m.gmm_states:
            s.parent = i
    ref_strings = [rec.path_to_models(z["gr_paths"][z["gr_paths_off"][u]:z["gr_paths_off"][u + 1]], seq)
                   for u in range(len(xs))]
    assert rec.recognize_strings(xs, models, n_layers=3) == ref_strings
    assert ref_strings == [list(s) for s in z["gr_strings"]]


def case_dtw_helpers(rec, G):
    z = G["dtw_helpers"]
    xs = unrag(z["dtw_X"], z["dtw_off"])
    for u, x in enumerate(xs):
        for tag, fun, var in (("e", rec.euclidean, None), ("m", rec.mahalanobis, z["dtw_var"])):
            costs, path = rec.dtw(x, z["dtw_y"], fun, z["dtw_trans"], var)
            o = z[f"dtw_{tag}_costs_off"]
            ref = z[f"dtw_{tag}_costs"][o[u]:o[u + 1]].T
            f = np.isfinite(ref)
            assert np.array_equal(np.isfinite(costs), f)
            close(costs[f], ref[f])
            po = z[f"dtw_{tag}_paths_off"]
            assert np.array_equal(path, z[f"dtw_{tag}_paths"][po[u]:po[u + 1]])
    # a user-supplied distance callable goes through the same DP
    costs, _ = rec.dtw(xs[0], z["dtw_y"], lambda a, b: np.linalg.norm(a - b), z["dtw_trans"])
    o = z["dtw_e_costs_off"]
    ref = z["dtw_e_costs"][o[0]:o[1]].T
    close(costs[np.isfinite(ref)], ref[np.isfinite(ref)])
    for a, b in (("tc_seg_lens", "tc_out"), ("tc_seg_lens2", "tc_out2")):
        out = rec.calc_transition_costs(z[a].shape[0], z[a])
        f = np.isfinite(z[b])
        assert np.array_equal(np.isfinite(out), f) and np.allclose(out[f], z[b][f], rtol=1e-12)
    assert np.array_equal(rec.get_segments_from_path(z["gs_path"], 5), z["gs_out"])
    temps = unrag(z["dtw_X"], z["dtw_off"])
    segs = rec.segment_data(temps, 3, 5, z["sd_seg_starts"])
    assert np.array_equal(np.concatenate(segs), z["sd_cat"]) and \
        np.array_equal(np.cumsum([0] + [len(s) for s in segs]), z["sd_off"])
    res, vars_ = rec.combine_templates(temps, 3, 5, z["sd_seg_starts"])
    close(res, z["ct_res"], atol=1e-6)
    close(vars_, z["ct_vars"])
    close(rec.calc_variance(temps[1].T), z["cv_out"])


def case_dtw_beam(rec, G):
    """dtw(..., beam=k) (decode.py:42-48,62-68): cost matrices incl. the -1 marks left in the last column, +inf at the
    cells the beam removed, and the paths the reference walks through its back-pointers."""
    z = G["beam"]
    xs = unrag(z["bm_X"], z["bm_off"])
    for k, (beam, maha, ti, u, plen) in enumerate(z["bm_cases"]):
        trans = z["bm_trans2"] if ti else z["bm_trans"]
        if maha:
            costs, path = rec.dtw(xs[u], z["bm_y"], rec.mahalanobis, trans, z["bm_var"], beam=int(beam))
        else:
            costs, path = rec.dtw(xs[u], z["bm_y"], rec.euclidean, trans, beam=int(beam))
        ref = z[f"bm_costs{k}"]
        f = np.isfinite(ref)
        assert np.array_equal(np.isfinite(costs), f), (k, beam)
        assert np.array_equal(costs == -1, ref == -1), (k, beam)
        close(costs[f], ref[f])
        assert np.array_equal(path, z[f"bm_path{k}"]), (k, beam)
    # no beam, long template (> 255 rows: past the graph kernel's 8-bit back-pointer codes) against a short one
    # embedded in it: rows 0..5 of a 300-row template whose other rows are unreachable behave like the 6-row template
    S, big = z["bm_y"].shape[0], 300
    y = np.concatenate([z["bm_y"], np.zeros((big - S, z["bm_y"].shape[1]))])
    tr = np.full((big, big), np.inf)
    tr[:S, :S] = z["bm_trans"]
    c_small, _ = rec.dtw(xs[1], z["bm_y"], rec.euclidean, z["bm_trans"])
    c_big, _ = rec.dtw(xs[1], y, rec.euclidean, tr)
    close(c_big[:S], c_small)
    assert np.isinf(c_big[S:]).all()
    try:
        rec.dtw(xs[0], z["bm_y"], rec.euclidean, z["bm_trans"], beam=2.5)
        raise AssertionError("a fractional beam must raise TypeError (sort_idx[beam:], decode.py:65)")
    except TypeError:
        pass


def case_end_points(rec, G):
    """decode_hmm_states with end points off the last frame, negative frames, a NES end (decode.py:126-146)."""
    z = G["ends"]
    models = load_models(rec, z["en_mu"], z["en_cov"], z["en_w"], z["en_wtrans"])
    seq, trans, ends = rec.build_state_sequences(models, [[0, 1]] * 2)
    assert list(ends) == list(z["en_ends"])
    x = z["en_X"]
    for k in range(int(z["en_n"])):
        costs, path = rec.decode_hmm_states(x, seq, trans, end_points=z[f"en_list{k}"].tolist())
        ref = z[f"en_costs{k}"]
        f = np.isfinite(ref)
        assert np.array_equal(np.isfinite(costs), f)
        close(costs[f], ref[f])
        assert np.array_equal(path.reshape(-1, 2), z[f"en_path{k}"]), k
    S = len(models[1].gmm_states)
    costs, path = rec.decode_hmm_states(x[:14], models[1].gmm_states, models[1].transitions,
                                        end_points=[[S - 1, 9], [S - 2, -2]])
    f = np.isfinite(z["en_iso_costs"])
    close(costs[f], z["en_iso_costs"][f])
    assert np.array_equal(path, z["en_iso_path"])
    try:
        rec.decode_hmm_states(x, seq, trans, end_points=[[0, len(x)]])
        raise AssertionError("an end frame past the utterance must raise IndexError")
    except IndexError:
        pass


# ------------------------------------------------------------------------------------------------ K-means / EM
def case_kmeans_em(rec, G):
    z = G["kmeans_em"]
    data = z["km_data"]
    np.random.seed(11)                                   # make_golden.py: the stream kmeans.py:171 consumes
    for k in (2, 4):
        clusters, cent, cov = rec.kmeans(data, k, z[f"km{k}_cent_in"], dist_fun=rec.mahalanobis)
        assert np.array_equal(clusters, z[f"km{k}_clusters"])
        close(cent, z[f"km{k}_cent"], atol=1e-6)
        close(cov, z[f"km{k}_cov"])
        close(rec.cluster_centroids(data, clusters, k), z[f"km{k}_cent"], atol=1e-6)
        for iters, tag in ((1, "it1"), (3, "it3"), (10000, "conv")):
            g = rec.GMM(z[f"em{k}_mu_in"][0].copy(), z[f"em{k}_cov_in"][0].copy(), 4)
            g.update_models(z[f"em{k}_mu_in"].copy(), z[f"em{k}_cov_in"].copy(), z[f"em{k}_w_in"].copy())
            _, stops = em_stops(g.em, data, k, iters)
            if iters == 10000:
                check_stops(stops, G["train_multi"][f"em{k}_conv_stop"])
            tol = RTOL if iters < 10000 else 2e-3        # the converged point depends on the stopping iteration
            close(g.w, z[f"em{k}_{tag}_w"], rtol=tol)
            close(np.array([d.mean for d in g.dists]), z[f"em{k}_{tag}_mu"], rtol=tol, atol=2e-5 if iters < 10000 else 1e-3)
            close(np.array([d.cov for d in g.dists]), z[f"em{k}_{tag}_cov"], rtol=tol)


# ------------------------------------------------------------------------------------------------ training
def _dump(models):
    mu = np.array([[[d.mean for d in s.dists] for s in m.gmm_states] for m in models])
    cov = np.array([[[d.cov for d in s.dists] for s in m.gmm_states] for m in models])
    w = np.array([[s.w for s in m.gmm_states] for m in models])
    tr = np.array([m.transitions for m in models])
    return mu, cov, w, tr


def _close_trans(a, b):
    f = np.isfinite(b)
    assert np.array_equal(np.isfinite(a), f)
    close(a[f], b[f], atol=1e-12)


def case_skmeans(rec, G):
    z = G["train"]
    utts = unrag(z["tr_X0"], z["tr_off0"])
    res, vars_, tc, segs = rec.skmeans(utts, 5, return_segmented_data=True)
    close(res, z["sk_res"], atol=1e-6)
    close(vars_, z["sk_vars"])
    _close_trans(tc, z["sk_trans"])
    assert np.array_equal(np.cumsum([0] + [len(s) for s in segs]), z["sk_seg_off"])
    assert np.array_equal(np.concatenate(segs), z["sk_seg_cat"])
    m1 = rec.HMM(5).fit(utts, n_gaussians=4, use_gmm=False)
    others = unrag(z["tr_X1"], z["tr_off1"])
    ev = [m1.evaluate(x) for x in utts[:3]] + [m1.evaluate(x) for x in others[:2]]
    close(ev, z["fit1_eval"])
    try:
        rec.skmeans([utts[0], utts[1][:4]], 5)
        raise AssertionError("a template with < 5 frames must raise NameError (kmeans.py:142-143)")
    except NameError:
        pass


def case_fit(rec, G):
    z = G["train"]
    W = 3
    utts = [unrag(z[f"tr_X{w}"], z[f"tr_off{w}"]) for w in range(W)]
    models = []
    for w in range(W):
        np.random.seed(1000 + w)
        m, stops = em_stops(rec.HMM(5).fit, utts[w], n_gaussians=4, use_gmm=True, use_em=True)
        check_stops(stops, G["train_multi"][f"fit_em_stop{w}"])
        models.append(m)
        assert np.array_equal(np.cumsum([0] + [len(s) for s in m.segments]), z[f"fit_seg_off{w}"])
        assert np.array_equal(np.concatenate(m.segments), z[f"fit_seg_cat{w}"])
    mu, cov, wt, tr = _dump(models)
    close(wt, z["fit_w"], rtol=2e-3)
    close(mu, z["fit_mu"], rtol=2e-3, atol=1e-3)
    close(cov, z["fit_cov"], rtol=2e-3)
    _close_trans(tr, z["fit_trans"])
    close(np.array([m.mu for m in models]), z["fit_hmm_mu"], atol=1e-6)
    close(np.array([m.sigma for m in models]), z["fit_hmm_sigma"])
    ev = np.array([[models[w].evaluate(utts[v][i]) for w in range(W)] for v in range(W) for i in range(2)])
    close(ev, z["fit_eval"], rtol=1e-3)
    assert models[0][1] is models[0].gmm_states[1] and models[0] == pickle.loads(pickle.dumps(models[0]))
    np.random.seed(2000)
    mk = rec.HMM(5).fit(utts[0], n_gaussians=4, use_gmm=True, use_em=False)
    mu1, cov1, w1, _ = _dump([mk])
    close(mu1, z["fitk_mu"], atol=1e-6)
    close(cov1, z["fitk_cov"])
    close(w1, z["fitk_w"])


def case_fit_words(rec, G, fused_align=True):
    """fit_words: the three golden words trained TOGETHER (one corpus, every phase batched over all states) must give
    what the reference's per-word HMM.fit gave -- same pooled segments, parameters, transitions and EM stop iterations."""
    z, zm = G["train"], G["train_multi"]
    W = 3
    utts = [unrag(z[f"tr_X{w}"], z[f"tr_off{w}"]) for w in range(W)]
    models = rec.fit_words(utts, 5, 4, use_em=True, seeds=[1000 + w for w in range(W)], fused_align=fused_align)
    ft = rec.fit_words.last
    for w, m in enumerate(models):
        assert np.array_equal(np.cumsum([0] + [len(s) for s in m.segments]), z[f"fit_seg_off{w}"])
        assert np.array_equal(np.concatenate(m.segments), z[f"fit_seg_cat{w}"])
        check_stops(ft.iterations["em"][0][w * 5:(w + 1) * 5], zm[f"fit_em_stop{w}"])
    mu, cov, wt, tr = _dump(models)
    close(wt, z["fit_w"], rtol=2e-3)
    close(mu, z["fit_mu"], rtol=2e-3, atol=1e-3)
    close(cov, z["fit_cov"], rtol=2e-3)
    _close_trans(tr, z["fit_trans"])
    close(np.array([m.mu for m in models]), z["fit_hmm_mu"], atol=1e-6)
    close(np.array([m.sigma for m in models]), z["fit_hmm_sigma"])
    ev = np.array([[models[w].evaluate(utts[v][i]) for w in range(W)] for v in range(W) for i in range(2)])
    close(ev, z["fit_eval"], rtol=1e-3)
    # k-means only (use_em=False): word 0 alone, the golden fitk_* run
    mk = rec.fit_words([utts[0]], 5, 4, use_em=False, seeds=[2000], fused_align=fused_align)
    mu1, cov1, w1, _ = _dump(mk)
    close(mu1, z["fitk_mu"], atol=1e-6)
    close(cov1, z["fitk_cov"])
    close(w1, z["fitk_w"])


def case_continuous_train(rec, G):
    z = G["train"]
    models = load_models(rec, z["fit_mu"], z["fit_cov"], z["fit_w"], z["fit_trans"])
    for i, m in enumerate(models):
        for s in m.gmm_states:
            s.parent = i
    data = unrag(z["ct_X"], z["ct_off"])
    strings = [list(map(int, s)) for s in z["ct_strings"]]
    with tempfile.TemporaryDirectory() as td:
        np.random.seed(3000)
        rec.continuous_train(data, models, strings, td, 4, 5, 1)
        new = [pickle.load(open(os.path.join(td, f"{i}.pkl"), "rb")) for i in range(len(models))]
    mu, cov, wt, tr = _dump(new)
    close(wt, z["ctr_w"], rtol=2e-3)
    close(mu, z["ctr_mu"], rtol=2e-3, atol=1e-3)
    close(cov, z["ctr_cov"], rtol=2e-3)
    _close_trans(tr, z["ctr_trans"])
    assert new[0].gmm_states[0].parent == 0


def case_continuous_train_multi(rec, G):
    """Three outer iterations (continuous_speech.py:73-179): from the second one on the reference looks the tied
    states up through `modelidx_state_map`, which still holds the FIRST copy's objects, in a dict keyed by the
    current copy's objects (hash by id, equality by value) -- so a state's arcs are re-estimated only while its
    mixture still equals the first iteration's, and every other state raises "No MFCC data for state".  The models,
    the number of those warnings, the outer iteration count and the EM stop iterations must follow the reference."""
    import contextlib
    import io
    import re
    import warnings
    z, zm = G["train"], G["train_multi"]
    data = unrag(z["ct_X"], z["ct_off"])
    strings = [list(map(int, s)) for s in z["ct_strings"]]
    for n_iter in (2, 3):
        models = load_models(rec, z["fit_mu"], z["fit_cov"], z["fit_w"], z["fit_trans"])
        for i, m in enumerate(models):
            for s in m.gmm_states:
                s.parent = i
        buf = io.StringIO()
        with tempfile.TemporaryDirectory() as td, warnings.catch_warnings(record=True) as wrn, \
                contextlib.redirect_stdout(buf):
            warnings.simplefilter("always")
            np.random.seed(3000)
            rec.continuous_train(data, models, strings, td, 4, 5, n_iter)
            new = [pickle.load(open(os.path.join(td, f"{i}.pkl"), "rb")) for i in range(len(models))]
        text = buf.getvalue()
        assert len(re.findall(r"Continuous training iteration:", text)) == int(zm[f"it{n_iter}_outer"])
        assert ("Continuous training converged" in text) == bool(zm[f"it{n_iter}_converged"])
        # Which states still "equal" their first-iteration mixture is an np.allclose(rtol 1e-5) decision on converged
        # EM parameters; the reference's own margin is read off the fixture.  A state whose largest relative change is
        # within a factor 5 of that threshold (state (0, 3) of iteration 3 sits at 1.39e-5) is decided by FP32 noise:
        # either outcome is accepted for it, every other state must follow the reference.
        lo, hi, undecided = 0, 0, set()
        for it in range(2, n_iter + 1):
            for wd in range(len(models)):
                for st in range(5):
                    ref1 = [zm["it1_w"][wd, st], zm["it1_mu"][wd, st], zm["it1_cov"][wd, st]]
                    cur = [zm[f"it{it}_w"][wd, st], zm[f"it{it}_mu"][wd, st], zm[f"it{it}_cov"][wd, st]]
                    stat = max(float(np.max(np.abs(c - r) / (1e-8 + 1e-5 * np.abs(r)))) for c, r in zip(cur, ref1))
                    if stat > 5.0:
                        lo, hi = lo + 1, hi + 1                      # decided: no match, one warning
                    elif stat > 0.2:
                        hi += 1                                      # inside the margin
                        if it == n_iter:
                            undecided.add((wd, st))
        n_warn = sum("No MFCC data" in str(x.message) for x in wrn)
        assert lo <= int(zm[f"it{n_iter}_no_data_warnings"]) <= hi
        assert lo <= n_warn <= hi, (n_warn, lo, hi)
        mu, cov, wt, tr = _dump(new)
        close(wt, zm[f"it{n_iter}_w"], rtol=2e-3)
        close(mu, zm[f"it{n_iter}_mu"], rtol=2e-3, atol=1e-3)
        close(cov, zm[f"it{n_iter}_cov"], rtol=2e-3)
        ref_tr = zm[f"it{n_iter}_trans"].copy()
        for wd, st in undecided:                                     # arcs of an undecided state: either version
            for r_, o_ in ((st, st), (min(st + 1, 4), st)):
                tr[wd][r_, o_] = ref_tr[wd][r_, o_]
        _close_trans(tr, ref_tr)
        # the 2e-3 tolerance of the converged points must not hide an early stop: the EM loops have to stop within a
        # few iterations of where the reference's did (np.allclose on float32-accumulated statistics)
        stops = np.array([int(k) for k in re.findall(r"EM converged at iteration: (\d+)", text)])
        ref = zm[f"it{n_iter}_em_stop"]
        assert len(stops) == len(ref), (len(stops), len(ref))
        assert np.all(np.abs(stops - ref) <= np.maximum(3, 0.1 * ref)), (stops.tolist(), ref.tolist())


ALL = [case_emission, case_isolated_decode, case_graph_decode, case_dtw_helpers, case_dtw_beam, case_end_points, case_kmeans_em, case_skmeans,
       case_fit, case_fit_words, case_continuous_train, case_continuous_train_multi]
