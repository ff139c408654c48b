"""Pin the CPU oracle (oracle/) against outputs of the UNMODIFIED reference (tests/golden/ref_*.npz).

CPU-only.  Tolerances are float64 round-off (the oracle restates the same IEEE arithmetic; only BLAS
summation order differs), integer outputs (paths, clusters, segment boundaries) are exact."""
import numpy as np
import pytest

from conftest import unragged
from oracle import oracle as orc

RT = 1e-10


def models_from(mu, cov, w, tr):
    return [(mu[i], cov[i], w[i], tr[i]) for i in range(mu.shape[0])]


def test_emission_linear_and_log(golden):
    g = golden["emission"]
    cost, comps = orc.gmm_evaluate(g["em_X"], g["em_mu"], g["em_cov"], g["em_w"])
    ref = g["em_cost"]
    assert np.array_equal(np.isinf(cost), np.isinf(ref)) and np.isinf(ref).sum() == 2   # A7 underflow rows
    fin = np.isfinite(ref)
    np.testing.assert_allclose(cost[fin], ref[fin], rtol=RT)
    np.testing.assert_allclose(comps, g["em_comps"], rtol=1e-9, atol=0)
    # log-domain restatement == reference wherever the reference is finite, and finite elsewhere
    lcost, lcomps = orc.gmm_evaluate(g["em_X"], g["em_mu"], g["em_cov"], g["em_w"], log_domain=True)
    np.testing.assert_allclose(lcost[fin], ref[fin], rtol=1e-12)
    assert np.isfinite(lcost).all()


def test_mahalanobis(golden):
    g = golden["emission"]
    out = [orc.mahalanobis(a, b, c) for a, b, c in zip(g["mh_v1"], g["mh_v2"], g["mh_var"])]
    np.testing.assert_allclose(out, g["mh_out"], rtol=1e-13)


def test_isolated_decode(golden):
    g = golden["decode"]
    models = models_from(g["iso_mu"], g["iso_cov"], g["iso_w"], g["iso_trans"])
    utts = unragged(g["iso_X"], g["iso_off"])
    rc = unragged(g["iso_costs"], g["iso_costs_off"])
    rp = unragged(g["iso_paths"], g["iso_paths_off"])
    k = 0
    for u, x in enumerate(utts):
        for w, m in enumerate(models):
            E = orc.emissions(x, [m], np.arange(5))
            costs, path, _ = orc.decode_hmm_states(E, m[3])
            ref = rc[k]
            assert np.array_equal(np.isinf(costs.T), np.isinf(ref))
            f = np.isfinite(ref)
            np.testing.assert_allclose(costs.T[f], ref[f], rtol=RT)
            assert np.array_equal(path, rp[k])
            np.testing.assert_allclose(orc.hmm_evaluate(x, m), g["iso_final"][u, w], rtol=RT)
            k += 1
    final, label = orc.classify(g["iso_X"], g["iso_off"], models)
    np.testing.assert_allclose(final, g["iso_final"], rtol=RT)
    assert np.array_equal(label, np.argmin(g["iso_final"], axis=1))


def test_graph_decode_and_forced_alignment(golden):
    g = golden["decode"]
    models = models_from(g["iso_mu"], g["iso_cov"], g["iso_w"], g["iso_trans"])
    emit, trans, ends = orc.build_state_sequences(5, [m[3] for m in models], [[0, 1, 2]] * 3)
    assert np.array_equal(emit, g["gr_emit"]) and np.array_equal(ends, g["gr_ends"])
    assert np.array_equal(trans, g["gr_trans"]) and np.array_equal(emit < 0, g["gr_is_nes"])
    utts = unragged(g["gr_X"], g["gr_off"])
    rc = unragged(g["gr_costs"], g["gr_costs_off"])
    rp = unragged(g["gr_paths"], g["gr_paths_off"])
    for i, x in enumerate(utts):
        E = orc.emissions(x, models, emit)
        costs, path, _ = orc.decode_hmm_states(E, trans, emit < 0, [[e, -1] for e in ends])
        f = np.isfinite(rc[i])
        assert np.array_equal(np.isfinite(costs.T), f)
        np.testing.assert_allclose(costs.T[f], rc[i][f], rtol=RT)
        assert np.array_equal(path, rp[i])
        assert orc.path_to_words(path, emit, 5) == list(g["gr_strings"][i])
    # forced-alignment chains (continuous_speech.py:80,89)
    fc = unragged(g["fa_costs"], g["fa_costs_off"])
    fp = unragged(g["fa_paths"], g["fa_paths_off"])
    for i, x in enumerate(utts):
        emit, trans, ends = orc.build_state_sequences(5, [m[3] for m in models], [[l] for l in g["gr_strings"][i]])
        assert np.array_equal(emit, g["fa_emit"][i]) and np.array_equal(trans, g["fa_trans"][i])
        E = orc.emissions(x, models, emit)
        costs, path, _ = orc.decode_hmm_states(E, trans, emit < 0, [[e, -1] for e in ends])
        f = np.isfinite(fc[i])
        np.testing.assert_allclose(costs.T[f], fc[i][f], rtol=RT)
        assert np.array_equal(path, fp[i])


def test_dtw_and_helpers(golden):
    g = golden["dtw_helpers"]
    xs = unragged(g["dtw_X"], g["dtw_off"])
    for nm, var in (("dtw_e", None), ("dtw_m", g["dtw_var"])):
        rc = unragged(g[nm + "_costs"], g[nm + "_costs_off"])
        rp = unragged(g[nm + "_paths"], g[nm + "_paths_off"])
        for i, x in enumerate(xs):
            costs, path = orc.dtw(x, g["dtw_y"], g["dtw_trans"], var)
            f = np.isfinite(rc[i])
            assert np.array_equal(np.isfinite(costs.T), f)
            np.testing.assert_allclose(costs.T[f], rc[i][f], rtol=RT)
            assert np.array_equal(path, rp[i])
    assert np.array_equal(orc.calc_transition_costs(3, g["tc_seg_lens"]), g["tc_out"], equal_nan=True)
    assert np.array_equal(orc.calc_transition_costs(3, g["tc_seg_lens2"]), g["tc_out2"])
    assert np.array_equal(orc.get_segments_from_path(g["gs_path"], 5), g["gs_out"])
    segs = orc.segment_data(xs, g["sd_seg_starts"])
    rs = unragged(g["sd_cat"], g["sd_off"])
    assert all(np.array_equal(a, b) for a, b in zip(segs, rs))
    res, var = orc.combine_templates(xs, g["sd_seg_starts"])
    np.testing.assert_allclose(res, g["ct_res"], rtol=1e-12)
    np.testing.assert_allclose(var, g["ct_vars"], rtol=1e-11)
    np.testing.assert_allclose(orc.calc_variance(xs[1]), g["cv_out"], rtol=1e-11)


def test_dtw_beam(golden):
    """oracle.dtw_beam against dtw(..., beam=k) of the reference (decode.py:42-48,62-68): costs incl. -1 marks, paths."""
    g = golden["beam"]
    xs = unragged(g["bm_X"], g["bm_off"])
    for k, (beam, maha, ti, u, plen) in enumerate(g["bm_cases"]):
        Dl = orc.local_dist(xs[u], g["bm_y"], g["bm_var"] if maha else None)
        costs, path = orc.dtw_beam(Dl, g["bm_trans2"] if ti else g["bm_trans"], int(beam))
        ref = g[f"bm_costs{k}"]
        f = np.isfinite(ref)
        assert np.array_equal(np.isfinite(costs), f)
        np.testing.assert_allclose(costs[f], ref[f], rtol=RT)
        assert np.array_equal(path, g[f"bm_path{k}"])
    # beam = inf is the C restatement's dtw
    Dl = orc.local_dist(xs[2], g["bm_y"])
    c0, p0 = orc.dtw(xs[2], g["bm_y"], g["bm_trans"])
    c1, p1 = orc.dtw_beam(Dl, g["bm_trans"])
    assert np.array_equal(c0, c1) and np.array_equal(p0, p1)


def test_end_points_off_the_last_frame(golden):
    """orc_decode with any [state, frame] end points (decode.py:126-146)."""
    g = golden["ends"]
    models = models_from(g["en_mu"], g["en_cov"], g["en_w"], g["en_wtrans"])
    emit, trans = g["en_emit"], g["en_trans"]
    E = orc.emissions(g["en_X"], models, emit)
    for k in range(int(g["en_n"])):
        costs, path, _ = orc.decode_hmm_states(E, trans, emit < 0, g[f"en_list{k}"])
        f = np.isfinite(g[f"en_costs{k}"])
        np.testing.assert_allclose(costs[f], g[f"en_costs{k}"][f], rtol=RT)
        assert np.array_equal(path.reshape(-1, 2), g[f"en_path{k}"])


def test_kmeans_and_em(golden):
    g = golden["kmeans_em"]
    data = g["km_data"]
    for k in (2, 4):
        clusters, cent, cov = orc.kmeans(data, k, g[f"km{k}_cent_in"], rng_randint=g[f"km{k}_partition"])
        assert np.array_equal(clusters, g[f"km{k}_clusters"])
        np.testing.assert_allclose(cent, g[f"km{k}_cent"], rtol=1e-12)
        np.testing.assert_allclose(cov, g[f"km{k}_cov"], rtol=1e-11)
        for iters in (1, 3):
            st = orc.GMMState(data.mean(axis=0), orc.calc_variance(data), 4)
            st.mu, st.cov, st.w = g[f"em{k}_mu_in"].copy(), g[f"em{k}_cov_in"].copy(), g[f"em{k}_w_in"].copy()
            st.em(data, k, iters)
            np.testing.assert_allclose(st.w, g[f"em{k}_it{iters}_w"], rtol=1e-10)
            np.testing.assert_allclose(st.mu, g[f"em{k}_it{iters}_mu"], rtol=1e-9, atol=1e-12)
            np.testing.assert_allclose(st.cov, g[f"em{k}_it{iters}_cov"], rtol=1e-9)


def test_kmeans_draws_global_rng_like_reference(golden):
    g = golden["kmeans_em"]
    np.random.seed(11)                                    # make_golden.py: same seed before the k=2 call
    clusters, cent, cov = orc.kmeans(g["km_data"], 2, g["km2_cent_in"])
    assert np.array_equal(clusters, g["km2_clusters"])
    np.testing.assert_allclose(cov, g["km2_cov"], rtol=1e-11)


def test_skmeans_and_fit(golden):
    g = golden["train"]
    utts = [unragged(g[f"tr_X{w}"], g[f"tr_off{w}"]) for w in range(3)]
    res, var, trans, segs = orc.skmeans(utts[0], 5)
    np.testing.assert_allclose(res, g["sk_res"], rtol=1e-11)
    np.testing.assert_allclose(var, g["sk_vars"], rtol=1e-10)
    assert np.array_equal(trans, g["sk_trans"])
    assert all(np.array_equal(a, b) for a, b in zip(segs, unragged(g["sk_seg_cat"], g["sk_seg_off"])))
    ev = [orc.hmm_evaluate_single(x, res, var, trans) for x in utts[0][:3] + utts[1][:2]]
    np.testing.assert_allclose(ev, g["fit1_eval"], rtol=1e-9)
    models = []
    for w in range(3):
        np.random.seed(1000 + w)
        model, hmu, hsig, segments = orc.hmm_fit(utts[w], 5, 4)
        models.append(model)
        np.testing.assert_allclose(hmu, g["fit_hmm_mu"][w], rtol=1e-11)
        # EM runs to np.allclose convergence (hundreds of iterations): round-off grows, so 1e-6
        np.testing.assert_allclose(model[0], g["fit_mu"][w], rtol=1e-6, atol=1e-8)
        np.testing.assert_allclose(model[1], g["fit_cov"][w], rtol=1e-6)
        np.testing.assert_allclose(model[2], g["fit_w"][w], rtol=1e-6)
        assert np.array_equal(model[3], g["fit_trans"][w])
        rs = unragged(g[f"fit_seg_cat{w}"], g[f"fit_seg_off{w}"])
        assert all(np.array_equal(a, b) for a, b in zip(segments, rs))
    ev = np.array([[orc.hmm_evaluate(utts[v][i], models[w]) for w in range(3)] for v in range(3) for i in range(2)])
    np.testing.assert_allclose(ev, g["fit_eval"], rtol=1e-6)
    np.random.seed(2000)
    mk, _, _, _ = orc.hmm_fit(utts[0], 5, 4, use_em=False)
    np.testing.assert_allclose(mk[0], g["fitk_mu"][0], rtol=1e-11)
    np.testing.assert_allclose(mk[1], g["fitk_cov"][0], rtol=1e-10)
    np.testing.assert_allclose(mk[2], g["fitk_w"][0], rtol=1e-12)


def test_continuous_train_iteration(golden):
    g = golden["train"]
    models = models_from(g["fit_mu"], g["fit_cov"], g["fit_w"], g["fit_trans"])
    data = unragged(g["ct_X"], g["ct_off"])
    np.random.seed(3000)
    new = orc.continuous_train_iteration(data, models, [list(s) for s in g["ct_strings"]])
    for w in range(3):
        np.testing.assert_allclose(new[w][3], g["ctr_trans"][w], rtol=1e-12)
        np.testing.assert_allclose(new[w][2], g["ctr_w"][w], rtol=1e-5)
        np.testing.assert_allclose(new[w][0], g["ctr_mu"][w], rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(new[w][1], g["ctr_cov"][w], rtol=1e-5)


def test_forward_backward_identities(golden):
    """a21 has no reference: tie it to decode_hmm_states and check the usual identities."""
    g = golden["decode"]
    models = models_from(g["iso_mu"], g["iso_cov"], g["iso_w"], g["iso_trans"])
    x = unragged(g["iso_X"], g["iso_off"])[2]
    for m in models:
        E = orc.emissions(x, [m], np.arange(5), log_domain=True)
        nlz, a, b, gam = orc.forward_backward(E, m[3])
        np.testing.assert_allclose(gam.sum(axis=1), 1.0, rtol=1e-10)          # sum_s gamma_t(s) = 1
        for t in (0, 7, len(x) - 1):                                          # log Z from any column
            col = a[t] + b[t]
            mn = col.min()
            np.testing.assert_allclose(mn - np.log(np.exp(-(col - mn)).sum()), nlz, rtol=1e-12)
        costs, _, _ = orc.decode_hmm_states(E, m[3])
        assert nlz <= costs[-1, -1] + 1e-9                                    # sum over paths >= best path
        # the SAME recursion in the (min, +) semiring is decode_hmm_states: bit for bit (SURVEY 8c asks "exactly")
        nlz_m, a_m, b_m, _ = orc.forward_backward(E, m[3], semiring="min")
        assert np.array_equal(a_m, costs.T)
        assert nlz_m == costs[-1, -1]
        # and its backward half: alpha + beta along the best path is the best cost
        path = orc.decode_hmm_states(E, m[3])[1]
        for s_, c_ in path:
            assert abs(a_m[c_, s_] + b_m[c_, s_] - nlz_m) <= 1e-9 * abs(nlz_m)
        # "hot" limit of the sum semiring: scaled costs -> soft-min tends to min
        s = 200.0
        nlz_s, a_s, _, _ = orc.forward_backward(E * s, np.where(np.isinf(m[3]), np.inf, m[3] * s))
        f = np.isfinite(costs.T)
        np.testing.assert_allclose(a_s[f] / s, costs.T[f], rtol=1e-6, atol=np.log(5 * len(x)) / s)


def test_forward_backward_against_path_enumeration():
    """Independent pin of the forward-backward restatement (a21 has no reference): every state sequence of small
    models enumerated explicitly -- log Z, state posteriors and summed edge posteriors to 1e-10."""
    rng = np.random.default_rng(77)
    for T, R in ((2, 2), (3, 3), (5, 2), (6, 3), (7, 4), (8, 4), (9, 3)):
        for trial in range(3):
            E = rng.uniform(0.5, 6.0, size=(T, R))
            tr = np.full((R, R), np.inf)
            for i in range(R):
                tr[i, i] = rng.uniform(0.05, 2.0)
                if i:
                    tr[i, i - 1] = rng.uniform(0.05, 2.5)
                if i > 1 and trial != 1:
                    tr[i, i - 2] = rng.uniform(0.5, 3.0)                      # skip arcs (kmeans.py:72-84)
            if trial == 2:
                tr = rng.uniform(0.1, 3.0, size=(R, R))                       # ergodic
            nlz_b, gam_b, xi_b = orc.forward_backward_bruteforce(E, tr)
            nlz, a, b, gam = orc.forward_backward(E, tr)
            if not np.isfinite(nlz_b):
                assert not np.isfinite(nlz)
                continue
            np.testing.assert_allclose(nlz, nlz_b, rtol=1e-12)
            np.testing.assert_allclose(gam, gam_b, rtol=1e-10, atol=1e-300)
            xi = np.zeros((R, R))
            for r in range(R):
                for o in range(R):
                    if np.isfinite(tr[r, o]):
                        xi[r, o] = np.exp(-(a[:-1, o] + tr[r, o] + E[1:, r] + b[1:, r] - nlz)).sum()
            np.testing.assert_allclose(xi, xi_b, rtol=1e-10, atol=1e-300)


def test_feature_oracle_matches_reference_front_end():
    """oracle/feature_oracle.py against the frozen outputs of the reference's mfcc_features / load_wav_as_mfcc1
    (tests/golden/make_golden_feature.py), and the product's host-built mel filter bank against the oracle's."""
    import os
    import sys
    from oracle import feature_oracle as fo
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_feature.npz"))
    for i in range(int(g["n"])):
        fb, mf = fo.mfcc_features(g[f"sig{i}"])
        np.testing.assert_allclose(fb, g[f"fbank{i}"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(mf, g[f"mfcc{i}"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(fo.load_features(g[f"sig{i}"]), g[f"feat{i}"], rtol=1e-9, atol=1e-9)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(__file__)), "speech-recognition_b200"))
    from srb200.frontend import mel_filterbank
    w, rng = mel_filterbank(16000)
    ref, _ = fo.mel_filterbank(16000)
    np.testing.assert_allclose(w, ref.astype(np.float32), rtol=0, atol=0)
    for m in range(40):
        nz = np.nonzero(ref[m])[0]
        assert rng[m, 0] <= nz.min() and nz.max() < rng[m, 1]
