#!/usr/bin/env python
"""Generate golden vectors by EXECUTING the unmodified reference (container only).

    python tests/golden/make_golden.py            # writes tests/golden/*.npz

The reference (`/root/reference/sr/recognition`) has no tests or fixtures of its own for this path
(SURVEY.md section 4), so parity is pinned by running its functions on seeded synthetic inputs and
freezing inputs + outputs here.  The fixtures travel to the GPU box; the reference does not.
Every array stored is float64 / int64 exactly as the reference produced it.

Reference entry points exercised (file:line in /root/reference):
  hmm_state.py:36-45   MultivariateNormal.pdf          hmm_state.py:48-58  mahalanobis
  hmm_state.py:114-120 GMM.evaluate                    hmm_state.py:122-159 GMM.em
  decode.py:7-77       dtw                             decode.py:80-146    decode_hmm_states
  kmeans.py:6-108      calc_variance/combine_templates/segment_data/calc_transition_costs/
                       get_segments_from_path          kmeans.py:111-155   skmeans
  kmeans.py:167-193    kmeans                          kmeans.py:196-205   align_gmm_states
  hmm.py:57-135        HMM.fit / HMM.evaluate          continuous_speech.py:13-53 build_state_sequences
  continuous_speech.py:56-179 continuous_train
"""
import contextlib
import io
import os
import pickle
import sys
import tempfile
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from _ref_shim import load_reference  # noqa: E402

warnings.simplefilter("ignore")
rec = load_reference()
D = 39


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


# ----------------------------------------------------------------------------- synthetic inputs
def make_word_templates(rng, n_words, n_states, dim=D):
    return rng.normal(0.0, 1.0, size=(n_words, n_states, dim))


def make_utterance(rng, base_w, T, noise=0.5):
    """T frames walking left-to-right through the S state means of one word (uniform segments)."""
    S = base_w.shape[0]
    seg = np.minimum((np.arange(T) * S) // T, S - 1)
    return base_w[seg] + noise * rng.normal(size=(T, base_w.shape[1]))


def make_gmm(rng, base_ws, M, weights=None):
    """A reference GMM object with M distinct components around base_ws."""
    g = rec.GMM(base_ws.copy(), np.ones(D), M)
    mus = base_ws[None, :] + 0.5 * rng.normal(size=(M, D))
    sig = rng.uniform(0.5, 1.5, size=(M, D))
    w = rng.dirichlet(np.ones(M)) if weights is None else np.asarray(weights, float)
    g.update_models(mus, sig, w)
    return g


def ltr_transitions(S, p_stay=0.9):
    tr = np.full((S, S), np.inf)
    for i in range(S):
        if i < S - 1:
            tr[i, i] = -np.log(p_stay)
            tr[i + 1, i] = -np.log(1 - p_stay)
        else:
            tr[i, i] = 0.0
    return tr


def make_model(rng, base_w, M, weights=None):
    S = base_w.shape[0]
    m = rec.HMM(S)
    m.gmm_states = [make_gmm(rng, base_w[s], M, weights) for s in range(S)]
    m.transitions = ltr_transitions(S)
    m.mu = base_w.copy()
    m.sigma = np.ones_like(base_w)
    return m


def dump_models(models):
    """(means[W,S,M,D], covs[W,S,M,D], w[W,S,M], trans[W,S,S])"""
    W, S, M = len(models), len(models[0].gmm_states), models[0].gmm_states[0].n_gaussians
    mu = np.zeros((W, S, M, D))
    cv = np.zeros((W, S, M, D))
    w = np.zeros((W, S, M))
    tr = np.zeros((W, S, S))
    for a, m in enumerate(models):
        tr[a] = m.transitions
        for b, g in enumerate(m.gmm_states):
            w[a, b] = g.w
            for c, dist in enumerate(g.dists):
                mu[a, b, c] = dist.mean
                cv[a, b, c] = dist.cov
    return mu, cv, w, tr


def ragged(arrs):
    """list of [T_i, ...] arrays -> (concatenated, offsets)"""
    off = np.zeros(len(arrs) + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(a) for a in arrs])
    if len(arrs) and sum(len(a) for a in arrs):
        cat = np.concatenate([np.asarray(a) for a in arrs if len(a)], axis=0)
    else:
        cat = np.zeros((0,))
    return cat, off


# ----------------------------------------------------------------------------- cases
def case_emission(out):
    rng = np.random.default_rng(101)
    base = rng.normal(size=D)
    # A2: reference-style weights that do not sum to 1
    g = make_gmm(rng, base, 4, weights=[0.552, 0.448, 0.25, 0.25])
    X = base[None, :] + rng.normal(size=(48, D))
    X[5] += 40.0  # A7: linear-domain pdf underflows -> cost inf
    X[17] = base + 9.0 * rng.normal(size=D)
    cost = np.array([g.evaluate(x) for x in X])
    comps = np.array([g.evaluate(x, return_neg_log_likelihood=False) for x in X])
    pdf0 = np.array([g.dists[0].pdf(x) for x in X])
    mu = np.array([d.mean for d in g.dists])
    cv = np.array([d.cov for d in g.dists])
    out.update(em_X=X, em_mu=mu, em_cov=cv, em_w=g.w.copy(), em_cost=cost, em_comps=comps, em_pdf0=pdf0)
    v1 = rng.normal(size=(8, D))
    v2 = rng.normal(size=(8, D))
    var = rng.uniform(0.3, 2.0, size=(8, D))
    out.update(mh_v1=v1, mh_v2=v2, mh_var=var,
               mh_out=np.array([rec.mahalanobis(a, b, c) for a, b, c in zip(v1, v2, var)]))


def case_isolated_decode(out):
    rng = np.random.default_rng(202)
    W, S, M = 3, 5, 4
    base = make_word_templates(rng, W, S)
    models = [make_model(rng, base[w], M) for w in range(W)]
    # one model with the isolated-training style transitions (A5) incl. a skip arc
    seg_lens = np.array([[8, 8, 8, 8, 8], [7, 7, 7, 7, 7]])
    models[1].transitions = rec.calc_transition_costs(2, seg_lens)
    lens = [20, 33, 40, 57, 6]
    words = [0, 1, 2, 1, 0]
    utts = [make_utterance(rng, base[w], T) for w, T in zip(words, lens)]
    costs_all, paths_all, final = [], [], np.zeros((len(utts), W))
    for u, x in enumerate(utts):
        for w in range(W):
            c, p = rec.decode_hmm_states(x, models[w].gmm_states, models[w].transitions)
            costs_all.append(c.T.copy())  # [T,S]
            paths_all.append(p)
            final[u, w] = models[w].evaluate(x)
    mu, cv, wt, tr = dump_models(models)
    X, off = ragged(utts)
    C, coff = ragged(costs_all)
    P, poff = ragged(paths_all)
    out.update(iso_mu=mu, iso_cov=cv, iso_w=wt, iso_trans=tr, iso_X=X, iso_off=off, iso_words=np.array(words),
               iso_costs=C, iso_costs_off=coff, iso_paths=P, iso_paths_off=poff, iso_final=final)
    return models, base


def case_graph_decode(out, models, base):
    rng = np.random.default_rng(303)
    W = len(models)
    # digit-loop style grammar (main.py:35): every word allowed in each of 3 layers
    seq, trans, ends = rec.build_state_sequences(models, [list(range(W))] * 3)
    is_nes = np.array([type(s) is rec.NES for s in seq])
    ids = {}
    for w, m in enumerate(models):
        for s, g in enumerate(m.gmm_states):
            ids[g.id] = w * len(m.gmm_states) + s
    emit = np.array([-1 if type(s) is rec.NES else ids[s.id] for s in seq])
    strings = [[2, 0, 1], [1, 1, 2]]
    utts = []
    for st in strings:
        utts.append(np.concatenate([make_utterance(rng, base[w], 18 + 3 * i) for i, w in enumerate(st)]))
    costs, paths = [], []
    for x in utts:
        c, p = rec.decode_hmm_states(x, seq, trans, end_points=[[e, -1] for e in ends])
        costs.append(c.T.copy())
        paths.append(p)
    X, off = ragged(utts)
    C, coff = ragged(costs)
    P, poff = ragged(paths)
    out.update(gr_trans=trans, gr_is_nes=is_nes, gr_emit=emit, gr_ends=np.array(ends), gr_X=X, gr_off=off,
               gr_costs=C, gr_costs_off=coff, gr_paths=P, gr_paths_off=poff, gr_strings=np.array(strings))
    # forced-alignment chains (continuous_speech.py:80,89)
    fa_costs, fa_paths, fa_trans, fa_emit, fa_nes, fa_ends = [], [], [], [], [], []
    for st, x in zip(strings, utts):
        seq, trans, ends = rec.build_state_sequences(models, [[l] for l in st])
        c, p = rec.decode_hmm_states(x, seq, trans, end_points=[[e, -1] for e in ends])
        fa_costs.append(c.T.copy())
        fa_paths.append(p)
        fa_trans.append(trans)
        fa_nes.append(np.array([type(s) is rec.NES for s in seq]))
        fa_emit.append(np.array([-1 if type(s) is rec.NES else ids[s.id] for s in seq]))
        fa_ends.append(np.array(ends))
    C, coff = ragged(fa_costs)
    P, poff = ragged(fa_paths)
    out.update(fa_costs=C, fa_costs_off=coff, fa_paths=P, fa_paths_off=poff, fa_trans=np.array(fa_trans),
               fa_is_nes=np.array(fa_nes), fa_emit=np.array(fa_emit), fa_ends=np.array(fa_ends))


def case_dtw_helpers(out):
    rng = np.random.default_rng(404)
    S = 5
    base = make_word_templates(rng, 1, S)[0]
    tr = rec.calc_transition_costs(3, np.array([[6] * S] * 3))
    xs = [make_utterance(rng, base, T) for T in (12, 30, 41)]
    cs, ps = [], []
    for x in xs:
        c, p = rec.dtw(x, base, lambda *a: np.linalg.norm(a[0] - a[1]), tr)
        cs.append(c.T.copy())
        ps.append(p)
    var = rng.uniform(0.4, 1.6, size=(S, D))
    cm, pm = [], []
    for x in xs:
        c, p = rec.dtw(x, base, rec.mahalanobis, tr, var)
        cm.append(c.T.copy())
        pm.append(p)
    X, off = ragged(xs)
    out.update(dtw_X=X, dtw_off=off, dtw_y=base, dtw_trans=tr, dtw_var=var)
    for nm, arrs in (("dtw_e_costs", cs), ("dtw_e_paths", ps), ("dtw_m_costs", cm), ("dtw_m_paths", pm)):
        a, o = ragged(arrs)
        out[nm] = a
        out[nm + "_off"] = o
    # helpers -----------------------------------------------------------------------------
    seg_lens = np.array([[4, 0, 5, 3, 6], [5, 0, 4, 4, 7], [6, 0, 3, 5, 4]])  # an empty segment: skip arc
    out.update(tc_seg_lens=seg_lens, tc_out=rec.calc_transition_costs(3, seg_lens),
               tc_seg_lens2=np.array([[6] * S] * 3), tc_out2=tr)
    out.update(gs_path=ps[1], gs_out=rec.get_segments_from_path(ps[1], S))
    seg_starts = np.array([[0, 2, 5, 7, 9], [0, 6, 12, 18, 24], [0, 8, 16, 24, 33]])
    segs = rec.segment_data(xs, 3, S, seg_starts)
    a, o = ragged(segs)
    res, vars_ = rec.combine_templates(xs, 3, S, seg_starts)
    out.update(sd_seg_starts=seg_starts, sd_cat=a, sd_off=o, ct_res=res, ct_vars=vars_,
               cv_out=rec.calc_variance(xs[1].T))


def case_kmeans_em(out):
    rng = np.random.default_rng(505)
    c0 = rng.normal(size=D)
    data = np.concatenate([c0 * 0.85 + 0.6 * rng.normal(size=(90, D)), c0 * 1.2 + 0.6 * rng.normal(size=(70, D))])
    rng.shuffle(data)
    cent = np.array([data.mean(axis=0)])
    out["km_data"] = data
    g = rec.GMM(data.mean(axis=0), rec.calc_variance(data.T), 4)
    weights = np.full(4, 1 / data.shape[0])
    # hmm.py:104: n_splits = int(np.log(4)) = 1 (A1); we also run the second split to cover k=4
    np.random.seed(11)
    for i in range(2):
        cent = np.concatenate([cent * 0.9, cent * 1.1], axis=0)
        k = 2 ** (i + 1)
        st = np.random.get_state()
        part = np.random.randint(0, k, data.shape[0])  # what kmeans.py:171 will draw
        np.random.set_state(st)
        out[f"km{k}_cent_in"] = cent.copy()
        clusters, cent, cov = rec.kmeans(data, k, cent, dist_fun=rec.mahalanobis)
        out[f"km{k}_partition"] = part
        out[f"km{k}_clusters"] = clusters
        out[f"km{k}_cent"] = cent.copy()
        out[f"km{k}_cov"] = cov.copy()
        cs, cc = np.unique(clusters, return_counts=True)
        for c in cs:
            weights[c] = cc[c] / data.shape[0]
        g.update_models(cent, cov, weights[:k])
        out[f"em{k}_w_in"] = g.w.copy()
        out[f"em{k}_mu_in"] = np.array([d.mean for d in g.dists])
        out[f"em{k}_cov_in"] = np.array([d.cov for d in g.dists])
        # one and three EM iterations from this state, then to convergence
        import copy
        for iters in (1, 3):
            h = copy.deepcopy(g)
            quiet(h.em, data, k, iters)
            out[f"em{k}_it{iters}_w"] = h.w.copy()
            out[f"em{k}_it{iters}_mu"] = np.array([d.mean for d in h.dists])
            out[f"em{k}_it{iters}_cov"] = np.array([d.cov for d in h.dists])
        quiet(g.em, data, k)
        out[f"em{k}_conv_w"] = g.w.copy()
        out[f"em{k}_conv_mu"] = np.array([d.mean for d in g.dists])
        out[f"em{k}_conv_cov"] = np.array([d.cov for d in g.dists])


def case_train(out):
    rng = np.random.default_rng(606)
    W, S, M = 3, 5, 4
    base = make_word_templates(rng, W, S)
    lens = [[34, 41, 30, 38, 45, 36], [31, 33, 44, 39, 35, 40], [30, 37, 42, 32, 43, 36]]
    utts = [[make_utterance(rng, base[w], T) for T in lens[w]] for w in range(W)]
    for w in range(W):
        X, off = ragged(utts[w])
        out[f"tr_X{w}"] = X
        out[f"tr_off{w}"] = off
    # skmeans (kmeans.py:111-155) on word 0
    res, vars_, tc, segs = rec.skmeans(utts[0], S, return_segmented_data=True)
    a, o = ragged(segs)
    out.update(sk_res=res, sk_vars=vars_, sk_trans=tc, sk_seg_cat=a, sk_seg_off=o)
    # single-gaussian model (hmm.py:74-75,134)
    m1 = quiet(rec.HMM(S).fit, utts[0], n_gaussians=4, use_gmm=False)
    out["fit1_eval"] = np.array([m1.evaluate(x) for x in utts[0][:3]] + [m1.evaluate(x) for x in utts[1][:2]])
    # full GMM-HMM fit per word, global RNG seeded per word (kmeans.py:171 consumes it)
    models = []
    for w in range(W):
        np.random.seed(1000 + w)
        m = quiet(rec.HMM(S).fit, utts[w], n_gaussians=M, use_gmm=True, use_em=True)
        models.append(m)
        a, o = ragged(m.segments)
        out[f"fit_seg_cat{w}"] = a
        out[f"fit_seg_off{w}"] = o
    mu, cv, wt, tr = dump_models(models)
    out.update(fit_mu=mu, fit_cov=cv, fit_w=wt, fit_trans=tr,
               fit_hmm_mu=np.array([m.mu for m in models]), fit_hmm_sigma=np.array([m.sigma for m in models]))
    ev = np.array([[models[w].evaluate(utts[v][i]) for w in range(W)] for v in range(W) for i in range(2)])
    out["fit_eval"] = ev
    # k-means-only variant (use_em=False)
    np.random.seed(2000)
    mk = quiet(rec.HMM(S).fit, utts[0], n_gaussians=M, use_gmm=True, use_em=False)
    mu1, cv1, wt1, tr1 = dump_models([mk])
    out.update(fitk_mu=mu1, fitk_cov=cv1, fitk_w=wt1)

    # continuous_train (continuous_speech.py:56-179), one outer iteration
    strings = [[0, 1, 2], [2, 0, 1], [1, 2, 0], [0, 2, 1], [1, 0, 2], [2, 1, 0], [0, 0, 1], [1, 1, 2], [2, 2, 0]]
    data = []
    for st in strings:
        data.append(np.concatenate([make_utterance(rng, base[w], 44 + 5 * i) for i, w in enumerate(st)]))
    X, off = ragged(data)
    out.update(ct_X=X, ct_off=off, ct_strings=np.array(strings))
    for hm_i, m in enumerate(models):
        for s in m.gmm_states:
            s.parent = hm_i  # core.py:107-109
    with tempfile.TemporaryDirectory() as td:
        np.random.seed(3000)
        quiet(rec.continuous_train, data, models, strings, td, 4, 5, 1)
        new = [pickle.load(open(os.path.join(td, f"{i}.pkl"), "rb")) for i in range(W)]
    mu2, cv2, wt2, tr2 = dump_models(new)
    out.update(ctr_mu=mu2, ctr_cov=cv2, ctr_w=wt2, ctr_trans=tr2)


def main():
    import time
    only = set(sys.argv[1:])
    for name, fn in (("emission", case_emission), ("decode", None), ("dtw_helpers", case_dtw_helpers),
                     ("kmeans_em", case_kmeans_em), ("train", case_train)):
        if only and name not in only:
            continue
        t0 = time.time()
        out = {}
        if name == "decode":
            models, base = case_isolated_decode(out)
            case_graph_decode(out, models, base)
        else:
            fn(out)
        path = os.path.join(HERE, f"ref_{name}.npz")
        np.savez_compressed(path, **out)
        print(f"{name}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB, {time.time() - t0:.1f}s")


if __name__ == "__main__":
    main()
