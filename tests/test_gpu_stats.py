"""GPU parity: K5/K6 statistics kernels (local distances, group moments, GMM E-step + M-step) against the
frozen reference outputs and the oracle, through the C ABI."""
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


def test_local_dist_golden(eng, golden):
    g = golden["dtw_helpers"]
    X = g["dtw_X"]
    for var in (None, g["dtw_var"]):
        ref = orc.local_dist(X, g["dtw_y"], var)
        out, am = eng.local_dist(X.astype(np.float32), g["dtw_y"], var, want_argmin=True)
        np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=RTOL)
        assert np.array_equal(am.cpu().numpy(), ref.argmin(axis=1))
    e = golden["emission"]                                  # mahalanobis() of the reference itself
    for v1, v2, var, ref in zip(e["mh_v1"], e["mh_v2"], e["mh_var"], e["mh_out"]):
        out, _ = eng.local_dist(v1[None].astype(np.float32), v2[None], var[None])
        np.testing.assert_allclose(out.cpu().numpy()[0, 0], ref, rtol=RTOL)


def test_group_stats_combine_templates_golden(eng, golden):
    """segment means / unbiased variances of combine_templates (kmeans.py:15-30)."""
    g = golden["dtw_helpers"]
    xs = unragged(g["dtw_X"], g["dtw_off"])
    starts = g["sd_seg_starts"]
    labels = []
    for r, x in enumerate(xs):
        b = list(starts[r]) + [len(x)]
        labels.append(np.repeat(np.arange(5), np.diff(b)))
    labels = np.concatenate(labels).astype(np.int32)
    X = g["dtw_X"].astype(np.float32)
    shift = np.tile(X.mean(axis=0), (5, 1))
    st = eng.group_stats(X, labels, 5, shift).cpu().numpy()
    n, mean, var = eng.moments_from_stats(st, shift, 39)
    assert np.array_equal(n, np.bincount(labels, minlength=5))
    np.testing.assert_allclose(mean, g["ct_res"], rtol=RTOL, atol=1e-5)
    np.testing.assert_allclose(var, g["ct_vars"], rtol=RTOL)


def test_group_stats_large_random(eng):
    rng = np.random.default_rng(0)
    N, K, D = 200_003, 7, 39
    X = (rng.normal(size=(N, D)) * 2 + 5).astype(np.float32)
    lab = rng.integers(-1, K, size=N).astype(np.int32)
    lab[1000:1400] = 3                                     # long runs too
    shift = rng.normal(size=(K, D)) + 5
    st = eng.group_stats(X, lab, K, shift).cpu().numpy()
    n, mean, var = eng.moments_from_stats(st, shift, D)
    for k in range(K):
        xk = X[lab == k].astype(np.float64)
        assert n[k] == len(xk)
        np.testing.assert_allclose(mean[k], xk.mean(axis=0), rtol=1e-5)
        np.testing.assert_allclose(var[k], xk.var(axis=0, ddof=1), rtol=RTOL)


@pytest.mark.parametrize("k", [2, 4])
def test_kmeans_pieces_golden(eng, golden, k):
    """The three device pieces of kmeans (kmeans.py:167-193): partition variances, assignment with cov[0],
    centroid update -- against the frozen reference run."""
    g = golden["kmeans_em"]
    data = g["km_data"]
    X = data.astype(np.float32)
    shift = np.tile(data.mean(axis=0), (k, 1))
    st = eng.group_stats(X, g[f"km{k}_partition"].astype(np.int32), k, shift).cpu().numpy()
    _, _, cov = eng.moments_from_stats(st, shift, 39)
    np.testing.assert_allclose(cov, g[f"km{k}_cov"], rtol=RTOL)
    # assignment of the converged centroids reproduces the converged clusters (fixed point)
    _, am = eng.local_dist(X, g[f"km{k}_cent"], g[f"km{k}_cov"][:1], shared_var=True, want_dist=False,
                           want_argmin=True)
    assert np.array_equal(am.cpu().numpy(), g[f"km{k}_clusters"])
    st = eng.group_stats(X, am, k, g[f"km{k}_cent"]).cpu().numpy()
    _, cent, _ = eng.moments_from_stats(st, g[f"km{k}_cent"], 39)
    np.testing.assert_allclose(cent, g[f"km{k}_cent"], rtol=RTOL, atol=1e-6)


@pytest.mark.parametrize("k", [2, 4])
def test_em_iteration_golden(eng, golden, k):
    """One and three GMM.em iterations (hmm_state.py:122-159) on the first k of 4 components."""
    g = golden["kmeans_em"]
    X = g["km_data"].astype(np.float32)
    lab = np.zeros(len(X), np.int32)
    pm = eng.PackedModel(g[f"em{k}_mu_in"][None], g[f"em{k}_cov_in"][None], g[f"em{k}_w_in"][None])
    for it in range(1, 4):
        st = eng.gmm_stats(X, lab, pm, k)
        eng.mstep(st, pm, k)
        if it in (1, 3):
            np.testing.assert_allclose(pm.w.cpu().numpy()[0], g[f"em{k}_it{it}_w"], rtol=RTOL)
            np.testing.assert_allclose(pm.mu.cpu().numpy()[0], g[f"em{k}_it{it}_mu"], rtol=RTOL, atol=1e-5)
            np.testing.assert_allclose(pm.var.cpu().numpy()[0], g[f"em{k}_it{it}_cov"], rtol=RTOL)


def test_gmm_stats_multi_state_vs_oracle(eng):
    """Frames hard-aligned to 55 tied states, ragged runs, skipped frames; statistics per state == oracle EM pass."""
    from srb200 import synth
    models, base = synth.make_models(51, W=11, S=5, M=4)
    X, off, words = synth.make_utterances(52, base, 300, T=60)
    rng = np.random.default_rng(3)
    N = X.shape[0]
    t_in = np.arange(N) - np.repeat(off[:-1], np.diff(off))
    lab = (np.repeat(words, np.diff(off)) * 5 + np.minimum(t_in // 12, 4)).astype(np.int32)
    lab[rng.random(N) < 0.05] = -1
    pm, _ = eng.PackedModel.from_words(models)
    st = eng.gmm_stats(X, lab, pm, 2)
    eng.mstep(st, pm, 2)
    mu, var, w = pm.mu.cpu().numpy(), pm.var.cpu().numpy(), pm.w.cpu().numpy()
    for s in (0, 7, 23, 54):
        m = models[s // 5]
        data = X[lab == s].astype(np.float64)
        gs = orc.GMMState(m[0][s % 5, 0], m[1][s % 5, 0], 4)
        gs.mu, gs.cov, gs.w = m[0][s % 5].copy(), m[1][s % 5].copy(), m[2][s % 5].copy()
        gs.em(data, 2, 1)
        np.testing.assert_allclose(mu[s], gs.mu, rtol=RTOL, atol=1e-5)
        np.testing.assert_allclose(var[s], gs.cov, rtol=RTOL)
        np.testing.assert_allclose(w[s], gs.w, rtol=RTOL)


def _pool_case(seed=52, B=300, T=60, ragged=None):
    from srb200 import synth
    models, base = synth.make_models(51, W=11, S=5, M=4)
    X, off, words = synth.make_utterances(seed, base, B, T=T) if ragged is None else \
        synth.make_utterances(seed, base, B, ragged=ragged)
    rng = np.random.default_rng(3)
    N = X.shape[0]
    t_in = np.arange(N) - np.repeat(off[:-1], np.diff(off))
    Tn = np.repeat(np.diff(off), np.diff(off))
    lab = (np.repeat(words, np.diff(off)) * 5 + np.minimum(t_in * 5 // Tn, 4)).astype(np.int32)
    lab[rng.random(N) < 0.05] = -1
    return models, X, lab


@pytest.mark.parametrize("k", [1, 2, 4])
def test_pool_em_stats_equals_gmm_stats(eng, k):
    """srb_pool_stats mode 0 == srb_gmm_stats (same arithmetic, other summation order), ragged tail chunk included, and
    two runs are bit-identical (private slices, program-order reductions)."""
    models, X, lab = _pool_case(ragged=(37, 81))
    pm, _ = eng.PackedModel.from_words(models)
    assert eng.pool_stats_supported(55, k, 39)
    ref = eng.gmm_stats(X, lab, pm, k).cpu().numpy()
    got = eng.pool_em_stats(X, lab, pm, k).cpu().numpy()
    again = eng.pool_em_stats(X, lab, pm, k).cpu().numpy()
    assert np.array_equal(got, again)
    D = 39
    scale = np.abs(ref[:, 1 + D:1 + 2 * D]).max(axis=1, keepdims=True) + 1e-30     # a row's second moments set its scale
    assert np.max(np.abs(got - ref)[:, :1 + 2 * D] / np.maximum(scale, np.abs(ref[:, :1]))) < 2e-5
    np.testing.assert_array_equal(got[:, 2 * D + 1], ref[:, 2 * D + 1])            # frame counts
    np.testing.assert_allclose(got[:, 2 * D + 2], ref[:, 2 * D + 2], rtol=1e-5)    # summed log-likelihood
    assert not got.reshape(55, 4, -1)[:, k:].any()                                 # untrained components untouched
    # a source that is not 16-byte aligned (row 1 of the device tensor) takes the plain-load path: same numbers as the
    # TMA path on the same frames
    import torch
    Xd, labd = torch.from_numpy(X).cuda(), torch.from_numpy(lab).cuda()
    a = eng.pool_em_stats(Xd[1:], labd[1:], pm, k).cpu().numpy()
    b = eng.pool_em_stats(Xd[1:].clone(), labd[1:].clone(), pm, k).cpu().numpy()
    assert Xd[1:].data_ptr() % 16 != 0 and np.array_equal(a, b)


def test_pool_stats_even_dimension(eng):
    """D = 24 (even: padded rows, no bulk copy) and k = 8 against srb_gmm_stats."""
    rng = np.random.default_rng(4)
    G, M, D, N = 7, 8, 24, 5000
    X = rng.normal(size=(N, D)).astype(np.float32)
    lab = np.sort(rng.integers(-1, G, N)).astype(np.int32)
    pm = eng.PackedModel(rng.normal(size=(G, M, D)), rng.uniform(0.5, 2.0, size=(G, M, D)), np.full((G, M), 1.0 / M))
    ref = eng.gmm_stats(X, lab, pm, 8).cpu().numpy()
    got = eng.pool_em_stats(X, lab, pm, 8).cpu().numpy()
    scale = np.abs(ref[:, 1 + D:1 + 2 * D]).max(axis=1, keepdims=True) + 1e-30
    assert np.max(np.abs(got - ref)[:, :1 + 2 * D] / np.maximum(scale, np.abs(ref[:, :1]))) < 2e-5


@pytest.mark.parametrize("k", [2, 4])
def test_pool_kmeans_stats_equals_assign_plus_group_stats(eng, k):
    """srb_pool_stats mode 1 == srb_kmeans_assign + srb_group_stats; assignments equal except where the two nearest
    centroids are closer than FP32 resolves (other summation order of the 39 terms)."""
    models, X, lab = _pool_case()
    rng = np.random.default_rng(9)
    G, D = 55, 39
    mu = np.concatenate([m[0] for m in models])[:, :k]                              # [G, k, D]
    var0 = np.concatenate([m[1] for m in models])[:, 0]
    shift = (mu + rng.normal(scale=0.05, size=mu.shape)).reshape(G * k, D)
    am_ref = eng.kmeans_assign(X, lab, mu, var0)
    st, am = eng.pool_kmeans_stats(X, lab, mu, var0, shift, want_argmin=True)
    am_ref, am = am_ref.cpu().numpy(), am.cpu().numpy()
    assert np.array_equal(am < 0, lab < 0) and np.array_equal(am_ref < 0, lab < 0)
    diff = np.flatnonzero(am != am_ref)
    assert len(diff) <= 2, len(diff)
    for n in diff:                                                                  # only near-ties may differ
        d = (0.5 * (X[n].astype(np.float64) - mu[lab[n]]) ** 2 / var0[lab[n]]).sum(axis=1)
        assert abs(d[am[n]] - d[am_ref[n]]) <= 1e-5 * d.min()
    ref = eng.group_stats(X, np.where(lab >= 0, lab * k + am, -1).astype(np.int32), G * k, shift).cpu().numpy()
    got = st.cpu().numpy()
    np.testing.assert_array_equal(got[:, 0], ref[:, 0])
    scale = np.abs(ref[:, 1 + D:1 + 2 * D]).max(axis=1, keepdims=True) + 1e-30
    assert np.max(np.abs(got - ref)[:, :1 + 2 * D] / scale) < 2e-5
    st2 = eng.pool_kmeans_stats(X, lab, mu, var0, shift).cpu().numpy()
    assert np.array_equal(got, st2)                                                 # the K-means stop test needs this


def test_pool_stats_rejects_tables_beyond_shared_memory(eng):
    from srb200 import lib as _l
    assert not eng.pool_stats_supported(400, 8, 39)
    assert not eng.pool_stats_supported(55, 3, 39)
    models, X, lab = _pool_case(B=20)
    big = np.zeros((400, 8, 39))
    with pytest.raises(_l.SrbError):
        eng.pool_kmeans_stats(X, lab, big, np.ones((400, 39)), big.reshape(-1, 39))


def test_fit_update_kernels_equal_numpy_semantics(eng):
    """srb_fit_em_update / srb_fit_kmeans_update against the float64 torch-on-CPU statement of the same updates
    (tests/engine_double.py): bit-identical parameters, same stop decisions, frozen states untouched."""
    import torch
    import engine_double as ed
    rng = np.random.default_rng(12)
    G, M, k, D = 9, 4, 2, 39
    row = 2 * D + 3
    mu, var = rng.normal(size=(G, M, D)), rng.uniform(0.5, 2.0, size=(G, M, D))
    w = np.full((G, M), 0.25)
    st = np.zeros((G, M, row))
    st[:, :k, 0] = rng.uniform(50, 500, size=(G, k))
    st[:, :k, 1:1 + D] = rng.normal(scale=1e-3, size=(G, k, D)) * st[:, :k, :1]
    st[:, :k, 1 + D:1 + 2 * D] = rng.uniform(0.5, 2.0, size=(G, k, D)) * st[:, :k, :1]
    st[:, 0, 2 * D + 1] = st[:, :k, 0].sum(axis=1)
    st[3, 1, 0] = 0.0                                                   # an empty component stays
    st[5, :k, 1:1 + D] = 0.0                                            # state 5: nothing moves -> converges when old == new
    active = np.ones(G, bool)
    active[7] = False                                                   # frozen before
    pm_d, pm_h = eng.PackedModel(mu, var, w), ed.PackedModel(mu, var, w)
    old_d = [pm_d.mu.clone(), pm_d.var.clone(), pm_d.w.clone()]
    old_h = [pm_h.mu.clone(), pm_h.var.clone(), pm_h.w.clone()]
    var5 = st[5, :k, 1 + D:1 + 2 * D] / st[5, :k, :1]                   # make state 5's old variances/weights the new ones
    for o in (old_d, old_h):
        o[1][5, :k] = torch.from_numpy(var5).to(o[1].device)
        o[2][5, :k] = torch.from_numpy(st[5, :k, 0] / st[5, 0, 2 * D + 1]).to(o[2].device)
    a_d, a_h = torch.from_numpy(active.copy()).cuda(), torch.from_numpy(active.copy())
    its_d, its_h = torch.zeros(G, dtype=torch.int64).cuda(), torch.zeros(G, dtype=torch.int64)
    lc_d, lc_h = torch.zeros(G * M, dtype=torch.float64).cuda(), torch.zeros(G * M, dtype=torch.float64)
    st_t = torch.from_numpy(st.reshape(G * M, row))
    eng.fit_em_update(st_t.cuda(), pm_d, k, old_d, a_d, its_d, 17, lc_d)
    ed.fit_em_update(st_t, pm_h, k, old_h, a_h, its_h, 17, lc_h)
    for x, y in zip([pm_d.mu, pm_d.var, pm_d.w] + old_d, [pm_h.mu, pm_h.var, pm_h.w] + old_h):
        assert np.array_equal(x.cpu().numpy(), y.numpy())
    assert np.array_equal(a_d.cpu().numpy(), a_h.numpy()) and not a_h[5] and not a_h[7] and a_h[0]
    assert np.array_equal(its_d.cpu().numpy(), its_h.numpy()) and its_h[7] == 0 and its_h[5] == 17
    np.testing.assert_allclose(lc_d.cpu().numpy(), lc_h.numpy(), rtol=1e-13)
    # K-means update
    cen = rng.normal(size=(G, k, D))
    shift = cen + rng.normal(scale=0.01, size=cen.shape)
    sk = np.zeros((G * k, row))
    sk[:, 0] = rng.integers(10, 90, G * k)
    sk[:, 1:1 + D] = rng.normal(size=(G * k, D))
    nc = shift.reshape(G * k, D) + sk[:, 1:1 + D] / sk[:, :1]
    cen[2] = nc.reshape(G, k, D)[2]                                     # pool 2 is at its fixed point
    outs = []
    for dev, mod in (("cuda", eng), ("cpu", ed)):
        c = torch.from_numpy(cen.copy()).to(dev)
        a = torch.from_numpy(active.copy()).to(dev)
        i_ = torch.zeros(G, dtype=torch.int64, device=dev)
        n_ = torch.zeros((G, k), dtype=torch.float64, device=dev)
        mod.fit_kmeans_update(torch.from_numpy(sk).to(dev), torch.from_numpy(shift.reshape(G * k, D)).to(dev), c, a, i_, 4, n_)
        outs.append([t.cpu().numpy() for t in (c, a, i_, n_)])
    for x, y in zip(*outs):
        assert np.array_equal(x, y)
    assert not outs[0][1][2] and not outs[0][1][7] and outs[0][1][0] and outs[0][3][7].sum() == 0


def test_pool_stats_more_than_64_pools(eng):
    """G = 70 pools: zeroed slices instead of the write-once slices with a 64-bit pool mask; short runs and every run
    length / alignment a chunk can hold (labels in random runs of 1..9 frames)."""
    rng = np.random.default_rng(8)
    G, M, D, N = 70, 2, 39, 20011
    X = rng.normal(size=(N, D)).astype(np.float32)
    runs = rng.integers(1, 10, N)
    lab = np.repeat(rng.integers(-1, G, N), runs)[:N].astype(np.int32)
    pm = eng.PackedModel(rng.normal(size=(G, M, D)), rng.uniform(0.5, 2.0, size=(G, M, D)), np.full((G, M), 0.5))
    for labels in (lab, np.minimum(lab, 40).astype(np.int32)):          # second pass: 41 pools in use -> mask path of G = 64
        Gx = G if labels is lab else 64
        pmx = pm if labels is lab else eng.PackedModel(pm.mu[:64].cpu().numpy(), pm.var[:64].cpu().numpy(),
                                                       pm.w[:64].cpu().numpy())
        ref = eng.gmm_stats(X, labels, pmx, 2).cpu().numpy()
        got = eng.pool_em_stats(X, labels, pmx, 2).cpu().numpy()
        assert got.shape == (Gx * M, 2 * D + 3)
        scale = np.abs(ref[:, 1 + D:1 + 2 * D]).max(axis=1, keepdims=True) + 1e-30
        assert np.max(np.abs(got - ref)[:, :1 + 2 * D] / np.maximum(scale, np.abs(ref[:, :1]))) < 2e-5
        np.testing.assert_array_equal(got[:, 2 * D + 1], ref[:, 2 * D + 1])
        np.testing.assert_allclose(got[:, 2 * D + 2], ref[:, 2 * D + 2], rtol=1e-5, atol=1e-3)


@pytest.mark.parametrize("k", [1, 2, 4])
def test_pool_stats_f64_equals_float64_statement(eng, k):
    """srb_pool_stats_f64 against the numpy float64 statement of the same statistics (tests/engine_double.py): 1e-11 of a
    row's scale for EM, identical assignments and 1e-12 for K-means; run-to-run bit-identical."""
    import engine_double as ed
    models, X, lab = _pool_case(ragged=(37, 81))
    G, D = (55, 39) if k < 4 else (25, 39)                 # the double tables of 55 x 4 x 39 exceed shared memory
    assert not eng.pool_stats_supported_f64(55, 4, 39)
    models, lab = models[:G // 5], np.where(lab < G, lab, -1).astype(np.int32)
    pm, _ = eng.PackedModel.from_words(models)
    pm_h = ed.PackedModel(pm.mu.cpu().numpy(), pm.var.cpu().numpy(), pm.w.cpu().numpy())
    assert eng.pool_stats_supported_f64(G, k, D)
    got = eng.pool_em_stats(X, lab, pm, k, f64=True).cpu().numpy()
    assert np.array_equal(got, eng.pool_em_stats(X, lab, pm, k, f64=True).cpu().numpy())
    ref = ed.gmm_stats(X, lab, pm_h, k).numpy()
    scale = np.abs(ref[:, 1 + D:1 + 2 * D]).max(axis=1, keepdims=True) + 1e-30
    assert np.max(np.abs(got - ref)[:, :1 + 2 * D] / np.maximum(scale, np.abs(ref[:, :1]))) < 1e-11
    np.testing.assert_array_equal(got[:, 2 * D + 1], ref[:, 2 * D + 1])
    np.testing.assert_allclose(got[:, 2 * D + 2], ref[:, 2 * D + 2], rtol=1e-11)
    if k > 1:
        rng = np.random.default_rng(9)
        mu = pm.mu.cpu().numpy()[:, :k].copy()
        var0 = pm.var.cpu().numpy()[:, 0].copy()
        shift = (mu + rng.normal(scale=0.05, size=mu.shape)).reshape(G * k, D)
        st, am = eng.pool_kmeans_stats(X, lab, mu, var0, shift, want_argmin=True, f64=True)
        st_h, am_h = ed.pool_kmeans_stats(X, lab, mu, var0, shift, want_argmin=True)
        assert np.array_equal(am.cpu().numpy(), am_h.numpy())
        st, st_h = st.cpu().numpy(), st_h.numpy()
        np.testing.assert_array_equal(st[:, 0], st_h[:, 0])
        scale = np.abs(st_h[:, 1 + D:1 + 2 * D]).max(axis=1, keepdims=True) + 1e-30
        assert np.max(np.abs(st - st_h)[:, :1 + 2 * D] / scale) < 1e-12
