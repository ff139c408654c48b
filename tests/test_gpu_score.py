"""GPU parity: K1 scoring and the fused classifier against the float64 oracle (through the C ABI)."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

RTOL = 1e-4          # north_star: log-likelihoods within 1e-4 relative (FP32)


@pytest.fixture(scope="module")
def eng():
    from srb200 import engine
    return engine


def _models(seed=1, **kw):
    from srb200 import synth
    return synth.make_models(seed, **kw)


def test_gmm_score_matches_oracle(eng):
    from srb200 import synth
    models, base = _models(3, W=11, S=5, M=4)
    X, off, _ = synth.make_utterances(4, base, 13, T=97)          # N = 1261: ragged last tile
    pm, _ = eng.PackedModel.from_words(models)
    out = eng.gmm_score(X, pm, want_E=True, want_best=True, want_comp=True)
    E = out["E"].cpu().numpy()
    comp = out["comp"].cpu().numpy()
    emit = np.arange(55)
    Eo = orc.emissions(X.astype(np.float64), models, emit, log_domain=True)
    np.testing.assert_allclose(E, Eo, rtol=RTOL)
    # per-Gaussian log(w N(x)) of state (word 2, state 3)
    mu, var, w, _ = models[2]
    _, co = orc.gmm_evaluate(X.astype(np.float64), mu[3], var[3], w[3], log_domain=True)
    np.testing.assert_allclose(comp[:, 13, :4], co, rtol=RTOL, atol=1e-3)
    bs = out["best_state"].cpu().numpy()
    bc = out["best_cost"].cpu().numpy()
    assert np.array_equal(bc, E.min(axis=1)) and np.array_equal(bs, E.argmin(axis=1))


def test_gmm_score_reference_linear_domain_and_unnormalised_weights(eng, golden):
    """Against the frozen outputs of the reference's own GMM.evaluate (hmm_state.py:114-120)."""
    g = golden["emission"]
    X = g["em_X"]
    pm = eng.PackedModel(g["em_mu"][None], g["em_cov"][None], g["em_w"][None])     # one state, weights sum != 1
    out = eng.gmm_score(X.astype(np.float32), pm, want_comp=True)
    E = out["E"].cpu().numpy()[:, 0]
    ref = g["em_cost"]
    f = np.isfinite(ref)
    assert f.any()
    np.testing.assert_allclose(E[f], ref[f], rtol=RTOL)
    comps = g["em_comps"]                                         # linear-domain w_m N_m(x)
    ok = comps > 1e-300
    np.testing.assert_allclose(out["comp"].cpu().numpy()[:, 0, :4][ok], np.log(comps[ok]), rtol=RTOL, atol=1e-3)


@pytest.mark.parametrize("W,S,B,T", [(11, 5, 13, 97), (3, 7, 300, 41), (1, 1, 1, 5), (57, 1, 9, 33), (11, 5, 700, 100)])
def test_gmm_score_tensor_core_matches_oracle(eng, W, S, B, T):
    """srb_gmm_score_tc (the tcgen05 classifier's scorer with the cost rows streamed out) against the oracle and the
    FFMA scorer: ragged last chunk, one frame per lane quarter, several CTAs, a strided E."""
    import torch
    from srb200 import synth
    models, base = _models(11, W=W, S=S, M=4)
    X, off, _ = synth.make_utterances(12, base, B, T=T)
    pm, _ = eng.PackedModel.from_words(models)
    assert pm.scorer_tensor_core_ok()
    E = eng.gmm_score(X, pm, tensor=True)["E"].cpu().numpy()
    Eo = orc.emissions(X.astype(np.float64), models, np.arange(W * S), log_domain=True)
    np.testing.assert_allclose(E, Eo, rtol=RTOL)
    Ef = eng.gmm_score(X, pm, tensor=False)["E"].cpu().numpy()
    np.testing.assert_allclose(E, Ef, rtol=2e-5)
    # strided output: only the R leading columns of each row are written
    wide = torch.full((X.shape[0], W * S + 3), -7.0, dtype=torch.float32, device="cuda")
    eng.gmm_score_tc(X, pm, wide[:, :W * S])
    wide = wide.cpu().numpy()
    assert np.array_equal(wide[:, :W * S], E) and np.all(wide[:, W * S:] == -7.0)
    if X.shape[0] >= eng.TC_SCORE_MIN_FRAMES:     # the default route picks the tensor-core scorer here
        assert np.array_equal(eng.gmm_score(X, pm)["E"].cpu().numpy(), E)
    with pytest.raises(ValueError):
        eng.gmm_score(X, pm, want_best=True, tensor=True)


@pytest.mark.parametrize("W,S,M,B,T", [(3, 4, 8, 40, 37), (1, 8, 16, 25, 90), (2, 7, 16, 9, 33), (4, 3, 12, 300, 41), (2, 5, 6, 20, 50)])
def test_gmm_score_tensor_core_sub_states(eng, W, S, M, B, T):
    """States of 4*J Gaussians on the tcgen05 scorer (J sub-states of 4 + srb_neg_lse_groups) against the oracle and
    the FFMA scorer."""
    from srb200 import synth
    models, base = _models(21, W=W, S=S, M=M)
    X, off, _ = synth.make_utterances(22, base, B, T=T)
    pm, _ = eng.PackedModel.from_words(models)
    assert pm.Mp == (M + 3) // 4 * 4 and pm.scorer_tensor_core_ok()
    E = eng.gmm_score(X, pm, tensor=True)["E"].cpu().numpy()
    Eo = orc.emissions(X.astype(np.float64), models, np.arange(W * S), log_domain=True)
    np.testing.assert_allclose(E, Eo, rtol=RTOL)
    np.testing.assert_allclose(E, eng.gmm_score(X, pm, tensor=False)["E"].cpu().numpy(), rtol=2e-5)
    old, eng.TC_MERGE_ROWS = eng.TC_MERGE_ROWS, 64             # several passes over the scratch
    try:
        assert np.array_equal(eng.gmm_score(X, pm, tensor=True)["E"].cpu().numpy(), E)
    finally:
        eng.TC_MERGE_ROWS = old


@pytest.mark.parametrize("R,M,N", [(70, 8, 3000), (20, 32, 1500), (5, 4, 700), (130, 16, 2100)])
def test_gmm_best_tensor_core_blocks(eng, R, M, N):
    """engine.gmm_best_tc (blocks of 57 // J states on the tcgen05 scorer + srb_merge_best) against the FFMA scorer's
    streaming best state: costs to FP32 noise, states equal wherever the two best costs differ by more than 1e-4."""
    import torch
    rng = np.random.default_rng(R + M)
    mu = rng.normal(size=(R, 1, 39)) + 0.5 * rng.normal(size=(R, M, 39))
    var = rng.uniform(0.5, 1.5, size=(R, M, 39))
    w = rng.dirichlet(np.ones(M), size=R)
    pm = eng.PackedModel(mu, var, w)
    X = torch.randn((N, 39), dtype=torch.float32, device="cuda", generator=torch.Generator("cuda").manual_seed(5))
    ref = eng.gmm_score(X, pm, want_E=True, want_best=True, tensor=False)
    E = ref["E"].cpu().numpy()
    old, eng.TC_BEST_ROWS = eng.TC_BEST_ROWS, 1024                 # several frame passes
    try:
        bc, bs = eng.gmm_best_tc(X, pm)
    finally:
        eng.TC_BEST_ROWS = old
    bc, bs = bc.cpu().numpy(), bs.cpu().numpy()
    np.testing.assert_allclose(bc, ref["best_cost"].cpu().numpy(), rtol=2e-5)
    so = np.sort(E, axis=1)
    decided = (so[:, 1] - so[:, 0]) > 1e-4 * np.abs(so[:, 0]) if R > 1 else np.ones(N, bool)
    assert decided.mean() > 0.9
    assert np.array_equal(bs[decided], ref["best_state"].cpu().numpy()[decided])
    np.testing.assert_allclose(E[np.arange(N), bs], bc, rtol=2e-5)   # the state named carries the cost named
    bc2, bs2 = eng.gmm_best_tc(X, pm)                                # cached block packing, one frame pass
    assert np.array_equal(bc2.cpu().numpy(), bc) and np.array_equal(bs2.cpu().numpy(), bs)
    out = eng.gmm_score(X, pm, want_E=False, want_best=True, tensor=True)       # the public route: this path for states
    if M % 16:                                                                  # of 4, 8, 12 Gaussians, else the GEMM kernel
        assert np.array_equal(out["best_cost"].cpu().numpy(), bc) and np.array_equal(out["best_state"].cpu().numpy(), bs)
    else:
        np.testing.assert_allclose(out["best_cost"].cpu().numpy(), bc, rtol=2e-5)
        assert np.array_equal(out["best_state"].cpu().numpy()[decided], bs[decided])


@pytest.mark.parametrize("R,M,N", [(20, 32, 1500), (130, 16, 2100), (1, 16, 130), (7, 48, 300), (2, 64, 129), (3, 160, 64),
                                   (2000, 32, 777), (401, 16, 20000)])
def test_gmm_best_gemm(eng, R, M, N):
    """engine.gmm_best_gemm (srb_gmm_best_gemm: one persistent tcgen05 GEMM, the frame tile's operand resident in tensor
    memory, the model streamed past it in 160-column blocks by TMA bulk copies, running best state in registers) against
    the float64 oracle's GMM.evaluate (hmm_state.py:114-120, log domain) and against the FFMA scorer: costs to 1e-4
    (FP32 rule), states equal wherever the two best costs differ by more than that.  Shapes: one block / many blocks,
    partial last block, a state per block (M = 160), partial last frame tile, more tiles than SMs (several tiles per
    CTA), and BASELINE config 5's 2 000 x 32."""
    import torch
    rng = np.random.default_rng(R + M)
    mu = rng.normal(size=(R, 1, 39)) + 0.5 * rng.normal(size=(R, M, 39))
    var = rng.uniform(0.5, 1.5, size=(R, M, 39))
    w = rng.dirichlet(np.ones(M), size=R)
    pm = eng.PackedModel(mu, var, w)
    X = torch.randn((N, 39), dtype=torch.float32, device="cuda", generator=torch.Generator("cuda").manual_seed(5))
    bc, bs = eng.gmm_best_gemm(X, pm)
    bc, bs = bc.cpu().numpy(), bs.cpu().numpy()
    n_or = min(N, 200)                                     # oracle sample (the scalar C restatement: R*M*39 per frame)
    idx = np.sort(rng.choice(N, size=n_or, replace=False))
    Xo = X.cpu().numpy()[idx].astype(np.float64)
    Eo = np.stack([orc.gmm_evaluate(Xo, mu[r], var[r], w[r], log_domain=True)[0] for r in range(R)], axis=1)
    np.testing.assert_allclose(bc[idx], Eo.min(axis=1), rtol=RTOL)
    so = np.sort(Eo, axis=1)
    decided = (so[:, 1] - so[:, 0]) > RTOL * np.abs(so[:, 0]) if R > 1 else np.ones(n_or, bool)
    assert decided.mean() > 0.9
    assert np.array_equal(bs[idx][decided], Eo.argmin(axis=1)[decided])
    np.testing.assert_allclose(Eo[np.arange(n_or), bs[idx]], bc[idx], rtol=RTOL)   # the state named carries the cost named
    if R * M <= 8192:
        ref = eng.gmm_score(X, pm, want_E=True, want_best=True, tensor=False)       # FFMA scorer, every frame
        E = ref["E"].cpu().numpy()
        np.testing.assert_allclose(bc, ref["best_cost"].cpu().numpy(), rtol=2e-5)
        s2 = np.sort(E, axis=1)
        dec = (s2[:, 1] - s2[:, 0]) > 1e-4 * np.abs(s2[:, 0]) if R > 1 else np.ones(N, bool)
        assert np.array_equal(bs[dec], ref["best_state"].cpu().numpy()[dec])
    bc2, bs2 = eng.gmm_best_gemm(X, pm)                    # cached blocks; bit-reproducible
    assert np.array_equal(bc2.cpu().numpy(), bc) and np.array_equal(bs2.cpu().numpy(), bs)
    out = eng.gmm_score(X, pm, want_E=False, want_best=True, tensor=True)           # the public route to the same path
    assert np.array_equal(out["best_cost"].cpu().numpy(), bc) and np.array_equal(out["best_state"].cpu().numpy(), bs)


@pytest.mark.parametrize("W,S,M", [(3, 8, 16), (11, 5, 32), (2, 3, 48), (4, 10, 16), (1, 1, 160)])
def test_gmm_score_gemm_rows_mode(eng, W, S, M):
    """srb_gmm_score_gemm: every word's frames against that word's states in ONE launch of the persistent GEMM kernel
    (a word = one 160-column block, tiles of <= 128 frames that never straddle words), against the oracle's
    GMM.evaluate and the FFMA scorer.  Ragged word blocks incl. an empty word and blocks shorter than a tile."""
    import torch
    from srb200 import lib as L
    rng = np.random.default_rng(W * 100 + S)
    R = W * S
    mu = rng.normal(size=(R, 1, 39)) + 0.5 * rng.normal(size=(R, M, 39))
    var = rng.uniform(0.5, 1.5, size=(R, M, 39))
    w = rng.dirichlet(np.ones(M), size=R)
    pm = eng.PackedModel(mu, var, w)
    rows = [int(v) for v in rng.integers(1, 700, size=W)]
    if W > 2:
        rows[1] = 0                                                       # a word without utterances
        rows[-1] = 37
    off = np.concatenate([[0], np.cumsum(rows)])
    N = int(off[-1])
    X = torch.randn((N, 39), dtype=torch.float32, device="cuda", generator=torch.Generator("cuda").manual_seed(11))
    lib = L.load()
    blocks = torch.empty(int(lib.srb_model_blocks_floats_sb(R, pm.Mp, 39, S)), dtype=torch.float32, device="cuda")
    L.check(lib.srb_pack_model_blocks_sb(L.ptr(pm.ab), L.ptr(pm.c), R, pm.Mp, 39, S, L.ptr(blocks), L.stream_ptr()))
    descs = []
    for wd in range(W):
        for f0 in range(int(off[wd]), int(off[wd + 1]), 128):
            descs.append([f0, min(128, int(off[wd + 1]) - f0), wd, 0])
    tiles = torch.tensor(descs, dtype=torch.int32, device="cuda")
    E = torch.full((N, S), -1.0, dtype=torch.float32, device="cuda")
    L.check(lib.srb_gmm_score_gemm(L.ptr(X), N, 39, L.ptr(blocks), L.ptr(pm.shift), R, pm.Mp, S, L.ptr(tiles), len(descs),
                                   L.ptr(E), S, L.stream_ptr()))
    E = E.cpu().numpy()
    ref = eng.gmm_score(X, pm, want_E=True, tensor=False)["E"].cpu().numpy()          # all states, FFMA
    Xh = X.cpu().numpy().astype(np.float64)
    for wd in range(W):
        a, b = int(off[wd]), int(off[wd + 1])
        if b == a:
            continue
        np.testing.assert_allclose(E[a:b], ref[a:b, wd * S:(wd + 1) * S], rtol=2e-5)
        k = min(b - a, 40)
        for s_ in range(S):
            co = orc.gmm_evaluate(Xh[a:a + k], mu[wd * S + s_], var[wd * S + s_], w[wd * S + s_], log_domain=True)[0]
            np.testing.assert_allclose(E[a:a + k, s_], co, rtol=RTOL)


def test_gmm_best_gemm_zero_weight_and_unsupported(eng):
    """Components of weight 0 (C = -inf) contribute nothing; a state whose components all have weight 0 never wins;
    shapes the kernel is not built for are refused loudly."""
    import torch
    rng = np.random.default_rng(3)
    R, M = 6, 16
    mu = rng.normal(size=(R, M, 39))
    var = rng.uniform(0.5, 1.5, size=(R, M, 39))
    w = rng.dirichlet(np.ones(M), size=R)
    w[:, 5:] = 0.0
    w[2, :] = 0.0
    pm = eng.PackedModel(mu, var, w)
    X = torch.randn((300, 39), dtype=torch.float32, device="cuda", generator=torch.Generator("cuda").manual_seed(6))
    bc, bs = eng.gmm_best_gemm(X, pm)
    ref = eng.gmm_score(X, pm, want_E=True, want_best=True, tensor=False)
    np.testing.assert_allclose(bc.cpu().numpy(), ref["best_cost"].cpu().numpy(), rtol=2e-5)
    assert (bs.cpu().numpy() != 2).all()
    assert np.isinf(ref["E"].cpu().numpy()[:, 2]).all()
    small = eng.PackedModel(mu[:, :4], var[:, :4], np.full((R, 4), 0.25))
    with pytest.raises(ValueError):
        eng.gmm_best_gemm(X, small)


def test_gmm_score_tensor_core_unsupported_is_loud(eng):
    import torch
    from srb200 import lib as L
    models, base = _models(5, W=3, S=4, M=8)                       # the C entry takes 4 Gaussians per state only (the
    pm, _ = eng.PackedModel.from_words(models)                     # engine scores 8 as 2 sub-states of 4, see below)
    assert pm.scorer_tensor_core_ok()
    big, _ = _models(5, W=6, S=5, M=8)                             # 240 Gaussians > 228 tensor-core columns
    assert not eng.PackedModel.from_words(big)[0].scorer_tensor_core_ok()
    assert L.load().srb_gmm_score_tc_supported(39, 58, 4) == 0     # packed model + cost ring > 227 KB of shared memory
    assert L.load().srb_gmm_score_tc_supported(39, 57, 4) == 1
    X = torch.zeros((64, 39), device="cuda")
    co = torch.tensor([0, 32, 64], dtype=torch.int32, device="cuda")
    E = torch.empty((64, 12), device="cuda")
    rc = L.load().srb_gmm_score_tc(X.data_ptr(), 64, 39, pm.ab.data_ptr(), pm.shift.data_ptr(), 12, 8, co.data_ptr(), 2,
                                   E.data_ptr(), 12, None)
    assert rc != 0 and b"not covered" in L.load().srb_last_error_string()


@pytest.mark.parametrize("M,S,W", [(4, 5, 11), (16, 8, 3), (2, 3, 4)])
def test_gmm_score_multi_block_states(eng, M, S, W):
    from srb200 import synth
    models, base = _models(5, W=W, S=S, M=M)
    X, off, _ = synth.make_utterances(6, base, 7, T=50)
    pm, _ = eng.PackedModel.from_words(models)
    E = eng.gmm_score(X, pm)["E"].cpu().numpy()
    Eo = orc.emissions(X.astype(np.float64), models, np.arange(W * S), log_domain=True)
    np.testing.assert_allclose(E, Eo, rtol=RTOL)


def test_gmm_score_param_groups(eng):
    """More Gaussians than fit shared memory at once: the kernel sweeps parameter groups."""
    from srb200 import synth
    models, base = _models(7, W=40, S=5, M=8)                     # 1600 Gaussians ~ 500 KB of parameters
    X, off, _ = synth.make_utterances(8, base, 6, T=100)
    pm, _ = eng.PackedModel.from_words(models)
    out = eng.gmm_score(X, pm, want_best=True)
    E = out["E"].cpu().numpy()
    Eo = orc.emissions(X.astype(np.float64), models, np.arange(200), log_domain=True)
    np.testing.assert_allclose(E, Eo, rtol=RTOL)
    assert np.array_equal(out["best_state"].cpu().numpy(), E.argmin(axis=1))


@pytest.mark.parametrize("tensor", [False, True], ids=["ffma", "tcgen05"])
@pytest.mark.parametrize("ragged", [None, (2, 40), (300, 1300)])
def test_classify_matches_oracle(eng, ragged, tensor):
    from srb200 import synth
    models, base = _models(11, W=11, S=5, M=4)
    X, off, words = synth.make_utterances(12, base, 64 if ragged != (300, 1300) else 9, T=100, ragged=ragged)
    pm, trans = eng.PackedModel.from_words(models)
    assert pm.tensor_core_ok(11, 5)
    final, label = eng.classify(X, off, pm, trans, 11, 5, tensor=tensor)
    final, label = final.cpu().numpy(), label.cpu().numpy()
    fo, lo = orc.classify(X.astype(np.float64), off.astype(np.int64), models, log_domain=True)
    fin = np.isfinite(fo)
    assert np.array_equal(np.isfinite(final), fin)
    np.testing.assert_allclose(final[fin], fo[fin], rtol=RTOL)
    # labels exact wherever the oracle's decision margin exceeds the tolerance
    so = np.sort(fo, axis=1)
    decided = (so[:, 1] - so[:, 0]) > RTOL * np.abs(so[:, 0])
    assert decided.mean() > 0.9
    assert np.array_equal(label[decided], lo[decided])
    if ragged is None:
        assert (label == words).mean() > 0.95


@pytest.mark.parametrize("tensor", [False, True], ids=["ffma", "tcgen05"])
def test_classify_golden_reference(eng, golden, tensor):
    """Frozen HMM.evaluate outputs of the reference (hmm.py:126-135) for 3 isolated models."""
    g = golden["decode"]
    models = [(g["iso_mu"][i], g["iso_cov"][i], g["iso_w"][i], g["iso_trans"][i]) for i in range(len(g["iso_mu"]))]
    pm, trans = eng.PackedModel.from_words(models)
    W, S = len(models), models[0][0].shape[0]
    if tensor and not pm.tensor_core_ok(W, S):
        pytest.skip("golden model outside the tensor-core classifier's shapes")
    final, label = eng.classify(g["iso_X"].astype(np.float32), g["iso_off"].astype(np.int32), pm, trans, W, S,
                                tensor=tensor)
    ref = g["iso_final"]                                          # [B, W]
    f = np.isfinite(ref)
    np.testing.assert_allclose(final.cpu().numpy()[f], ref[f], rtol=RTOL)


@pytest.mark.parametrize("tensor", [False, True], ids=["ffma", "tcgen05"])
def test_classify_dense_transitions(eng, tensor):
    """Non left-to-right (ergodic) transitions take the dense DP branch."""
    from srb200 import synth
    rng = np.random.default_rng(5)
    models, base = _models(21, W=4, S=4, M=4)
    models = [(m[0], m[1], m[2], rng.uniform(0.5, 3.0, size=(4, 4))) for m in models]
    X, off, _ = synth.make_utterances(22, base, 10, T=30)
    pm, trans = eng.PackedModel.from_words(models)
    final, _ = eng.classify(X, off, pm, trans, 4, 4, tensor=tensor)
    fo, _ = orc.classify(X.astype(np.float64), off.astype(np.int64), models, log_domain=True)
    np.testing.assert_allclose(final.cpu().numpy(), fo, rtol=RTOL)


@pytest.mark.parametrize("W,S,ragged,n", [(11, 5, (2, 300), 6000), (10, 5, (90, 110), 3000), (3, 8, (5, 700), 800),
                                          (1, 2, (2, 9), 50)])
def test_classify_tensor_core_equals_ffma(eng, W, S, ragged, n):
    """srb_classify_tc (tcgen05 3xTF32, TMEM accumulators) against srb_classify (FFMA2) on corpora that fill the whole
    grid: costs within FP32 accumulation noise, labels equal wherever the two best words are further apart than
    that; utterances straddle tiles, chunks and CTA ranges."""
    from srb200 import synth
    models, base = _models(31 + W, W=W, S=S, M=4)
    X, off, words = synth.make_utterances(32, base, n, ragged=ragged, noise=0.9)
    pm, trans = eng.PackedModel.from_words(models)
    assert pm.tensor_core_ok(W, S)
    fa, la = eng.classify(X, off, pm, trans, W, S, tensor=False)
    fb, lb = eng.classify(X, off, pm, trans, W, S, tensor=True)
    fa, la, fb, lb = fa.cpu().numpy(), la.cpu().numpy(), fb.cpu().numpy(), lb.cpu().numpy()
    assert np.array_equal(np.isfinite(fa), np.isfinite(fb))
    fin = np.isfinite(fa)
    np.testing.assert_allclose(fb[fin], fa[fin], rtol=2e-5)
    so = np.sort(np.where(fin, fa, np.inf), axis=1)
    if W > 1:
        decided = ~(so[:, 1] - so[:, 0] <= 1e-4 * np.abs(so[:, 0]))
        assert decided.mean() > 0.9
        assert np.array_equal(la[decided], lb[decided])
    else:
        assert np.array_equal(la, lb)
