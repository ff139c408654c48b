"""CPU float64 oracle -- TEST INFRASTRUCTURE ONLY.

Restates, function by function, the reference hot path (`/root/reference/sr/recognition`, cited as
file:line below).  The inner loops live in `sr_oracle.c` (compiled to `oracle/_build/libsroracle.so`);
this module holds the ctypes bindings and the orchestration-level restatements (skmeans / kmeans / EM
loops, HMM.fit, continuous_train) on plain arrays.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may import
this module.  The product (`speech-recognition_b200/`) never does.

Parity status: PINNED by tests/test_oracle_golden.py against tests/golden/ref_*.npz (outputs of the
unmodified reference, made by tests/golden/make_golden.py).  `forward_backward` is the exception: the
reference has no such function ("parity unpinned"); it is tied to `decode_hmm_states` by identities.

Model layout used throughout: a word model is (mu[S,M,D], cov[S,M,D], w[S,M], trans[S,S]) float64.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libsroracle.so")
_SRC = os.path.join(_HERE, "sr_oracle.c")
_lib = None

c_dp = ctypes.POINTER(ctypes.c_double)
c_ip = ctypes.POINTER(ctypes.c_int64)
c_bp = ctypes.POINTER(ctypes.c_uint8)


def build(force=False):
    """gcc the C restatement (building the checker is not using it)."""
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-o", _SO, _SRC, "-lm"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.orc_mahalanobis.restype = ctypes.c_double
        _lib.orc_forward_backward.restype = ctypes.c_double
        _lib.orc_decode.restype = ctypes.c_int64
        _lib.orc_dtw.restype = ctypes.c_int64
    return _lib


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(c_dp)


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int64)
    return a, a.ctypes.data_as(c_ip)


def num_threads():
    return int(lib().orc_num_threads())


def set_num_threads(n=None):
    """Use n (default: all host cores) OpenMP threads regardless of OMP_NUM_THREADS (torchrun sets it to 1)."""
    import os
    lib().orc_set_num_threads(int(n or os.cpu_count() or 1))
    return num_threads()


# ------------------------------------------------------------------------------------ emission (a2,a3,a6)
def gmm_evaluate(X, mu, cov, w, log_domain=False):
    """hmm_state.py:114-120 for a batch of frames -> (cost[N], comps[N,M]).

    comps = w_m*pdf_m(x) (or its log when log_domain)."""
    X, Xp = _d(np.atleast_2d(X))
    mu, mup = _d(mu)
    cov, covp = _d(cov)
    w, wp = _d(w)
    N, D = X.shape
    M = mu.shape[0]
    assert M <= 1024, "orc_gmm_evaluate keeps the component log-densities of a frame in a 1024-entry array"
    cost = np.empty(N)
    comps = np.empty((N, M))
    lib().orc_gmm_evaluate(Xp, ctypes.c_int64(N), D, mup, covp, wp, M, int(log_domain), cost.ctypes.data_as(c_dp),
                           comps.ctypes.data_as(c_dp))
    return cost, comps


def mahalanobis(v1, v2, variance):
    """hmm_state.py:48-58"""
    v1, p1 = _d(v1)
    v2, p2 = _d(v2)
    var, p3 = _d(variance)
    return float(lib().orc_mahalanobis(p1, p2, p3, len(var)))


def emissions(X, word_models, emit, log_domain=False):
    """E[T,R]: cost of flat state r at frame t.  emit[r] = flat (word*S+s) index, or -1 for a NES
    (hmm_state.py:90-91 -> 0)."""
    X = np.atleast_2d(np.asarray(X, np.float64))
    T = X.shape[0]
    S = word_models[0][0].shape[0]
    emit = np.asarray(emit)
    E = np.zeros((T, len(emit)))
    cache = {}
    for r, e in enumerate(emit):
        if e < 0:
            continue
        if e not in cache:
            mu, cov, w, _ = word_models[e // S]
            cache[e] = gmm_evaluate(X, mu[e % S], cov[e % S], w[e % S], log_domain)[0]
        E[:, r] = cache[e]
    return E


# ------------------------------------------------------------------------------------ DP (a9,a10)
def decode_hmm_states(E, trans, is_nes=None, end_points=None):
    """decode.py:80-146 on a precomputed emission matrix E[T,R] -> (costs[R,T], path[L,2], best_end)."""
    E, Ep = _d(E)
    T, R = E.shape
    trans, tp = _d(trans)
    nes = np.zeros(R, np.uint8) if is_nes is None else np.ascontiguousarray(is_nes, dtype=np.uint8)
    if end_points is None:
        end_points = [[R - 1, T - 1]]                      # decode.py:126-127
    ends, ep = _i(np.asarray(end_points).reshape(-1, 2))
    costs = np.empty((R, T))
    pm = np.empty((R, T, 2), np.int64)
    max_path = 4 * (T + R) + 8
    path = np.empty((max_path, 2), np.int64)
    best = np.zeros(2, np.int64)
    n = lib().orc_decode(Ep, ctypes.c_int64(T), R, tp, nes.ctypes.data_as(c_bp), ep, len(ends),
                         costs.ctypes.data_as(c_dp), pm.ctypes.data_as(c_ip), path.ctypes.data_as(c_ip),
                         ctypes.c_int64(max_path), best.ctypes.data_as(c_ip))
    if n == -1:
        raise NameError("origin_point == [r, c]")           # decode.py:120-121
    if n < 0:
        raise RuntimeError(f"orc_decode failed ({n})")
    return costs, path[:n].copy(), best


def local_dist(X, Y, var=None):
    """dist_fun of dtw: Euclidean norm (kmeans.py:111) or mahalanobis with per-row variance (hmm.py:134)."""
    X, Xp = _d(X)
    Y, Yp = _d(Y)
    T, D = X.shape
    S = Y.shape[0]
    out = np.empty((T, S))
    if var is None:
        lib().orc_local_dist(Xp, ctypes.c_int64(T), D, Yp, None, S, 0, out.ctypes.data_as(c_dp))
    else:
        var, vp = _d(var)
        lib().orc_local_dist(Xp, ctypes.c_int64(T), D, Yp, vp, S, 1, out.ctypes.data_as(c_dp))
    return out


def dtw(X, Y, trans, var=None):
    """decode.py:7-77 (beam=inf) -> (costs[S,T], path)."""
    Dl, Dp = _d(local_dist(X, Y, var))
    T, S = Dl.shape
    assert T > 1 and S > 1                                  # decode.py:22
    trans, tp = _d(trans)
    costs = np.empty((S, T))
    path = np.empty((2 * (T + S) + 8, 2), np.int64)
    n = lib().orc_dtw(Dp, ctypes.c_int64(T), S, tp, costs.ctypes.data_as(c_dp), path.ctypes.data_as(c_ip),
                      ctypes.c_int64(len(path)))
    if n < 0:
        raise RuntimeError(f"orc_dtw failed ({n})")
    return costs, path[:n].copy()


def dtw_beam(Dl, trans, beam=np.inf):
    """decode.py:19-77 WITH the beam (pure-Python loops, small cases only) on a precomputed local-distance matrix
    Dl[T,S] -> (costs[S,T], path).  Follows the reference statement by statement: -1 marks cells excluded by the beam
    (:62-68); the first row swept in the next column skips marked origins and resets the mark to inf (:44-51)."""
    Dl = np.asarray(Dl, np.float64)
    T, S = Dl.shape
    assert T > 1 and S > 1                                                      # :22
    costs = np.full((S, T), np.inf)
    pm = np.zeros((S, T, 2), np.int64)
    for j in range(T):
        for i in range(S):
            if i == 0 and j == 0:
                costs[0, 0] = Dl[0, 0]                                          # :34-40
                continue
            prev_costs, from_pts = [], []
            for origin in range(trans.shape[1]):                               # :44-51
                if costs[origin, j - 1] == -1:
                    costs[origin, j - 1] = np.inf
                else:
                    prev_costs.append(trans[i, origin] + costs[origin, j - 1])
                    from_pts.append([origin, j - 1])
            min_i = int(np.argmin(prev_costs))                                  # :53 (ValueError when empty)
            pm[i, j] = from_pts[min_i]
            costs[i, j] = min(costs[i, j], prev_costs[min_i] + Dl[j, i])        # :56-60
        if not np.isinf(beam):                                                  # :62-68
            sort_idx = np.argsort(costs[:, j].flatten(), kind="stable")
            for i in sort_idx[beam:]:
                if not np.isinf(costs[i, j]):
                    costs[i, j] = -1
    i, j = S - 1, T - 1
    path = []
    while i != 0 or j != 0:                                                     # :71-77
        if j == 0:
            raise RuntimeError("walk back left the first column (the reference wraps to column -1 here)")
        i, j = pm[i, j]
        path.append([i, j])
    return costs, np.array(path, np.int64).reshape(-1, 2)


def forward_backward(E, trans, semiring="sum"):
    """No reference counterpart (parity unpinned).  -> (neg_log_Z, alpha[T,R], beta[T,R], gamma[T,R]).
    semiring="min" runs the SAME recursion with min in place of -log-sum-exp: alpha is then decode_hmm_states' cost
    matrix bit for bit, the identity that ties this restatement to the reference (decode.py:97-124)."""
    E, Ep = _d(E)
    T, R = E.shape
    trans, tp = _d(trans)
    a = np.empty((T, R))
    b = np.empty((T, R))
    g = np.empty((T, R))
    fn = lib().orc_forward_backward_sr
    fn.restype = ctypes.c_double
    nlz = fn(Ep, ctypes.c_int64(T), R, tp, a.ctypes.data_as(c_dp), b.ctypes.data_as(c_dp), g.ctypes.data_as(c_dp),
             int(semiring == "min"))
    return float(nlz), a, b, g


def forward_backward_bruteforce(E, trans, end_states=None):
    """Forward-backward by ENUMERATING every state sequence (independent of any recursion; T <= 10, R <= 4 or so):
    weight(path) = exp(-(sum_c E[c, s_c] + sum_c trans[s_c, s_{c-1}])), paths start in state 0 at frame 0
    (decode.py:99-100) and end in one of end_states (default: the last state, decode.py:126-127) at the last frame.
    -> (neg_log_Z, gamma[T,R], xi[R,R] summed over frames: xi[r, o] = posterior mass of the arcs o -> r)."""
    import itertools
    E = np.asarray(E, np.float64)
    T, R = E.shape
    trans = np.asarray(trans, np.float64)
    ends = [R - 1] if end_states is None else list(end_states)
    Z = 0.0
    gamma = np.zeros((T, R))
    xi = np.zeros((R, R))
    for tail in itertools.product(range(R), repeat=T - 1):
        path = (0,) + tail
        if path[-1] not in ends:
            continue
        cost = E[0, 0]
        for c in range(1, T):
            cost += trans[path[c], path[c - 1]] + E[c, path[c]]
        if not np.isfinite(cost):
            continue
        w = np.exp(-cost)
        Z += w
        for c in range(T):
            gamma[c, path[c]] += w
            if c:
                xi[path[c], path[c - 1]] += w
    return -np.log(Z), gamma / Z, xi / Z


# ------------------------------------------------------------------------------------ kmeans.py helpers
def calc_variance(data_rows):
    """kmeans.py:6-12 on [N,D] rows (callers pass seg.T to the reference)."""
    X, Xp = _d(data_rows)
    out = np.empty(X.shape[1])
    with np.errstate(all="ignore"):
        lib().orc_calc_variance(Xp, ctypes.c_int64(X.shape[0]), X.shape[1], out.ctypes.data_as(c_dp))
    return out


def segment_data(templates, seg_starts):
    """kmeans.py:33-50: pool the frames of segment s over all templates."""
    n, S = seg_starts.shape
    out = []
    for s in range(S):
        parts = []
        for r in range(n):
            lo = seg_starts[r, s]
            hi = len(templates[r]) if s == S - 1 else seg_starts[r, s + 1]
            parts.append(templates[r][lo:hi])
        out.append(np.concatenate(parts, axis=0))
    return out


def combine_templates(templates, seg_starts):
    """kmeans.py:15-30: per-segment mean and unbiased variance."""
    segs = segment_data(templates, seg_starts)
    res = np.array([s.mean(axis=0) for s in segs])
    var = np.array([calc_variance(s) for s in segs])
    return res, var


def calc_transition_costs(n_temps, seg_lens, max_jump_dist=2):
    """kmeans.py:53-95 (incl. its look-ahead at segment s+1)."""
    seg_lens = np.asarray(seg_lens)
    S = seg_lens.shape[1]
    empty = seg_lens == 0
    res = np.full((S, S), np.inf)
    for i in range(S):
        jump = 1
        n_jump = 0 if i == S - 1 else n_temps
        s = i + 1
        while s < S - 1:
            if empty[:, s + 1].sum() == 0:
                break
            jump += 1
            if jump > max_jump_dist:
                break
            s += 1
        n_all = np.float64(seg_lens[:, i].sum())
        with np.errstate(all="ignore"):
            if n_jump:
                res[i + jump, i] = -np.log(np.float64(n_jump) / n_all)
            res[i, i] = -np.log((n_all - n_jump) / n_all)
    return res


def get_segments_from_path(path, n_segments):
    """kmeans.py:98-108: occupancy histogram of path[:,0] -> start frames of segments 1..S-1."""
    counts = np.bincount(np.asarray(path)[:, 0], minlength=n_segments)
    return np.cumsum(counts)[:-1]


def skmeans(templates, n_segments, max_iteration=1000):
    """kmeans.py:111-155 -> (res, vars, transition_costs, segmented_data)."""
    assert max_iteration > 0
    n = len(templates)
    S = n_segments
    seg_lens = np.array([[len(t) // S] * S for t in templates], dtype=np.int64)          # :126-132
    seg_starts = np.concatenate([np.zeros((n, 1), np.int64), np.cumsum(seg_lens, axis=1)[:, :-1]], axis=1)
    res, var = combine_templates(templates, seg_starts)
    trans = None
    for _ in range(max_iteration):
        seg_starts = np.zeros((n, S), np.int64)
        trans = calc_transition_costs(n, seg_lens)                                          # :139 (A5)
        for r in range(n):
            if templates[r].shape[0] < 5:
                raise NameError("template is too small, cannot do dtw on it")               # :142-143
            _, path = dtw(templates[r], res, trans)
            seg_starts[r, 1:] = get_segments_from_path(path, S)
        new_res, var = combine_templates(templates, seg_starts)
        if np.allclose(res, new_res):                                                       # :148
            break
        res = new_res
    return res, var, trans, segment_data(templates, seg_starts)


def kmeans(data, k, centroids, max_iteration=1000, rng_randint=None):
    """kmeans.py:167-193 with dist_fun=mahalanobis -> (clusters, centroids, cov).

    Draws the initial partition from the GLOBAL numpy RNG like kmeans.py:171 unless `rng_randint` supplies it."""
    data, dp = _d(data)
    N, D = data.shape
    assert k == centroids.shape[0]
    part = np.random.randint(0, k, N) if rng_randint is None else np.asarray(rng_randint)
    with np.errstate(all="ignore"):
        cov = np.array([calc_variance(data[part == c]) for c in range(k)])                  # :173-177 (A3)
    centroids = np.array(centroids, dtype=np.float64)
    clusters = np.empty(N, np.int64)
    for _ in range(max(max_iteration, 1)):
        cent, cp = _d(centroids)
        cov0, c0p = _d(cov[0])                                                              # :183 cov[0] for all
        new = np.empty((k, D))
        lib().orc_kmeans_iteration(dp, ctypes.c_int64(N), D, cp, k, c0p, clusters.ctypes.data_as(c_ip),
                                   new.ctypes.data_as(c_dp))
        if np.array_equal(new, centroids):                                                  # :190
            break
        centroids = new
    return clusters.copy(), centroids, cov


class GMMState:
    """hmm_state.py:100-176 as plain arrays (one HMM state)."""

    def __init__(self, mu, sigma, n_gaussians):             # :105-112
        self.n = n_gaussians
        self.w = np.full(n_gaussians, 1 / n_gaussians)
        self.mu = np.tile(np.asarray(mu, float), (n_gaussians, 1))
        self.cov = np.tile(np.asarray(sigma, float), (n_gaussians, 1))
        self.mu_old = self.mu.copy()
        self.sigma_old = self.cov.copy()
        self.w_old = self.w.copy()

    def update_models(self, mus, sigmas, weights):          # :161-165
        k = mus.shape[0]
        self.w[:k] = weights
        self.mu[:k] = mus
        self.cov[:k] = sigmas

    def em(self, data, n, max_iteration=10000):             # :122-159
        data, dp = _d(data)
        N, D = data.shape
        iters = 0
        for it in range(max_iteration):
            mu, mp = _d(self.mu)
            cov, cp = _d(self.cov)
            w, wp = _d(self.w)
            nm = np.empty((n, D))
            ns = np.empty((n, D))
            nw = np.empty(n)
            lib().orc_em_iteration(dp, ctypes.c_int64(N), D, mp, cp, wp, self.n, n, nm.ctypes.data_as(c_dp),
                                   ns.ctypes.data_as(c_dp), nw.ctypes.data_as(c_dp))
            self.update_models(nm, ns, nw)
            iters = it + 1
            if np.allclose(nm, self.mu_old[:n]) and np.allclose(ns, self.sigma_old[:n]) and \
                    np.allclose(nw, self.w_old[:n]):        # :150-152
                break
            self.mu_old[:n] = nm
            self.sigma_old[:n] = ns
            self.w_old[:n] = nw
        return iters


def fit_gmm_state(state, data, n_gaussians, centroid0, weight_fill, weight_den, use_em=True):
    """hmm.py:97-124 (_fit_GMM) / continuous_speech.py:121-142: binary split + kmeans + EM for one state."""
    n_splits = int(np.log(n_gaussians))                     # natural log (A1)
    assert n_splits > 0
    centroids = np.asarray(centroid0, float).reshape(1, -1)
    weights = np.full(n_gaussians, weight_fill)
    for i in range(n_splits):
        centroids = np.concatenate([centroids * 0.9, centroids * 1.1], axis=0)
        k = 2 ** (i + 1)
        clusters, centroids, variance = kmeans(data, k, centroids)
        cs, cc = np.unique(clusters, return_counts=True)
        for c in cs:
            weights[c] = cc[c] / weight_den                 # (A4: counts indexed by label)
        state.update_models(centroids, variance, weights[:k])
        if use_em:
            state.em(data, k)


def align_gmm_states(templates, word_model, n_segments):
    """kmeans.py:196-205"""
    mu, cov, w, trans = word_model
    seg_starts = np.zeros((len(templates), n_segments), np.int64)
    for r, t in enumerate(templates):
        E = emissions(t, [word_model], np.arange(n_segments))
        _, path, _ = decode_hmm_states(E, trans)
        seg_starts[r, 1:] = get_segments_from_path(path, n_segments)
    return segment_data(templates, seg_starts)


def hmm_fit(ys, n_segments, n_gaussians, use_em=True):
    """hmm.py:57-95 with use_gmm=True -> ((mu,cov,w,trans), hmm_mu, hmm_sigma, segments)."""
    hmu, hsig, trans, segments = skmeans(ys, n_segments)
    states = [GMMState(hmu[i], hsig[i], n_gaussians) for i in range(n_segments)]            # :78-79
    for i in range(n_segments):
        fit_gmm_state(states[i], segments[i], n_gaussians, hmu[i], 1 / segments[i].shape[0],
                      segments[i].shape[0], use_em)
    model = (np.array([s.mu for s in states]), np.array([s.cov for s in states]),
             np.array([s.w for s in states]), trans)
    segments = align_gmm_states(ys, model, n_segments)                                       # :95
    return model, hmu, hsig, segments


def hmm_evaluate(x, word_model):
    """hmm.py:126-135 (use_gmm=True)"""
    S = word_model[0].shape[0]
    E = emissions(x, [word_model], np.arange(S))
    costs, _, _ = decode_hmm_states(E, word_model[3])
    return costs[-1, -1]


def hmm_evaluate_single(x, mu, sigma, trans):
    """hmm.py:134 (use_gmm=False): dtw with mahalanobis local distance."""
    costs, _ = dtw(x, mu, trans, sigma)
    return costs[-1, -1]


def classify(X, off, word_models, log_domain=False):
    """core.py:79-87 over hmm.py:126-135 -> (final[B,W], label[B]).  OpenMP over utterances."""
    X, Xp = _d(X)
    off, op = _i(off)
    B = len(off) - 1
    W = len(word_models)
    S, M, D = word_models[0][0].shape
    mu, mp = _d(np.array([m[0] for m in word_models]))
    cov, cp = _d(np.array([m[1] for m in word_models]))
    w, wp = _d(np.array([m[2] for m in word_models]))
    tr, tp = _d(np.array([m[3] for m in word_models]))
    final = np.empty((B, W))
    label = np.empty(B, np.int64)
    lib().orc_classify(Xp, op, ctypes.c_int64(B), D, mp, cp, wp, tp, W, S, M, int(log_domain),
                       final.ctypes.data_as(c_dp), label.ctypes.data_as(c_ip))
    return final, label


# ------------------------------------------------------------------------------------ connected speech (a19,a20,a22)
def build_state_sequences(n_states_per_word, word_trans, label_matrix):
    """continuous_speech.py:13-53 -> (emit[R] flat state index or -1 for NES, trans[R,R], end_indices)."""
    S = n_states_per_word
    R = 1 + sum(len(labels) * S + 1 for labels in label_matrix)
    trans = np.full((R, R), np.inf)
    emit = [-1]
    starts, ends, nes = [], [], [0]
    for labels in label_matrix:
        st, en = [], []
        for l in labels:
            o = len(emit)
            st.append(o)
            en.append(o + S - 1)
            emit += [l * S + s for s in range(S)]
            trans[o:o + S, o:o + S] = word_trans[l]
        emit.append(-1)
        nes.append(len(emit) - 1)
        starts.append(st)
        ends.append(en)
    for layer in range(len(label_matrix)):
        for s in starts[layer]:
            trans[s, nes[layer]] = 0
        for e in ends[layer]:
            trans[nes[layer + 1], e] = 0
    return np.array(emit), trans, ends[-1]


def path_to_words(path, emit, n_states_per_word):
    """main.py:61-67: forward state sequence -> drop consecutive duplicates -> split on NES -> first state of
    each run -> word index."""
    st = np.asarray(path)[:, 0][::-1]
    if len(st) == 0:
        return []
    keep = np.insert(np.diff(st) != 0, 0, True)
    st = st[keep]
    words, in_run = [], False
    for s in st:
        if emit[s] < 0:
            in_run = False
        elif not in_run:
            words.append(int(emit[s]) // n_states_per_word)
            in_run = True
    return words


def pool_segments(path_forward, is_nes):
    """The literal pooling loop of continuous_speech.py:91-106 on one forward path [[state, frame], ...]
    -> list of (state_row, start_frame, end_frame) in flush order (quirk A12: the first frame after a state change is
    skipped, the last open segment is never flushed)."""
    out = []
    start_idx = None
    current_state = None
    for r, c in path_forward:
        r, c = int(r), int(c)
        if start_idx is None and not is_nes[r]:                       # :97-99
            start_idx = c
            current_state = r
        if r != current_state and start_idx is not None and start_idx < c:   # :100-106
            out.append((current_state, start_idx, c))
            start_idx = None
            current_state = None
    return out


def continuous_pool(data, word_models, label_seqs):
    """continuous_speech.py:80-107: forced alignment + the reference's frame pooling (A12).
    -> ordered dict {flat_state: [frame arrays]} in first-flush order."""
    S = word_models[0][0].shape[0]
    wt = [m[3] for m in word_models]
    pools = {}
    for x, labels in zip(data, label_seqs):
        emit, trans, ends = build_state_sequences(S, wt, [[l] for l in labels])
        E = emissions(x, word_models, emit)
        _, path, _ = decode_hmm_states(E, trans, emit < 0, [[e, -1] for e in ends])
        start, cur = None, None
        for r, c in path[::-1]:
            if start is None and emit[r] >= 0:
                start, cur = c, r
            if r != cur and start is not None and start < c:
                pools.setdefault(int(emit[cur]), []).append(x[start:c])
                start, cur = None, None
    return pools


def continuous_train_iteration(data, word_models, label_seqs, n_gaussians=4, n_segments=5):
    """One outer iteration of continuous_speech.py:73-164 -> new word models."""
    S = word_models[0][0].shape[0]
    new = [tuple(np.array(a, dtype=np.float64, copy=True) for a in m) for m in word_models]
    pools = continuous_pool(data, new, label_seqs)
    for flat, parts in pools.items():
        seg = np.vstack(parts)
        wd, s = divmod(flat, S)
        mu, cov, w, _ = new[wd]
        st = GMMState(mu[s, 0], cov[s, 0], mu.shape[1])
        st.mu, st.cov, st.w = mu[s], cov[s], w[s]          # views: em() updates the model in place
        # mu_old etc. are whatever the pickled state carried; the golden models come straight from HMM.fit,
        # whose last em() call left *_old at the converged values (hmm_state.py:157-159)
        st.mu_old, st.sigma_old, st.w_old = mu[s].copy(), cov[s].copy(), w[s].copy()
        fit_gmm_state(st, seg, n_gaussians, np.mean(seg, axis=0), 1 / n_segments, n_segments, True)
    for wd in range(len(new)):                              # :146-164
        tr = new[wd][3]
        for si in range(S):
            parts = pools.get(wd * S + si)
            if parts is None:
                continue
            with np.errstate(all="ignore"):
                p_jump = np.float64(len(parts)) / sum(p.shape[0] for p in parts)
                if si < S - 1:
                    tr[si + 1, si] = -np.log(p_jump)
                tr[si, si] = -np.log(1 - p_jump)
    return new


# ------------------------------------------------------------------------------------ batched training pass (CPU baseline)
def align_labels(X, off, words, word_models, log_domain=False):
    """kmeans.py:196-205 for a labelled batch -> (labels[N] tied state per frame, cost[B]).  OpenMP over utterances."""
    X, Xp = _d(X)
    off, op = _i(off)
    words, wp_ = _i(words)
    S, M, D = word_models[0][0].shape
    mu, mp = _d(np.array([m[0] for m in word_models]))
    cov, cp = _d(np.array([m[1] for m in word_models]))
    w, wp = _d(np.array([m[2] for m in word_models]))
    tr, tp = _d(np.array([m[3] for m in word_models]))
    B = len(off) - 1
    labels = np.empty(X.shape[0], np.int64)
    cost = np.empty(B)
    lib().orc_align_labels(Xp, op, ctypes.c_int64(B), D, wp_, mp, cp, wp, tp, S, M, int(log_domain),
                           labels.ctypes.data_as(c_ip), cost.ctypes.data_as(c_dp))
    return labels, cost


def train_iteration(X, off, words, word_models, ncomp, log_domain=False):
    """One segmental-K-means + EM pass over all words: Viterbi re-alignment (align_gmm_states) then ONE GMM.em
    iteration (hmm_state.py:124-148) per tied state on its pooled frames.  -> (new models, mean cost)."""
    X = np.ascontiguousarray(X, dtype=np.float64)
    labels, cost = align_labels(X, off, words, word_models, log_domain)
    S, M, D = word_models[0][0].shape
    order = np.argsort(labels, kind="stable")
    bounds = np.searchsorted(labels[order], np.arange(len(word_models) * S + 1))
    new = []
    for wd, m in enumerate(word_models):
        mu, cov, w = m[0].copy(), m[1].copy(), m[2].copy()
        for s in range(S):
            t = wd * S + s
            idx = np.sort(order[bounds[t]:bounds[t + 1]])
            if len(idx) == 0:
                continue
            st = GMMState(mu[s, 0], cov[s, 0], M)
            st.mu, st.cov, st.w = mu[s], cov[s], w[s]
            st.em(X[idx], ncomp, 1)
        new.append((mu, cov, w, m[3]))
    return new, float(cost.mean())
