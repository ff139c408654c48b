"""TEST INFRASTRUCTURE ONLY -- a numpy float64 stand-in for `srb200.engine`, installed by the `host_double`
fixture so that the HOST LOGIC of the sr.recognition mirror (segment bookkeeping, pooling state machine, RNG order,
convergence loops, graph building) can be exercised by `-m "not gpu"` tests in a container without a GPU.

It is never imported by the product tree (tests/test_abi.py checks that), it is not a fallback -- the product
engine raises without CUDA -- and no parity claim rests on it: the `-m gpu` tests run the SAME cases
(tests/mirror_cases.py) through libsrb200 on the B200.
"""
import numpy as np
import torch

LOG2PI = 1.8378770664093454836
CPU = torch.device("cpu")


def device():
    return CPU


def _np(a, dtype=None):
    if torch.is_tensor(a):
        a = a.detach().cpu().numpy()
    return np.asarray(a, dtype=dtype)


class PackedModel:
    def __init__(self, mu, var, w, shift=None):
        self.mu = torch.as_tensor(_np(mu, np.float64)).clone()
        self.var = torch.as_tensor(_np(var, np.float64)).clone()
        self.w = torch.as_tensor(_np(w, np.float64)).clone()
        self.R, self.M, self.D = (int(v) for v in self.mu.shape)
        self.Mp = (self.M + 3) // 4 * 4
        self.ab = torch.zeros(1)

    def repack(self):
        return self


def as_frames(X):
    return torch.as_tensor(np.ascontiguousarray(_np(X, np.float32)))


def as_offsets(off):
    return torch.as_tensor(np.ascontiguousarray(_np(off, np.int32)))


def _log_comps(X, pm):
    X = _np(X, np.float64)
    mu, var, w = pm.mu.numpy(), pm.var.numpy(), pm.w.numpy()
    with np.errstate(divide="ignore"):
        lc = np.log(w) - 0.5 * (pm.D * LOG2PI + np.log(var).sum(-1))                       # [R, M]
    q = ((X[:, None, None, :] - mu[None]) ** 2 / var[None]).sum(-1)                         # [N, R, M]
    return lc[None] - 0.5 * q


def _neg_lse(l, axis):
    m = l.max(axis=axis, keepdims=True)
    m = np.where(np.isfinite(m), m, 0.0)
    with np.errstate(divide="ignore"):
        return -(np.squeeze(m, axis) + np.log(np.exp(l - m).sum(axis=axis)))


def gmm_score(X, pm, want_E=True, want_best=False, want_comp=False):
    l = _log_comps(X, pm)
    E = _neg_lse(l, 2)
    out = {}
    if want_E:
        out["E"] = torch.from_numpy(E.astype(np.float32))
    if want_best:
        out["best_cost"] = torch.from_numpy(E.min(1).astype(np.float32))
        out["best_state"] = torch.from_numpy(E.argmin(1).astype(np.int32))
    if want_comp:
        comp = np.full((l.shape[0], pm.R, pm.Mp), -np.inf)
        comp[:, :, :pm.M] = l
        out["comp"] = torch.from_numpy(comp.astype(np.float32))
    return out


def _viterbi_one(E, g, want_costs):
    """Sequential sweep in state-index order with the same-column rule for NES edges (decode.py:97-124)."""
    T, R = E.shape[0], g.R
    emit, pp, pi, pc = g.emit, g.pred_ptr, g.pred_idx, g.pred_cost.astype(np.float64)
    costs = np.full((T, R), np.inf)
    bp = np.zeros((T, R), np.int64)
    for c in range(T):
        prev = costs[c - 1] if c > 0 else np.full(R, np.inf)
        cur = costs[c]
        for r in range(R):
            e = E[c, emit[r]] if emit[r] >= 0 else 0.0
            if c == 0 and r == 0:
                cur[0] = e
                continue
            best, bj = np.inf, 0
            for j in range(pp[r], pp[r + 1]):
                o = pi[j]
                if emit[o] < 0 or emit[r] < 0:
                    v = cur[o] if o < r else np.inf
                else:
                    v = prev[o]
                if pc[j] + v < best:
                    best, bj = pc[j] + v, j - pp[r]
            if pp[r + 1] > pp[r]:
                cur[r] = best + e
                bp[c, r] = bj
    return costs, bp


class _VRes:
    pass


def viterbi(E, off, gb, utt_graph=None, want_costs=False, want_path=True, frame_label=None):
    E = _np(E, np.float64)
    off = _np(off, np.int64)
    B = len(off) - 1
    bc, be = np.full(B, np.inf), np.full(B, -1, np.int64)
    paths, costs_flat, ws = [], [], np.zeros(B + 1, np.int64)
    fl = np.full(E.shape[0], -1, np.int32)
    for u in range(B):
        gi = 0 if utt_graph is None else int(utt_graph[u])
        g, s0 = gb.graphs[gi], int(gb.state_off[gi])
        Eu = E[off[u]:off[u + 1]]
        T = Eu.shape[0]
        costs, bp = _viterbi_one(Eu, g, want_costs)
        ws[u + 1] = ws[u] + T * g.R
        costs_flat.append(costs.ravel())
        best, bi, bj = np.inf, -1, T - 1
        frames = g.end_frames if getattr(g, "end_frames", None) is not None else [T - 1] * len(g.end_states)
        for s, c in zip(g.end_states, frames):
            c = int(c) + T if c < 0 else int(c)
            if best >= costs[c, s]:
                best, bi, bj = costs[c, s], int(s), c
        bc[u], be[u] = best, bi
        path = []
        if np.isfinite(best):
            i, j = bi, bj
            if g.emit[i] >= 0:
                fl[off[u] + j] = s0 + i
            while j > 0:
                o = int(g.pred_idx[g.pred_ptr[i] + bp[j, i]])
                if not (g.emit[o] < 0 or g.emit[i] < 0):
                    j -= 1
                i = o
                path.append([i, j])
                if g.emit[i] >= 0:
                    fl[off[u] + j] = s0 + i
        paths.append(np.array(path, np.int64).reshape(-1, 2))
    res = _VRes()
    res.best_cost = torch.from_numpy(bc.astype(np.float32))
    res.best_end = torch.from_numpy(be.astype(np.int32))
    res.ws_off = ws
    res.costs = torch.from_numpy(np.concatenate(costs_flat).astype(np.float32)) if want_costs else None
    res.paths = lambda: paths

    def words(state_word, max_words):
        # main.py:59-67 in numpy: forward order, drop consecutive duplicates, split at NES (-1), first state per piece
        out = []
        for u, pth in enumerate(paths):
            gi = 0 if utt_graph is None else int(utt_graph[u])
            sw = np.asarray(state_word)[int(gb.state_off[gi]):]
            rows = pth[:, 0][::-1]
            ws_, prev = [], None
            for r in rows:
                if prev is not None and r == prev:
                    continue
                if sw[r] >= 0 and (prev is None or sw[prev] < 0):
                    ws_.append(int(sw[r]))
                prev = r
            assert len(ws_) <= max_words
            out.append(ws_)
        return out

    res.words = words
    res.frame_label = None
    if frame_label is not None:
        frame_label[:len(fl)] = torch.from_numpy(fl)
        res.frame_label = frame_label
    return res


def dtw_beam(dist, off, trans, beam=None):
    """numpy stand-in of engine.dtw_beam: the column rules of decode.py:42-68 (marked origins skipped by row 0, +inf
    for the other rows; finite cells of rank >= beam marked)."""
    Dl = _np(dist, np.float64)
    off = _np(off, np.int64)
    tr = _np(trans, np.float64)
    S = Dl.shape[1]
    B = len(off) - 1
    ws = np.zeros(B + 1, np.int64)
    flat, paths = [], []
    for u in range(B):
        d = Dl[off[u]:off[u + 1]]
        T = d.shape[0]
        ws[u + 1] = ws[u] + T * S
        costs = np.full((T, S), np.inf)
        bp = np.zeros((T, S), np.int64)
        mark = np.zeros(S, bool)
        for j in range(T):
            for i in range(S):
                if j == 0:
                    costs[0, i] = d[0, 0] if i == 0 else np.inf
                    continue
                cand = np.where(mark, np.inf, costs[j - 1]) + tr[i]
                idx = np.flatnonzero(~mark) if i == 0 else np.arange(S)
                if len(idx) == 0:
                    raise ValueError("attempt to get argmin of an empty sequence")
                bo = idx[np.argmin(cand[idx])]
                bp[j, i] = bo
                costs[j, i] = cand[bo] + d[j, i]
            cj = costs[j]
            fin = np.isfinite(cj)
            if beam is not None:
                rank = np.empty(S, np.int64)
                rank[np.argsort(cj, kind="stable")] = np.arange(S)
                mark = fin & (rank >= beam)
            else:
                mark = np.zeros(S, bool)
            if j > 0:
                costs[j - 1][prev_mark] = np.inf
            prev_mark = mark
        costs[T - 1][mark] = -1.0
        i, j, path = S - 1, T - 1, []
        while i != 0 or j != 0:
            if j == 0:
                path = None
                break
            i, j = int(bp[j, i]), j - 1
            path.append([i, j])
        flat.append(costs.ravel())
        paths.append(None if path is None else np.array(path, np.int64).reshape(-1, 2))
    status = np.array([-3 if p is None else len(p) for p in paths], np.int32)
    return torch.from_numpy(np.concatenate(flat).astype(np.float32)), ws, paths, status


def classify(X, off, pm, trans, W, S, final=None, label=None):
    from srb200.graph import StateGraph
    E = _neg_lse(_log_comps(X, pm), 2)
    off = _np(off, np.int64)
    tr = _np(trans, np.float64)
    B = len(off) - 1
    fin = np.empty((B, W))
    for w in range(W):
        g = StateGraph(tr[w], np.arange(S))
        for u in range(B):
            fin[u, w] = _viterbi_one(E[off[u]:off[u + 1], w * S:(w + 1) * S], g, False)[0][-1, -1]
    return torch.from_numpy(fin.astype(np.float32)), torch.from_numpy(fin.argmin(1).astype(np.int32))


def stats_row(D):
    return 2 * D + 3


def group_stats(X, label, K, shift, out=None, accumulate=False):
    X, lab, sh = _np(X, np.float64), _np(label, np.int64), _np(shift, np.float64)
    D = X.shape[1]
    st = np.zeros((K, stats_row(D)))
    lab = lab[:X.shape[0]]
    for k in range(K):
        e = X[lab == k] - sh[k]
        st[k, 0] = e.shape[0]
        st[k, 1:1 + D] = e.sum(0)
        st[k, 1 + D:1 + 2 * D] = (e * e).sum(0)
    return torch.from_numpy(st)


def moments_from_stats(stats, shift, D, unbiased=True):
    S0 = stats[:, 0]
    S1, S2 = stats[:, 1:1 + D], stats[:, 1 + D:1 + 2 * D]
    m1 = S1 / S0[:, None]
    den = (S0 - 1.0) if unbiased else S0
    return S0, shift + m1, (S2 - S1 * m1) / den[:, None]


def gmm_stats(X, label, pm, ncomp, out=None, accumulate=False):
    X, lab = _np(X, np.float64), _np(label, np.int64)[:len(X)]
    D, M = pm.D, pm.M
    l = _log_comps(X, pm)                                                       # [N, R, M]
    st = np.zeros((pm.R * M, stats_row(D)))
    mu = pm.mu.numpy()
    for r in range(pm.R):
        sel = lab == r
        if not sel.any():
            continue
        lr = l[sel, r, :ncomp]
        mx = lr.max(1, keepdims=True)
        p = np.exp(lr - mx)
        den = p.sum(1, keepdims=True)
        p /= den
        for m in range(ncomp):
            e = X[sel] - mu[r, m]
            row = st[r * M + m]
            row[0] = p[:, m].sum()
            row[1:1 + D] = (p[:, [m]] * e).sum(0)
            row[1 + D:1 + 2 * D] = (p[:, [m]] * e * e).sum(0)
        st[r * M, 2 * D + 1] = sel.sum()
        st[r * M, 2 * D + 2] = (mx[:, 0] + np.log(den[:, 0])).sum()
    return torch.from_numpy(st)


def mstep(stats, pm, ncomp, var_floor=0.0):
    st = _np(stats, np.float64)
    D, M = pm.D, pm.M
    for r in range(pm.R):
        cnt = st[r * M, 2 * D + 1]
        for m in range(ncomp):
            row = st[r * M + m]
            if not row[0] > 0:
                continue
            m1 = row[1:1 + D] / row[0]
            pm.mu[r, m] += torch.from_numpy(m1)
            pm.var[r, m] = torch.from_numpy(np.maximum(row[1 + D:1 + 2 * D] / row[0] - m1 * m1, var_floor))
            if cnt > 0:
                pm.w[r, m] = row[0] / cnt
    return pm


def local_dist(X, Y, var=None, shared_var=False, want_dist=True, want_argmin=False):
    X, Y = _np(X, np.float64), _np(Y, np.float64)
    e = X[:, None, :] - Y[None]
    if var is None:
        d = np.sqrt((e * e).sum(-1))
    else:
        v = _np(var, np.float64)
        v = np.broadcast_to(v[:1] if shared_var else v, Y.shape)
        d = 0.5 * (Y.shape[1] * LOG2PI + np.log(v).sum(-1))[None] + 0.5 * (e * e / v[None]).sum(-1)
    return (torch.from_numpy(d.astype(np.float32)) if want_dist else None,
            torch.from_numpy(d.argmin(1).astype(np.int32)) if want_argmin else None)


def kmeans_assign(X, group, centroids, var0):
    X, g = _np(X, np.float64), _np(group, np.int64)
    c, v = _np(centroids, np.float64), _np(var0, np.float64)
    out = np.full(X.shape[0], -1, np.int32)
    ok = (g >= 0) & (g < c.shape[0])
    e = X[ok, None, :] - c[g[ok]]                                        # [n, k, D]
    d = (0.5 * e * e / v[g[ok]][:, None, :]).sum(-1)
    d = np.where(np.isnan(d), np.inf, d)
    out[ok] = d.argmin(1)
    return torch.from_numpy(out)


def pool_stats_supported(G, k, D):
    return k in (1, 2, 4, 8)


def pool_stats_supported_f64(G, k, D):
    return k in (1, 2, 4)


def pool_em_stats(X, state, pm, k, lc=None, f64=False):
    return gmm_stats(X, state, pm, k)


def pool_kmeans_stats(X, state, centroids, var0, shift, want_argmin=False, f64=False):
    c = _np(centroids, np.float64)
    G, k, D = c.shape
    am = kmeans_assign(X, state, c, var0)
    st_ = _np(state, np.int64)
    lab = np.where(st_ >= 0, st_ * k + am.numpy(), -1)
    out = group_stats(X, lab, G * k, _np(shift, np.float64).reshape(G * k, D))
    return (out, am) if want_argmin else out


def gauss_consts(var, w):
    v, ww = _np(var, np.float64), _np(w, np.float64)
    G = ww.size
    v = v.reshape(G, -1)
    with np.errstate(divide="ignore"):
        return torch.from_numpy(np.log(ww.reshape(G)) - 0.5 * (v.shape[1] * np.log(2 * np.pi) + np.log(v).sum(1)))


def _allclose_rows(a, b):
    return ((a - b).abs() <= 1e-8 + 1e-5 * b.abs()).flatten(1).all(dim=1)


def fit_em_update(stats, pm, k, old, active, its, it, lc):
    """numpy-semantics stand-in of srb_fit_em_update (float64 torch on the CPU), in place."""
    G, M, D = pm.mu.shape
    st = stats.reshape(G, M, 2 * D + 3)
    mu, var, w = pm.mu[:, :k], pm.var[:, :k], pm.w[:, :k]
    mu_old, var_old, w_old = (t[:, :k] for t in old)
    S0 = st[:, :k, 0]
    cntg = st[:, 0, 2 * D + 1]
    m1 = st[:, :k, 1:1 + D] / S0[:, :, None]
    nmu = mu + m1
    nvar = st[:, :k, 1 + D:1 + 2 * D] / S0[:, :, None] - m1 * m1
    nw = S0 / cntg[:, None]
    ok = (S0 > 0) & active[:, None]
    mu.copy_(torch.where(ok[:, :, None], nmu, mu))
    var.copy_(torch.where(ok[:, :, None], nvar, var))
    w.copy_(torch.where(ok & (cntg > 0)[:, None], nw, w))
    conv = _allclose_rows(mu, mu_old) & _allclose_rows(var, var_old) & _allclose_rows(w, w_old)
    its.copy_(torch.where(active, torch.full_like(its, it), its))
    active.copy_(active & ~conv)
    mu_old.copy_(torch.where(active[:, None, None], mu, mu_old))
    var_old.copy_(torch.where(active[:, None, None], var, var_old))
    w_old.copy_(torch.where(active[:, None], w, w_old))
    lc.copy_(gauss_consts(pm.var, pm.w).reshape(lc.shape))


def fit_kmeans_update(stats, shift, cen, active, its, it, cnt):
    G, k, D = cen.shape
    c = stats[:, 0].reshape(G, k)
    new_c = (shift + stats[:, 1:1 + D] / stats[:, :1]).reshape(G, k, D)
    same = (new_c == cen).flatten(1).all(dim=1)
    its.copy_(torch.where(active, torch.full_like(its, it), its))
    cnt.copy_(torch.where(active[:, None], c, cnt))
    active.copy_(active & ~same)
    cen.copy_(torch.where(active[:, None, None], new_c, cen))


def install(monkeypatch):
    """Replace the device entry points of srb200.engine for the duration of one test."""
    from srb200 import engine
    for name in ("device", "PackedModel", "as_frames", "as_offsets", "gmm_score", "viterbi", "classify",
                 "group_stats", "moments_from_stats", "gmm_stats", "mstep", "local_dist", "dtw_beam", "kmeans_assign",
                 "pool_stats_supported", "pool_stats_supported_f64", "pool_em_stats", "pool_kmeans_stats",
                 "gauss_consts", "fit_em_update", "fit_kmeans_update"):
        monkeypatch.setattr(engine, name, globals()[name])
