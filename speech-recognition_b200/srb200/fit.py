"""Batched `HMM.fit` for ALL words of a labelled corpus (BASELINE config 2 in full: segmental K-means -> binary-split
K-means -> per-state EM -> Viterbi re-alignment), device resident and sharded by utterance.

The reference trains one word at a time and one state at a time (hmm.py:57-124): `skmeans` (kmeans.py:111-155), then per
state `kmeans` on the x0.9 / x1.1 split centroids (kmeans.py:166-193) and `GMM.em` (hmm_state.py:122-159), then
`align_gmm_states` (kmeans.py:196-205).  The mirror's `HMM.fit` follows it call by call (a Python loop of small launches per
state).  Here every phase runs for all W*S states at once:

    phase                      per iteration                                               exchange (N GPUs)
    skmeans                    W x srb_local_dist + one srb_viterbi_graph + srb_group_stats  fp64 allreduce of [W*S, 2D+3]
    K-means (per split)        srb_kmeans_assign + srb_group_stats                           fp64 allreduce of [W*S*k, 2D+3]
    EM (per split)             srb_gmm_stats (all states, label = state)                     fp64 allreduce of [W*S*M, 2D+3]
    re-alignment               srb_em_isolated (alignment only)                              none

The statistics are summed over the ranks (`torch.distributed.all_reduce`), every rank applies the same update: the
"local partials -> one fp64 allreduce -> replicated update" pattern of SURVEY 8e for the two phases the round-1 trainer
left single-GPU.  What stays on the host is O(states) bookkeeping in float64: the convergence tests of the reference
(np.allclose / np.array_equal per word or per state, with converged words / states frozen at the iteration where the
reference's own loop would have stopped), the M-step on the 200-odd reduced rows, and the global numpy RNG stream of
kmeans.py:171, drawn state by state in the reference's order and indexed by a frame's rank in its state's pool.
"""
import sys

import numpy as np
import torch

from . import engine
from .graph import GraphBatch, StateGraph
from .trainer import sort_by_word


def _dist_on(group):
    return group is not None or (torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1)


def _allreduce(t, group):
    if _dist_on(group):
        torch.distributed.all_reduce(t, group=group)
    return t


def transition_costs(n_temps, n_all, S):
    """calc_transition_costs (kmeans.py:53-95) for uniform segment lengths without empty segments: stay / advance costs
    from the pooled segment lengths n_all[S] of n_temps utterances; the last state stays with probability 1."""
    tr = np.full((S, S), np.inf)
    with np.errstate(divide="ignore", invalid="ignore"):
        for i in range(S):
            n_jump = 0 if i == S - 1 else n_temps
            if n_jump:
                tr[i + 1, i] = -np.log(n_jump / n_all[i])
            tr[i, i] = -np.log((n_all[i] - n_jump) / n_all[i])
    return tr


class WordFitter:
    """All W word models of this rank's shard of a labelled corpus, trained together.  X [N, D] float32 (numpy or CUDA,
    word-sorted when on the device), off [B+1], words [B]."""

    def __init__(self, X, off, words, W, n_segments, n_gaussians, use_em=True, group=None, seeds=None, verbose=False,
                 fused_align=True, precision="f64"):
        X, off, words, self.order = sort_by_word(X if torch.is_tensor(X) else np.asarray(X), np.asarray(off), words)
        self.dev = engine.device()
        self.X = engine.as_frames(X)
        self.W, self.S, self.M, self.D = int(W), int(n_segments), int(n_gaussians), int(self.X.shape[1])
        self.use_em, self.group, self.seeds, self.verbose = use_em, group, seeds, verbose
        self.fused_align = fused_align
        # "f64": the K-means / EM passes score and accumulate in float64 (srb_pool_stats_f64) so that the iterated loops
        # follow the reference's float64 trajectories to its fixed points and stop iterations; "f32": srb_pool_stats
        assert precision in ("f64", "f32")
        self.f64 = precision == "f64"
        self.words = np.asarray(words, np.int64)
        self.off = np.asarray(off, np.int64)
        self.B, self.N = len(self.off) - 1, int(self.off[-1])
        self.T = np.diff(self.off)
        uptr = np.searchsorted(self.words, np.arange(self.W + 1), side="left")
        self.word_utts = [(int(uptr[w]), int(uptr[w + 1])) for w in range(self.W)]
        self.word_rows = [(int(self.off[a]), int(self.off[b])) for a, b in self.word_utts]
        self.off_d = torch.from_numpy(self.off).to(self.dev)
        self.T_d = torch.from_numpy(self.T).to(self.dev)
        self.utt_of = torch.repeat_interleave(torch.arange(self.B, device=self.dev), self.T_d)        # [N]
        self.t_in = torch.arange(self.N, device=self.dev) - self.off_d[self.utt_of]
        self.word_of_frame = torch.from_numpy(self.words).to(self.dev)[self.utt_of]
        self.iterations = {}
        self.converged = {}                                   # per EM pass: which states met the stop test (not the cap)
        world = torch.distributed.get_world_size(group) if _dist_on(group) else 1
        self.rank = torch.distributed.get_rank(group) if _dist_on(group) else 0
        self.world = world

    # ------------------------------------------------------------------------------------------ helpers
    def _labels_from_starts(self, starts):
        """frame -> w*S + segment for seg_starts [B, S] (segment_data's slicing, kmeans.py:41-50)."""
        st = torch.from_numpy(np.ascontiguousarray(starts[:, 1:], dtype=np.int64)).to(self.dev)       # [B, S-1]
        seg = (self.t_in[:, None] >= st[self.utt_of]).sum(dim=1)
        return (self.word_of_frame * self.S + seg).to(torch.int32)

    def _moments(self, label, K, shift=None, unbiased=True):
        """(count[K], mean[K, D], var[K, D]) of the frames of every group over ALL ranks; two passes (mean, then moments
        about it) unless a shift is given."""
        D = self.D
        if shift is None:
            z = torch.zeros((K, D), dtype=torch.float64, device=self.dev)
            st = _allreduce(engine.group_stats(self.X, label, K, z), self.group)
            shift_t = st[:, 1:1 + D] / st[:, :1]
            shift_t = torch.where(torch.isfinite(shift_t), shift_t, torch.zeros_like(shift_t))
        else:
            shift_t = torch.as_tensor(shift, dtype=torch.float64).to(self.dev).reshape(K, D)
        st = _allreduce(engine.group_stats(self.X, label, K, shift_t), self.group)
        with np.errstate(all="ignore"):
            cnt, mean, var = engine.moments_from_stats(st, shift_t, D, unbiased=unbiased)
        return cnt.cpu().numpy(), mean.cpu().numpy(), var.cpu().numpy()

    def _global_sum(self, a):
        t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(self.dev)
        return _allreduce(t, self.group).cpu().numpy()

    # ------------------------------------------------------------------------------------------ phase 1: skmeans
    def skmeans(self, max_iteration=1000):
        """kmeans.py:111-155 for every word at once -> (mu [W,S,D], var [W,S,D], trans [W,S,S], seg_starts [B,S])."""
        W, S = self.W, self.S
        if (self.T < 5).any():
            raise NameError('template is too small, cannot do dtw on it')                 # kmeans.py:142-143
        seg_len = self.T // S
        starts = np.concatenate([np.zeros((self.B, 1), np.int64),
                                 np.cumsum(np.repeat(seg_len[:, None], S, axis=1), axis=1)[:, :-1]], axis=1)
        _, res, var = self._moments(self._labels_from_starts(starts), W * S)
        res, var = res.reshape(W, S, -1), var.reshape(W, S, -1)
        # transition costs come from the UNIFORM lengths in every iteration (A5): computed once
        n_temps = self._global_sum(np.bincount(self.words, minlength=W).astype(np.float64))
        n_all = self._global_sum(np.bincount(self.words, weights=seg_len.astype(np.float64), minlength=W))
        trans = np.stack([transition_costs(n_temps[w], np.full(S, n_all[w]), S) for w in range(W)])
        graphs = [StateGraph(trans[w], np.arange(S), end_states=[S - 1], keep_inf=True) for w in range(W)]
        gb = GraphBatch(graphs, self.dev)
        fl = torch.empty(max(self.N, 1), dtype=torch.int32, device=self.dev)
        dist = torch.empty((max(self.N, 1), S), dtype=torch.float32, device=self.dev)
        state_off = torch.from_numpy(np.arange(W, dtype=np.int64) * S).to(self.dev)
        active = np.ones(W, bool)
        its = np.zeros(W, np.int64)
        for it in range(max_iteration):
            for w, (r0, r1) in enumerate(self.word_rows):                     # Euclidean distances to the word's means
                if r1 > r0 and active[w]:
                    dist[r0:r1] = engine.local_dist(self.X[r0:r1], res[w])[0]
            engine.viterbi(dist, self.off.astype(np.int32), gb, utt_graph=self.words.astype(np.int32), want_path=False,
                           frame_label=fl)
            lab = fl[:self.N]                                                 # graph offset + state = w*S + s
            cnt = torch.bincount((self.utt_of * S + (lab.long() - state_off[self.word_of_frame])).clamp_(min=0),
                                 weights=(lab >= 0).double(), minlength=self.B * S).reshape(self.B, S)
            new_starts = np.zeros((self.B, S), np.int64)
            new_starts[:, 1:] = np.cumsum(cnt.cpu().numpy().astype(np.int64), axis=1)[:, :-1]
            _, new_res, new_var = self._moments(self._labels_from_starts(new_starts), W * S)
            new_res, new_var = new_res.reshape(W, S, -1), new_var.reshape(W, S, -1)
            for w in np.flatnonzero(active):
                a, b = self.word_utts[w]
                starts[a:b] = new_starts[a:b]
                var[w] = new_var[w]
                its[w] = it
                if np.allclose(res[w], new_res[w]):
                    active[w] = False                                         # `break` before res = new_res (:148-150)
                else:
                    res[w] = new_res[w]
            if not active.any():
                break
        self.iterations["skmeans"] = its
        self.hmm_mu, self.hmm_sigma, self.trans, self.seg_starts = res, var, trans, starts
        return res, var, trans, starts

    # ------------------------------------------------------------------------------------------ phase 2: per-state GMMs
    def _pool_ranks(self, state):
        """Rank of every frame inside its state's pool (segment_data order: utterance by utterance), offset by the frames
        the lower ranks contribute to the pool.  -> (rank_in_pool int64[N], n_local[G], n_global[G], offset[G])"""
        G = self.W * self.S
        order = torch.argsort(state, stable=True)
        sorted_state = state[order].long()
        n_local = torch.bincount(sorted_state, minlength=G)
        start = torch.cumsum(n_local, 0) - n_local
        rank = torch.empty(self.N, dtype=torch.int64, device=self.dev)
        rank[order] = torch.arange(self.N, device=self.dev) - start[sorted_state]
        n_loc = n_local.cpu().numpy().astype(np.int64)
        if self.world > 1:
            allc = [torch.empty_like(n_local) for _ in range(self.world)]
            torch.distributed.all_gather(allc, n_local, group=self.group)
            allc = np.stack([c.cpu().numpy() for c in allc]).astype(np.int64)
            offset, n_glob = allc[:self.rank].sum(axis=0), allc.sum(axis=0)
        else:
            offset, n_glob = np.zeros(G, np.int64), n_loc
        return rank, n_loc, n_glob, offset

    @classmethod
    def from_pools(cls, X, state, G, n_gaussians, use_em=True, group=None, seeds=None, verbose=False, precision="f64"):
        """A fitter over ready-made pools instead of a labelled corpus: X [N, D] frames (numpy or CUDA), state int[N] the
        pool of every frame (frames of a pool in the order the reference would concatenate them).  For callers that pool
        themselves (continuous_train, continuous_speech.py:101-142); use fit_states(state=..., init=..., ...)."""
        self = cls.__new__(cls)
        self.dev = engine.device()
        self.X = engine.as_frames(X)
        self.W, self.S, self.M, self.D = 1, int(G), int(n_gaussians), int(self.X.shape[1])
        self.use_em, self.group, self.seeds, self.verbose = use_em, group, seeds, verbose
        self.fused_align, self.f64 = False, precision == "f64"
        self.N = int(self.X.shape[0])
        self.iterations, self.converged = {}, {}
        self.world = torch.distributed.get_world_size(group) if _dist_on(group) else 1
        self.rank = torch.distributed.get_rank(group) if _dist_on(group) else 0
        self.pool_state = torch.as_tensor(state).to(self.dev, torch.int32)
        return self

    def fit_states(self, kmeans_max_iteration=1000, em_max_iteration=10000, *, state=None, init=None, old=None,
                   centroid0=None, weight_fill=None, weight_den=None):
        """hmm.py:97-124 for every state of every word at once (after skmeans).  The parameters, the stop tests and the
        iteration counters live on the device (float64 torch tensors, same IEEE operations as the reference's numpy);
        the host only learns, every few iterations, how many states are still running.
        Keyword arguments for pools that do not come from skmeans (continuous_speech.py:121-142): state int[N] labels,
        init = (mu, var, w) [G, M, ...] the states' current mixtures, old = (mu_old, var_old, w_old) what their first EM
        stop test compares with, centroid0 [G, D] the first centroid (the pool mean there), weight_fill / weight_den [G]
        the value the weight vector starts from and the denominator of the cluster counts."""
        W, S, M, D = self.W, self.S, self.M, self.D
        G = W * S
        n_splits = int(np.log(M))
        assert n_splits > 0
        if state is None:
            state = self._labels_from_starts(self.seg_starts)                  # pool of every frame
        rank, n_loc, n_glob, offset = self._pool_ranks(state)
        if init is None:
            mu = np.repeat(self.hmm_mu.reshape(G, 1, D), M, axis=1).copy()      # GMM.__init__: M copies (hmm_state.py:105-112)
            var = np.repeat(self.hmm_sigma.reshape(G, 1, D), M, axis=1).copy()
            wgt = np.full((G, M), 1.0 / M)
        else:
            mu, var, wgt = (np.array(a, dtype=np.float64) for a in init)
        pm = engine.PackedModel(mu, var, wgt)
        self._old = [pm.mu.clone(), pm.var.clone(), pm.w.clone()] if old is None else [self._t(a) for a in old]
        centroids = self._t((self.hmm_mu if centroid0 is None else np.asarray(centroid0)).reshape(G, 1, D))
        weight_den = n_glob if weight_den is None else np.broadcast_to(np.asarray(weight_den, np.float64), (G,))
        if weight_fill is None:
            weight_fill = 1.0 / np.maximum(n_glob, 1)                           # np.full(n_gaussians, 1 / n) (hmm.py:108)
        weights = np.repeat(np.broadcast_to(np.asarray(weight_fill, np.float64), (G,)).reshape(G, 1), M, axis=1)
        self.iterations["kmeans"], self.iterations["em"] = [], []
        self._n_loc = torch.from_numpy(n_loc).to(self.dev)
        # the global numpy RNG stream of kmeans.py:171 in the reference's order: word by word (optionally re-seeded per
        # word, like a caller who seeds before each HMM.fit), state by state, split by split
        draws = [[None] * G for _ in range(n_splits)]
        for w in range(W):
            if self.seeds is not None:
                np.random.seed(self.seeds[w])
            for s_ in range(S):
                for i in range(n_splits):
                    draws[i][w * S + s_] = np.random.randint(0, 2 ** (i + 1), int(n_glob[w * S + s_]))
        has = torch.from_numpy(n_glob > 0).to(self.dev)
        for i in range(n_splits):
            k = 2 ** (i + 1)
            centroids = torch.cat([centroids * 0.9, centroids * 1.1], dim=1)                  # [G, k, D]  (hmm.py:110-112)
            # ---- K-means: the random initial partition (kmeans.py:171) only gives the variances (A3)
            parts = draws[i]
            part_off = np.concatenate([[0], np.cumsum([len(p) for p in parts])])
            flat = torch.from_numpy(np.concatenate(parts) if parts else np.zeros(0, np.int64)).to(self.dev)
            gidx = torch.from_numpy(part_off[:-1] + offset).to(self.dev)[state.long()] + rank
            part = flat[gidx]
            _, _, cov = self._moments((state.long() * k + part).to(torch.int32), G * k)
            cov = cov.reshape(G, k, D)
            centroids, cnt, km_its = self._kmeans(state, k, centroids, self._t(cov[:, 0]), has.clone(), kmeans_max_iteration)
            self.iterations["kmeans"].append(km_its)
            for g in range(G):                                                   # weights from cluster counts (A4, sic)
                if n_glob[g] == 0:
                    continue
                present = np.flatnonzero(cnt[g] > 0)
                c_counts = cnt[g][present]
                for cl in present:
                    weights[g, cl] = c_counts[cl] / weight_den[g]                # IndexError like hmm.py:116-118 if a cluster is empty
            pm.mu[:, :k], pm.var[:, :k], pm.w[:, :k] = centroids, self._t(cov), self._t(weights[:, :k])
            pm.repack()
            if self.use_em:
                self.iterations["em"].append(self._em(pm, state, k, has.clone(), em_max_iteration))
        self.pm = pm
        self.mu, self.var, self.w = (t.cpu().numpy().copy() for t in (pm.mu, pm.var, pm.w))
        self.mu_old, self.var_old, self.w_old = (t.cpu().numpy().copy() for t in self._old)
        return self.mu, self.var, self.w

    def _t(self, a):
        return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(self.dev)

    class _WorkingSet:
        """The frames of the states that are still iterating.  Once fewer than half of the frames in hand belong to
        running states the set is compacted (one gather), so the tail of the slowest states does not re-read the corpus."""

        def __init__(self, fitter, state):
            self.f, self.X, self.state = fitter, fitter.X, state
            self.n = int(state.shape[0])

        def maybe_compact(self, active):
            need = int(self.f._n_loc[active].sum().item())
            if need * 2 < self.n:
                idx = torch.nonzero(active[self.state.long()]).squeeze(1)
                self.X, self.state, self.n = self.X[idx].contiguous(), self.state[idx].contiguous(), int(idx.numel())

    CHECK_EVERY = 8          # iterations between two host reads of the running-state count

    def _kmeans(self, state, k, cen, cov0, active, max_iteration):
        """kmeans.py:178-193 for every pool at once: a pool stops at the iteration where its new centroids equal the
        previous ones (np.array_equal, :190) and keeps those.  -> (centroids [G,k,D] device, counts [G,k] numpy, its[G])"""
        G, D = cen.shape[0], self.D
        cen = cen.contiguous().clone()
        shift = cen.clone().reshape(G * k, D)
        fused = engine.pool_stats_supported(G, k, D)          # one pass over X per iteration instead of two
        f64 = self.f64 and engine.pool_stats_supported_f64(G, k, D)
        ws = self._WorkingSet(self, state)
        its = torch.zeros(G, dtype=torch.int64, device=self.dev)
        cnt = torch.zeros((G, k), dtype=torch.float64, device=self.dev)
        zero = torch.zeros((G * k, 2 * D + 3), dtype=torch.float64, device=self.dev)
        for it in range(max(max_iteration, 1)):
            if ws.n and fused:
                st = engine.pool_kmeans_stats(ws.X, ws.state, cen, cov0, shift, f64=f64)
            elif ws.n:
                am = engine.kmeans_assign(ws.X, ws.state, cen, cov0)
                st = engine.group_stats(ws.X, (ws.state.long() * k + am.long()).to(torch.int32), G * k, shift)
            else:
                st = zero.clone()
            st = _allreduce(st, self.group)
            engine.fit_kmeans_update(st, shift, cen, active, its, it, cnt)
            if it % self.CHECK_EVERY == self.CHECK_EVERY - 1:
                n_run = int(active.sum().item())
                if self.verbose:
                    print(f"[fit] k-means k={k} iteration {it}: {n_run} pools running, {ws.n} frames in hand", file=sys.stderr)
                if not n_run:
                    break
                ws.maybe_compact(active)
        return cen, cnt.cpu().numpy(), its.cpu().numpy()

    def _em(self, pm, state, k, active, max_iteration):
        """GMM.em (hmm_state.py:122-159) over the first k components of every state at once; a state stops at the
        iteration where the reference's loop would (np.allclose of the new parameters with the previous ones) and is
        frozen from then on.  Updates pm.mu / pm.var / pm.w in place.  -> its[G] (last iteration index of every state)"""
        G, M, D = pm.mu.shape
        ws = self._WorkingSet(self, state)
        its = torch.zeros(G, dtype=torch.int64, device=self.dev)
        row = 2 * D + 3
        fused = engine.pool_stats_supported(G, k, D)
        f64 = self.f64 and engine.pool_stats_supported_f64(G, k, D)
        zero = torch.zeros((G * M, row), dtype=torch.float64, device=self.dev)
        lc = engine.gauss_consts(pm.var, pm.w)
        for it in range(max_iteration):
            if not ws.n:
                st = zero.clone()
            elif fused:
                st = engine.pool_em_stats(ws.X, ws.state, pm, k, lc=lc, f64=f64)
            else:
                st = engine.gmm_stats(ws.X, ws.state, pm, k)
            st = _allreduce(st, self.group)
            engine.fit_em_update(st, pm, k, self._old, active, its, it, lc)       # M-step, stop test, freeze, new lc
            if it % self.CHECK_EVERY == self.CHECK_EVERY - 1:
                n_run = int(active.sum().item())
                if self.verbose and (it < 64 or it % 256 == 255 or not n_run):
                    print(f"[fit] EM k={k} iteration {it}: {n_run} states running, {ws.n} frames in hand", file=sys.stderr)
                if not n_run:
                    break
                ws.maybe_compact(active)
        pm.repack()
        self.converged.setdefault("em", []).append(~active.cpu().numpy())
        return its.cpu().numpy()

    # ------------------------------------------------------------------------------------------ phase 3: re-alignment
    def align(self):
        """align_gmm_states (kmeans.py:196-205): Viterbi re-alignment with the trained states -> seg_starts [B, S]."""
        if self.fused_align and self.dev.type != "cuda":
            self.fused_align = False                           # host-logic runs on the numpy stand-in: K1 + K2 path
        if self.fused_align:
            from .trainer import IsolatedTrainer
            tr = IsolatedTrainer(self.X, self.off.astype(np.int32), self.words, self.pm, self.trans, self.S, 1, fused=None,
                                 peer=False)
        if self.fused_align and tr.fused:
            seg = torch.zeros((self.B, self.S), dtype=torch.int32, device=self.dev)
            tr.fused_pass(seg_start=seg, stats=False)
            starts = seg.cpu().numpy().astype(np.int64)
        else:                                                  # K1 on all states + K2 per word graph
            E = engine.gmm_score(self.X, self.pm, want_E=True)["E"]
            graphs = [StateGraph(self.trans[w], np.arange(self.S) + w * self.S) for w in range(self.W)]
            fl = torch.empty(max(self.N, 1), dtype=torch.int32, device=self.dev)
            engine.viterbi(E, self.off.astype(np.int32), GraphBatch(graphs, self.dev), utt_graph=self.words.astype(np.int32),
                           want_path=False, frame_label=fl)
            lab = fl[:self.N].long() - (self.word_of_frame * self.S)
            cnt = torch.bincount((self.utt_of * self.S + lab).clamp_(min=0), weights=(lab >= 0).double(),
                                 minlength=self.B * self.S).reshape(self.B, self.S).cpu().numpy().astype(np.int64)
            starts = np.zeros((self.B, self.S), np.int64)
            starts[:, 1:] = np.cumsum(cnt, axis=1)[:, :-1]
        self.final_starts = starts
        return starts

    def fit(self):
        """-> list of W word models (mu[S,M,D], var[S,M,D], w[S,M], trans[S,S])."""
        self.skmeans()
        self.fit_states()
        self.align()
        S = self.S
        return [(self.mu[w * S:(w + 1) * S], self.var[w * S:(w + 1) * S], self.w[w * S:(w + 1) * S], self.trans[w])
                for w in range(self.W)]
