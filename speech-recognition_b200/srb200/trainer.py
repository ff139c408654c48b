"""Device-resident batched training of isolated-word GMM-HMMs (BASELINE config 2).

The reference trains one word at a time (`HMM.fit(ys)`, hmm.py:57-95): Viterbi re-alignment of every utterance
with the current GMM states (align_gmm_states, kmeans.py:196-205), pooled frames per state, per-state EM
(GMM.em, hmm_state.py:122-159).  Here all words are trained together from one word-sorted corpus that stays in
HBM; one `em_iteration()` is

    score (K1, each word's block against its own S states) -> Viterbi + traceback (K2, frame labels)
    -> E-step statistics (K5) -> [allreduce over GPUs] -> M-step -> repack

Utterances are sharded over ranks; the only collective is one allreduce of the float64 statistics buffer
(`[W*S*M, 2D+3]` + total cost) per iteration, after which every rank applies the same M-step.
"""
import os

import numpy as np
import torch

from . import engine
from . import lib as _l
from .graph import GraphBatch, StateGraph


def shard_range(n_items, rank, world):
    """Contiguous, balanced [lo, hi) of items for `rank` (utterance sharding)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def sort_by_word(X, off, words):
    """Reorder utterances so that each word's frames are contiguous.  -> (X, off, words, order)
    X may be a CUDA tensor when the corpus is already word-sorted (it is then passed through untouched)."""
    words = np.asarray(words)
    order = np.argsort(words, kind="stable")
    lens = np.diff(off)
    if np.array_equal(order, np.arange(len(order))):
        return X, np.asarray(off), words, order
    if torch.is_tensor(X):
        X = X.cpu().numpy()
    new_off = np.zeros(len(off), np.int64)
    np.cumsum(lens[order], out=new_off[1:])
    idx = np.concatenate([np.arange(off[u], off[u + 1]) for u in order]) if len(order) else np.zeros(0, np.int64)
    return X[idx], new_off.astype(off.dtype), words[order], order


class PeerExchange:
    """Exchange buffers of the fused training tail (srb_em_push / srb_em_update): one device buffer per rank that every
    other process of the box maps through CUDA IPC, so that a rank's statistics are STORED into its peers' memory over
    NVLink by the kernel that forms them -- no collective call, no host in the loop."""

    def __init__(self, length, group=None):
        import ctypes
        import torch.distributed as dist
        L = _l.load()
        multi = dist.is_initialized() and dist.get_world_size(group) > 1
        self.world = dist.get_world_size(group) if multi else 1
        self.rank = dist.get_rank(group) if multi else 0
        self.length, self.epoch, self.opened = int(length), 0, []
        nbytes = int(L.srb_em_sync_bytes(self.world, self.length))
        if nbytes <= 0:
            raise _l.SrbError(-2, f"no exchange buffer layout for {self.world} ranks")
        own = ctypes.c_void_p()
        _l.check(L.srb_peer_alloc(nbytes, ctypes.byref(own)))
        self.own = own.value
        ptrs = [None] * self.world
        ptrs[self.rank] = self.own
        if multi:
            dev = engine.device()
            handle = (ctypes.c_uint8 * 64)()
            ok = L.srb_peer_export(self.own, handle) == 0
            mine = torch.tensor(list(handle) + [int(ok)], dtype=torch.uint8, device=dev)
            allh = [torch.empty_like(mine) for _ in range(self.world)]
            dist.all_gather(allh, mine, group=group)
            allh = [h.cpu().numpy() for h in allh]
            ok = all(int(h[64]) for h in allh)
            for r in range(self.world):
                if ok and r != self.rank:
                    p = ctypes.c_void_p()
                    hb = (ctypes.c_uint8 * 64)(*[int(v) for v in allh[r][:64]])
                    if L.srb_peer_open(hb, ctypes.byref(p)) == 0:
                        ptrs[r] = p.value
                        self.opened.append(p.value)
                    else:
                        ok = False
            flag = torch.tensor([int(ok)], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)          # every rank mapped every peer, or nobody
            if not int(flag.item()):
                self.close()
                raise _l.SrbError(-3, "CUDA IPC mapping of the peers' exchange buffers failed: " +
                                  L.srb_last_error_string().decode())
        self.ptr_array = (ctypes.c_void_p * self.world)(*ptrs)

    def close(self):
        L = _l.load()
        for p in self.opened:
            L.srb_peer_close(p)
        self.opened = []
        if getattr(self, "own", None):
            L.srb_peer_free(self.own)
            self.own = None


class ViterbiPlan:
    """Pre-planned K2 launch for a fixed batch (offsets, graphs): all buffers allocated once."""

    def __init__(self, off, gb, utt_graph=None, want_path=False, want_label=True, banded=None):
        self.banded = banded
        dev = gb["emit"].device
        self.gb = gb
        self.off = engine.as_offsets(off)
        self.B = self.off.numel() - 1
        offn = np.asarray(off.cpu() if torch.is_tensor(off) else off, dtype=np.int64)
        ws, po = engine._utt_layout(offn, gb, utt_graph)
        self.ws, self.po = ws, po
        self.ws_d = torch.from_numpy(ws).to(dev)
        self.ug = None if utt_graph is None else torch.as_tensor(np.asarray(utt_graph, np.int32)).to(dev)
        self.bp = torch.empty(max(int(ws[-1]), 1), dtype=torch.uint8, device=dev)
        self.best_cost = torch.empty(self.B, dtype=torch.float32, device=dev)
        self.best_end = torch.empty(self.B, dtype=torch.int32, device=dev)
        self.N = int(offn[-1])
        self.label = torch.empty(max(self.N, 1), dtype=torch.int32, device=dev) if want_label else None
        self.po_d = self.ps = self.pf = self.pl = None
        if want_path:
            self.po_d = torch.from_numpy(po).to(dev)
            self.ps = torch.empty(max(int(po[-1]), 1), dtype=torch.int32, device=dev)
            self.pf = torch.empty(max(int(po[-1]), 1), dtype=torch.int32, device=dev)
            self.pl = torch.zeros(self.B, dtype=torch.int32, device=dev)

    def run(self, E):
        engine.viterbi_launch(self.gb, E, self.off, self.B, self.ug, self.ws_d, self.bp, None, self.best_cost,
                              self.best_end, self.po_d, self.ps, self.pf, self.pl, self.label, banded=self.banded)
        return self


class IsolatedTrainer:
    """All W word models trained together on this rank's shard of a word-sorted corpus."""

    def __init__(self, X, off, words, pm, trans, S, ncomp, group=None, var_floor=1e-6, fused=None, peer=None):
        """X [N,D] float32, off [B+1], words [B] (numpy, this rank's shard); pm: PackedModel with W*S states;
        trans: [W,S,S] float64 numpy (decode.py:86-87 convention).
        peer: None = the two-launch tail over NVLink peer memory (srb_em_push / srb_em_update) when the fused pass runs
        and the peers' buffers can be mapped, else reduce + NCCL allreduce + M-step + packing; True / False force one."""
        X, off, words, self.order = sort_by_word(X if torch.is_tensor(X) else np.asarray(X), np.asarray(off), words)
        self.dev = engine.device()
        self.X = engine.as_frames(X)
        self.pm, self.S, self.ncomp, self.group, self.var_floor = pm, S, ncomp, group, var_floor
        self.W = pm.R // S
        self.words = words
        off = np.asarray(off, np.int64)
        self.N, self.B = int(off[-1]), len(off) - 1
        uptr = np.searchsorted(words, np.arange(self.W + 1), side="left")
        self.word_rows = [(int(off[uptr[w]]), int(off[uptr[w + 1]])) for w in range(self.W)]
        graphs = [StateGraph(trans[w], np.arange(S)) for w in range(self.W)]
        self.gb = GraphBatch(graphs, self.dev)
        self.plan = ViterbiPlan(off.astype(np.int32), self.gb, utt_graph=words.astype(np.int32))
        self.E = torch.empty((max(self.N, 1), S), dtype=torch.float32, device=self.dev)
        L = _l.load()
        self.row = L.srb_stats_row(pm.D)
        self.G = pm.R * pm.M
        self.stats = torch.zeros(self.G * self.row + 2, dtype=torch.float64, device=self.dev)   # + cost, utterances
        self.partial = torch.empty(L.srb_stats_parts() * self.G * self.row, dtype=torch.float32, device=self.dev)
        self.lc = torch.empty(self.G, dtype=torch.float64, device=self.dev)
        self.launches_per_iteration = self.W + 1 + 1 + 3 + 2 + 1 + 1
        # ---- fused pass (srb_em_isolated): one kernel scores, aligns and accumulates; needs left-to-right models
        # with skips <= 2 (what calc_transition_costs produces, kmeans.py:53-95) and utterances that fit a tile
        self.fused = False
        if fused is not False:
            self.fused = self._plan_fused(off.astype(np.int32), words.astype(np.int32), np.asarray(trans, np.float64))
            if fused is True and not self.fused:
                raise ValueError("fused training pass not applicable: " + self.fused_why)
        if self.fused:
            self.launches_per_iteration = 1 + 1 + 1 + 2 + 1 + 1   # memset+em+reduce, cost sum (2), mstep, pack
        self.peer = None
        if self.fused and peer is not False and pm.D == 39:
            try:
                self.peer = PeerExchange(self.G * self.row, group)
                self.launches_per_iteration = 2 + 1 + 1            # memset + em kernel, push, update
            except _l.SrbError:
                if peer:
                    raise
        elif peer:
            raise ValueError("the peer-memory tail needs the fused pass and D = 39")

    def _plan_fused(self, off32, words32, trans):
        import ctypes
        L = _l.load()
        S, W = self.S, self.W
        r, o = np.indices((S, S))
        if S > 8 or self.ncomp > 8 or np.isfinite(trans[:, (o > r) | (r - o > 2)]).any():
            self.fused_why = "transitions are not left-to-right with skips <= 2 (or S, ncomp > 8)"
            return False
        tiles = np.empty((max(self.B, 1), 4), np.int32)
        nt, tf = ctypes.c_int32(0), ctypes.c_int32(0)
        off32, words32 = np.ascontiguousarray(off32), np.ascontiguousarray(words32)
        rc = L.srb_em_plan(off32.ctypes.data, words32.ctypes.data, self.B, tiles.ctypes.data, ctypes.byref(nt),
                           ctypes.byref(tf))
        if rc != 0:
            self.fused_why = L.srb_last_error_string().decode()
            return False
        self.ntiles, self.tile_frames = int(nt.value), int(tf.value)
        self.tiles_host = tiles[:max(self.ntiles, 1)].copy()
        self.tiles = torch.from_numpy(self.tiles_host).to(self.dev)
        self._copy_stream = torch.cuda.Stream(self.dev)
        self.off_d = torch.from_numpy(off32).to(self.dev)
        self.trans_d = torch.from_numpy(trans.astype(np.float32)).to(self.dev).contiguous()
        self.em_partial = torch.empty(int(L.srb_em_partial_floats(W, S, self.pm.M, self.pm.D)), dtype=torch.float32,
                                      device=self.dev)
        self.best_cost = torch.empty(max(self.B, 1), dtype=torch.float32, device=self.dev)
        return True

    def em_iteration_from_host(self, Xh, chunks=8):
        """One iteration with the shard coming from PINNED host memory (the end-to-end path bench.py times): the
        shard is cut at tile boundaries into `chunks` pieces; piece k+1 is copied host->device on a copy stream
        while srb_em_isolated runs on piece k (statistics accumulate across the pieces)."""
        assert self.fused, "pipelined host path needs the fused pass"
        pm, p, L = self.pm, _l.ptr, _l.load()
        cur, cp = torch.cuda.current_stream(), self._copy_stream
        cp.wait_stream(cur)                                   # the previous iteration is done reading self.X
        body = self.stats[:self.G * self.row]
        cuts = np.linspace(0, self.ntiles, min(chunks, max(self.ntiles, 1)) + 1).astype(np.int64)
        for k in range(len(cuts) - 1):
            t0, t1 = int(cuts[k]), int(cuts[k + 1])
            if t1 <= t0:
                continue
            f0 = int(self.tiles_host[t0, 1])
            f1 = int(self.tiles_host[t1, 1]) if t1 < self.ntiles else self.N
            with torch.cuda.stream(cp):
                self.X[f0:f1].copy_(Xh[f0:f1], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cp)
            cur.wait_event(ev)
            _l.check(L.srb_em_isolated(
                p(self.X), self.N, pm.D, p(self.off_d), self.B, self.tiles[t0:].data_ptr(), t1 - t0, self.tile_frames,
                p(pm.ab), p(pm.c), p(pm.shift), self.W, self.S, pm.M, pm.Mp, p(self.trans_d), self.ncomp, p(pm.mu),
                p(self.em_partial), p(body), int(k > 0), p(self.best_cost), None, None, cur.cuda_stream))
        self.stats[self.G * self.row] = self.best_cost[:self.B].sum(dtype=torch.float64)
        self.stats[self.G * self.row + 1] = float(self.B)
        self._allreduce()                                      # the chunks accumulate through stats_out: collective tail
        self._mstep()
        return self.stats

    def fused_pass(self, seg_start=None, frame_label=None, stats=True, keep_slices=False):
        """srb_em_isolated over the shard: statistics body of self.stats (+ best_cost, optional boundaries/labels).
        keep_slices: leave the per-CTA slices for reduce() = srb_em_push (what em_iteration does with the peer tail)
        instead of folding them into self.stats here."""
        pm, p = self.pm, _l.ptr
        body = self.stats[:self.G * self.row]
        keep = stats and keep_slices and self.peer is not None
        _l.check(_l.load().srb_em_isolated(
            p(self.X), self.N, pm.D, p(self.off_d), self.B, p(self.tiles), self.ntiles, self.tile_frames, p(pm.ab),
            p(pm.c), p(pm.shift), self.W, self.S, pm.M, pm.Mp, p(self.trans_d), self.ncomp, p(pm.mu),
            p(self.em_partial) if stats else None, p(body) if stats else None, -1 if keep else 0, p(self.best_cost),
            p(seg_start), p(frame_label), _l.stream_ptr()))
        if stats and not keep:
            self.stats[self.G * self.row] = self.best_cost[:self.B].sum(dtype=torch.float64)
            self.stats[self.G * self.row + 1] = float(self.B)
        return self.stats

    def score(self):
        """K1: each word's contiguous block of frames against that word's S states -> E[N, S]."""
        pm, S, L, p = self.pm, self.S, _l.load(), _l.ptr
        st = _l.stream_ptr()
        blk = S * pm.Mp * pm.D * 2
        for w, (r0, r1) in enumerate(self.word_rows):
            if r1 > r0:
                _l.check(L.srb_gmm_score_rows(p(self.X), self.N, r0, r1 - r0, pm.D, pm.ab[w * blk:].data_ptr(),
                                              pm.c[w * S * pm.Mp:].data_ptr(), p(pm.shift), S, pm.Mp, p(self.E), S,
                                              None, None, None, st))
        return self.E

    def align(self):
        """K1 per word block + K2: frame labels (tied state per frame) and per-utterance best cost."""
        self.score()
        self.plan.run(self.E)
        return self.plan.label

    def accumulate(self):
        pm, L, p = self.pm, _l.load(), _l.ptr
        st = _l.stream_ptr()
        _l.check(L.srb_gauss_consts(p(pm.var), p(pm.w), self.G, pm.D, p(self.lc), st))
        body = self.stats[:self.G * self.row]
        _l.check(L.srb_gmm_stats(p(self.X), self.N, pm.D, p(self.plan.label), pm.R, pm.M, self.ncomp, p(pm.mu),
                                 p(pm.var), p(self.lc), p(self.partial), p(body), 0, st))
        self.stats[self.G * self.row] = self.plan.best_cost.sum(dtype=torch.float64)
        self.stats[self.G * self.row + 1] = float(self.B)
        return self.stats

    def reduce(self):
        """The one exchange of an iteration: the sum of the statistics over the ranks.  Peer tail: srb_em_push folds
        the per-CTA slices and stores the result into every rank's exchange buffer over NVLink; else NCCL allreduce."""
        if self.peer is not None:
            pe = self.peer
            pe.epoch += 1
            pm = self.pm
            _l.check(_l.load().srb_em_push(_l.ptr(self.em_partial), _l.load().srb_em_parts(), self.W, self.S, pm.M,
                                           self.ncomp, pm.D, _l.ptr(self.best_cost), self.B, pe.ptr_array, pe.world,
                                           pe.rank, pe.epoch, _l.stream_ptr()))
            return self.stats
        return self._allreduce()

    def _allreduce(self):
        if self.group is not None or (torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1):
            torch.distributed.all_reduce(self.stats, group=self.group)
        return self.stats

    def update(self):
        """Replicated M-step + repacking.  Peer tail: srb_em_update (waits for the peers' flags, sums the slots in rank
        order, M-step and both packings, one CTA per Gaussian)."""
        if self.peer is None:
            return self._mstep()
        pm, pe, p = self.pm, self.peer, _l.ptr
        _l.check(_l.load().srb_em_update(pe.own, pe.world, pe.epoch, pm.R, pm.M, pm.Mp, self.ncomp, pm.D, p(pm.mu),
                                         p(pm.var), p(pm.w), p(pm.shift64), float(self.var_floor), p(pm.ab), p(pm.c),
                                         p(pm.tcb), p(self.stats), _l.stream_ptr()))
        pm.version += 1                                        # caches of derived packings key on it

    def _mstep(self):
        pm = self.pm
        _l.check(_l.load().srb_mstep(_l.ptr(self.stats), pm.R, pm.M, self.ncomp, pm.D, _l.ptr(pm.mu), _l.ptr(pm.var),
                                     _l.ptr(pm.w), float(self.var_floor), _l.stream_ptr()))
        pm.repack()

    def em_iteration(self):
        """One segmental-K-means + EM pass.  Returns the device statistics buffer (last two entries: summed
        Viterbi cost and utterance count over all ranks) without synchronising."""
        if self.fused:
            self.fused_pass(keep_slices=True)
        else:
            self.align()
            self.accumulate()
        self.reduce()
        self.update()
        return self.stats

    def mean_cost(self):
        return float(self.stats[-2] / self.stats[-1])


class BaumWelchTrainer:
    """Forward-backward (soft-count) EM over this rank's shard of a word-sorted corpus (BASELINE config 3).

    The reference has no Baum-Welch; the iteration is defined as the soft version of IsolatedTrainer's:
        score (K1, per word block) -> forward-backward (K4: gamma, xi, log Z) -> soft statistics (srb_gmm_stats_soft)
        -> [allreduce over GPUs] -> M-step (means / variances / weights of the first ncomp components as in
        hmm_state.py:138-148 with gamma-weighted posteriors; transitions a[o->r] = xi[o->r] / sum_r' xi[o->r'])
    All words must share the state count S (continuous_speech.py:27-28 makes the same assumption)."""

    def __init__(self, X, off, words, pm, trans, S, ncomp, group=None, var_floor=1e-6, gmin=1e-7,
                 update_transitions=True, tensor=None):
        X, off, words, self.order = sort_by_word(X if torch.is_tensor(X) else np.asarray(X), np.asarray(off), words)
        self.dev = engine.device()
        self.X = engine.as_frames(X)
        self.pm, self.S, self.ncomp, self.group = pm, S, ncomp, group
        self.var_floor, self.gmin, self.update_transitions = var_floor, gmin, update_transitions
        self.W = pm.R // S
        self.words = np.asarray(words).astype(np.int32)
        off = np.asarray(off, np.int64)
        self.off = off.astype(np.int32)
        self.N, self.B = int(off[-1]), len(off) - 1
        uptr = np.searchsorted(words, np.arange(self.W + 1), side="left")
        self.word_rows = [(int(off[uptr[w]]), int(off[uptr[w + 1]])) for w in range(self.W)]
        self._trans0 = np.array(trans, dtype=np.float64)
        self.fb_plan = None
        self._build_graphs()
        self.E = torch.empty((max(self.N, 1), S), dtype=torch.float32, device=self.dev)
        L = _l.load()
        self.row = L.srb_stats_row(pm.D)
        self.G = pm.R * pm.M
        self.nnz = int(self.gb.edge_off[-1])
        # tensor-core scoring (srb_gmm_score_tc, 4 Gaussians per column group): a state of Mp Gaussians is scored as
        # J = Mp/4 sub-states whose costs srb_neg_lse_groups merges.  tensor: None = when covered and worth the
        # per-CTA model staging, True = required, False = FFMA scorer.
        self.J = pm.Mp // 4 if pm.Mp % 4 == 0 else 0
        ok = self.J > 0 and bool(L.srb_gmm_score_tc_supported(pm.D, S * max(self.J, 1), 4))
        if tensor and not ok:
            raise ValueError("BaumWelchTrainer: the tensor-core scorer does not cover this model")
        self.tensor = bool(ok and (tensor or (tensor is None and self.N >= 8 * engine.TC_SCORE_MIN_FRAMES)))
        use_gemm = bool(self.tensor and pm.Mp % 16 == 0 and S * pm.Mp <= 160 and
                        L.srb_gmm_best_gemm_supported(pm.D, pm.R, pm.Mp) and not os.environ.get("SRB_BW_NO_GEMM"))
        if self.tensor and not use_gemm:
            self.tc_floats = int(L.srb_tc_model_floats(S * self.J, 4, pm.D))
            self.tcb = torch.empty(self.W * self.tc_floats, dtype=torch.float32, device=self.dev)
            self.P = self.E if self.J == 1 else torch.empty((max(self.N, 1), S * self.J), dtype=torch.float32,
                                                            device=self.dev)
        # one-launch tensor-core scoring (srb_gmm_score_gemm, rows mode): a word's states are one 160-column block of
        # the persistent GEMM kernel, tiles of 128 frames never straddle words; E is written directly (no sub-state
        # scratch, no merge pass).  Needs Mp a multiple of 16 and S * Mp <= 160.
        self.gemm = use_gemm
        if self.gemm:
            self.blocks = torch.empty(int(L.srb_model_blocks_floats_sb(pm.R, pm.Mp, pm.D, S)), dtype=torch.float32,
                                      device=self.dev)
            descs, self.word_tiles = [], [0]
            for w, (r0, r1) in enumerate(self.word_rows):
                f0 = np.arange(r0, r1, 128, dtype=np.int64)
                d = np.zeros((len(f0), 4), np.int32)
                d[:, 0], d[:, 1], d[:, 2] = f0, np.minimum(128, r1 - f0), w
                descs.append(d)
                self.word_tiles.append(self.word_tiles[-1] + len(f0))
            tiles = np.concatenate(descs) if descs else np.zeros((0, 4), np.int32)
            self.score_tiles = torch.from_numpy(np.ascontiguousarray(tiles if len(tiles) else np.zeros((1, 4), np.int32))
                                                ).to(self.dev)
        # [statistics | xi per edge | sum of -log Z, utterances]
        self.stats = torch.zeros(self.G * self.row + self.nnz + 2, dtype=torch.float64, device=self.dev)
        self.partial = torch.empty(L.srb_stats_parts() * S * pm.M * self.row, dtype=torch.float32, device=self.dev)
        self.lc = torch.empty(self.G, dtype=torch.float64, device=self.dev)

    def _build_graphs(self):
        self.graphs = [StateGraph(self._trans0[w], np.arange(self.S)) for w in range(self.W)]
        self.gb = GraphBatch(self.graphs, self.dev)

    @property
    def trans(self):
        """[W, S, S] float64 transition costs (decode.py:86-87 convention) read back from the device graphs, where
        srb_update_transitions re-estimates them in place."""
        pc = self.gb["pred_cost"].cpu().numpy().astype(np.float64)
        out = np.full((self.W, self.S, self.S), np.inf)
        for w, g in enumerate(self.graphs):
            e0 = int(self.gb.edge_off[w])
            out[w][g.pred_tgt, g.pred_idx] = pc[e0:e0 + g.nnz]
        return out

    def em_iteration_from_host(self, Xh):
        """One iteration with the shard coming from PINNED host memory (the end-to-end path bench.py times): every
        word's block of frames is copied host->device on a copy stream while the blocks before it are scored."""
        cur = torch.cuda.current_stream()
        if not hasattr(self, "_copy_stream"):
            self._copy_stream = torch.cuda.Stream(self.dev)
        cp = self._copy_stream
        cp.wait_stream(cur)                                   # the previous iteration is done reading self.X
        ready = []
        for r0, r1 in self.word_rows:
            with torch.cuda.stream(cp):
                if r1 > r0:
                    self.X[r0:r1].copy_(Xh[r0:r1], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cp)
            ready.append(ev)
        self.e_step(ready=ready)
        self.accumulate()
        self.reduce()
        self.update()
        return self.stats

    def e_step(self, ready=None):
        pm, S, L, p = self.pm, self.S, _l.load(), _l.ptr
        st = _l.stream_ptr()
        blk = S * pm.Mp * pm.D * 2
        RJ = S * self.J if self.tensor else 0
        if self.gemm:
            _l.check(L.srb_pack_model_blocks_sb(p(pm.ab), p(pm.c), pm.R, pm.Mp, pm.D, S, p(self.blocks), st))
            spans = [(0, self.word_tiles[-1])] if ready is None else list(zip(self.word_tiles[:-1], self.word_tiles[1:]))
            for k, (t0, t1) in enumerate(spans):
                if ready is not None:
                    torch.cuda.current_stream().wait_event(ready[k])      # this word's frames have arrived
                if t1 > t0:
                    _l.check(L.srb_gmm_score_gemm(p(self.X), self.N, pm.D, p(self.blocks), p(pm.shift), pm.R, pm.Mp, S,
                                                  self.score_tiles[t0:].data_ptr(), t1 - t0, p(self.E), S, st))
        for w, (r0, r1) in enumerate(self.word_rows):
            if self.gemm:
                break
            if ready is not None:
                torch.cuda.current_stream().wait_event(ready[w])
            if r1 <= r0:
                continue
            ab_w, c_w = pm.ab[w * blk:].data_ptr(), pm.c[w * S * pm.Mp:].data_ptr()
            ra = min(r1, (r0 + 3) & ~3) if self.tensor else r1        # the tensor-core scorer reads X in 16-byte units
            if ra > r0:
                _l.check(L.srb_gmm_score_rows(p(self.X), self.N, r0, ra - r0, pm.D, ab_w, c_w, p(pm.shift), S, pm.Mp,
                                              p(self.E), S, None, None, None, st))
            if r1 > ra:
                tcb_w = self.tcb[w * self.tc_floats:]
                _l.check(L.srb_pack_model_tc(ab_w, c_w, RJ, 4, pm.D, tcb_w.data_ptr(), st))
                co = engine.score_chunk_offsets(r1 - ra, self.dev)
                _l.check(L.srb_gmm_score_tc(self.X[ra:].data_ptr(), r1 - ra, pm.D, tcb_w.data_ptr(), p(pm.shift), RJ, 4,
                                            p(co), co.numel() - 1, self.P[ra:].data_ptr(), RJ, st))
                if self.J > 1:
                    _l.check(L.srb_neg_lse_groups(self.P[ra:].data_ptr(), r1 - ra, S, self.J, self.E[ra:].data_ptr(), S,
                                                  st))
        if self.fb_plan is None:                               # offsets, layout, order and buffers: once
            self.fb_plan = engine.ForwardBackwardPlan(self.off, self.gb, utt_graph=self.words, want_xi=True)
        self.gamma, _, self.nlz, self.xi = self.fb_plan.run(self.E)
        return self.gamma

    def accumulate(self):
        pm, S, L, p = self.pm, self.S, _l.load(), _l.ptr
        st = _l.stream_ptr()
        _l.check(L.srb_gauss_consts(p(pm.var), p(pm.w), self.G, pm.D, p(self.lc), st))
        body = self.stats[:self.G * self.row]
        body.zero_()
        mu, var = pm.mu.reshape(-1, pm.D), pm.var.reshape(-1, pm.D)
        for w, (r0, r1) in enumerate(self.word_rows):
            if r1 > r0:
                g0 = w * S * pm.M
                _l.check(L.srb_gmm_stats_soft(p(self.X), self.N, r0, r1 - r0, pm.D, p(self.gamma), S, pm.M, self.ncomp,
                                              mu[g0:].data_ptr(), var[g0:].data_ptr(), self.lc[g0:].data_ptr(),
                                              float(self.gmin), p(self.partial), body[g0 * self.row:].data_ptr(), 0,
                                              st))
        o = self.G * self.row
        self.stats[o:o + self.nnz] = self.xi[:self.nnz]
        fin = torch.isfinite(self.nlz)
        self.stats[o + self.nnz] = torch.where(fin, self.nlz, torch.zeros_like(self.nlz)).sum()
        self.stats[o + self.nnz + 1] = fin.sum().double()
        return self.stats

    def reduce(self):
        if self.group is not None or (torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1):
            torch.distributed.all_reduce(self.stats, group=self.group)
        return self.stats

    def update(self):
        pm = self.pm
        _l.check(_l.load().srb_mstep(_l.ptr(self.stats), pm.R, pm.M, self.ncomp, pm.D, _l.ptr(pm.mu), _l.ptr(pm.var),
                                     _l.ptr(pm.w), float(self.var_floor), _l.stream_ptr()))
        pm.repack()
        if self.update_transitions:
            # a[o -> r] = xi[o -> r] / sum_r' xi[o -> r'], rewritten in place in the device graphs (no host round trip)
            o, gb, p = self.G * self.row, self.gb, _l.ptr
            xi = self.stats[o:o + self.nnz]
            _l.check(_l.load().srb_update_transitions(
                xi.data_ptr(), p(gb["succ_ptr"]), p(gb["succ_edge"]), p(gb["band_slot"]) if gb.banded else None,
                int(gb.state_off[-1]), p(gb["pred_cost"]), p(gb["succ_cost"]), p(gb["band"]) if gb.banded else None,
                _l.stream_ptr()))

    def em_iteration(self):
        self.e_step()
        self.accumulate()
        self.reduce()
        self.update()
        return self.stats

    def mean_neg_log_z(self):
        return float(self.stats[-2] / self.stats[-1])
