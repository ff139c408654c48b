"""Batched device-side operators over libsrb200 (torch is used for memory and streams only)."""
import os
import numpy as np
import torch

from . import lib as _l


def device():
    if not torch.cuda.is_available():
        raise RuntimeError("srb200 needs a CUDA device (B200, sm_100a); there is no CPU path")
    return torch.device("cuda", torch.cuda.current_device())


def fma_peak(packed=True, iters=4096):
    """Measured FP32 FMA rate (FMA/s) of the current device with plain FFMA or packed fma.rn.f32x2."""
    import ctypes
    ms, n = ctypes.c_float(), ctypes.c_double()
    _l.check(_l.load().srb_fma_peak(int(packed), iters, ctypes.byref(ms), ctypes.byref(n), _l.stream_ptr()))
    return n.value / (ms.value * 1e-3)


class PackedModel:
    """R tied states x M diagonal Gaussians in the scorer's layout (see include/srb200.h, srb_pack_model).

    Keeps the float64 parameters (mu, var, w: [R, M, D] / [R, M]) on the device next to the packed fp32 copy;
    `repack()` refreshes the latter after an M-step wrote the former."""

    def __init__(self, mu, var, w, shift=None):
        dev = device()
        mu = torch.as_tensor(np.asarray(mu, np.float64)) if not torch.is_tensor(mu) else mu
        var = torch.as_tensor(np.asarray(var, np.float64)) if not torch.is_tensor(var) else var
        w = torch.as_tensor(np.asarray(w, np.float64)) if not torch.is_tensor(w) else w
        assert mu.dim() == 3 and var.shape == mu.shape and w.shape == mu.shape[:2]
        self.R, self.M, self.D = (int(v) for v in mu.shape)
        self.Mp = (self.M + 3) // 4 * 4
        self.mu = mu.to(dev, torch.float64).contiguous()
        self.var = var.to(dev, torch.float64).contiguous()
        self.w = w.to(dev, torch.float64).contiguous()
        if shift is None:
            shift = self.mu.reshape(-1, self.D).mean(dim=0)
        self.shift64 = torch.as_tensor(shift).to(dev, torch.float64).contiguous()
        L = _l.load()
        self.ab = torch.empty(L.srb_packed_ab_floats(self.R, self.M, self.D), dtype=torch.float32, device=dev)
        self.c = torch.empty(L.srb_packed_c_floats(self.R, self.M), dtype=torch.float32, device=dev)
        self.shift = torch.empty(self.D, dtype=torch.float32, device=dev)
        n_tc = int(L.srb_tc_model_floats(self.R, self.Mp, self.D))      # TF32-split copy for the tensor-core classifier
        self.tcb = torch.empty(n_tc, dtype=torch.float32, device=dev) if n_tc > 0 else None
        self.repack()

    def repack(self):
        self.version = getattr(self, "version", 0) + 1               # caches of derived packings key on it
        _l.check(_l.load().srb_pack_model(_l.ptr(self.mu), _l.ptr(self.var), _l.ptr(self.w), _l.ptr(self.shift64),
                                          self.R, self.M, self.D, _l.ptr(self.ab), _l.ptr(self.c),
                                          _l.ptr(self.shift), _l.stream_ptr()))
        if self.tcb is not None:
            _l.check(_l.load().srb_pack_model_tc(_l.ptr(self.ab), _l.ptr(self.c), self.R, self.Mp, self.D, _l.ptr(self.tcb),
                                                 _l.stream_ptr()))
        return self

    def tensor_core_ok(self, W, S):
        """True when srb_classify_tc covers this model (D = 39, 4 Gaussians per state, <= 256 Gaussians)."""
        return self.tcb is not None and bool(_l.load().srb_classify_tc_supported(self.D, W, S, self.Mp))

    def scorer_tensor_core_ok(self):
        """True when srb_gmm_score_tc covers this model: D = 39, 4*J Gaussians per state (a state is scored as J
        sub-states of 4 -- the packed model [R, Mp/4, D, 4, 2] IS the packed model of R*J states x 4 -- and
        srb_neg_lse_groups merges them), at most 57 sub-states = 228 Gaussians (the packed model and the cost ring
        share 227 KB of shared memory)."""
        return (self.tcb is not None and self.Mp % 4 == 0 and
                bool(_l.load().srb_gmm_score_tc_supported(self.D, self.R * (self.Mp // 4), 4)))

    @classmethod
    def from_words(cls, word_models, shift=None):
        """word_models: list of (mu[S,M,D], cov[S,M,D], w[S,M], trans[S,S]) -> (PackedModel, trans[W,S,S] fp32)."""
        mu = np.concatenate([np.asarray(m[0], np.float64) for m in word_models], axis=0)
        var = np.concatenate([np.asarray(m[1], np.float64) for m in word_models], axis=0)
        w = np.concatenate([np.asarray(m[2], np.float64) for m in word_models], axis=0)
        pm = cls(mu, var, w, shift)
        trans = torch.as_tensor(np.stack([np.asarray(m[3], np.float64) for m in word_models])).to(pm.ab.device,
                                                                                               torch.float32)
        return pm, trans.contiguous()


def as_frames(X):
    """[N, D] float32 contiguous CUDA tensor (16-byte aligned by the torch allocator)."""
    if not torch.is_tensor(X):
        X = torch.as_tensor(np.ascontiguousarray(X, dtype=np.float32))
    return X.to(device(), torch.float32).contiguous()


def as_offsets(off):
    if not torch.is_tensor(off):
        off = torch.as_tensor(np.ascontiguousarray(off, dtype=np.int32))
    return off.to(device(), torch.int32).contiguous()


_scratch = {}                   # caller-owned workspaces of the kernels, grown on demand and reused (per device)

TC_SCORE_MIN_FRAMES = 4096      # below this the FFMA scorer's launch is cheaper than staging the packed model per CTA
_chunk_off_cache = {}


def score_chunk_offsets(N, dev):
    """int32[ceil(N/32)+1] partition of N frames into 32-frame chunks (srb_gmm_score_tc's work split), cached per N."""
    key = (int(N), str(dev))
    t = _chunk_off_cache.get(key)
    if t is None:
        if len(_chunk_off_cache) > 16:
            _chunk_off_cache.clear()
        t = torch.arange(0, N + 32, 32, dtype=torch.int64, device=dev).clamp_(max=N).to(torch.int32)
        if t.numel() >= 2 and int(t[-1]) == int(t[-2]):
            t = t[:-1].contiguous()
        _chunk_off_cache[key] = t
    return t


TC_MERGE_ROWS = 1 << 20         # frames per pass when sub-state costs are merged (bounds the scratch to 228 floats x 1 M)


def gmm_score_tc(X, pm, E=None):
    """K1 on the tensor cores (3xTF32, srb_gmm_score_tc): E[N,R] only.  States of 4*J Gaussians go through a scratch
    P[rows, R*J] of sub-state costs and srb_neg_lse_groups."""
    X = as_frames(X)
    N = X.shape[0]
    if E is None:
        E = torch.empty((N, pm.R), dtype=torch.float32, device=X.device)
    assert E.is_cuda and E.dtype == torch.float32 and tuple(E.shape) == (N, pm.R) and (N == 0 or E.stride(1) == 1)
    L, J = _l.load(), pm.Mp // 4
    if J <= 1:
        co = score_chunk_offsets(N, X.device)
        _l.check(L.srb_gmm_score_tc(_l.ptr(X), N, pm.D, _l.ptr(pm.tcb), _l.ptr(pm.shift), pm.R, pm.Mp, _l.ptr(co),
                                    co.numel() - 1, E.data_ptr(), E.stride(0) if N else pm.R, _l.stream_ptr()))
        return E
    RJ = pm.R * J
    P = torch.empty((min(N, TC_MERGE_ROWS), RJ), dtype=torch.float32, device=X.device)
    for r0 in range(0, N, TC_MERGE_ROWS):                    # r0 % 4 == 0: X[r0:] stays 16-byte aligned
        n = min(TC_MERGE_ROWS, N - r0)
        co = score_chunk_offsets(n, X.device)
        _l.check(L.srb_gmm_score_tc(X[r0:].data_ptr(), n, pm.D, _l.ptr(pm.tcb), _l.ptr(pm.shift), RJ, 4, _l.ptr(co),
                                    co.numel() - 1, P.data_ptr(), RJ, _l.stream_ptr()))
        _l.check(L.srb_neg_lse_groups(P.data_ptr(), n, pm.R, J, E[r0:].data_ptr(), E.stride(0), _l.stream_ptr()))
    return E


# frames per pass of gmm_best_tc.  Measured on config 5 (frames/s): 2^17 6.4e6, 2^18 6.9e6, 2^20 7.9e6, 2^22 8.4e6,
# one pass 8.5e6 -- the per-launch cost of staging the block's model outweighs keeping the scratch in L2.
TC_BEST_ROWS = int(os.environ.get("SRB_TC_BEST_ROWS", 1 << 22))


def gmm_best_tc(X, pm, best_cost=None, best_state=None):
    """Per-frame best state and its cost (srb_gmm_score's best_cost / best_state) for a model of ANY number of states
    on the tensor cores: blocks of 57 // J states are scored as sub-states of 4 Gaussians (srb_gmm_score_tc on the
    block's slice of the packed model, split once per model update and cached on pm) and folded into the running
    minimum by srb_merge_best.  -> (best_cost[N] float32, best_state[N] int32)"""
    X = as_frames(X)
    N, dev, L = X.shape[0], X.device, _l.load()
    J = pm.Mp // 4
    SB = 57 // J if J else 0
    if pm.D != 39 or pm.Mp % 4 or SB < 1 or not L.srb_gmm_score_tc_supported(pm.D, min(SB, pm.R) * J, 4):
        raise ValueError("gmm_best_tc: D = 39 and 4*J <= 228 Gaussians per state needed")
    SB = min(SB, pm.R)
    nb = (pm.R + SB - 1) // SB
    nf = int(L.srb_tc_model_floats(SB * J, 4, pm.D))
    cache = getattr(pm, "_tc_blocks", None)
    if cache is None or cache[0] != pm.version:
        tcb = torch.empty(nb * nf, dtype=torch.float32, device=dev)
        for b in range(nb):
            r0, rb = b * SB, min(SB, pm.R - b * SB)
            _l.check(L.srb_pack_model_tc(pm.ab[r0 * J * pm.D * 8:].data_ptr(), pm.c[r0 * pm.Mp:].data_ptr(), rb * J, 4,
                                         pm.D, tcb[b * nf:].data_ptr(), _l.stream_ptr()))
        cache = pm._tc_blocks = (pm.version, tcb)
    tcb = cache[1]
    bc = torch.empty(N, dtype=torch.float32, device=dev) if best_cost is None else best_cost
    bs = torch.empty(N, dtype=torch.int32, device=dev) if best_state is None else best_state
    P = torch.empty((min(N, TC_BEST_ROWS), SB * J), dtype=torch.float32, device=dev)
    st = _l.stream_ptr()
    for f0 in range(0, N, TC_BEST_ROWS):
        n = min(TC_BEST_ROWS, N - f0)
        co = score_chunk_offsets(n, dev)
        for b in range(nb):
            r0, rb = b * SB, min(SB, pm.R - b * SB)
            _l.check(L.srb_gmm_score_tc(X[f0:].data_ptr(), n, pm.D, tcb[b * nf:].data_ptr(), _l.ptr(pm.shift), rb * J, 4,
                                        _l.ptr(co), co.numel() - 1, P.data_ptr(), rb * J, st))
            _l.check(L.srb_merge_best(P.data_ptr(), n, rb, J, r0, int(b == 0), bc[f0:].data_ptr(), bs[f0:].data_ptr(), st))
    return bc, bs


def gmm_best_gemm(X, pm, best_cost=None, best_state=None):
    """Per-frame best state and its cost for a model of ANY number of states as one persistent tensor-core GEMM
    (srb_gmm_best_gemm: the frame tile's operand stays in tensor memory, the model streams past it, nothing but the
    result leaves the SM).  Mp a multiple of 16.  -> (best_cost[N] float32, best_state[N] int32)"""
    X = as_frames(X)
    N, dev, L = X.shape[0], X.device, _l.load()
    if not L.srb_gmm_best_gemm_supported(pm.D, pm.R, pm.Mp):
        raise ValueError("gmm_best_gemm: D = 39 and a multiple of 16 Gaussians per state (at most 160) needed")
    cache = getattr(pm, "_gemm_blocks", None)
    if cache is None or cache[0] != pm.version:
        blocks = torch.empty(int(L.srb_model_blocks_floats(pm.R, pm.Mp, pm.D)), dtype=torch.float32, device=dev)
        _l.check(L.srb_pack_model_blocks(_l.ptr(pm.ab), _l.ptr(pm.c), pm.R, pm.Mp, pm.D, _l.ptr(blocks), _l.stream_ptr()))
        cache = pm._gemm_blocks = (pm.version, blocks)
    bc = torch.empty(N, dtype=torch.float32, device=dev) if best_cost is None else best_cost
    bs = torch.empty(N, dtype=torch.int32, device=dev) if best_state is None else best_state
    _l.check(L.srb_gmm_best_gemm(_l.ptr(X), N, pm.D, _l.ptr(cache[1]), _l.ptr(pm.shift), pm.R, pm.Mp, _l.ptr(bc),
                                 _l.ptr(bs), _l.stream_ptr()))
    return bc, bs


TC_BEST_MIN_FRAMES = 1 << 16    # best-state-only requests of at least this many frames go to gmm_best_tc


def gmm_score(X, pm, want_E=True, want_best=False, want_comp=False, tensor=None):
    """K1.  -> dict(E[N,R], best_cost[N], best_state[N], comp[N,R,Mp]) for the requested outputs.
    tensor: None = the tcgen05 (3xTF32) scorer for E-only requests of >= TC_SCORE_MIN_FRAMES frames when it covers
    the model and for best-state-only requests of >= TC_BEST_MIN_FRAMES frames (any number of states), else the FFMA
    scorer; True / False force one."""
    X = as_frames(X)
    N = X.shape[0]
    assert X.shape[1] == pm.D
    dev = X.device
    out = {}
    e_only = want_E and not want_best and not want_comp
    if tensor is None:
        tensor = e_only and N >= TC_SCORE_MIN_FRAMES and pm.scorer_tensor_core_ok()
    best_only = want_best and not want_E and not want_comp
    if best_only and tensor is not False and pm.D == 39 and pm.Mp % 4 == 0 and 4 <= pm.Mp <= 228 and (
            tensor or N >= TC_BEST_MIN_FRAMES):
        if _l.load().srb_gmm_best_gemm_supported(pm.D, pm.R, pm.Mp):
            bc, bs = gmm_best_gemm(X, pm)                # one persistent GEMM, the model streams past the frame tile
        else:
            bc, bs = gmm_best_tc(X, pm)                  # block by block through the classifier kernel's scorer mode
        return {"best_cost": bc, "best_state": bs}
    if tensor:
        if not e_only:
            raise ValueError("the tensor-core scorer returns E only, or the best state only")
        return {"E": gmm_score_tc(X, pm)}
    E = torch.empty((N, pm.R), dtype=torch.float32, device=dev) if want_E else None
    bc = torch.empty(N, dtype=torch.float32, device=dev) if want_best else None
    bs = torch.empty(N, dtype=torch.int32, device=dev) if want_best else None
    comp = torch.empty((N, pm.R, pm.Mp), dtype=torch.float32, device=dev) if want_comp else None
    _l.check(_l.load().srb_gmm_score(_l.ptr(X), N, pm.D, _l.ptr(pm.ab), _l.ptr(pm.c), _l.ptr(pm.shift), pm.R, pm.Mp,
                                     _l.ptr(E), pm.R, _l.ptr(bc), _l.ptr(bs), _l.ptr(comp), _l.stream_ptr()))
    if want_E:
        out["E"] = E
    if want_best:
        out["best_cost"], out["best_state"] = bc, bs
    if want_comp:
        out["comp"] = comp
    return out


def classify(X, off, pm, trans, W, S, final=None, label=None, tensor=None):
    """Fused score + Viterbi over W isolated-word models (core.py:79-87).  -> (final[B,W] f32, label[B] i32).
    tensor: None = the tcgen05 (3xTF32) kernel when it covers the model, else the FFMA kernel; True / False force one."""
    X = as_frames(X)
    off = as_offsets(off)
    B = off.numel() - 1
    assert pm.R == W * S and tuple(trans.shape) == (W, S, S)
    if final is None:
        final = torch.empty((B, W), dtype=torch.float32, device=X.device)
    if label is None:
        label = torch.empty(B, dtype=torch.int32, device=X.device)
    if tensor is None:
        tensor = pm.tensor_core_ok(W, S)
    if tensor:
        _l.check(_l.load().srb_classify_tc(_l.ptr(X), _l.ptr(off), B, X.shape[0], pm.D, _l.ptr(pm.tcb),
                                           _l.ptr(pm.shift), W, S, pm.Mp, _l.ptr(trans), _l.ptr(final),
                                           _l.ptr(label), _l.stream_ptr()))
    else:
        _l.check(_l.load().srb_classify(_l.ptr(X), _l.ptr(off), B, X.shape[0], pm.D, _l.ptr(pm.ab), _l.ptr(pm.c),
                                        _l.ptr(pm.shift), W, S, pm.Mp, _l.ptr(trans), _l.ptr(final), _l.ptr(label),
                                        _l.stream_ptr()))
    return final, label


class HostClassifier:
    """Host-facing isolated-word classifier (core.py:63-94's inner loop): numpy / pinned host frames in, labels out.

    The batch is cut into chunks of whole utterances; chunk k+1 is copied host->device on one stream while chunk k
    is scored and decoded on the other (two device staging buffers), and the labels come back with one D2H copy
    per chunk.  All device work goes through srb_classify."""

    def __init__(self, pm, trans, W, S, max_utts=8192, max_frames=8192 * 100):
        dev = device()
        self.pm, self.trans, self.W, self.S = pm, trans, W, S
        self.max_utts, self.max_frames = int(max_utts), int(max_frames)
        self.streams = [torch.cuda.Stream(dev) for _ in range(2)]
        self.xbuf = [torch.empty((self.max_frames, pm.D), dtype=torch.float32, device=dev) for _ in range(2)]
        self.obuf = [torch.empty(self.max_utts + 1, dtype=torch.int32, device=dev) for _ in range(2)]
        self.final = [torch.empty((self.max_utts, W), dtype=torch.float32, device=dev) for _ in range(2)]
        self.label = [torch.empty(self.max_utts, dtype=torch.int32, device=dev) for _ in range(2)]
        self.off_pin = [torch.empty(self.max_utts + 1, dtype=torch.int32).pin_memory() for _ in range(2)]
        self.done = [None, None]

    def _chunks(self, off):
        """Greedy split of utterances into chunks bounded by max_utts / max_frames."""
        B = len(off) - 1
        u = 0
        while u < B:
            ue = min(B, u + self.max_utts)
            lim = int(off[u]) + self.max_frames
            if int(off[ue]) > lim:
                ue = int(np.searchsorted(off, lim, side="right")) - 1
                if ue <= u:
                    raise ValueError(f"utterance {u} has more than max_frames={self.max_frames} frames")
            yield u, ue
            u = ue

    def run(self, X, off, labels_out=None, final_out=None, sync=True):
        """X: [N, D] float32 host tensor/array (pinned for full speed); off: [B+1] int32 -> labels [B] int32 (host).
        The labels are valid on return (sync=True); with sync=False the caller must synchronise the current stream
        before reading them."""
        Xt = X if torch.is_tensor(X) else torch.from_numpy(np.ascontiguousarray(X, dtype=np.float32))
        offn = off.numpy() if torch.is_tensor(off) else np.asarray(off)
        B = len(offn) - 1
        if labels_out is None:
            labels_out = torch.empty(B, dtype=torch.int32).pin_memory()
        cur = torch.cuda.current_stream()
        start = torch.cuda.Event()
        start.record(cur)
        L = _l.load()
        for k, (u0, u1) in enumerate(self._chunks(offn)):
            s = k & 1
            st = self.streams[s]
            nb, f0, f1 = u1 - u0, int(offn[u0]), int(offn[u1])
            if self.done[s] is not None:
                self.done[s].synchronize()          # pinned offsets / staging buffers of slot s are free again
            np.subtract(offn[u0:u1 + 1], f0, out=self.off_pin[s].numpy()[:nb + 1])
            with torch.cuda.stream(st):
                st.wait_event(start)
                self.xbuf[s][:f1 - f0].copy_(Xt[f0:f1], non_blocking=True)
                self.obuf[s][:nb + 1].copy_(self.off_pin[s][:nb + 1], non_blocking=True)
                if self.pm.tensor_core_ok(self.W, self.S):
                    par = (_l.ptr(self.pm.tcb),)
                    fn = L.srb_classify_tc
                else:
                    par = (_l.ptr(self.pm.ab), _l.ptr(self.pm.c))
                    fn = L.srb_classify
                _l.check(fn(self.xbuf[s].data_ptr(), self.obuf[s].data_ptr(), nb, f1 - f0, self.pm.D,
                            *par, _l.ptr(self.pm.shift), self.W, self.S,
                            self.pm.Mp, _l.ptr(self.trans), self.final[s].data_ptr(),
                            self.label[s].data_ptr(), st.cuda_stream))
                labels_out[u0:u1].copy_(self.label[s][:nb], non_blocking=True)
                if final_out is not None:
                    final_out[u0:u1].copy_(self.final[s][:nb], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(st)
                self.done[s] = ev
        for s in range(2):
            if self.done[s] is not None:
                cur.wait_event(self.done[s])
        if sync:
            for s in range(2):                      # the D2H copies are asynchronous: the HOST must wait as well before
                if self.done[s] is not None:        # it reads labels_out (sync=False: the caller synchronises the stream)
                    self.done[s].synchronize()
        return labels_out


class HostBestState:
    """Host-facing likelihood sweep (BASELINE config 5): pinned host frames in, per-frame best state and cost out.
    The frames are cut into chunks; chunk k+1 is copied host->device on one stream while srb_gmm_best_gemm runs on
    chunk k on the other, results come back with one D2H copy per chunk."""

    def __init__(self, pm, chunk_frames=1 << 20):
        dev = device()
        self.pm, self.chunk = pm, int(chunk_frames) // 4 * 4            # chunk starts stay 16-byte aligned
        self.streams = [torch.cuda.Stream(dev) for _ in range(2)]
        self.xbuf = [torch.empty((self.chunk, pm.D), dtype=torch.float32, device=dev) for _ in range(2)]
        self.bc = [torch.empty(self.chunk, dtype=torch.float32, device=dev) for _ in range(2)]
        self.bs = [torch.empty(self.chunk, dtype=torch.int32, device=dev) for _ in range(2)]
        gmm_best_gemm(self.xbuf[0][:4], pm)                             # packs the model's blocks once

    def run(self, Xh, cost_out, state_out):
        """Xh [N, D] float32 pinned host tensor; cost_out [N] float32 / state_out [N] int32 pinned host tensors."""
        N = Xh.shape[0]
        cur = torch.cuda.current_stream()
        start = torch.cuda.Event()
        start.record(cur)
        done = [None, None]
        for k, f0 in enumerate(range(0, N, self.chunk)):
            s, n = k & 1, min(self.chunk, N - f0)
            st = self.streams[s]
            with torch.cuda.stream(st):
                st.wait_event(start)
                self.xbuf[s][:n].copy_(Xh[f0:f0 + n], non_blocking=True)
                gmm_best_gemm(self.xbuf[s][:n], self.pm, self.bc[s][:n], self.bs[s][:n])
                cost_out[f0:f0 + n].copy_(self.bc[s][:n], non_blocking=True)
                state_out[f0:f0 + n].copy_(self.bs[s][:n], non_blocking=True)
                done[s] = torch.cuda.Event()
                done[s].record(st)
        for ev in done:
            if ev is not None:
                cur.wait_event(ev)
        return cost_out, state_out


# ------------------------------------------------------------------------------------------------ K2 / K4
from .graph import GraphBatch, StateGraph  # noqa: E402


def _utt_layout(off, gb, utt_graph):
    """Per-utterance workspace offsets: ws_off[u] = sum_{v<u} T_v * R_{graph(v)} (int64), path capacity offsets."""
    offn = off.cpu().numpy().astype(np.int64) if torch.is_tensor(off) else np.asarray(off, np.int64)
    T = np.diff(offn)
    R = np.diff(gb.state_off.astype(np.int64))
    nn = np.diff(gb.host["g_nes_off"].astype(np.int64))
    if utt_graph is None:
        Ru, Nu = np.full(len(T), R[0]), np.full(len(T), nn[0])
    else:
        ug = np.asarray(utt_graph)
        Ru, Nu = R[ug], nn[ug]
    ws = np.zeros(len(T) + 1, np.int64)
    np.cumsum(T * Ru, out=ws[1:])
    po = np.zeros(len(T) + 1, np.int64)
    np.cumsum(T + 2 * Nu + 2, out=po[1:])
    return ws, po


class ViterbiResult:
    """Device outputs of srb_viterbi_graph plus host-side views in the reference's formats."""

    def __init__(self, best_cost, best_end, path_state, path_frame, path_len, path_off, costs, ws_off, bp):
        self.best_cost, self.best_end = best_cost, best_end
        self.path_state, self.path_frame, self.path_len, self.path_off = path_state, path_frame, path_len, path_off
        self.costs, self.ws_off, self.bp = costs, ws_off, bp

    def paths(self):
        """list of reversed [[state, frame], ...] int64 arrays (decode.py:142-146 format)."""
        ps, pf = self.path_state.cpu().numpy(), self.path_frame.cpu().numpy()
        ln = self.path_len.cpu().numpy()
        out = []
        for u, n in enumerate(ln):
            if n < 0:
                raise RuntimeError("path capacity exhausted")
            o = self.path_off[u]
            out.append(np.stack([ps[o:o + n], pf[o:o + n]], axis=1).astype(np.int64))
        return out

    def words(self, state_word, max_words):
        """Word strings straight from the device paths (main.py:59-67; srb_path_words).
        state_word: int32 per packed state (graphs back to back), -1 for a non-emitting state.
        -> list of int lists.  Raises when a string is longer than max_words."""
        dev = self.path_len.device
        B = self.path_len.numel()
        sw = torch.as_tensor(np.ascontiguousarray(state_word, dtype=np.int32)).to(dev)
        words = torch.empty((B, max(max_words, 1)), dtype=torch.int32, device=dev)
        nw = torch.empty(B, dtype=torch.int32, device=dev)
        _l.check(_l.load().srb_path_words(_l.ptr(self.path_state), _l.ptr(self.path_off_dev), _l.ptr(self.path_len), B,
                                          _l.ptr(self.utt_graph_dev), _l.ptr(self.gb["g_state_off"]), _l.ptr(sw),
                                          max_words, _l.ptr(words), _l.ptr(nw), _l.stream_ptr()))
        wn, nn = words.cpu().numpy(), nw.cpu().numpy()
        if (nn > max_words).any():
            raise RuntimeError("srb_path_words: a decoded string has more than max_words words")
        return [wn[u, :nn[u]].tolist() for u in range(B)]


def viterbi_launch(gb, E, offd, B, ug, ws_d, bp, costs, bc, be, po_d, ps, pf, pl, frame_label, banded=None):
    """srb_viterbi_banded when every graph of the batch is in banded form (banded=False forces the generic kernel;
    both give identical outputs), else srb_viterbi_graph."""
    p, L = _l.ptr, _l.load()
    use_banded = gb.banded if banded is None else (banded and gb.banded)
    end_frame = gb.dev.get("end_frame")                   # end points off the last frame: generic kernel only
    if end_frame is not None:
        if banded:
            raise ValueError("end points off the last frame run through srb_viterbi_graph (banded=False)")
        use_banded = False
    head = (p(E), E.stride(0), p(offd), B, p(ug), p(gb["g_state_off"]), p(gb["g_nes_off"]), p(gb["g_end_off"]),
            p(gb["emit"]), p(gb["pred_ptr"]), p(gb["pred_idx"]), p(gb["pred_cost"]), p(gb["nes_idx"]), p(gb["end_state"]))
    tail = (gb.Rmax, p(ws_d), p(bp), p(costs), p(bc), p(be), p(po_d), p(ps), p(pf), p(pl), p(frame_label),
            _l.stream_ptr())
    if use_banded:
        _l.check(L.srb_viterbi_banded(*head, p(gb["band"]), p(gb["codes"]), p(gb["entry_nes"]), p(gb["entry_cost"]),
                                      p(gb["g_entry_off"]), p(gb["entry_state"]), *tail))
    else:
        _l.check(L.srb_viterbi_graph(*head, p(end_frame), *tail))
    return use_banded


def viterbi(E, off, gb, utt_graph=None, want_costs=False, want_path=True, frame_label=None, banded=None):
    """K2.  E: [N, ldE] float32 CUDA emission costs; off: [B+1] int32; gb: GraphBatch."""
    dev = E.device
    offd = as_offsets(off)
    B = offd.numel() - 1
    ws, po = _utt_layout(off, gb, utt_graph)
    ws_d = torch.from_numpy(ws).to(dev)
    po_d = torch.from_numpy(po).to(dev)
    ug = None if utt_graph is None else torch.as_tensor(np.asarray(utt_graph, np.int32)).to(dev)
    bp = torch.empty(max(int(ws[-1]), 1), dtype=torch.uint8, device=dev)
    costs = torch.empty(max(int(ws[-1]), 1), dtype=torch.float32, device=dev) if want_costs else None
    bc = torch.empty(B, dtype=torch.float32, device=dev)
    be = torch.empty(B, dtype=torch.int32, device=dev)
    ps = torch.empty(max(int(po[-1]), 1), dtype=torch.int32, device=dev) if want_path else None
    pf = torch.empty(max(int(po[-1]), 1), dtype=torch.int32, device=dev) if want_path else None
    pl = torch.zeros(B, dtype=torch.int32, device=dev)
    viterbi_launch(gb, E, offd, B, ug, ws_d, bp, costs, bc, be, po_d, ps, pf, pl, frame_label, banded=banded)
    res = ViterbiResult(bc, be, ps, pf, pl, po, costs, ws, bp)
    res.frame_label = frame_label
    res.path_off_dev, res.utt_graph_dev, res.gb = po_d, ug, gb
    return res


def dtw_beam(dist, off, trans, beam=None):
    """dtw with the reference's beam pruning (decode.py:7-77; srb_dtw_beam) for a ragged batch of utterances against
    ONE template.  dist: [N, S] float32 CUDA local distances; trans: [S, S] (cost o -> i at [i, o], inf entries stay
    candidates); beam: None / inf = no pruning.
    -> (costs flat float32 [sum T*S] with costs[i, j] at ws[u] + j*S + i, ws int64[B+1], paths: list of [T-1, 2] int64
        or None, status int32[B]: path length, -2 where the reference's np.argmin gets an empty list (ValueError), -3
        where its walk back would leave the first column off row 0 (the reference does not terminate there))."""
    dev = dist.device
    S = int(dist.shape[1])
    offd = as_offsets(off)
    offn = offd.cpu().numpy().astype(np.int64)
    B = len(offn) - 1
    T = np.diff(offn)
    ws = np.zeros(B + 1, np.int64)
    np.cumsum(T * S, out=ws[1:])
    po = np.zeros(B + 1, np.int64)
    np.cumsum(np.maximum(T - 1, 0), out=po[1:])
    tr = torch.as_tensor(np.ascontiguousarray(np.asarray(trans, np.float64))).to(dev, torch.float32).contiguous()
    assert tuple(tr.shape) == (S, S)
    bp = torch.empty(max(int(ws[-1]), 1), dtype=torch.int16, device=dev)
    costs = torch.empty(max(int(ws[-1]), 1), dtype=torch.float32, device=dev)
    ps = torch.empty(max(int(po[-1]), 1), dtype=torch.int32, device=dev)
    pf = torch.empty(max(int(po[-1]), 1), dtype=torch.int32, device=dev)
    pl = torch.zeros(max(B, 1), dtype=torch.int32, device=dev)
    b = -1 if beam is None or beam == float("inf") else int(beam)
    ws_d, po_d = torch.from_numpy(ws).to(dev), torch.from_numpy(po).to(dev)    # named: they must outlive the launch
    _l.check(_l.load().srb_dtw_beam(_l.ptr(dist), dist.stride(0), _l.ptr(offd), B, _l.ptr(tr), S, b, _l.ptr(ws_d),
                                    _l.ptr(bp), _l.ptr(costs), _l.ptr(po_d), _l.ptr(ps), _l.ptr(pf), _l.ptr(pl),
                                    _l.stream_ptr()))
    ln = pl.cpu().numpy()[:B]
    psn, pfn = ps.cpu().numpy(), pf.cpu().numpy()
    paths = [np.stack([psn[po[u]:po[u] + ln[u]], pfn[po[u]:po[u] + ln[u]]], axis=1).astype(np.int64) if ln[u] >= 0
             else None for u in range(B)]
    return costs, ws, paths, ln


def _xi_part(nnz_total, dev):
    """Caller-owned workspace of the forward-backward kernels: srb_fb_xi_parts() rows of nnz_total doubles."""
    need = _l.load().srb_fb_xi_parts() * nnz_total
    key = (dev.index, "xi_part")
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.empty(need, dtype=torch.float64, device=dev)
        _scratch[key] = buf
    return buf


def forward_backward(E, off, gb, utt_graph=None, want_xi=False, banded=None):
    """K4.  -> (gamma flat float32 [sum T*R], ws_off, neg_log_z[B] float64, xi_sum[nnz] float64 or None).
    Left-to-right models of at most 32 states (gb.banded) run srb_forward_backward_banded, one lane per state with the
    utterances ordered by falling length; banded=False forces the generic CSR kernel (same outputs to rounding)."""
    dev = E.device
    offd = as_offsets(off)
    B = offd.numel() - 1
    if gb.n_nes_max:
        raise ValueError("forward_backward is defined for NES-free graphs only")
    ws, _ = _utt_layout(off, gb, utt_graph)
    ws_d = torch.from_numpy(ws).to(dev)
    ug = None if utt_graph is None else torch.as_tensor(np.asarray(utt_graph, np.int32)).to(dev)
    gamma = torch.empty(max(int(ws[-1]), 1), dtype=torch.float32, device=dev)
    nlz = torch.empty(B, dtype=torch.float64, device=dev)
    nnz_total = max(int(gb.edge_off[-1]), 1)
    xi = torch.zeros(nnz_total, dtype=torch.float64, device=dev) if want_xi else None
    xi_part = _xi_part(nnz_total, dev) if want_xi else None          # per-warp rows: edge posteriors without atomics
    fits = gb.banded and gb.Rmax <= 32
    if banded and not fits:
        raise ValueError("forward_backward: banded=True needs left-to-right graphs of at most 32 states")
    if fits and banded is not False:
        order = torch.argsort(offd[1:] - offd[:-1], descending=True, stable=True).to(torch.int32) if B > 1 else None
        _l.check(_l.load().srb_forward_backward_banded(
            _l.ptr(E), E.stride(0), _l.ptr(offd), B, _l.ptr(ug), _l.ptr(gb["g_state_off"]), _l.ptr(gb["g_end_off"]),
            _l.ptr(gb["emit"]), _l.ptr(gb["pred_ptr"]), _l.ptr(gb["end_state"]), _l.ptr(gb["band"]), _l.ptr(gb["codes"]),
            _l.ptr(order), gb.Rmax, _l.ptr(ws_d), _l.ptr(gamma), _l.ptr(nlz), _l.ptr(xi), _l.ptr(xi_part), nnz_total,
            _l.stream_ptr()))
        return gamma, ws, nlz, xi
    _l.check(_l.load().srb_forward_backward(
        _l.ptr(E), E.stride(0), _l.ptr(offd), B, _l.ptr(ug), _l.ptr(gb["g_state_off"]), _l.ptr(gb["g_end_off"]),
        _l.ptr(gb["emit"]), _l.ptr(gb["pred_ptr"]), _l.ptr(gb["pred_idx"]), _l.ptr(gb["pred_cost"]),
        _l.ptr(gb["pred_tgt"]), _l.ptr(gb["succ_ptr"]), _l.ptr(gb["succ_idx"]), _l.ptr(gb["succ_cost"]),
        _l.ptr(gb["end_state"]), gb.Rmax, gb.NNZmax, _l.ptr(ws_d), _l.ptr(gamma), _l.ptr(nlz), _l.ptr(xi),
        _l.ptr(xi_part), nnz_total, _l.stream_ptr()))
    return gamma, ws, nlz, xi


class ForwardBackwardPlan:
    """Pre-planned K4 launch for a fixed batch (offsets, graphs): offsets, workspace layout, the length-sorted utterance
    order and every output buffer live on the device and are built once -- an iteration of BaumWelchTrainer then
    enqueues the kernel without touching the host (forward_backward() re-derives all of this per call, which for 1 M
    utterances is tens of milliseconds of host time per iteration)."""

    def __init__(self, off, gb, utt_graph=None, want_xi=True):
        if gb.n_nes_max:
            raise ValueError("forward_backward is defined for NES-free graphs only")
        dev = gb["emit"].device
        self.gb = gb
        self.off = as_offsets(off)
        self.B = self.off.numel() - 1
        ws, _ = _utt_layout(off, gb, utt_graph)
        self.ws = ws
        self.ws_d = torch.from_numpy(ws).to(dev)
        self.ug = None if utt_graph is None else torch.as_tensor(np.asarray(utt_graph, np.int32)).to(dev)
        self.gamma = torch.empty(max(int(ws[-1]), 1), dtype=torch.float32, device=dev)
        self.nlz = torch.empty(max(self.B, 1), dtype=torch.float64, device=dev)
        self.nnz_total = max(int(gb.edge_off[-1]), 1)
        self.xi = torch.zeros(self.nnz_total, dtype=torch.float64, device=dev) if want_xi else None
        self.xi_part = _xi_part(self.nnz_total, dev) if want_xi else None
        self.banded = bool(gb.banded and gb.Rmax <= 32)
        self.order = None
        if self.banded and self.B > 1:
            self.order = torch.argsort(self.off[1:] - self.off[:-1], descending=True, stable=True).to(torch.int32)

    def run(self, E):
        gb, L, p = self.gb, _l.load(), _l.ptr
        if self.xi is not None:
            self.xi.zero_()
        if self.banded:
            _l.check(L.srb_forward_backward_banded(
                p(E), E.stride(0), p(self.off), self.B, p(self.ug), p(gb["g_state_off"]), p(gb["g_end_off"]), p(gb["emit"]),
                p(gb["pred_ptr"]), p(gb["end_state"]), p(gb["band"]), p(gb["codes"]), p(self.order), gb.Rmax, p(self.ws_d),
                p(self.gamma), p(self.nlz), p(self.xi), p(self.xi_part), self.nnz_total, _l.stream_ptr()))
        else:
            _l.check(L.srb_forward_backward(
                p(E), E.stride(0), p(self.off), self.B, p(self.ug), p(gb["g_state_off"]), p(gb["g_end_off"]), p(gb["emit"]),
                p(gb["pred_ptr"]), p(gb["pred_idx"]), p(gb["pred_cost"]), p(gb["pred_tgt"]), p(gb["succ_ptr"]),
                p(gb["succ_idx"]), p(gb["succ_cost"]), p(gb["end_state"]), gb.Rmax, gb.NNZmax, p(self.ws_d), p(self.gamma),
                p(self.nlz), p(self.xi), p(self.xi_part), self.nnz_total, _l.stream_ptr()))
        return self.gamma, self.ws, self.nlz, self.xi


# ------------------------------------------------------------------------------------------------ K5 / K6


def _partial(groups, D, dev):
    """Caller-owned scratch for the per-warp private slices (grown on demand, reused across calls)."""
    L = _l.load()
    need = L.srb_stats_parts() * groups * L.srb_stats_row(D)
    key = (dev.index, "partial")
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.empty(need, dtype=torch.float32, device=dev)
        _scratch[key] = buf
    return buf


def _f64(a, dev):
    if torch.is_tensor(a):
        return a.to(dev, torch.float64).contiguous()
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)


def _i32(a, dev):
    if torch.is_tensor(a):
        return a.to(dev, torch.int32).contiguous()
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.int32)).to(dev)


def group_stats(X, label, K, shift, out=None, accumulate=False):
    """-> double [K, 2D+3] rows [S0, S1, S2, -, -] about shift[K, D]."""
    X = as_frames(X)
    dev, D = X.device, X.shape[1]
    label, shift = _i32(label, dev), _f64(shift, dev)
    assert shift.shape == (K, D)
    L = _l.load()
    if out is None:
        out = torch.zeros((K, L.srb_stats_row(D)), dtype=torch.float64, device=dev)
    _l.check(L.srb_group_stats(_l.ptr(X), X.shape[0], D, _l.ptr(label), K, _l.ptr(shift), _l.ptr(_partial(K, D, dev)),
                               _l.ptr(out), int(accumulate), _l.stream_ptr()))
    return out


def moments_from_stats(stats, shift, D, unbiased=True):
    """(count[K], mean[K,D], var[K,D]) from group_stats rows (calc_variance: unbiased, kmeans.py:12)."""
    S0 = stats[:, 0]
    S1, S2 = stats[:, 1:1 + D], stats[:, 1 + D:1 + 2 * D]
    m1 = S1 / S0[:, None]
    mean = shift + m1
    den = (S0 - 1.0) if unbiased else S0
    var = (S2 - S1 * m1) / den[:, None]
    return S0, mean, var


def gauss_consts(var, w):
    G, D = var.reshape(-1, var.shape[-1]).shape
    lc = torch.empty(G, dtype=torch.float64, device=var.device)
    _l.check(_l.load().srb_gauss_consts(_l.ptr(var), _l.ptr(w), G, D, _l.ptr(lc), _l.stream_ptr()))
    return lc


def gmm_stats(X, label, pm, ncomp, out=None, accumulate=False):
    """E-step statistics of frames hard-aligned to tied states of PackedModel pm -> double [R*M, 2D+3]."""
    X = as_frames(X)
    dev = X.device
    label = _i32(label, dev)
    L = _l.load()
    lc = gauss_consts(pm.var, pm.w)
    if out is None:
        out = torch.zeros((pm.R * pm.M, L.srb_stats_row(pm.D)), dtype=torch.float64, device=dev)
    _l.check(L.srb_gmm_stats(_l.ptr(X), X.shape[0], pm.D, _l.ptr(label), pm.R, pm.M, ncomp, _l.ptr(pm.mu),
                             _l.ptr(pm.var), _l.ptr(lc), _l.ptr(_partial(pm.R * pm.M, pm.D, dev)), _l.ptr(out),
                             int(accumulate), _l.stream_ptr()))
    return out


def kmeans_assign(X, group, centroids, var0):
    """K-means assignment of many pools at once (srb_kmeans_assign): group int32[N] (pool of every frame, < 0 = none),
    centroids [G, k, D], var0 [G, D] (the one variance row a pool uses, kmeans.py:183).  -> int32[N] cluster (-1 = none)"""
    X = as_frames(X)
    dev = X.device
    c, v = _f64(centroids, dev), _f64(var0, dev)
    G, k, D = (int(t) for t in c.shape)
    g = _i32(group, dev)
    out = torch.empty(X.shape[0], dtype=torch.int32, device=dev)
    _l.check(_l.load().srb_kmeans_assign(_l.ptr(X), X.shape[0], D, _l.ptr(g), _l.ptr(c), _l.ptr(v), G, k, _l.ptr(out),
                                         _l.stream_ptr()))
    return out


def pool_stats_supported(G, k, D):
    return k in (1, 2, 4, 8) and bool(_l.load().srb_pool_stats_supported(int(G), int(k), int(D)))


def _pool_partial(G, k, D, dev):
    need = _l.load().srb_pool_stats_partial_floats(int(G), int(k), int(D))
    key = (dev.index, "pool_partial")
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.empty(need, dtype=torch.float32, device=dev)
        _scratch[key] = buf
    return buf


def pool_stats_supported_f64(G, k, D):
    return k in (1, 2, 4) and bool(_l.load().srb_pool_stats_f64_supported(int(G), int(k), int(D)))


def _pool_partial64(G, k, D, dev):
    need = _l.load().srb_pool_stats_f64_partial_doubles(int(G), int(k), int(D))
    key = (dev.index, "pool_partial64")
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.empty(need, dtype=torch.float64, device=dev)
        _scratch[key] = buf
    return buf


def pool_em_stats(X, state, pm, k, lc=None, f64=False):
    """srb_pool_stats mode 0: E-step statistics of GMM.em for every state at once, frames hard-assigned to states
    (same rows as gmm_stats: double [R*M, 2D+3]).  lc: the model's constants when the caller keeps them current."""
    X = as_frames(X)
    dev = X.device
    state = _i32(state, dev)
    L = _l.load()
    if lc is None:
        lc = gauss_consts(pm.var, pm.w)
    out = torch.empty((pm.R * pm.M, L.srb_stats_row(pm.D)), dtype=torch.float64, device=dev)
    if f64:
        _l.check(L.srb_pool_stats_f64(_l.ptr(X), X.shape[0], pm.D, _l.ptr(state), pm.R, pm.M, int(k), 0, _l.ptr(pm.mu),
                                      _l.ptr(pm.var), _l.ptr(lc), None, _l.ptr(_pool_partial64(pm.R, k, pm.D, dev)),
                                      _l.ptr(out), None, 0, _l.stream_ptr()))
        return out
    _l.check(L.srb_pool_stats(_l.ptr(X), X.shape[0], pm.D, _l.ptr(state), pm.R, pm.M, int(k), 0, _l.ptr(pm.mu),
                              _l.ptr(pm.var), _l.ptr(lc), None, _l.ptr(_pool_partial(pm.R, k, pm.D, dev)), _l.ptr(out),
                              None, 0, _l.stream_ptr()))
    return out


def pool_kmeans_stats(X, state, centroids, var0, shift, want_argmin=False, f64=False):
    """srb_pool_stats mode 1: one K-means iteration for every pool at once -- assignment to the nearest of the pool's k
    centroids [G, k, D] under the pool's variance row var0 [G, D], moments of every cluster about shift [G*k, D].
    -> double [G*k, 2D+3] (and the int32[N] assignment)"""
    X = as_frames(X)
    dev = X.device
    state = _i32(state, dev)
    c, v, sh = _f64(centroids, dev), _f64(var0, dev), _f64(shift, dev)
    G, k, D = (int(t) for t in c.shape)
    L = _l.load()
    out = torch.empty((G * k, L.srb_stats_row(D)), dtype=torch.float64, device=dev)
    am = torch.empty(X.shape[0], dtype=torch.int32, device=dev) if want_argmin else None
    if f64:
        _l.check(L.srb_pool_stats_f64(_l.ptr(X), X.shape[0], D, _l.ptr(state), G, k, k, 1, _l.ptr(c), _l.ptr(v), None,
                                      _l.ptr(sh), _l.ptr(_pool_partial64(G, k, D, dev)), _l.ptr(out), _l.ptr(am), 0,
                                      _l.stream_ptr()))
    else:
        _l.check(L.srb_pool_stats(_l.ptr(X), X.shape[0], D, _l.ptr(state), G, k, k, 1, _l.ptr(c), _l.ptr(v), None, _l.ptr(sh),
                                  _l.ptr(_pool_partial(G, k, D, dev)), _l.ptr(out), _l.ptr(am), 0, _l.stream_ptr()))
    return (out, am) if want_argmin else out


def fit_em_update(stats, pm, k, old, active, its, it, lc):
    """srb_fit_em_update: M-step + np.allclose stop test of GMM.em for every active state, in place on pm.mu / pm.var /
    pm.w (double), old = [mu_old, var_old, w_old], active bool[G], its int64[G], lc double[G*M]."""
    _l.check(_l.load().srb_fit_em_update(_l.ptr(stats), pm.R, pm.M, int(k), pm.D, _l.ptr(pm.mu), _l.ptr(pm.var), _l.ptr(pm.w),
                                         _l.ptr(old[0]), _l.ptr(old[1]), _l.ptr(old[2]), _l.ptr(active), _l.ptr(its), int(it),
                                         _l.ptr(lc), _l.stream_ptr()))


def fit_kmeans_update(stats, shift, cen, active, its, it, cnt):
    """srb_fit_kmeans_update: centroid update + np.array_equal stop test for every active pool, in place on cen [G,k,D]."""
    G, k, D = (int(t) for t in cen.shape)
    _l.check(_l.load().srb_fit_kmeans_update(_l.ptr(stats), G, k, D, _l.ptr(shift), _l.ptr(cen), _l.ptr(active), _l.ptr(its),
                                             int(it), _l.ptr(cnt), _l.stream_ptr()))


def mstep(stats, pm, ncomp, var_floor=0.0):
    """In-place M-step on pm.mu / pm.var / pm.w (double), then refresh the packed fp32 copy."""
    _l.check(_l.load().srb_mstep(_l.ptr(stats), pm.R, pm.M, ncomp, pm.D, _l.ptr(pm.mu), _l.ptr(pm.var), _l.ptr(pm.w),
                                 float(var_floor), _l.stream_ptr()))
    return pm.repack()


def local_dist(X, Y, var=None, shared_var=False, want_dist=True, want_argmin=False):
    """kind 0 (var None): Euclidean norm; kind 1: mahalanobis with var[S,D] (or one shared row, kmeans.py:183)."""
    X = as_frames(X)
    dev, D = X.device, X.shape[1]
    Y = _f64(Y, dev)
    S = Y.shape[0]
    v = None if var is None else _f64(var, dev)
    out = torch.empty((X.shape[0], S), dtype=torch.float32, device=dev) if want_dist else None
    am = torch.empty(X.shape[0], dtype=torch.int32, device=dev) if want_argmin else None
    _l.check(_l.load().srb_local_dist(_l.ptr(X), X.shape[0], D, _l.ptr(Y), _l.ptr(v), S, 0 if var is None else 1,
                                      0 if shared_var else D, _l.ptr(out), S, _l.ptr(am), _l.stream_ptr()))
    return out, am


# ------------------------------------------------------------------------------------------------ lextree (f4)
class DeviceLexTree:
    """A flattened lextree (sr/langmodel/lextree.py FlatTree) resident on the device for srb_text_viterbi."""

    def __init__(self, flat):
        dev = device()
        self.flat = flat
        self.n_rows, self.n_levels, self.n_ends = int(flat.n_rows), int(flat.n_levels), int(len(flat.word_ends))
        self.val = torch.from_numpy(flat.val).to(dev)
        self.parent = torch.from_numpy(flat.parent).to(dev)
        self.level_nodes = torch.from_numpy(flat.level_nodes if len(flat.level_nodes) else np.zeros(1, np.int32)).to(dev)
        self.level_off = torch.from_numpy(flat.level_off).to(dev)
        self.word_ends = torch.from_numpy(flat.word_ends).to(dev)


def text_viterbi(codes, tree, dist=None):
    """srb_text_viterbi for a batch of strings against one tree.
    codes: list of int32 arrays = character codes of '*' + text (spellchecker.py:26); tree: DeviceLexTree;
    dist: optional [sum n_cols, n_rows] float32 matrix of a caller-supplied dist_fun.
    -> (best_cost float64 [B], list of node-index arrays in FORWARD order = the matched strings' nodes)."""
    dev = tree.val.device
    B = len(codes)
    n_cols = np.array([len(c) for c in codes], np.int64)
    if (n_cols < 2).any():
        raise ValueError("text_viterbi needs a non-empty text (the reference's walk back does not terminate on '')")
    x_off = np.zeros(B + 1, np.int64)
    np.cumsum(n_cols, out=x_off[1:])
    out_off = np.zeros(B + 1, np.int64)
    np.cumsum(n_cols * (tree.n_levels + 2), out=out_off[1:])
    x = torch.from_numpy(np.concatenate(codes).astype(np.int32)).to(dev)
    col_ws = torch.empty(B * 2 * tree.n_rows, dtype=torch.float32, device=dev)
    bp = torch.empty(int(x_off[-1]) * tree.n_rows, dtype=torch.uint8, device=dev)
    jump = torch.empty((int(x_off[-1]), 2), dtype=torch.int32, device=dev)
    best = torch.empty(B, dtype=torch.float32, device=dev)
    out = torch.empty(int(out_off[-1]), dtype=torch.int32, device=dev)
    out_len = torch.empty(B, dtype=torch.int32, device=dev)
    d = None if dist is None else dist.to(dev, torch.float32).contiguous()
    assert d is None or tuple(d.shape) == (int(x_off[-1]), tree.n_rows)
    x_off_d, bp_off_d = torch.from_numpy(x_off).to(dev), torch.from_numpy(x_off * tree.n_rows).to(dev)
    out_off_d = torch.from_numpy(out_off).to(dev)                   # named: they must outlive the launch
    _l.check(_l.load().srb_text_viterbi(
        _l.ptr(x), _l.ptr(x_off_d), B, _l.ptr(tree.val), _l.ptr(tree.parent), _l.ptr(tree.level_nodes),
        _l.ptr(tree.level_off), tree.n_levels, tree.n_rows, _l.ptr(tree.word_ends), tree.n_ends, _l.ptr(d), _l.ptr(col_ws),
        _l.ptr(bp), _l.ptr(bp_off_d), _l.ptr(jump), _l.ptr(best), _l.ptr(out), _l.ptr(out_off_d), _l.ptr(out_len),
        _l.stream_ptr()))
    ln, on = out_len.cpu().numpy(), out.cpu().numpy()
    if (ln == -3).any() or (ln == -2).any():
        raise RuntimeError("text_viterbi: the walk back does not reach column 1 (the reference loops forever here)")
    if (ln < 0).any():
        raise RuntimeError("text_viterbi: output capacity exhausted")
    return best.double().cpu().numpy(), [on[out_off[u]:out_off[u] + ln[u]][::-1].copy() for u in range(B)]
