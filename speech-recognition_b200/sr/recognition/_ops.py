"""Device plumbing shared by the sr.recognition mirror: every numeric step of the hot path goes through
srb200.engine (libsrb200.so, sm_100a); this module only shapes the reference's Python objects (lists of ragged
utterances, GMM state objects, dense transition matrices) into the batched device form and back.

There is no CPU path here: `engine.device()` raises without a CUDA device and `lib.load()` raises without the .so.
"""
import numpy as np
import torch

from srb200 import engine
from srb200.graph import GraphBatch, StateGraph


def cat_frames(seqs):
    """list of [T_r, D] arrays -> (X [N, D] float32 C-contiguous, off [B+1] int32)."""
    lens = np.fromiter((len(s) for s in seqs), dtype=np.int64, count=len(seqs))
    off = np.zeros(len(seqs) + 1, np.int64)
    np.cumsum(lens, out=off[1:])
    if len(seqs) == 0:
        return np.zeros((0, 0), np.float32), off.astype(np.int32)
    X = np.ascontiguousarray(np.concatenate([np.asarray(s, dtype=np.float32).reshape(len(s), -1) for s in seqs]))
    return X, off.astype(np.int32)


def to_device(data):
    """numpy [N, D] (any float) or CUDA tensor -> float32 CUDA tensor."""
    return engine.as_frames(data)


def pack_gmms(gmms):
    """list of GMM state objects (hmm_state.GMM) -> engine.PackedModel with one tied state per object.
    Objects with fewer components than the widest one are padded with weight-0 Gaussians (C = -inf)."""
    R = len(gmms)
    M = max(g.n_gaussians for g in gmms)
    D = int(np.asarray(gmms[0].dists[0].mean).shape[0])
    mu = np.zeros((R, M, D))
    var = np.ones((R, M, D))
    w = np.zeros((R, M))
    for r, g in enumerate(gmms):
        for m in range(g.n_gaussians):
            mu[r, m] = g.dists[m].mean
            var[r, m] = g.dists[m].diag_cov()
            w[r, m] = g.w[m]
    return engine.PackedModel(mu, var, w)


def group_moments(Xd, label, K, shift=None):
    """Per-group frame count, mean and UNBIASED variance (np.cov, kmeans.py:12) of rows of Xd by label.
    Two passes when no shift is given: the first finds the means, the second takes moments about them so the
    variance does not cancel in FP32.  -> numpy (count[K], mean[K,D], var[K,D]) float64."""
    D = Xd.shape[1]
    lab = label if torch.is_tensor(label) else torch.from_numpy(np.ascontiguousarray(label, dtype=np.int32))
    lab = lab.to(Xd.device, torch.int32)
    if shift is None:
        z = torch.zeros((K, D), dtype=torch.float64, device=Xd.device)
        st = engine.group_stats(Xd, lab, K, z)
        shift_t = st[:, 1:1 + D] / st[:, :1]
        shift_t = torch.where(torch.isfinite(shift_t), shift_t, torch.zeros_like(shift_t))
    else:
        shift_t = torch.as_tensor(shift, dtype=torch.float64).to(Xd.device).reshape(K, D)
    st = engine.group_stats(Xd, lab, K, shift_t)
    with np.errstate(all="ignore"):
        cnt, mean, var = engine.moments_from_stats(st, shift_t, D, unbiased=True)
    return cnt.cpu().numpy(), mean.cpu().numpy(), var.cpu().numpy()


def segment_labels(off, seg_starts, n_segments):
    """Frame labels of segment_data's slicing (kmeans.py:41-50): frames [seg_starts[r,s], seg_starts[r,s+1]) of
    utterance r belong to segment s, the last segment runs to the end.  -> int32 [N]"""
    off = np.asarray(off, np.int64)
    T = np.diff(off)
    st = np.asarray(seg_starts, np.int64)
    ends = np.concatenate([st[:, 1:], T[:, None]], axis=1)
    lens = np.clip(ends - st, 0, None)
    lab = np.repeat(np.tile(np.arange(n_segments, dtype=np.int32), len(T)), lens.ravel())
    if lab.shape[0] != int(off[-1]):
        # non-monotone starts (cannot come from a left-to-right alignment): fall back to explicit slices
        lab = np.full(int(off[-1]), -1, np.int32)
        for r in range(len(T)):
            for s in range(n_segments):
                lab[off[r] + st[r, s]: off[r] + ends[r, s]] = s
    return lab


def starts_from_labels(frame_label, off, n_segments, state_off=None):
    """get_segments_from_path (kmeans.py:98-108) for every utterance at once: the path visits every frame but the
    end cell once, so the start of segment s is the number of frames aligned to segments < s.
    frame_label: int [N] (state index, optionally offset by the utterance's graph offset state_off[B])."""
    off = np.asarray(off, np.int64)
    B = len(off) - 1
    T = np.diff(off)
    u_of = np.repeat(np.arange(B), T)
    lab = np.asarray(frame_label, np.int64)
    if state_off is not None:
        lab = lab - np.asarray(state_off, np.int64)[u_of]
    ok = lab >= 0
    cnt = np.bincount(u_of[ok] * n_segments + lab[ok], minlength=B * n_segments).reshape(B, n_segments)
    starts = np.zeros((B, n_segments), dtype=int)
    starts[:, 1:] = np.cumsum(cnt, axis=1)[:, :-1]
    return starts


def viterbi_batch(E, off, graphs, utt_graph=None, want_costs=False, want_path=True, want_label=False):
    """K2 over a list of StateGraphs.  -> engine.ViterbiResult (frame_label attribute when asked)."""
    gb = GraphBatch(graphs, E.device)
    fl = torch.empty(max(E.shape[0], 1), dtype=torch.int32, device=E.device) if want_label else None
    return engine.viterbi(E, off, gb, utt_graph=utt_graph, want_costs=want_costs, want_path=want_path, frame_label=fl)


__all__ = ["cat_frames", "to_device", "pack_gmms", "group_moments", "segment_labels", "starts_from_labels",
           "viterbi_batch", "StateGraph", "GraphBatch", "engine"]
