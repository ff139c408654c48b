"""Dynamic-programming layer of the sr.recognition mirror (reference: sr/recognition/decode.py:7-146).

`dtw` and `decode_hmm_states` keep the reference's signatures, conventions (transitions[i, j] = cost j -> i,
inf = none; reversed [[state, frame], ...] paths that exclude the end cell) and warnings.  The recursion, the
back-pointers and the traceback run in srb_viterbi_graph (K2); emissions in srb_gmm_score (K1) or
srb_local_dist.  `decode_batch` is the batched form the trainers use (one launch for many utterances).
"""
import warnings
from math import isinf

import numpy as np
import torch

from . import _ops
from ._ops import StateGraph, engine
from .hmm_state import GMM, NES, mahalanobis


def euclidean(*args):
    """Default local distance of dtw / skmeans / kmeans (kmeans.py:111): ||a - b||_2, extra arguments ignored."""
    return np.linalg.norm(args[0] - args[1])


def _local_distances(Xd, X_host, y, dist_fun, variance):
    """[N, S] float32 device matrix of dist_fun(x_n, y_s[, variance_s]).  The two distance functions the reference
    uses run on the device; any other callable is evaluated by the caller's own Python function."""
    y = np.asarray(y, np.float64)
    if dist_fun is mahalanobis and variance is not None:
        return engine.local_dist(Xd, y, var=np.asarray(variance, np.float64))[0]
    if dist_fun is euclidean and variance is None:
        return engine.local_dist(Xd, y)[0]
    X_host = np.asarray(X_host)
    out = np.empty((X_host.shape[0], y.shape[0]), np.float32)
    for n in range(X_host.shape[0]):
        for s in range(y.shape[0]):
            out[n, s] = dist_fun(X_host[n], y[s]) if variance is None else dist_fun(X_host[n], y[s], variance[s])
    return torch.from_numpy(out).to(Xd.device)


def dtw(x, y, dist_fun, transitions, variance=None, beam=np.inf):
    """Min-cost alignment of x [T, D] to the template y [S, D] (decode.py:7-77).
    -> (costs [S, T] float64, path [T-1, 2] int64 reversed)."""
    col_count, row_count = len(x), len(y)
    assert col_count > 1 and row_count > 1
    Xd = _ops.to_device(np.asarray(x))
    dist = _local_distances(Xd, x, y, dist_fun, variance)
    off = np.array([0, col_count], np.int32)
    if not isinf(beam) or row_count > 255:
        # beam pruning (decode.py:42-48,62-68: -1 sentinels, pruned origins skipped by row 0 and +inf for the other
        # rows), and templates too long for the graph kernel's 8-bit back-pointer codes: srb_dtw_beam
        if not isinf(beam) and int(beam) != beam:
            raise TypeError("slice indices must be integers or None or have an __index__ method")   # sort_idx[beam:]
        costs, _, paths, status = engine.dtw_beam(dist, off, transitions, None if isinf(beam) else int(beam))
        if status[0] == -2:
            raise ValueError("attempt to get argmin of an empty sequence")        # np.argmin([]) at decode.py:53
        if status[0] == -3:
            raise RuntimeError("dtw: the walk back left the first column off row 0 (the reference indexes column -1 "
                               "from there and does not terminate)")
        costs = costs[:col_count * row_count].reshape(col_count, row_count).t().double().cpu().numpy()
        return costs, paths[0]
    graph = StateGraph(transitions, np.arange(row_count), end_states=[row_count - 1], keep_inf=True)
    res = _ops.viterbi_batch(dist, off, [graph], want_costs=True, want_path=True)
    costs = res.costs[:col_count * row_count].reshape(col_count, row_count).t().double().cpu().numpy()
    return costs, res.paths()[0]


class DecodeResult:
    """Batched decode output: per-utterance best cost / end state, paths, optional cost matrices and frame labels."""

    def __init__(self, vres, off, R_of_utt, emit_cols):
        self.vres, self.off, self.R_of_utt, self.emit_cols = vres, np.asarray(off, np.int64), R_of_utt, emit_cols
        self.best_cost = vres.best_cost.double().cpu().numpy()
        self.best_end = vres.best_end.cpu().numpy()

    def paths(self):
        return self.vres.paths()

    def costs(self, u):
        """costs[R, T] float64 of utterance u (needs want_costs)."""
        T, R = int(self.off[u + 1] - self.off[u]), int(self.R_of_utt[u])
        o = int(self.vres.ws_off[u])
        return self.vres.costs[o:o + T * R].reshape(T, R).t().double().cpu().numpy()

    def frame_labels(self):
        return self.vres.frame_label.cpu().numpy()


def _emission_matrix(Xd, X_host, state_lists):
    """Scores every distinct emitting state object once.  -> (E [N, C] device, {id(state): column})."""
    uniq, col = [], {}
    others = []
    for states in state_lists:
        for s in states:
            if type(s) == NES or id(s) in col:
                continue
            if isinstance(s, GMM):
                col[id(s)] = len(uniq)
                uniq.append(s)
            else:
                col[id(s)] = -2 - len(others)          # user-defined state: column filled from its own evaluate()
                others.append(s)
    N = Xd.shape[0]
    C = len(uniq) + len(others)
    E = torch.empty((N, max(C, 1)), dtype=torch.float32, device=Xd.device)
    if uniq:
        pm = _ops.pack_gmms(uniq)
        E[:, :len(uniq)] = engine.gmm_score(Xd, pm, want_E=True)["E"]
    for k, s in enumerate(others):
        vals = np.array([s.evaluate(xx) for xx in np.asarray(X_host)], np.float32)
        E[:, len(uniq) + k] = torch.from_numpy(vals).to(Xd.device)
        col[id(s)] = len(uniq) + k
    return E, col


def _graph_of(states, transitions, end_states, col):
    emit = np.array([-1 if type(s) == NES else col[id(s)] for s in states], np.int32)
    end_frames = None
    if isinstance(end_states, tuple):              # (states, frames): end points off the last frame
        end_states, end_frames = end_states
    g = StateGraph(transitions, emit, end_states=end_states, end_frames=end_frames)
    if g.has_nes_self_loop:
        raise NameError("a non-emitting state transits to itself in the same column (decode.py:120-121)")
    return g


def decode_batch(xs, graphs, utt_graph=None, want_costs=False, want_path=True, want_label=False):
    """Viterbi-decode many utterances in one K1 + one K2 launch.
    xs: list of [T, D];  graphs: list of (states, transitions, end_states or None);  utt_graph[u]: graph index."""
    X, off = _ops.cat_frames(xs)
    Xd = _ops.to_device(X)
    E, col = _emission_matrix(Xd, X, [g[0] for g in graphs])
    sg = [_graph_of(st, tr, None if ends is None else (ends if isinstance(ends, tuple) else list(ends)), col)
          for st, tr, ends in graphs]
    ug = None if utt_graph is None else np.asarray(utt_graph, np.int32)
    res = _ops.viterbi_batch(E, off, sg, utt_graph=ug, want_costs=want_costs, want_path=want_path,
                             want_label=want_label)
    R = np.array([g.R for g in sg])
    R_of = R[ug] if ug is not None else np.full(len(xs), R[0])
    return DecodeResult(res, off, R_of, col)


def decode_hmm_states(x, states, transitions, end_points=None):
    """Viterbi over an arbitrary state list with non-emitting states (decode.py:80-146).
    -> (costs [R, T] float64, path [L, 2] int64 reversed, end cell excluded, down to frame 0)."""
    T = len(x)
    ends = None
    if end_points is not None:
        frames = [int(e[1]) for e in end_points]
        for c in frames:
            if not -T <= c < T:
                raise IndexError(f"index {c} is out of bounds for axis 1 with size {T}")     # costs[end[0], end[1]]
        ends = [int(e[0]) for e in end_points]
        if any(c not in (-1, T - 1) for c in frames):
            ends = (ends, frames)                  # any [state, frame] (decode.py:126-134): srb_viterbi_graph end_frame
    res = decode_batch([np.asarray(x)], [(states, np.asarray(transitions), ends)], want_costs=True, want_path=True)
    if isinf(res.best_cost[0]):
        warnings.warn("decode_hmm_states: Cannot find a path when decoding sequence")
    return res.costs(0), res.paths()[0].reshape(-1, 2)


__all__ = ["dtw", "decode_hmm_states", "decode_batch", "euclidean"]
