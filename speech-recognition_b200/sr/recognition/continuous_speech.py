"""Connected-word layer of the sr.recognition mirror (reference: sr/recognition/continuous_speech.py:13-179).

`build_state_sequences` builds the same flat state list / dense transition matrix / end indices.
`continuous_train` runs the embedded Viterbi training loop with ONE batched scoring launch and ONE batched
graph-Viterbi launch per outer iteration for all utterances (one forced-alignment chain graph per distinct label
sequence), then the per-state split / K-means / EM of the reference on device-resident pools.  The tied-state lookup
of the transition update follows the reference's `modelidx_state_map` semantics exactly (see continuous_train).
"""
import copy
import os
import pickle
import warnings

import numpy as np
import torch

from . import _ops
from .decode import decode_batch
from .hmm_state import NES


def build_state_sequences(hmms, label_matrix):
    """Flat graph [NES, layer 0 words, NES, layer 1 words, NES, ...] (continuous_speech.py:13-53).
    Every word of layer k is entered from NES k (cost 0) and leaves to NES k+1 (cost 0); inside a word block the
    word's own transitions apply.  -> (states, transitions [R, R], end-state indices of the last layer)"""
    S = len(hmms[0].gmm_states)                       # all models share the state count (asserted below)
    R = 1 + sum(len(hmms[l].gmm_states) for labels in label_matrix for l in labels) + len(label_matrix)
    trans = np.full((R, R), np.inf)
    seq = [NES()]
    entry = 0                                          # index of the NES feeding the current layer
    last_ends = []
    for labels in label_matrix:
        starts, ends = [], []
        for l in labels:
            o = len(seq)
            assert S == len(hmms[l].gmm_states)
            seq += hmms[l].gmm_states
            trans[o:o + S, o:o + S] = hmms[l].transitions
            starts.append(o)
            ends.append(o + S - 1)
        seq.append(NES())
        exit_ = len(seq) - 1
        trans[starts, entry] = 0
        trans[exit_, ends] = 0
        entry, last_ends = exit_, ends
    return seq, trans, last_ends


def _pool_segments(path, is_nes):
    """The reference's pooling state machine (continuous_speech.py:90-106, quirks A12) on one forward path:
    -> list of (state_row, start_frame, end_frame).  A segment opens at the first point whose state emits, closes
    at the first later point with a different row AND a later frame (that point's frame is excluded and the point
    does not open a segment itself); the last open segment is never flushed.

    The loop runs over RUNS of equal rows, not over points: inside a run the machine can only (1) close the segment
    a previous run left open -- at the run's first point whose frame lies past the segment's start, which is the
    second point when the open state began on this very frame (a one-frame state followed by the same-column hop
    through a NES, decode.py:109-111) -- and then (2) reopen at the following point.  Frames never decrease along
    a path, so (1) is a binary search."""
    r, c = np.asarray(path[:, 0]), np.asarray(path[:, 1])
    n = len(r)
    if n == 0:
        return []
    run_first = np.flatnonzero(np.concatenate([[True], r[1:] != r[:-1]]))
    run_end = np.concatenate([run_first[1:], [n]])
    out = []
    start = cur = None
    for p, e in zip(run_first.tolist(), run_end.tolist()):
        ri = int(r[p])
        if start is not None:
            if ri == cur:
                continue
            q = p + int(np.searchsorted(c[p:e], start, side="right"))     # first point of the run with frame > start
            if q >= e:
                continue
            out.append((cur, start, int(c[q])))
            start = cur = None
            p = q + 1
        if p < e and not is_nes[ri]:
            start, cur = int(c[p]), ri
    return out


def continuous_train(data, models, label_seqs, output_path, n_gaussians=4, n_segments=5, max_iteration=1000, *,
                     state_lookup="reference"):
    """Embedded training on connected utterances (continuous_speech.py:56-179).

    state_lookup="reference" (default) reproduces the reference's transition update exactly: `modelidx_state_map`
    holds the state objects of the FIRST deep copy for the whole run (continuous_speech.py:66-71), and the pools are
    a dict keyed by the CURRENT copy's state objects, which hash by `id` (kept by deepcopy) and compare by VALUE
    (GMM.__eq__, np.allclose).  So from the second outer iteration on a state's transitions are re-estimated only
    while its freshly trained mixture still equals the one the first iteration produced; otherwise the reference
    warns "No MFCC data for state" and the arcs keep their previous costs.  state_lookup="current" looks the states
    up in the current copy instead (every state with data gets its arcs re-estimated in every iteration)."""
    assert state_lookup in ("reference", "current")
    old_models = models
    new_models = copy.deepcopy(models)
    modelidx_state_map = {i: list(m.gmm_states) for i, m in enumerate(new_models)}     # :66-71, never refreshed
    X, off = _ops.cat_frames(data)
    Xd = _ops.to_device(X)
    off64 = off.astype(np.int64)
    for it in range(max_iteration):
        print('=' * 25)
        print('Continuous training iteration:', it)
        # ---- alignment: one chain graph per distinct label sequence, all utterances in one launch
        keys, graphs, utt_graph = {}, [], []
        for labels in label_seqs:
            k = tuple(labels)
            if k not in keys:
                keys[k] = len(graphs)
                seq, trans, ends = build_state_sequences(new_models, [[l] for l in labels])
                graphs.append((seq, trans, ends))
            utt_graph.append(keys[k])
        res = decode_batch(data, graphs, utt_graph=utt_graph, want_path=True)
        paths = res.paths()
        # ---- pooling per tied state: a dict keyed by the state objects like the reference's gmm_data (:76,101-104);
        # its insertion order = first flush, which fixes the order kmeans consumes the global RNG stream in
        gmm_data = {}
        for u, path in enumerate(paths):
            seq = graphs[utt_graph[u]][0]
            is_nes = np.array([type(s) is NES for s in seq])
            for row, a, b in _pool_segments(path[::-1], is_nes):
                spans = gmm_data.get(seq[row])
                if spans is None:
                    gmm_data[seq[row]] = spans = []
                spans.append((int(off64[u]) + a, int(off64[u]) + b))
        # ---- per-state split / K-means / EM (continuous_speech.py:110-142), all pooled states at once: the pools are
        # gathered in the dict's order (= the order the reference's loop consumes the global RNG stream in) and go through
        # the batched loops of srb200.fit (one pass over the pooled frames per K-means / EM iteration)
        states = list(gmm_data.keys())
        assert all(type(st) is not NES for st in states)
        if states:
            from srb200.fit import WordFitter
            rows = [np.concatenate([np.arange(a, b) for a, b in gmm_data[st]]) for st in states]
            idx = torch.from_numpy(np.concatenate(rows)).to(Xd.device)
            Xp = Xd.index_select(0, idx)
            lab = torch.from_numpy(np.repeat(np.arange(len(states)), [len(r) for r in rows])).to(Xd.device)
            Gp, Dp = len(states), Xp.shape[1]
            cnt = torch.bincount(lab, minlength=Gp).double()
            mean = torch.zeros((Gp, Dp), dtype=torch.float64, device=Xd.device).index_add_(0, lab, Xp.double()) / cnt[:, None]
            ft = WordFitter.from_pools(Xp, lab, Gp, n_gaussians, use_em=True)
            init = (np.stack([np.stack([d.mean for d in st.dists]) for st in states]),
                    np.stack([np.stack([np.asarray(d.cov, np.float64) for d in st.dists]) for st in states]),
                    np.stack([st.w for st in states]))
            old = (np.stack([st.mu_old for st in states]), np.stack([st.sigma_old for st in states]),
                   np.stack([st.w_old for st in states]))
            mu, var, wgt = ft.fit_states(state=lab.to(torch.int32), init=init, old=old, centroid0=mean.cpu().numpy(),
                                         weight_fill=1 / n_segments, weight_den=n_segments)   # continuous_speech.py:135-137 (sic)
            for g, st in enumerate(states):
                st.update_models(mu[g].copy(), var[g].copy(), wgt[g].copy())
                st.mu_old, st.sigma_old, st.w_old = ft.mu_old[g].copy(), ft.var_old[g].copy(), ft.w_old[g].copy()
                for its, conv in zip(ft.iterations.get("em", []), ft.converged.get("em", [])):
                    if conv[g]:
                        print("EM converged at iteration:", int(its[g]))
        # ---- transitions from segment counts (continuous_speech.py:144-164)
        for mi, m in enumerate(new_models):
            states = modelidx_state_map[mi] if state_lookup == "reference" else m.gmm_states
            for si in range(len(states)):
                spans = gmm_data.get(states[si])                # hash by id, then stored_key == states[si] by value
                if spans is None:
                    warnings.warn("No MFCC data for state", UserWarning)
                    continue
                p_jump = len(spans) / sum(b - a for a, b in spans)
                if si < len(states) - 1:
                    m.transitions[si + 1, si] = -np.log(p_jump)
                m.transitions[si, si] = -np.log(1 - p_jump)
        for i, m in enumerate(new_models):
            with open(os.path.join(output_path, str(i) + '.pkl'), 'wb') as f:
                pickle.dump(m, f)
        if all(nm == om for nm, om in zip(new_models, old_models)):
            print('Continuous training converged')
            break
        old_models = new_models
        new_models = copy.deepcopy(old_models)


def path_to_models(path, seq):
    """main.py:61-67: reversed path -> forward state rows, consecutive duplicates removed, split at the
    non-emitting states, first state of every piece mapped to the index of the model that owns it
    (`state.parent`, set by core.py:107-109, else looked up by position in `seq`)."""
    rows = np.asarray(path)[:, 0][::-1]
    if rows.size == 0:
        return []
    rows = rows[np.concatenate([[True], rows[1:] != rows[:-1]])]
    nes = np.array([type(seq[r]) is NES for r in rows])
    first = ~nes & np.concatenate([[True], nes[:-1]])           # first emitting state after a NES (or at the start)
    return [seq[r].parent for r in rows[first]]


def recognize_strings(xs, models, n_layers=7):
    """Connected-word recognition of many utterances (main.py:35,59-67): every model allowed in each of
    `n_layers` layers, one batched scoring + graph-Viterbi launch, word strings from the paths."""
    for i, m in enumerate(models):
        for s in m.gmm_states:
            if s.parent is None:
                s.parent = i
    seq, trans, ends = build_state_sequences(models, [list(range(len(models)))] * n_layers)
    res = decode_batch(xs, [(seq, trans, ends)], want_path=True)
    # the paths stay on the device: srb_path_words turns them into word strings (path_to_models is the host form)
    state_word = [-1 if type(s) is NES else int(s.parent) for s in seq]
    return res.vres.words(state_word, n_layers)


__all__ = ["build_state_sequences", "continuous_train", "path_to_models", "recognize_strings"]
