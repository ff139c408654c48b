"""State graphs in the packed CSR form srb_viterbi_graph / srb_forward_backward take (include/srb200.h).

Host-side (numpy) packing of what the reference passes around as a dense `transitions[R, R]` matrix
(`transitions[i, j]` = cost j -> i, inf = none, decode.py:86-87) plus a list of state objects.
"""
import numpy as np
import torch


class StateGraph:
    """One graph: emit[r] (column of the emission matrix, -1 = NES), CSR predecessors in ascending origin order
    (decode.py:104-107 iterates origins ascending, so 'first minimum' is preserved), end states."""

    def __init__(self, transitions, emit, end_states=None, keep_inf=False, end_frames=None):
        tr = np.asarray(transitions, dtype=np.float64)
        R = tr.shape[0]
        assert tr.shape == (R, R)
        self.R = R
        self.emit = np.asarray(emit, dtype=np.int32).reshape(R)
        # dtw (decode.py:44-51) keeps inf entries as candidates; decode_hmm_states (:106) skips them
        fin = np.ones_like(tr, dtype=bool) if keep_inf else ~np.isposinf(tr)
        tgt, org = np.nonzero(fin)                      # row-major: targets ascending, origins ascending within
        self.pred_ptr = np.zeros(R + 1, np.int32)
        np.cumsum(np.bincount(tgt, minlength=R), out=self.pred_ptr[1:])
        self.pred_idx = org.astype(np.int32)
        self.pred_tgt = tgt.astype(np.int32)
        self.pred_cost = tr[tgt, org].astype(np.float32)
        if (np.diff(self.pred_ptr) > 255).any():
            raise ValueError("a state has more than 255 predecessors (back-pointer codes are 8 bit)")
        self.nes_idx = np.nonzero(self.emit < 0)[0].astype(np.int32)
        if end_states is None:
            end_states = [R - 1]                         # decode.py:126-127
        self.end_states = np.asarray(end_states, dtype=np.int32).reshape(-1)
        # frames of the end points when some lie off the last frame (decode.py:126-134 takes any [state, frame])
        self.end_frames = None if end_frames is None else np.asarray(end_frames, dtype=np.int32).reshape(-1)
        assert self.end_frames is None or self.end_frames.shape == self.end_states.shape
        # successor CSR of the same edges (forward-backward's beta recursion)
        order = np.lexsort((tgt, org))
        self.succ_ptr = np.zeros(R + 1, np.int32)
        np.cumsum(np.bincount(org, minlength=R), out=self.succ_ptr[1:])
        self.succ_idx = tgt[order].astype(np.int32)
        self.succ_cost = self.pred_cost[order]
        self.succ_edge = order.astype(np.int32)          # predecessor-CSR index of every successor entry
        self.has_nes_self_loop = bool(any(np.isfinite(tr[n, n]) for n in self.nes_idx))

    @property
    def nnz(self):
        return int(self.pred_idx.shape[0])

    def banded_form(self):
        """Arrays of srb_viterbi_banded (include/srb200.h) or None when the graph is not in banded form: emitting
        predecessors of an emitting state r inside {r-2, r-1, r}, at most one NES predecessor n < r (its entry),
        NES fed only by emitting states n' < n that have no entry themselves."""
        R, emit = self.R, self.emit
        is_nes = emit < 0
        band = np.full((R, 3), np.inf, np.float32)
        codes = np.zeros((R, 4), np.int64)
        entry_nes = np.full(R, -1, np.int32)
        entry_cost = np.zeros(R, np.float32)
        feeds_nes = np.zeros(R, bool)
        for r in range(R):
            p0, p1 = int(self.pred_ptr[r]), int(self.pred_ptr[r + 1])
            for pos in range(p1 - p0):
                o, c = int(self.pred_idx[p0 + pos]), self.pred_cost[p0 + pos]
                if is_nes[r]:
                    if is_nes[o] or o >= r:
                        return None
                    feeds_nes[o] = True
                elif is_nes[o]:
                    if o >= r or entry_nes[r] >= 0:
                        return None
                    entry_nes[r], entry_cost[r], codes[r, 3] = o, c, pos
                else:
                    k = r - o
                    if k < 0 or k > 2:
                        return None
                    band[r, 2 - k], codes[r, 2 - k] = c, pos
        if (feeds_nes & (entry_nes >= 0)).any():
            return None
        packed = (codes[:, 0] | (codes[:, 1] << 8) | (codes[:, 2] << 16) | (codes[:, 3] << 24)).astype(np.int32)
        return {"band": band.reshape(-1), "codes": packed, "entry_nes": entry_nes, "entry_cost": entry_cost,
                "entry_state": np.nonzero(entry_nes >= 0)[0].astype(np.int32)}


class GraphBatch:
    """Several StateGraphs packed back to back on the device."""

    def __init__(self, graphs, device):
        self.graphs = graphs
        R = [g.R for g in graphs]
        self.Rmax = max(R)
        self.NNZmax = max(g.nnz for g in graphs)
        self.n_nes_max = max(len(g.nes_idx) for g in graphs)
        self.state_off = np.zeros(len(graphs) + 1, np.int32)
        np.cumsum(R, out=self.state_off[1:])
        edge_off = np.zeros(len(graphs) + 1, np.int64)
        np.cumsum([g.nnz for g in graphs], out=edge_off[1:])
        self.edge_off = edge_off

        def cat(parts, dtype):
            return np.concatenate(parts).astype(dtype) if parts else np.zeros(0, dtype)

        def offs(lens):
            o = np.zeros(len(lens) + 1, np.int32)
            np.cumsum(lens, out=o[1:])
            return o

        pred_ptr = cat([g.pred_ptr[:-1].astype(np.int64) + edge_off[i] for i, g in enumerate(graphs)] +
                       [np.array([edge_off[-1]])], np.int32)
        succ_ptr = cat([g.succ_ptr[:-1].astype(np.int64) + edge_off[i] for i, g in enumerate(graphs)] +
                       [np.array([edge_off[-1]])], np.int32)
        host = {
            "g_state_off": self.state_off,
            "g_nes_off": offs([len(g.nes_idx) for g in graphs]),
            "g_end_off": offs([len(g.end_states) for g in graphs]),
            "emit": cat([g.emit for g in graphs], np.int32),
            "pred_ptr": pred_ptr,
            "pred_idx": cat([g.pred_idx for g in graphs], np.int32),
            "pred_tgt": cat([g.pred_tgt for g in graphs], np.int32),
            "pred_cost": cat([g.pred_cost for g in graphs], np.float32),
            "nes_idx": cat([g.nes_idx for g in graphs], np.int32),
            "end_state": cat([g.end_states for g in graphs], np.int32),
            "succ_ptr": succ_ptr,
            "succ_idx": cat([g.succ_idx for g in graphs], np.int32),
            "succ_cost": cat([g.succ_cost for g in graphs], np.float32),
            "succ_edge": cat([g.succ_edge.astype(np.int64) + edge_off[i] for i, g in enumerate(graphs)], np.int32),
        }
        if any(g.end_frames is not None for g in graphs):
            host["end_frame"] = cat([g.end_frames if g.end_frames is not None else np.full(len(g.end_states), -1)
                                     for g in graphs], np.int32)
        forms = [g.banded_form() for g in graphs]
        self.banded = all(f is not None for f in forms)
        if self.banded:
            host.update({
                "band": cat([f["band"] for f in forms], np.float32),
                "codes": cat([f["codes"] for f in forms], np.int32),
                "entry_nes": cat([f["entry_nes"] for f in forms], np.int32),
                "entry_cost": cat([f["entry_cost"] for f in forms], np.float32),
                "g_entry_off": offs([len(f["entry_state"]) for f in forms]),
                "entry_state": cat([f["entry_state"] for f in forms], np.int32),
            })
            # index into `band` of every predecessor-CSR edge (-1: the arc from a non-emitting entry state), so that
            # srb_update_transitions can rewrite a re-estimated cost wherever the kernels read it
            slots = []
            for i, g in enumerate(graphs):
                k = g.pred_tgt.astype(np.int64) - g.pred_idx
                sl = (int(self.state_off[i]) + g.pred_tgt.astype(np.int64)) * 3 + (2 - k)
                slots.append(np.where((k >= 0) & (k <= 2) & (g.emit[g.pred_idx] >= 0), sl, -1))
            host["band_slot"] = cat(slots, np.int32)
        self.host = host
        self.dev = {}
        for k, v in host.items():
            if v.size == 0:                              # keep a valid pointer for empty arrays
                v = np.zeros(1, v.dtype)
            self.dev[k] = torch.from_numpy(v).to(device)

    def __getitem__(self, k):
        return self.dev[k]


def word_loop_graph(word_trans, n_layers):
    """The connected-word grammar of main.py:35 / continuous_speech.py:13-53 as a StateGraph: layers of all W words
    between non-emitting states; emit[r] = w*S + s (column of the tied state), -1 for the NES.
    -> (StateGraph, word_of_state[R] with -1 at the NES)"""
    word_trans = np.asarray(word_trans, np.float64)
    W, S, _ = word_trans.shape
    R = 1 + n_layers * (W * S + 1)
    tr = np.full((R, R), np.inf)
    emit = np.full(R, -1, np.int32)
    word_of = np.full(R, -1, np.int32)
    entry, ends = 0, []
    for layer in range(n_layers):
        base = 1 + layer * (W * S + 1)
        exit_ = base + W * S
        ends = []
        for w in range(W):
            o = base + w * S
            tr[o:o + S, o:o + S] = word_trans[w]
            emit[o:o + S] = w * S + np.arange(S)
            word_of[o:o + S] = w
            tr[o, entry] = 0.0
            tr[exit_, o + S - 1] = 0.0
            ends.append(o + S - 1)
        entry = exit_
    return StateGraph(tr, emit, end_states=ends), word_of
