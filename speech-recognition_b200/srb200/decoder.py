"""Connected-word decoding of batches of utterances (main.py:35-67 of the reference): every word model allowed in each
of `n_layers` layers between non-emitting states (continuous_speech.py:13-53), K1 scoring of the tied states, one-graph
banded Viterbi with back-pointers and traceback (srb_viterbi_banded), decoded paths turned into word strings on the
device (srb_path_words).  All buffers of a batch shape are allocated once."""
import numpy as np
import torch

from . import engine
from . import lib as _l
from .graph import GraphBatch, word_loop_graph
from .trainer import ViterbiPlan


class StringDecoder:
    """Device-resident decoder for a fixed batch shape (utterance offsets `off`)."""

    def __init__(self, pm, word_trans, n_layers, off, tensor=None):
        self.pm, self.n_layers = pm, int(n_layers)
        dev = engine.device()
        g, word_of = word_loop_graph(np.asarray(word_trans, np.float64), n_layers)
        self.word_of = word_of
        self.gb = GraphBatch([g], dev)
        self.plan = ViterbiPlan(np.asarray(off, np.int32), self.gb, want_path=True, want_label=False)
        self.N, self.B = self.plan.N, self.plan.B
        self.E = torch.empty((max(self.N, 1), pm.R), dtype=torch.float32, device=dev)
        self.state_word = torch.from_numpy(word_of.astype(np.int32)).to(dev)
        self.words = torch.empty((max(self.B, 1), self.n_layers), dtype=torch.int32, device=dev)
        self.nwords = torch.empty(max(self.B, 1), dtype=torch.int32, device=dev)
        self.tensor = pm.scorer_tensor_core_ok() if tensor is None else bool(tensor)

    def score(self, Xd):
        pm, L = self.pm, _l.load()
        if self.tensor:
            engine.gmm_score_tc(Xd, pm, self.E[:Xd.shape[0]])
        else:
            _l.check(L.srb_gmm_score(_l.ptr(Xd), Xd.shape[0], pm.D, _l.ptr(pm.ab), _l.ptr(pm.c), _l.ptr(pm.shift), pm.R,
                                     pm.Mp, _l.ptr(self.E), pm.R, None, None, None, _l.stream_ptr()))
        return self.E

    def decode(self):
        self.plan.run(self.E)
        p = self.plan
        _l.check(_l.load().srb_path_words(_l.ptr(p.ps), _l.ptr(p.po_d), _l.ptr(p.pl), self.B, None,
                                          _l.ptr(self.gb["g_state_off"]), _l.ptr(self.state_word), self.n_layers,
                                          _l.ptr(self.words), _l.ptr(self.nwords), _l.stream_ptr()))
        return self.words, self.nwords

    def run(self, Xd):
        """Xd [N, D] float32 CUDA frames of the planned batch -> (words [B, n_layers] int32, nwords [B]) on the device."""
        self.score(Xd)
        return self.decode()


class HostStringDecoder:
    """Host-facing form: pinned host frames in, word strings out.  The batch is cut into `chunks` groups of whole
    utterances, each with its own StringDecoder; group k+1 is copied host->device on one stream while group k is
    scored and decoded on the other."""

    def __init__(self, pm, word_trans, n_layers, off, chunks=4):
        off = np.asarray(off, np.int64)
        B = len(off) - 1
        cuts = np.linspace(0, B, min(chunks, max(B, 1)) + 1).astype(np.int64)
        dev = engine.device()
        self.parts = []
        for k in range(len(cuts) - 1):
            u0, u1 = int(cuts[k]), int(cuts[k + 1])
            if u1 <= u0:
                continue
            f0, f1 = int(off[u0]), int(off[u1])
            dec = StringDecoder(pm, word_trans, n_layers, (off[u0:u1 + 1] - f0).astype(np.int32))
            xbuf = torch.empty((f1 - f0, pm.D), dtype=torch.float32, device=dev)
            self.parts.append((u0, u1, f0, f1, dec, xbuf))
        self.streams = [torch.cuda.Stream(dev) for _ in range(2)]

    def run(self, Xh, words_out, nwords_out):
        """Xh [N, D] pinned host frames; words_out [B, n_layers] int32 / nwords_out [B] int32 pinned host tensors."""
        cur = torch.cuda.current_stream()
        start = torch.cuda.Event()
        start.record(cur)
        evs = []
        for k, (u0, u1, f0, f1, dec, xbuf) in enumerate(self.parts):
            st = self.streams[k & 1]
            with torch.cuda.stream(st):
                st.wait_event(start)
                xbuf.copy_(Xh[f0:f1], non_blocking=True)
                words, nwords = dec.run(xbuf)
                words_out[u0:u1].copy_(words[:u1 - u0], non_blocking=True)
                nwords_out[u0:u1].copy_(nwords[:u1 - u0], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(st)
            evs.append(ev)
        for ev in evs:
            cur.wait_event(ev)
        return words_out, nwords_out
