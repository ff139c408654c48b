"""Text Viterbi spell-checker of the sr.langmodel mirror (reference: sr/langmodel/spellchecker.py:7-149).

`text_viterbi(x, lextree, dist_fun)` keeps the reference's signature and result `(best_cost, matched_string)`; the
dynamic programme and the walk back run in srb_text_viterbi (libsrb200).  `text_viterbi_batch` checks many strings
against one tree in one launch (one CTA per string).
"""
import numpy as np
import torch

from srb200 import engine

from .lextree import FlatTree, LexNode


def get_nodes(node_list, lexnode):
    """Depth-first (pre-order) list of the tree's nodes (spellchecker.py:7-12)."""
    node_list.extend(FlatTree(lexnode).nodes)


def _default_dist(*args):
    return int(args[0] != args[1])


_tree_cache = {}


def _device_tree(lextree):
    """Flattened device copy of a tree, cached per tree object (flattening 27 k nodes is host work worth keeping)."""
    key = id(lextree)
    hit = _tree_cache.get(key)
    if hit is None or hit[0] is not lextree:
        if len(_tree_cache) > 8:
            _tree_cache.clear()
        hit = _tree_cache[key] = (lextree, engine.DeviceLexTree(FlatTree(lextree)))
    return hit[1]


def text_viterbi_batch(xs, lextree, dist_fun=_default_dist):
    """text_viterbi of every string of `xs` against one tree -> list of (best_cost, matched_string)."""
    assert type(lextree) is LexNode
    tree = _device_tree(lextree)
    flat = tree.flat
    texts = ['*' + x for x in xs]
    if dist_fun is _default_dist and (flat.val >= 0).all():
        codes = [np.array([ord(ch) for ch in t], np.int32) for t in texts]
        dist = None
    else:
        # a caller-supplied distance: evaluated by the caller's own function, once per distinct (character, node value)
        codes = [np.zeros(len(t), np.int32) for t in texts]
        memo = {}
        rows = []
        for t in texts:
            for ch in t:
                row = memo.get(ch)
                if row is None:
                    per_val = {}
                    row = memo[ch] = np.array([per_val.setdefault(v, dist_fun(ch, v)) for v in flat.vals], np.float32)
                rows.append(row)
        dist = torch.from_numpy(np.stack(rows))
    best, nodes = engine.text_viterbi(codes, tree, dist)
    return [(float(b), ''.join(flat.vals[r] for r in nd)) for b, nd in zip(best, nodes)]


def text_viterbi(x, lextree, dist_fun=_default_dist):
    """Best match of the string x against the dictionary tree with word loop-back (spellchecker.py:15-135).
    -> (best_cost, matched_string)"""
    return text_viterbi_batch([x], lextree, dist_fun)[0]


class SpellChecker:
    """Stub kept like the reference's (spellchecker.py:138-149)."""

    def __init__(self, beam):
        self.dictionary = None
        self.beam = beam

    def fit(self, dictionary):
        self.dictionary = dictionary

    def spell_check(self, text):
        pass


__all__ = ["get_nodes", "text_viterbi", "text_viterbi_batch", "SpellChecker"]
