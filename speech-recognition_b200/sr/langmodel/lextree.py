"""Lexical tree of the sr.langmodel mirror (reference: sr/langmodel/lextree.py:5-76).

Host-side data structure only: `LexNode` keeps the reference's attributes (`val`, `children`, `property`: 0 normal,
1 root, 2 end of word) so that trees built by either side are interchangeable; `flatten` turns a tree into the arrays
srb_text_viterbi takes.
"""
from collections import Counter

import numpy as np


class LexNode:
    def __init__(self, val):
        self.val = val
        self.children = []
        self.property = 0          # 0: normal node, 1: start (root) node, 2: end-of-word node

    def pretty_str(self, level=0):
        return "\t" * level + repr(self.val) + "\n" + "".join(c.pretty_str(level + 1) for c in self.children)

    def __str__(self):
        return self.val

    def get_max_level(self, level=0):
        if not self.children:
            return level
        return max(c.get_max_level(level=level + 1) for c in self.children)


def append_lex_node(parent, child):
    assert type(parent) is LexNode and type(child) is LexNode
    parent.children.append(child)


def _build_lextree(node, words, i, max_word_len):
    """Children of `node` for the padded `words` that share the prefix leading to it (lextree.py:38-57).

    A word whose last character sits at position i becomes a leaf of its own (property 2), so every leaf stands for
    exactly one word.  The reference finds those words by REMOVING from `words` while it iterates over it, which makes
    the iteration skip the element after each removal; a skipped word stays in the list and continues as an ordinary
    branch whose leaf is then the padding blank after it.  The scan below is that iteration spelled out (an index
    that advances once per visit while the list shrinks under it)."""
    if i >= max_word_len:
        return
    last = i == max_word_len - 1
    k = 0
    while k < len(words):
        w = words[k]
        k += 1
        if last or w[i + 1] == ' ':
            leaf = LexNode(w[i])
            leaf.property = 2
            append_lex_node(node, leaf)
            words.remove(w)                    # first equal element; the scan index has already moved on
    for ch in Counter(w[i] for w in words):    # first-appearance order, like Counter(chars).items()
        child = LexNode(ch)
        append_lex_node(node, child)
        _build_lextree(child, [w for w in words if w[i] == ch], i + 1, max_word_len)


def lextree_from_words(words):
    """Lexical tree of a word list (lextree.py:60-76).  Like the reference it pads the caller's list in place with
    blanks to the longest word."""
    tree = LexNode('*')
    tree.property = 1
    max_word_len = max(len(w) for w in words)
    for n in range(len(words)):
        words[n] = words[n].ljust(max_word_len, ' ')
    _build_lextree(tree, words, 0, max_word_len)
    return tree


class FlatTree:
    """A lextree in the array form of srb_text_viterbi (include/srb200.h): nodes in the depth-first order of
    spellchecker.py:7-12 plus the blank node the spell-checker appends (spellchecker.py:37), parent of every node
    (spellchecker.py:52-63), the word-end list (blank node first, spellchecker.py:54-59) and the nodes grouped by depth
    (a node's same-column predecessor is its parent, so a column is swept level by level)."""

    def __init__(self, root):
        assert type(root) is LexNode
        nodes, parent, depth = [], [], []
        stack = [(root, -1, 0)]
        while stack:                                   # iterative pre-order: children in list order
            n, p, d = stack.pop()
            idx = len(nodes)
            nodes.append(n)
            parent.append(p)
            depth.append(d)
            for c in reversed(n.children):
                stack.append((c, idx, d + 1))
        self.nodes = nodes
        self.n_rows = len(nodes) + 1                   # + the blank node
        self.vals = [n.val for n in nodes] + [' ']
        self.val = np.array([ord(v) if len(v) == 1 else -1 for v in self.vals], np.int32)
        self.parent = np.array(parent + [-1], np.int32)
        depth = np.array(depth, np.int64)
        self.word_ends = np.array([self.n_rows - 1] + [i for i, n in enumerate(nodes) if n.property == 2], np.int32)
        order = np.argsort(depth, kind="stable")
        order = order[depth[order] >= 1]
        self.level_nodes = order.astype(np.int32)
        self.n_levels = int(depth.max()) if len(depth) else 0
        self.level_off = np.searchsorted(depth[order], np.arange(1, self.n_levels + 2)).astype(np.int32)


__all__ = ["LexNode", "append_lex_node", "lextree_from_words"]
