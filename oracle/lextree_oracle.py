"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's lexical tree and text Viterbi spell-checker
(/root/reference/sr/langmodel/lextree.py:38-76, spellchecker.py:15-135), SURVEY 8 row f4.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module; the product path
(speech-recognition_b200/) never does.  Pinned: tests/test_oracle_golden.py checks it against tests/golden/ref_lextree.npz,
frozen from the UNMODIFIED reference by tests/golden/make_golden_r2.py.

Two forms of the same recursion:
  text_viterbi_cells   cell by cell in the reference's loop order (pure Python; small trees only)
  text_viterbi         numpy, one vector operation per tree level and column (a node's only same-column predecessor is
                       its parent) -- checked against the cell-by-cell form, used for the larger parity sweeps
Trees are plain tuples here: (val, property, [children]).
"""
from collections import Counter

import numpy as np

DELETION, INSERTION, MATCH, SPACE, LOOP = 1.0, 1.0, 0.0, 0.0, 1.0          # spellchecker.py:28-32


def build_lextree(words):
    """lextree.py:60-76 -> nested (val, property, children) tuples; the list-mutation-while-iterating of
    lextree.py:41-48 is spelled out as an index scan over a shrinking list."""
    max_len = max(len(w) for w in words)
    words = [w.ljust(max_len, ' ') for w in words]

    def build(ws, i):
        if i >= max_len:
            return []
        children = []
        k = 0
        while k < len(ws):                                                  # `for w in words:` + words.remove(w)
            w = ws[k]
            k += 1
            if (i + 1 < max_len and w[i + 1] == ' ') or i == max_len - 1:    # :44
                children.append((w[i], 2, []))
                ws.remove(w)
        for ch in Counter(w[i] for w in ws):                                 # :51-57
            children.append((ch, 0, build([w for w in ws if w[i] == ch], i + 1)))
        return children

    return ('*', 1, build(words, 0))


def flatten(tree):
    """spellchecker.py:35-63 -> (vals, parent, props, word_ends) with the blank node appended last."""
    vals, parent, props = [], [], []

    def rec(n, p):
        idx = len(vals)
        vals.append(n[0])
        parent.append(p)
        props.append(n[1])
        for c in n[2]:
            rec(c, idx)

    import sys
    sys.setrecursionlimit(max(10000, sys.getrecursionlimit()))
    rec(tree, -1)
    vals.append(' ')
    parent.append(-1)
    props.append(0)
    n_rows = len(vals)
    word_ends = [n_rows - 1] + [i for i in range(n_rows) if props[i] == 2]   # :54-59
    return vals, np.array(parent), np.array(props), word_ends


def _walk(path_matrix, vals, r, c):
    s = vals[r]                                                              # :128
    while True:
        if c == 1:                                                           # :130-131
            break
        if c < 1:
            raise RuntimeError("walk back passed column 1 (the reference loops forever here)")
        r, c = path_matrix[r, c]
        if r != 0:
            s += vals[r]
    return s[::-1]


def text_viterbi_cells(x, tree, dist_fun=lambda a, b: int(a != b)):
    """spellchecker.py:15-135 cell by cell -> (best_cost, matched_string, costs[n_rows, n_cols])."""
    x = '*' + x
    vals, parent, props, word_ends = flatten(tree)
    n_cols, n_rows = len(x), len(vals)
    costs = np.full((n_rows, n_cols), np.inf)
    costs[0, :] = 0
    pm = np.zeros((n_rows, n_cols, 2), np.int64)
    for c in range(n_cols):
        for r in range(n_rows):
            if r == 0 and c == 0:
                continue
            sub = [np.inf] * 4
            frm = [[0, 0] for _ in range(4)]
            nd = dist_fun(x[c], vals[r])
            par = parent[r] if parent[r] >= 0 else None
            if r == n_rows - 1:                                              # blank node :80-82
                sub = [nd + SPACE + costs[we, c - 1] for we in word_ends[1:]]
                frm = [[we, c - 1] for we in word_ends[1:]]
            elif par == 0:                                                   # children of the root :83-91
                if c > 0:
                    sub[0] = nd + DELETION + costs[r, c - 1]
                    frm[0] = [r, c - 1]
                sub[2] = nd + INSERTION + costs[par, c]
                frm[2] = [par, c]
            elif r == 0:                                                     # root: loop back :92-100
                if c > 0:
                    loop = [nd + LOOP + costs[we, c - 1] for we in word_ends]
                    k = int(np.argmin(loop))
                    sub[3] = loop[k]
                    frm[3] = [word_ends[k], c - 1]
            else:                                                            # :101-113
                if c > 0:
                    sub[0] = nd + DELETION + costs[r, c - 1]
                    frm[0] = [r, c - 1]
                if par is not None and c > 0:
                    sub[1] = nd + MATCH + costs[par, c - 1]
                    frm[1] = [par, c - 1]
                if par is not None:
                    sub[2] = nd + INSERTION + costs[par, c]
                    frm[2] = [par, c]
            k = int(np.argmin(sub))                                          # :116-118
            pm[r, c] = frm[k]
            costs[r, c] = sub[k]
    fin = [costs[we, n_cols - 1] for we in word_ends]                        # :120-123
    k = int(np.argmin(fin))
    return float(fin[k]), _walk(pm, vals, word_ends[k], n_cols - 1), costs


def text_viterbi(x, tree, dist_fun=None):
    """The same recursion with one numpy operation per tree level and column -> (best_cost, matched_string)."""
    x = '*' + x
    vals, parent, props, word_ends = flatten(tree)
    n_cols, n_rows = len(x), len(vals)
    depth = np.zeros(n_rows, np.int64)
    for r in range(1, n_rows - 1):
        depth[r] = depth[parent[r]] + 1
    levels = [np.flatnonzero(depth == d)[:] for d in range(1, int(depth.max()) + 1)]
    levels = [l[l != n_rows - 1] for l in levels]
    vals_arr = np.array(vals)
    we = np.array(word_ends)
    space = n_rows - 1
    pm = np.zeros((n_rows, n_cols, 2), np.int64)
    prev = np.full(n_rows, np.inf)
    memo = {}
    for c in range(n_cols):
        if dist_fun is None:
            nd = (vals_arr != x[c]).astype(np.float64)
        else:
            if x[c] not in memo:
                memo[x[c]] = np.array([dist_fun(x[c], v) for v in vals], np.float64)
            nd = memo[x[c]]
        cur = np.full(n_rows, np.inf)
        if c == 0:
            cur[0] = 0.0
            pm[space, 0] = [we[1], -1]
        else:
            sub = (nd[space] + SPACE) + prev[we[1:]]
            k = int(np.argmin(sub))
            cur[space], pm[space, c] = sub[k], [we[1 + k], c - 1]
            loop = (nd[0] + LOOP) + prev[we]
            k = int(np.argmin(loop))
            cur[0] = loop[k]
            pm[0, c] = [we[k], c - 1] if np.isfinite(loop[k]) else [0, 0]
        for lv in levels:
            par = parent[lv]
            dele = (nd[lv] + DELETION) + prev[lv] if c > 0 else np.full(len(lv), np.inf)
            mat = np.where(par != 0, (nd[lv] + MATCH) + prev[par], np.inf) if c > 0 else np.full(len(lv), np.inf)
            ins = (nd[lv] + INSERTION) + cur[par]
            sub = np.stack([dele, mat, ins])
            k = np.argmin(sub, axis=0)
            cur[lv] = sub[k, np.arange(len(lv))]
            pm[lv, c, 0] = np.where(k == 0, lv, par)
            pm[lv, c, 1] = np.where(k == 2, c, c - 1)
        prev = cur
    fin = prev[we]
    k = int(np.argmin(fin))
    return float(fin[k]), _walk(pm, vals, int(we[k]), n_cols - 1)
