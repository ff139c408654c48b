"""lexical tree + text Viterbi spell-checker (sr/langmodel/lextree.py:38-76, spellchecker.py:15-135).

CPU (`-m "not gpu"`): the oracle restatement and the mirror's host-side tree builder against the golden vectors frozen
from the unmodified reference (tests/golden/make_golden_r2.py lextree).
GPU (`-m gpu`): the drop-in sr.langmodel.text_viterbi (srb_text_viterbi through the C ABI) against the same golden
vectors, and against the oracle on dictionaries of the reference test's size (6 k words, 27 k tree nodes)."""
import numpy as np
import pytest

from oracle import lextree_oracle as lo


def soft_dist(a, b):
    if a == b:
        return 0.0
    return 0.5 if (a in "aeiou" and b in "aeiou") else 1.0


def dump_flat(vals, props, parent):
    return [str(v) for v in vals], [int(p) for p in props], [int(p) for p in parent]


def test_oracle_tree_and_viterbi_match_reference(golden):
    g = golden["lextree"]
    for d in range(int(g["lt_n"])):
        words = [str(w) for w in g[f"lt{d}_words"]]
        tree = lo.build_lextree(words)
        vals, parent, props, _ = lo.flatten(tree)
        assert (vals[:-1], list(props[:-1]), list(parent[:-1])) == dump_flat(g[f"lt{d}_vals"], g[f"lt{d}_props"],
                                                                            g[f"lt{d}_parent"])
        for t, cost, match in zip(g[f"lt{d}_texts"], g[f"lt{d}_cost"], g[f"lt{d}_match"]):
            c1, m1 = lo.text_viterbi(str(t), tree)
            assert (c1, m1) == (float(cost), str(match))
            if d < 2:
                c0, m0, _ = lo.text_viterbi_cells(str(t), tree)
                assert (c0, m0) == (float(cost), str(match))
        if f"lt{d}_soft_cost" in g.files:
            for t, cost, match in zip(g[f"lt{d}_texts"], g[f"lt{d}_soft_cost"], g[f"lt{d}_soft_match"]):
                c1, m1 = lo.text_viterbi(str(t), tree, soft_dist)
                assert (c1, m1) == (float(cost), str(match))


def test_mirror_tree_builder_matches_reference(golden):
    """Host logic only (no device call): lextree_from_words incl. the skipped-element quirk of lextree.py:41-48, and the
    flattening srb_text_viterbi consumes."""
    from sr.langmodel.lextree import FlatTree, lextree_from_words
    g = golden["lextree"]
    for d in range(int(g["lt_n"])):
        words = [str(w) for w in g[f"lt{d}_words"]]
        tree = lextree_from_words(words)
        assert all(len(w) == len(words[0]) for w in words)            # padded in place like the reference (:73-74)
        ft = FlatTree(tree)
        assert ft.vals[:-1] == [str(v) for v in g[f"lt{d}_vals"]] and ft.vals[-1] == ' '
        assert np.array_equal(ft.parent[:-1], g[f"lt{d}_parent"])
        assert [n.property for n in ft.nodes] == list(g[f"lt{d}_props"])
        _, _, _, we = lo.flatten(lo.build_lextree([str(w) for w in g[f"lt{d}_words"]]))
        assert list(ft.word_ends) == list(we)
        depth = np.zeros(ft.n_rows, int)
        for r in range(1, ft.n_rows - 1):
            depth[r] = depth[ft.parent[r]] + 1
        for lv in range(ft.n_levels):
            assert (depth[ft.level_nodes[ft.level_off[lv]:ft.level_off[lv + 1]]] == lv + 1).all()
        assert ft.level_off[-1] == ft.n_rows - 2


@pytest.mark.gpu
def test_text_viterbi_golden(golden):
    import sr.langmodel as lm
    g = golden["lextree"]
    for d in range(int(g["lt_n"])):
        tree = lm.lextree_from_words([str(w) for w in g[f"lt{d}_words"]])
        texts = [str(t) for t in g[f"lt{d}_texts"]]
        for t, cost, match in zip(texts, g[f"lt{d}_cost"], g[f"lt{d}_match"]):
            assert lm.text_viterbi(t, tree) == (float(cost), str(match))
        got = lm.text_viterbi_batch(texts, tree)
        assert got == [(float(c), str(m)) for c, m in zip(g[f"lt{d}_cost"], g[f"lt{d}_match"])]
        if f"lt{d}_soft_cost" in g.files:
            got = lm.text_viterbi_batch(texts, tree, soft_dist)
            assert [m for _, m in got] == [str(m) for m in g[f"lt{d}_soft_match"]]
            np.testing.assert_allclose([c for c, _ in got], g[f"lt{d}_soft_cost"], rtol=1e-6)
    with pytest.raises(ValueError):
        lm.text_viterbi("", tree)


@pytest.mark.gpu
def test_text_viterbi_large_dictionary_against_oracle():
    """A dictionary of the reference test's size (tests/lextree_tests.py: 6 249 words) and misspelt sentences drawn from
    it: costs and matched strings equal to the oracle's."""
    import sr.langmodel as lm
    rng = np.random.default_rng(12)
    letters = np.array(list("abcdefghijklmnopqrstuvwxyz"))
    words = sorted({"".join(rng.choice(letters, size=int(rng.integers(1, 12)))) for _ in range(6500)})
    texts = []
    for _ in range(6):
        ws = [words[i] for i in rng.integers(0, len(words), size=int(rng.integers(3, 9)))]
        chars = list(" ".join(ws))
        for k in rng.integers(0, len(chars), size=max(1, len(chars) // 6)):     # typos: substitute / delete / insert
            op = rng.integers(0, 3)
            if op == 0:
                chars[k] = str(rng.choice(letters))
            elif op == 1 and len(chars) > 2:
                del chars[min(k, len(chars) - 1)]
            else:
                chars.insert(k, str(rng.choice(letters)))
        texts.append("".join(chars))
    tree = lm.lextree_from_words(list(words))
    otree = lo.build_lextree(list(words))
    got = lm.text_viterbi_batch(texts, tree)
    for t, (c, m) in zip(texts, got):
        assert (c, m) == lo.text_viterbi(t, otree), t
