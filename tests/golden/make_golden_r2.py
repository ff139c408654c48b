#!/usr/bin/env python
"""Round-2 golden vectors, produced by EXECUTING the unmodified reference (container only; shim in _ref_shim.py).

    python tests/golden/make_golden_r2.py [beam] [ends] [lextree]      # writes tests/golden/ref_<name>.npz

  beam     dtw(..., beam=k) (decode.py:42-48,62-68): costs incl. the -1 sentinels left in the last column, paths
  ends     decode_hmm_states with end points off the last frame (decode.py:126-137)
  lextree  sr/langmodel: build_lextree (lextree.py:38-76) and text_viterbi (spellchecker.py:15-135)
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from _ref_shim import load_reference  # noqa: E402
from make_golden import (D, dump_models, make_model, make_utterance, make_word_templates, ragged)  # noqa: E402

warnings.simplefilter("ignore")
rec = load_reference()


def case_beam(out):
    rng = np.random.default_rng(707)
    S = 6
    base = make_word_templates(rng, 1, S)[0]
    tr = rec.calc_transition_costs(3, np.array([[6] * S] * 3))
    tr2 = tr.copy()                                             # hand-made skip arcs (finite; calc_transition_costs
    tr2[2, 0] = -np.log(0.05)                                   # only makes them next to NaN rows of empty segments)
    tr2[4, 2] = -np.log(0.2)
    var = rng.uniform(0.4, 1.6, size=(S, D))
    xs = [make_utterance(rng, base, T) for T in (9, 23, 40)]
    X, off = ragged(xs)
    out.update(bm_X=X, bm_off=off, bm_y=base, bm_trans=tr, bm_trans2=tr2, bm_var=var)
    euclid = lambda *a: np.linalg.norm(a[0] - a[1])  # noqa: E731
    k = 0
    cases = []
    for beam in (1, 2, 3, 5, 6, 50):
        for kind in ("e", "m"):
            for ti, trans in enumerate((tr, tr2)):
                for u, x in enumerate(xs):
                    try:
                        if kind == "e":
                            c, p = rec.dtw(x, base, euclid, trans, beam=beam)
                        else:
                            c, p = rec.dtw(x, base, rec.mahalanobis, trans, var, beam=beam)
                    except Exception as e:  # noqa: BLE001
                        cases.append((beam, kind == "m", ti, u, -1))
                        out[f"bm_err{k}"] = np.array(type(e).__name__)
                        k += 1
                        continue
                    cases.append((beam, kind == "m", ti, u, len(p)))
                    out[f"bm_costs{k}"] = c
                    out[f"bm_path{k}"] = p
                    k += 1
    out["bm_cases"] = np.array(cases, np.int64)


def case_ends(out):
    rng = np.random.default_rng(808)
    W, S, M = 2, 4, 2
    base = make_word_templates(rng, W, S)
    models = [make_model(rng, base[w], M) for w in range(W)]
    mu, cv, wt, tr = dump_models(models)
    seq, trans, ends = rec.build_state_sequences(models, [[0, 1]] * 2)
    is_nes = np.array([type(s) is rec.NES for s in seq])
    ids = {g.id: w * S + s for w, m in enumerate(models) for s, g in enumerate(m.gmm_states)}
    emit = np.array([-1 if type(s) is rec.NES else ids[s.id] for s in seq])
    x = np.concatenate([make_utterance(rng, base[1], 14), make_utterance(rng, base[0], 17)])
    out.update(en_mu=mu, en_cov=cv, en_w=wt, en_wtrans=tr, en_trans=trans, en_is_nes=is_nes, en_emit=emit, en_X=x,
               en_ends=np.array(ends))
    T = len(x)
    lists = [
        [[ends[0], T - 5], [ends[1], T - 5]],                    # both ends a few frames early
        [[ends[0], -1], [ends[1], -3], [ends[0], 20]],           # mixed, negative frames (python indexing)
        [[S, 13], [2 * S, 13]],                                  # first-layer end states (rows S, 2S) mid-utterance
        [[S, 9], [len(seq) - 1, -1]],                            # a NES as end point at the last frame
        [[3, 0]],                                                # frame 0: no walk at all
    ]
    # iso: a plain left-to-right model with an early end point
    c0, p0 = rec.decode_hmm_states(x[:14], models[1].gmm_states, models[1].transitions, end_points=[[S - 1, 9], [S - 2, -2]])
    out.update(en_iso_costs=c0, en_iso_path=p0)
    for k, el in enumerate(lists):
        c, p = rec.decode_hmm_states(x, seq, trans, end_points=el)
        out[f"en_list{k}"] = np.array(el, np.int64)
        out[f"en_costs{k}"] = c
        out[f"en_path{k}"] = np.asarray(p, np.int64).reshape(-1, 2)
    out["en_n"] = np.int64(len(lists))


VOWELS = set("aeiou")


def soft_dist(a, b):
    """A caller-supplied dist_fun for text_viterbi: vowels are half as far from each other as other characters."""
    if a == b:
        return 0.0
    return 0.5 if (a in VOWELS and b in VOWELS) else 1.0


def case_lextree(out):
    import random
    import pandas as pd
    import sr.langmodel as lm
    assert lm.__file__.startswith("/root/reference")
    sys.setrecursionlimit(20000)
    words_all = [str(w) for w in pd.read_csv("/root/reference/tests/test_data/dict1.txt")["words"].values.tolist()]
    rnd = random.Random(4242)
    dicts = [
        ["once", "upon", "a", "time", "on", "one", "up", "tim", "at", "an", "and", "ant"],
        ["a", "a", "ab", "ab", "abc", "b", "ba", "bab", "c"],                       # duplicates: the skipped-element quirk
        rnd.sample(words_all, 120),
        rnd.sample(words_all, 400) + ["the", "a", "of", "to", "monkey", "crocodile", "said", "heart", "tree", "river"],
    ]
    texts = [
        ["onse apon a tyme", "a", "at an ant", "xyz", "oneup on", "tim e", "uponn a a tme"],
        ["abc ab a", "bab c", "cab", "aabbcc", "b"],
        ["tha quick brwn fox", "helo wrld", "zzzz", "e"],
        ["the munkey sed to the krocodile", "harte of the tre", "rever", "a", "onse upon a tyme the monkey said"],
    ]

    def flat(tree):
        nodes = []
        lm.get_nodes(nodes, tree)
        idx = {id(n): i for i, n in enumerate(nodes)}
        parent = np.full(len(nodes), -1, np.int64)
        for n in nodes:
            for c in n.children:
                parent[idx[id(c)]] = idx[id(n)]
        return np.array([n.val for n in nodes]), np.array([n.property for n in nodes], np.int64), parent

    for d, (words, txts) in enumerate(zip(dicts, texts)):
        out[f"lt{d}_words"] = np.array(words)
        tree = lm.lextree_from_words(list(words))
        v, pr, pa = flat(tree)
        out[f"lt{d}_vals"], out[f"lt{d}_props"], out[f"lt{d}_parent"] = v, pr, pa
        out[f"lt{d}_texts"] = np.array(txts)
        res = [lm.text_viterbi(t, tree) for t in txts]
        out[f"lt{d}_cost"] = np.array([r[0] for r in res], np.float64)
        out[f"lt{d}_match"] = np.array([r[1] for r in res])
        if d in (0, 2):
            res = [lm.text_viterbi(t, tree, soft_dist) for t in txts]
            out[f"lt{d}_soft_cost"] = np.array([r[0] for r in res], np.float64)
            out[f"lt{d}_soft_match"] = np.array([r[1] for r in res])
    out["lt_n"] = np.int64(len(dicts))


def main():
    import time
    only = set(sys.argv[1:])
    for name, fn in (("beam", case_beam), ("ends", case_ends), ("lextree", case_lextree)):
        if only and name not in only:
            continue
        t0 = time.time()
        out = {}
        fn(out)
        path = os.path.join(HERE, f"ref_{name}.npz")
        np.savez_compressed(path, **out)
        print(f"{name}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB, {time.time() - t0:.1f}s")


if __name__ == "__main__":
    main()
