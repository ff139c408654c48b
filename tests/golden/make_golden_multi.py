#!/usr/bin/env python
"""Golden vectors for continuous_train over SEVERAL outer iterations (continuous_speech.py:56-179), produced by
EXECUTING the unmodified reference (container only; see make_golden.py for the shim).

    python tests/golden/make_golden_multi.py       # writes tests/golden/ref_train_multi.npz

The starting models are rebuilt from the arrays of ref_train.npz (fit_mu / fit_cov / fit_w / fit_trans) with the
reference's own constructors, exactly as tests/mirror_cases.py:load_models does on the mirror side, so that both
sides start from the same objects (including the `*_old` convergence buffers of hmm_state.py:110-112).
For max_iteration = 1, 2, 3 (same seed, so run n+1 repeats run n and adds one outer iteration) it stores the models
the reference pickled, the number of "No MFCC data for state" warnings (continuous_speech.py:149-153: from the second
iteration on `modelidx_state_map` still holds the FIRST copy's states and the dict lookup compares them by value), the
outer iteration count and every "EM converged at iteration: k" the reference printed (hmm_state.py:153).
It also records the EM stop iterations of the converged cases frozen by make_golden.py (GMM.em to convergence,
HMM.fit), so that the parity tests can check that the mirror's loops stop where the reference's do.
"""
import contextlib
import io
import os
import pickle
import re
import sys
import tempfile
import time
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from _ref_shim import load_reference  # noqa: E402
from make_golden import dump_models  # noqa: E402

rec = load_reference()


def build_models(z):
    mu, cov, w, trans = z["fit_mu"], z["fit_cov"], z["fit_w"], z["fit_trans"]
    models = []
    for wd in range(mu.shape[0]):
        S, M = mu.shape[1], mu.shape[2]
        m = rec.HMM(S)
        m.transitions = trans[wd].copy()
        m.gmm_states = []
        for s in range(S):
            g = rec.GMM(mu[wd, s, 0].copy(), cov[wd, s, 0].copy(), M)
            g.update_models(mu[wd, s].copy(), cov[wd, s].copy(), w[wd, s].copy())
            g.parent = wd
            m.gmm_states.append(g)
        models.append(m)
    return models


def main():
    z = np.load(os.path.join(HERE, "ref_train.npz"))
    off = z["ct_off"]
    data = [z["ct_X"][off[i]:off[i + 1]] for i in range(len(off) - 1)]
    strings = [list(map(int, s)) for s in z["ct_strings"]]
    out = {}
    for n_iter in (1, 2, 3):
        t0 = time.time()
        models = build_models(z)
        buf = io.StringIO()
        with tempfile.TemporaryDirectory() as td, warnings.catch_warnings(record=True) as wrn, \
                contextlib.redirect_stdout(buf):
            warnings.simplefilter("always")
            np.random.seed(3000)
            rec.continuous_train(data, models, strings, td, 4, 5, n_iter)
            new = [pickle.load(open(os.path.join(td, f"{i}.pkl"), "rb")) for i in range(len(models))]
        text = buf.getvalue()
        mu, cv, wt, tr = dump_models(new)
        out[f"it{n_iter}_mu"], out[f"it{n_iter}_cov"], out[f"it{n_iter}_w"], out[f"it{n_iter}_trans"] = mu, cv, wt, tr
        out[f"it{n_iter}_no_data_warnings"] = np.int64(sum("No MFCC data" in str(x.message) for x in wrn))
        out[f"it{n_iter}_outer"] = np.int64(len(re.findall(r"Continuous training iteration:", text)))
        out[f"it{n_iter}_converged"] = np.int64("Continuous training converged" in text)
        out[f"it{n_iter}_em_stop"] = np.array([int(k) for k in re.findall(r"EM converged at iteration: (\d+)", text)],
                                              np.int64)
        print(f"max_iteration={n_iter}: {time.time() - t0:.0f}s, warnings {out[f'it{n_iter}_no_data_warnings']}, "
              f"em stops {out[f'it{n_iter}_em_stop'][:8]}...", file=sys.stderr)
    # ---- stop iterations of the converged EM loops whose RESULTS are frozen in ref_kmeans_em.npz / ref_train.npz
    # (same inputs, same seeds as make_golden.py; only the printed "EM converged at iteration: k" is added here)
    zk = np.load(os.path.join(HERE, "ref_kmeans_em.npz"))
    for k in (2, 4):
        g = rec.GMM(zk[f"em{k}_mu_in"][0].copy(), zk[f"em{k}_cov_in"][0].copy(), 4)
        g.update_models(zk[f"em{k}_mu_in"].copy(), zk[f"em{k}_cov_in"].copy(), zk[f"em{k}_w_in"].copy())
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            g.em(zk["km_data"], k)
        out[f"em{k}_conv_stop"] = np.array([int(x) for x in re.findall(r"EM converged at iteration: (\d+)",
                                                                        buf.getvalue())], np.int64)
        assert np.allclose(np.array([d.mean for d in g.dists]), zk[f"em{k}_conv_mu"], rtol=1e-6, atol=1e-9)
    for w in range(3):
        o = z[f"tr_off{w}"]
        utts = [z[f"tr_X{w}"][o[i]:o[i + 1]] for i in range(len(o) - 1)]
        buf = io.StringIO()
        np.random.seed(1000 + w)
        with contextlib.redirect_stdout(buf):
            m = rec.HMM(5).fit(utts, n_gaussians=4, use_gmm=True, use_em=True)
        out[f"fit_em_stop{w}"] = np.array([int(x) for x in re.findall(r"EM converged at iteration: (\d+)",
                                                                       buf.getvalue())], np.int64)
        assert np.array_equal(dump_models([m])[0][0], z["fit_mu"][w])
        print(f"fit word {w}: em stops {out[f'fit_em_stop{w}']}", file=sys.stderr)
    path = os.path.join(HERE, "ref_train_multi.npz")
    np.savez_compressed(path, **out)
    print(f"train_multi: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
