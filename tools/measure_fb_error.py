#!/usr/bin/env python
"""Measures the FP32 error of the forward-backward kernels and of one Baum-Welch iteration against the float64 oracle
(run on the GPU box; writes gpurun_out/fb_error.json).  The tolerances of tests/test_gpu_dp.py and tests/test_gpu_bw.py
are set from these numbers."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "speech-recognition_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

from oracle import oracle as orc  # noqa: E402
from srb200 import engine, synth  # noqa: E402
from srb200.graph import GraphBatch, StateGraph  # noqa: E402

out = {}
for S, M, banded, bs in ((5, 4, True, 1.0), (5, 4, True, 0.12), (5, 4, False, 0.12), (8, 16, True, 0.12), (8, 16, False, 0.12),
                         (20, 2, True, 0.12), (20, 2, False, 0.12), (8, 16, True, 0.05)):
    models, base = synth.make_models(41, W=3, S=S, M=M, base_scale=bs, spread=0.5 * bs)
    X, off, words = synth.make_utterances(42, base, 40, ragged=(S + 3, 300), noise=1.0 if bs < 1 else 0.7)
    pm, _ = engine.PackedModel.from_words(models)
    E = engine.gmm_score(X.astype(np.float32), pm)["E"]
    Eh = E.double().cpu().numpy()
    graphs = [StateGraph(m[3], np.arange(S) + S * w) for w, m in enumerate(models)]
    gb = GraphBatch(graphs, E.device)
    gamma, ws, nlz, xi = engine.forward_backward(E, off, gb, utt_graph=words.astype(np.int32), want_xi=True, banded=banded)
    gamma, nlz, xi = gamma.cpu().numpy(), nlz.cpu().numpy(), xi.cpu().numpy()
    xi_ref = np.zeros_like(xi)
    g_err, g_err_same, z_err, z_err_same = [], [], [], []
    for u in range(len(off) - 1):
        m = models[words[u]]
        x = X[off[u]:off[u + 1]].astype(np.float64)
        Eo = orc.emissions(x, [m], np.arange(S), log_domain=True)                 # float64 emissions
        Es = Eh[off[u]:off[u + 1], S * words[u]:S * words[u] + S]                 # the kernel's own fp32 emissions
        got = gamma[ws[u]:ws[u + 1]].reshape(-1, S)
        for tag, Em, ge, ze in (("f64", Eo, g_err, z_err), ("same", Es, g_err_same, z_err_same)):
            z, a, b, go = orc.forward_backward(Em, m[3])
            big = go > 1e-3
            ge.append(np.max(np.abs(got[big] - go[big]) / go[big]))
            ze.append(abs(nlz[u] - z) / abs(z))
            if tag == "same":
                sg, e0 = graphs[words[u]], gb.edge_off[words[u]]
                for j in range(sg.nnz):
                    o, r = sg.pred_idx[j], sg.pred_tgt[j]
                    xi_ref[e0 + j] += np.exp(-(a[:-1, o] + m[3][r, o] + Em[1:, r] + b[1:, r] - z)).sum()
    big = xi_ref > 1e-2
    key = f"S{S}_M{M}_{'banded' if banded else 'generic'}_scale{bs}"
    out[key] = {
        "gamma_entries_between_0.05_and_0.95": float(np.mean((gamma > 0.05) & (gamma < 0.95))),
        "gamma_rel_err_vs_f64_emissions(gamma>1e-3)": float(np.max(g_err)),
        "gamma_rel_err_same_emissions(gamma>1e-3)": float(np.max(g_err_same)),
        "nlz_rel_err_vs_f64_emissions": float(np.max(z_err)), "nlz_rel_err_same_emissions": float(np.max(z_err_same)),
        "xi_rel_err_same_emissions(xi>1e-2)": float(np.max(np.abs(xi[big] - xi_ref[big]) / xi_ref[big])),
    }
    # run-to-run reproducibility of the edge posteriors (no atomics)
    xi2 = engine.forward_backward(E, off, gb, utt_graph=words.astype(np.int32), want_xi=True, banded=banded)[3]
    out[key]["xi_bit_reproducible"] = bool(np.array_equal(xi2.cpu().numpy(), xi))

# ---- one Baum-Welch iteration: parameter errors
import test_gpu_bw as tb  # noqa: E402
from srb200.trainer import BaumWelchTrainer, sort_by_word  # noqa: E402

for S, M, ncomp, tensor, bs in ((8, 16, 4, True, 1.0), (8, 16, 4, True, 0.12), (8, 16, 4, False, 0.12), (5, 4, 2, False, 0.12),
                                (4, 8, 4, True, 0.12)):
    W = 2
    models, base = synth.make_models(81, W=W, S=S, M=M, base_scale=bs, spread=0.5 * bs)
    X, off, words = synth.make_utterances(82, base, 48, ragged=(12, 80), noise=1.1)
    X, off, words, _ = sort_by_word(X, off, words)
    pm, _ = engine.PackedModel.from_words(models)
    tr = BaumWelchTrainer(X, off, words, pm, np.stack([m[3] for m in models]), S, ncomp, gmin=0.0, tensor=tensor)
    tr.em_iteration()
    mu, var, w, trans, nlz = tb.numpy_iteration(X.astype(np.float64), off.astype(np.int64), words, models, S, ncomp)
    gm, gv, gw = pm.mu.cpu().numpy(), pm.var.cpu().numpy(), pm.w.cpu().numpy()
    sd = np.sqrt(var)
    sl = np.s_[:, :ncomp]
    fin = np.isfinite(trans)
    out[f"bw_S{S}_M{M}_n{ncomp}_{'tc' if tensor else 'ffma'}_scale{bs}"] = {
        "w_rel": float(np.max(np.abs(gw[sl] - w[sl]) / w[sl])),
        "mu_err_over_sd": float(np.max(np.abs(gm[sl] - mu[sl]) / sd[sl])),
        "mu_err_over_max(|mu|,sd)": float(np.max(np.abs(gm[sl] - mu[sl]) / np.maximum(np.abs(mu[sl]), sd[sl]))),
        "var_rel": float(np.max(np.abs(gv[sl] - var[sl]) / var[sl])),
        "trans_abs": float(np.max(np.abs(tr.trans[fin] - trans[fin]))),
        "trans_prob_rel": float(np.max(np.abs(np.exp(-tr.trans[fin]) - np.exp(-trans[fin])) / np.exp(-trans[fin]))),
        "nlz_rel": float(abs(tr.mean_neg_log_z() - nlz) / abs(nlz)),
    }
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "fb_error.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
