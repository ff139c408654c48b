"""Where one K-means / EM iteration of WordFitter spends its time (config 2: 100 k utterances x 100 frames, 55 states)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "speech-recognition_b200")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from srb200 import engine, synth  # noqa: E402
from srb200.fit import WordFitter  # noqa: E402

W, S, M = 11, 5, 4
_, base = synth.make_models(1, W=W, S=S, M=M)
X, off, words = synth.make_utterances(100, base, int(os.environ.get("UTTS", 100000)), T=100)
ft = WordFitter(X, off, words, W, S, M, seeds=[1000 + w for w in range(W)])
ft.skmeans()


def timed(name, fn, n=20):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    print(f"{name:28s} {(time.perf_counter() - t0) / n * 1e3:8.3f} ms", flush=True)


G, D, k = W * S, ft.D, 2
state = ft._labels_from_starts(ft.seg_starts)
cen = torch.cat([ft._t(ft.hmm_mu.reshape(G, 1, D)) * 0.9, ft._t(ft.hmm_mu.reshape(G, 1, D)) * 1.1], dim=1)
cov0 = ft._t(ft.hmm_sigma.reshape(G, D))
shift = cen.clone().reshape(G * k, D)
am = engine.kmeans_assign(ft.X, state, cen, cov0)
lab = (state.long() * k + am.long()).to(torch.int32)
timed("kmeans_assign", lambda: engine.kmeans_assign(ft.X, state, cen, cov0))
timed("label arithmetic (torch)", lambda: (state.long() * k + am.long()).to(torch.int32))
timed("group_stats", lambda: engine.group_stats(ft.X, lab, G * k, shift))
mu = np.repeat(ft.hmm_mu.reshape(G, 1, D), M, axis=1) * np.array([0.9, 1.1, 1, 1])[None, :, None]
var = np.repeat(ft.hmm_sigma.reshape(G, 1, D), M, axis=1)
pm = engine.PackedModel(mu, var, np.full((G, M), 0.25))
timed("gmm_stats", lambda: engine.gmm_stats(ft.X, state, pm, k))
timed("repack", pm.repack)
timed("gauss_consts", lambda: engine.gauss_consts(pm.var, pm.w))
ft._old = [pm.mu.clone(), pm.var.clone(), pm.w.clone()]
ft._n_loc = torch.bincount(state.long(), minlength=G)
has = torch.ones(G, dtype=torch.bool, device=ft.dev)
for n in (8, 64):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ft._em(pm, state, k, has.clone(), n)
    torch.cuda.synchronize()
    print(f"_em, {n} iterations: {(time.perf_counter() - t0) / n * 1e3:.3f} ms / iteration", flush=True)
    t0 = time.perf_counter()
    ft._kmeans(state, k, cen, cov0, has.clone(), n)
    torch.cuda.synchronize()
    print(f"_kmeans, {n} iterations: {(time.perf_counter() - t0) / n * 1e3:.3f} ms / iteration", flush=True)
timed("pool_em_stats", lambda: engine.pool_em_stats(ft.X, state, pm, k))
timed("pool_kmeans_stats", lambda: engine.pool_kmeans_stats(ft.X, state, cen, cov0, shift))
pm4 = engine.PackedModel(np.repeat(ft.hmm_mu.reshape(G, 1, D), M, axis=1) * np.array([0.9, 1.1, 0.95, 1.05])[None, :, None],
                         var, np.full((G, M), 0.25))
timed("pool_em_stats k=4", lambda: engine.pool_em_stats(ft.X, state, pm4, 4))
timed("pool_em_stats f64", lambda: engine.pool_em_stats(ft.X, state, pm, k, f64=True))
timed("pool_kmeans_stats f64", lambda: engine.pool_kmeans_stats(ft.X, state, cen, cov0, shift, f64=True))
