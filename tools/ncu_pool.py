"""A few launches of srb_pool_stats (EM, k = 2; K-means, k = 2) on the config-2 corpus, for ncu."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "speech-recognition_b200")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from srb200 import engine, synth  # noqa: E402

W, S, M, k = 11, 5, 4, 2
models, base = synth.make_models(1, W=W, S=S, M=M)
X, off, words = synth.make_utterances(100, base, int(os.environ.get("UTTS", 100000)), T=100)
N = X.shape[0]
t_in = np.arange(N) % 100
state = torch.from_numpy((np.repeat(words, 100) * S + t_in // 20).astype(np.int32)).cuda()
Xd = torch.from_numpy(X).cuda()
pm, _ = engine.PackedModel.from_words(models)
cen = pm.mu[:, :k].contiguous()
for _ in range(3):
    engine.pool_em_stats(Xd, state, pm, k)
    engine.pool_em_stats(Xd, state, pm, k, f64=True)
    engine.pool_kmeans_stats(Xd, state, cen, pm.var[:, 0].contiguous(), cen.reshape(-1, 39))
torch.cuda.synchronize()
