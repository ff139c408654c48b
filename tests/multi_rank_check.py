"""Run by tests/test_gpu_multi.py, once as a plain process (writes the 1-rank result) and once under torchrun (N ranks):
three fused training iterations of the 11-word digit model on the SAME corpus, sharded by utterance over the ranks
(trainer.shard_range), through the peer-memory tail (srb_em_push / srb_em_update) and through the NCCL tail.

Checks: (1) all ranks hold bit-identical models after every iteration (the slots are added in rank order on every rank);
(2) the N-rank statistics and models equal the 1-rank ones to 1e-5 of a row's scale (measured 3e-6 at 2 ranks) -- not to
fp64 round-off: a CTA folds the runs of its tiles in FP32 (registers, per-warp shared partials, its private slice) before
the fixed-order fp64 reduction, and which tiles a CTA gets depends on the shard; the bound is sqrt(n) * 6e-8 per slice;
(3) the peer tail and the NCCL tail give the same model to the same tolerance."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "speech-recognition_b200")):
    sys.path.insert(0, p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from srb200 import engine, synth  # noqa: E402
from srb200.trainer import IsolatedTrainer, shard_range, sort_by_word  # noqa: E402

W, S, M, B, ITER = 11, 5, 4, 6000, 3


def run(peer, rank, world):
    models, base = synth.make_models(1, W=W, S=S, M=M)
    X, off, words = synth.make_utterances(100, base, B, ragged=(60, 110))
    lo, hi = shard_range(B, rank, world)
    X, words = X[int(off[lo]):int(off[hi])], words[lo:hi]
    off = (off[lo:hi + 1] - off[lo]).astype(np.int32)
    X, off, words, _ = sort_by_word(X, off, words)
    pm, _ = engine.PackedModel.from_words(models)
    tr = IsolatedTrainer(X, off, words, pm, np.stack([m[3] for m in models]), S, 2, peer=peer)
    assert tr.fused and (tr.peer is not None) == peer
    out = []
    for _ in range(ITER):
        tr.em_iteration()
        torch.cuda.synchronize()
        out.append((tr.stats.cpu().numpy().copy(), pm.mu.cpu().numpy().copy(), pm.var.cpu().numpy().copy(),
                    pm.w.cpu().numpy().copy(), pm.ab.cpu().numpy().copy(), pm.tcb.cpu().numpy().copy()))
    return out


def run_fit(rank, world):
    """The complete HMM.fit of 5 words (WordFitter: skmeans, split K-means, EM, re-alignment) on this rank's shard."""
    from srb200.fit import WordFitter
    Wf, Bf = 5, 1500
    _, base = synth.make_models(3, W=Wf, S=S, M=M)
    X, off, words = synth.make_utterances(101, base, Bf, ragged=(60, 110))
    lo, hi = shard_range(Bf, rank, world)
    ft = WordFitter(X[int(off[lo]):int(off[hi])], off[lo:hi + 1] - off[lo], words[lo:hi], Wf, S, M, use_em=True,
                    seeds=[50 + w for w in range(Wf)])
    models = ft.fit()
    starts = np.empty_like(ft.final_starts)
    starts[ft.order] = ft.final_starts
    return {"mu": np.stack([m[0] for m in models]), "var": np.stack([m[1] for m in models]),
            "w": np.stack([m[2] for m in models]), "trans": np.stack([m[3] for m in models]), "starts": starts,
            "lo": lo, "hi": hi}


def close(a, b, tol):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = np.maximum(np.abs(b).max(axis=-1, keepdims=True), 1e-30) if b.ndim > 1 else np.maximum(np.abs(b), 1e-30)
    assert np.max(np.abs(a - b) / scale) < tol, float(np.max(np.abs(a - b) / scale))


def main():
    path = sys.argv[1]
    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if world == 1:
        res = run(True, 0, 1)
        res_nccl = run(False, 0, 1)
        for a, b in zip(res, res_nccl):                       # one rank: the two tails on the same slices
            close(a[0][:-2].reshape(-1, 81), b[0][:-2].reshape(-1, 81), 1e-12)     # fp64 sums of the same slices, other order
            np.testing.assert_allclose(a[0][-2:], b[0][-2:], rtol=1e-12)
            close(a[1], b[1], 1e-12)
            np.testing.assert_allclose(a[4], b[4], rtol=1e-6, atol=1e-30)         # the FFMA packing
            ta, tb = a[5].reshape(2, -1).sum(axis=0), b[5].reshape(2, -1).sum(axis=0)   # tensor-core operand: hi + lo
            np.testing.assert_allclose(ta, tb, rtol=1e-6, atol=1e-30)             # (the constant's fp64 sum order differs)
        fit = run_fit(0, 1)
        np.savez(path, **{f"it{i}_{k}": v for i, r in enumerate(res) for k, v in zip(("stats", "mu", "var", "w"), r)},
                 **{f"fit_{k}": v for k, v in fit.items()})
        print("single-rank reference written")
        return
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))))
    ref = np.load(path)
    for peer in (True, False):
        res = run(peer, rank, world)
        for i, (stats, mu, var, w, ab, tcb) in enumerate(res):
            # (1) bit-identical replicas
            for arr in (stats, mu, var, w, ab, tcb):
                t = torch.from_numpy(np.ascontiguousarray(arr)).cuda().view(torch.uint8).to(torch.int32)
                lo_, hi_ = t.clone(), t.clone()
                dist.all_reduce(lo_, op=dist.ReduceOp.MIN)
                dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
                assert torch.equal(lo_, hi_), f"replicas differ (peer={peer}, iteration {i})"
            # (2) N ranks == 1 rank
            close(stats[:-2].reshape(-1, 81), ref[f"it{i}_stats"][:-2].reshape(-1, 81), 1e-5)
            np.testing.assert_allclose(stats[-2:], ref[f"it{i}_stats"][-2:], rtol=1e-6)
            close(mu, ref[f"it{i}_mu"], 1e-5)
            close(var, ref[f"it{i}_var"], 1e-5)
            close(w, ref[f"it{i}_w"], 1e-5)
    # (4) the sharded full fit: same models (every stop test of the reference is taken on allreduced statistics, so the
    # ranks agree on it; against one rank the fp32 in-kernel partials may move a borderline stop by an iteration)
    fit = run_fit(rank, world)
    for k in ("mu", "var", "w"):
        t = torch.from_numpy(fit[k]).cuda().view(torch.uint8).to(torch.int32)
        lo_, hi_ = t.clone(), t.clone()
        dist.all_reduce(lo_, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
        assert torch.equal(lo_, hi_), f"fit: replicas differ in {k}"
        close(fit[k], ref[f"fit_{k}"], 1e-3)
    np.testing.assert_array_equal(fit["trans"], ref["fit_trans"])
    same = (fit["starts"] == ref["fit_starts"][fit["lo"]:fit["hi"]]).all(axis=1).mean()
    assert same > 0.995, same
    dist.barrier()
    if rank == 0:
        print(f"{world} ranks == 1 rank (peer-memory tail, NCCL tail, sharded HMM.fit)")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
