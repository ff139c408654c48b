"""N>1 host path on the CPU: two gloo ranks shard the utterances with trainer.shard_range, accumulate their shard's
statistics buffer (numpy stand-in of the device kernels, tests/engine_double.py), allreduce it exactly as
IsolatedTrainer.reduce does and apply the M-step; every rank must end with the model a single process gets from
the whole corpus (one fp64 allreduce per iteration, replicated M-step)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _shard_stats(X, off, words, models, S, ncomp, lo, hi):
    """statistics buffer [G*row + 2] of utterances [lo, hi): labels from a uniform segmentation (host logic only)."""
    import engine_double as ed
    mu = np.concatenate([m[0] for m in models])
    var = np.concatenate([m[1] for m in models])
    w = np.concatenate([m[2] for m in models])
    pm = ed.PackedModel(mu, var, w)
    lab = []
    for u in range(lo, hi):
        T = off[u + 1] - off[u]
        lab.append(words[u] * S + np.minimum(np.arange(T) * S // T, S - 1))
    lab = np.concatenate(lab) if lab else np.zeros(0, np.int64)
    st = ed.gmm_stats(X[off[lo]:off[hi]], lab, pm, ncomp).reshape(-1)
    return torch.cat([st, torch.tensor([float(len(lab)), float(hi - lo)], dtype=torch.float64)]), pm


def _worker(rank, world, port, out):
    sys.path[:0] = [HERE, os.path.dirname(HERE), os.path.join(os.path.dirname(HERE), "speech-recognition_b200")]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import engine_double as ed
    from srb200 import synth
    from srb200.trainer import shard_range
    W, S, M, ncomp = 3, 5, 4, 2
    models, base = synth.make_models(5, W=W, S=S, M=M)
    X, off, words = synth.make_utterances(6, base, 41, ragged=(20, 60))
    lo, hi = shard_range(len(words), rank, world)
    stats, pm = _shard_stats(X, off, words, models, S, ncomp, lo, hi)
    dist.all_reduce(stats)                                   # the one collective of an iteration
    G = W * S * M
    ed.mstep(stats[:G * 81].reshape(G, 81), pm, ncomp)
    torch.save({"mu": pm.mu, "var": pm.var, "w": pm.w, "tail": stats[-2:], "range": (lo, hi)},
               os.path.join(out, f"rank{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_allreduce_equals_single_process(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(tmp_path / f"rank{k}.pt") for k in range(world)]
    assert r[0]["range"][0] == 0 and r[0]["range"][1] == r[1]["range"][0] and r[1]["range"][1] == 41
    for k in ("mu", "var", "w", "tail"):
        assert torch.equal(r[0][k], r[1][k]), k             # replicated M-step: identical on every rank
    # single process over the whole corpus
    import engine_double as ed
    from srb200 import synth
    models, base = synth.make_models(5, W=3, S=5, M=4)
    X, off, words = synth.make_utterances(6, base, 41, ragged=(20, 60))
    stats, pm = _shard_stats(X, off, words, models, 5, 2, 0, 41)
    ed.mstep(stats[:60 * 81].reshape(60, 81), pm, 2)
    np.testing.assert_allclose(r[0]["mu"].numpy(), pm.mu.numpy(), rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(r[0]["var"].numpy(), pm.var.numpy(), rtol=1e-9)
    np.testing.assert_allclose(r[0]["w"].numpy(), pm.w.numpy(), rtol=1e-10)
    assert r[0]["tail"][1].item() == 41


def test_shard_range_partitions():
    from srb200.trainer import shard_range
    for n in (0, 1, 7, 100):
        for world in (1, 2, 3, 8):
            parts = [shard_range(n, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1


# ---------------------------------------------------------------------------------------------- sharded HMM.fit
class _Patch:
    """monkeypatch stand-in for engine_double.install inside a spawned worker."""

    @staticmethod
    def setattr(obj, name, value):
        setattr(obj, name, value)


def _fit_corpus():
    from srb200 import synth
    _, base = synth.make_models(7, W=3, S=5, M=4)
    return synth.make_utterances(8, base, 46, ragged=(30, 70))


def _fit(X, off, words, group=None):
    from srb200.fit import WordFitter
    ft = WordFitter(X, off, words, 3, 5, 4, use_em=True, seeds=[11, 12, 13], fused_align=False, group=group)
    models = ft.fit()
    return ft, models


def _fit_worker(rank, world, port, out):
    sys.path[:0] = [HERE, os.path.dirname(HERE), os.path.join(os.path.dirname(HERE), "speech-recognition_b200")]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import engine_double as ed
    ed.install(_Patch)
    from srb200.trainer import shard_range
    X, off, words = _fit_corpus()
    lo, hi = shard_range(len(words), rank, world)
    ft, models = _fit(X[off[lo]:off[hi]], off[lo:hi + 1] - off[lo], words[lo:hi])
    # final segment starts back in the shard's original utterance order
    starts = np.empty_like(ft.final_starts)
    starts[ft.order] = ft.final_starts
    torch.save({"models": models, "starts": starts, "range": (lo, hi), "its": ft.iterations}, os.path.join(out, f"fit{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_rank_fit_equals_single_process(tmp_path, monkeypatch):
    """WordFitter on two gloo ranks (each its contiguous shard of the utterances) = WordFitter on the whole corpus: same
    pools in the same order (rank offsets into the reference's RNG draws), same stop iterations, same models."""
    world = 2
    mp.spawn(_fit_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(tmp_path / f"fit{k}.pt", weights_only=False) for k in range(world)]
    import engine_double as ed
    ed.install(monkeypatch)
    X, off, words = _fit_corpus()
    ft, models = _fit(X, off, words)
    starts = np.empty_like(ft.final_starts)
    starts[ft.order] = ft.final_starts
    assert np.array_equal(np.concatenate([r[0]["starts"], r[1]["starts"]]), starts)
    for k in range(world):
        assert np.array_equal(r[k]["its"]["skmeans"], ft.iterations["skmeans"])
        for a, b in zip(r[k]["its"]["kmeans"] + r[k]["its"]["em"], ft.iterations["kmeans"] + ft.iterations["em"]):
            assert np.array_equal(a, b)
        for (mu, var, w, tr), (mu1, var1, w1, tr1) in zip(r[k]["models"], models):
            np.testing.assert_allclose(mu, mu1, rtol=1e-9, atol=1e-11)
            np.testing.assert_allclose(var, var1, rtol=1e-8)
            np.testing.assert_allclose(w, w1, rtol=1e-9)
            np.testing.assert_array_equal(tr, tr1)
    for a, b in zip(r[0]["models"], r[1]["models"]):
        for x, y in zip(a, b):
            assert np.array_equal(x, y)                      # replicated updates: identical on every rank
