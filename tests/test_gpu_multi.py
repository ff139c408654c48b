"""Multi-GPU check on real hardware (needs >= 2 GPUs; the driver's one-GPU run skips it, `gpurun --gpus 2` runs it):
the N-rank training pass equals the 1-rank one on the same corpus and all replicas stay bit-identical
(tests/multi_rank_check.py).  The host-side sharding logic is covered on the CPU by tests/test_dist_gloo.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_single_rank_peer_tail_equals_collective_tail(tmp_path):
    """One rank: srb_em_push + srb_em_update against reduce_slices + srb_mstep + srb_pack_model(_tc)."""
    r = subprocess.run([sys.executable, os.path.join(HERE, "multi_rank_check.py"), str(tmp_path / "ref.npz")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_n_ranks_equal_one_rank(tmp_path, world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ref = str(tmp_path / "ref.npz")
    r = subprocess.run([sys.executable, os.path.join(HERE, "multi_rank_check.py"), ref], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29500 + world),
                        os.path.join(HERE, "multi_rank_check.py"), ref], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-12000:]
