"""srb_group_stats_static (include/srb200.h): the allocation-free variant of srb_group_stats, its private slices in a
file-scope `static __device__` array.  Collected last (file name) so that a failure here hides no other test.  Checked
through the C-ABI against srb_group_stats on two calls in a row (the array keeps the first call's contents: the second
call uses fewer groups and other labels, so stale slices would show) and for the size limit."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(L, _l, X, label, K, shift):
    dev, D = X.device, X.shape[1]
    row = L.srb_stats_row(D)
    ref = torch.zeros((K, row), dtype=torch.float64, device=dev)
    out = torch.full((K, row), 7.0, dtype=torch.float64, device=dev)        # accumulate = 0 must overwrite
    partial = torch.empty(L.srb_stats_parts() * K * row, dtype=torch.float32, device=dev)
    st = _l.stream_ptr()
    _l.check(L.srb_group_stats(_l.ptr(X), X.shape[0], D, _l.ptr(label), K, _l.ptr(shift), _l.ptr(partial), _l.ptr(ref), 0, st))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    _l.check(L.srb_group_stats_static(_l.ptr(X), X.shape[0], D, _l.ptr(label), K, _l.ptr(shift), _l.ptr(out), 0, st))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return ref.cpu().numpy(), out.cpu().numpy()


def test_group_stats_static_matches_group_stats_twice():
    from srb200 import lib as _l
    L = _l.load()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(11)
    D = 39
    for K, N in ((55, 300_000), (7, 41_003)):                      # second call: fewer groups, ragged size, same array
        X = torch.randn((N, D), generator=g).to(dev)
        label = torch.randint(-1, K, (N,), generator=g, dtype=torch.int32).to(dev)
        shift = torch.randn((K, D), generator=g, dtype=torch.float64).to(dev)
        ref, out = _both(L, _l, X, label, K, shift)
        # same kernels, same slices, same fixed-order double reduction: the results are bit-identical
        assert np.array_equal(ref, out)
        lab = label.cpu().numpy()
        assert np.array_equal(out[:, 0], np.bincount(lab[lab >= 0], minlength=K).astype(np.float64))
    # accumulate != 0 adds to `out`
    row = L.srb_stats_row(D)
    acc = torch.from_numpy(out).to(dev).clone()
    _l.check(L.srb_group_stats_static(_l.ptr(X), N, D, _l.ptr(label), K, _l.ptr(shift), _l.ptr(acc), 1, _l.stream_ptr()))
    torch.cuda.synchronize()
    np.testing.assert_allclose(acc.cpu().numpy(), 2 * out, rtol=1e-12)
    assert acc.shape == (K, row)


def test_group_stats_static_rejects_oversize():
    from srb200 import lib as _l
    L = _l.load()
    dev = torch.device("cuda", 0)
    D = 39
    cap = L.srb_static_scratch_floats() // (L.srb_stats_parts() * L.srb_stats_row(D))
    assert cap >= 55
    K = int(cap) + 1
    X = torch.zeros((8, D), device=dev)
    label = torch.zeros(8, dtype=torch.int32, device=dev)
    shift = torch.zeros((K, D), dtype=torch.float64, device=dev)
    out = torch.zeros((K, L.srb_stats_row(D)), dtype=torch.float64, device=dev)
    rc = L.srb_group_stats_static(_l.ptr(X), 8, D, _l.ptr(label), K, _l.ptr(shift), _l.ptr(out), 0, _l.stream_ptr())
    assert rc == -2                                                # SRB_EUNSUPPORTED
    assert "static scratch" in L.srb_last_error_string().decode()
    with pytest.raises(_l.SrbError):
        _l.check(rc)
