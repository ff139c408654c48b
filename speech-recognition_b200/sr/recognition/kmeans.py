"""Segmental K-means / K-means layer of the sr.recognition mirror (reference: sr/recognition/kmeans.py:6-205).

All utterances of a call are concatenated once into one device matrix; alignment (K2 on local distances or on
GMM emissions), assignment (srb_local_dist argmin) and the segment / cluster moments (srb_group_stats) are
batched launches over it.  The host keeps only the [n_utterances, n_segments] boundary bookkeeping.
"""
import numpy as np
import torch

from . import _ops
from ._ops import StateGraph, engine
from .decode import _local_distances, decode_batch, euclidean
from .hmm_state import mahalanobis


def calc_variance(data):
    """Unbiased per-row variance of data [D, N] (np.cov(data).diagonal(), kmeans.py:6-12)."""
    Xd = _ops.to_device(np.ascontiguousarray(np.asarray(data).T))
    lab = torch.zeros(max(Xd.shape[0], 1), dtype=torch.int32, device=Xd.device)
    return _ops.group_moments(Xd, lab, 1)[2][0]


def segment_data(templates, n_temps, n_segments, seg_starts):
    """Pools the frames of segment s over all templates (kmeans.py:33-50) -> list of n_segments arrays."""
    segments = []
    for s in range(n_segments):
        parts = []
        for r in range(n_temps):
            hi = None if s == n_segments - 1 else seg_starts[r, s + 1]
            parts.append(np.asarray(templates[r])[seg_starts[r, s]:hi])
        segments.append(np.concatenate(parts, axis=0))
    return segments


def _segment_moments(Xd, off, seg_starts, n_segments):
    lab = _ops.segment_labels(off, seg_starts, n_segments)
    _, mean, var = _ops.group_moments(Xd, lab, n_segments)
    return mean, var


def combine_templates(templates, n_temps, n_segments, seg_starts):
    """Per-segment mean and unbiased variance over all templates (kmeans.py:15-30)."""
    X, off = _ops.cat_frames(templates[:n_temps])
    return _segment_moments(_ops.to_device(X), off, np.asarray(seg_starts), n_segments)


def calc_transition_costs(n_temps, seg_lens, max_jump_dist=2):
    """Count-based left-to-right costs (kmeans.py:53-95): res[i, i] = -log p_stay, res[i+jump, i] = -log p_jump,
    where a jump skips following segments that are empty in some template (up to max_jump_dist)."""
    seg_lens = np.asarray(seg_lens)
    S = seg_lens.shape[1]
    has_empty = (seg_lens == 0).any(axis=0)
    n_all = seg_lens[:n_temps].sum(axis=0)
    res = np.full((S, S), np.inf)
    with np.errstate(divide="ignore", invalid="ignore"):
        for i in range(S):
            n_jump = 0 if i == S - 1 else n_temps
            jump, s = 1, i + 1
            while s < S - 1 and has_empty[s + 1]:
                jump += 1
                if jump > max_jump_dist:
                    break
                s += 1
            if n_jump:
                res[i + jump, i] = -np.log(n_jump / n_all[i])
            res[i, i] = -np.log((n_all[i] - n_jump) / n_all[i])
    return res


def get_segments_from_path(path, n_segments):
    """Start frames of segments 1..S-1 from a reversed alignment path (kmeans.py:98-108)."""
    counts = np.bincount(np.asarray(path)[:, 0].astype(np.int64), minlength=n_segments)[:n_segments]
    return np.cumsum(counts)[:-1]


def _align_to_templates(Xd, X, off, res, dist_fun, transition_costs, n_segments):
    """dtw of every utterance against the segment means in one launch -> seg_starts [n, S]."""
    dist = _local_distances(Xd, X, res, dist_fun, None)
    graph = StateGraph(transition_costs, np.arange(n_segments), end_states=[n_segments - 1], keep_inf=True)
    v = _ops.viterbi_batch(dist, off, [graph], want_path=False, want_label=True)
    return _ops.starts_from_labels(v.frame_label.cpu().numpy(), off, n_segments)


def skmeans(templates, n_segments, dist_fun=euclidean, return_segmented_data=False, max_iteration=1000):
    """Segmental K-means (kmeans.py:111-155): uniform initial segmentation, then alternate batched DTW alignment
    against the segment means and re-estimation until np.allclose.  -> (means, variances, transitions[, segments])"""
    assert max_iteration > 0
    n_temps = len(templates)
    X, off = _ops.cat_frames(templates)
    Xd = _ops.to_device(X)
    T = np.diff(off.astype(np.int64))
    seg_lens = np.repeat((T // n_segments)[:, None], n_segments, axis=1)
    seg_starts = np.concatenate([np.zeros((n_temps, 1), dtype=int), np.cumsum(seg_lens, axis=1)[:, :-1]], axis=1)
    res, vars_ = _segment_moments(Xd, off, seg_starts, n_segments)
    transition_costs = None
    for _ in range(max_iteration):
        transition_costs = calc_transition_costs(n_temps, seg_lens)     # from the uniform lengths every time (A5)
        if (T < 5).any():
            raise NameError('template is too small, cannot do dtw on it')
        seg_starts = _align_to_templates(Xd, X, off, res, dist_fun, transition_costs, n_segments)
        new_res, vars_ = _segment_moments(Xd, off, seg_starts, n_segments)
        if np.allclose(res, new_res):
            break
        res = new_res
    if return_segmented_data:
        return res, vars_, transition_costs, segment_data(templates, n_temps, n_segments, seg_starts)
    return res, vars_, transition_costs


def cluster_centroids(data, clusters, k):
    """Mean of each cluster (kmeans.py:158-164); an empty cluster gives NaN like np.mean of nothing."""
    Xd = _ops.to_device(data)
    return _ops.group_moments(Xd, np.asarray(clusters), k)[1]


def _kmeans_device(Xd, X_host, k, centroids, dist_fun, max_iteration):
    n = Xd.shape[0]
    clusters = np.random.randint(0, k, n)                                # kmeans.py:171 (global RNG stream)
    cov = _ops.group_moments(Xd, clusters, k)[2]                         # variances of the RANDOM partition (A3)
    centroids = np.asarray(centroids, np.float64)
    shift = centroids.copy()                                             # fixed for the whole call: equal clusters
    for _ in range(max(max_iteration, 1)):                               # then give bit-equal centroids
        if dist_fun is mahalanobis:
            _, am = engine.local_dist(Xd, centroids, var=cov[:1], shared_var=True, want_dist=False, want_argmin=True)
        elif dist_fun is euclidean:
            _, am = engine.local_dist(Xd, centroids, want_dist=False, want_argmin=True)
        else:
            d = np.array([[dist_fun(centroids[c], xx, cov[0]) for c in range(k)] for xx in np.asarray(X_host)])
            am = torch.from_numpy(np.argmin(d, axis=1).astype(np.int32)).to(Xd.device)
        new_centroids = _ops.group_moments(Xd, am, k, shift=shift)[1]
        clusters = am.cpu().numpy().astype(np.int64)
        if np.array_equal(new_centroids, centroids):
            break
        centroids = new_centroids
    return clusters, centroids, cov


def kmeans(data, k, centroids, dist_fun=euclidean, max_iteration=1000):
    """The reference's modified K-means (kmeans.py:166-193): distances use the variance of cluster 0 of a random
    initial partition for every centroid; stops when the centroids repeat exactly.
    -> (clusters [N], centroids [k, D], cov [k, D] of the initial partition)"""
    assert k == centroids.shape[0]
    return _kmeans_device(_ops.to_device(data), data, k, centroids, dist_fun, max_iteration)


def _align_starts(templates, gmm_states, transition_costs, n_segments):
    res = decode_batch(templates, [(gmm_states, np.asarray(transition_costs), None)], want_path=False,
                       want_label=True)
    return _ops.starts_from_labels(res.frame_labels(), res.off, n_segments)


def align_gmm_states(templates, gmm_states, transition_costs, n_segments):
    """Viterbi re-alignment of all templates with the GMM states, pooled per state (kmeans.py:196-205)."""
    seg_starts = _align_starts(templates, gmm_states, transition_costs, n_segments)
    return segment_data(templates, len(templates), n_segments, seg_starts)


__all__ = ["calc_variance", "combine_templates", "segment_data", "calc_transition_costs", "get_segments_from_path",
           "skmeans", "cluster_centroids", "kmeans", "align_gmm_states"]
