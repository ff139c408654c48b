"""Emission layer of the sr.recognition mirror (reference: sr/recognition/hmm_state.py:5-179).

Same class names, attributes and call signatures as the reference so pickled models and callers keep working;
the arithmetic runs on the B200 through libsrb200 (K1 scorer, K5 statistics, M-step), in the log domain.
Objects hold only numpy / Python data (picklable, deep-copyable); device buffers are built per call or per
batched operation and never stored on the object.
"""
import uuid

import numpy as np
import torch

from . import _ops
from ._ops import engine


def _as_diag(cov):
    """The scorer is diagonal-covariance (BASELINE north_star).  A full matrix is accepted when it is diagonal."""
    cov = np.asarray(cov, dtype=np.float64)
    if cov.ndim == 1:
        return cov
    if cov.ndim == 2 and np.count_nonzero(cov - np.diag(np.diagonal(cov))) == 0:
        return np.diagonal(cov).copy()
    raise NotImplementedError("only diagonal covariances are supported by the B200 scorer")


class MultivariateNormal:
    """Diagonal Gaussian (hmm_state.py:5-45).  `cov` is the 1-D variance vector (or a diagonal matrix)."""

    def __init__(self, mean, cov):
        self.mean = mean
        self._cov = cov

    @property
    def cov(self):
        return self._cov

    @cov.setter
    def cov(self, val):
        self._cov = val

    @property
    def inv_cov(self):
        """Dense inverse kept for API compatibility (hmm_state.py:13-17); the kernels never use it."""
        return np.diag(1.0 / _as_diag(self._cov))

    def diag_cov(self):
        return _as_diag(self._cov)

    def __getstate__(self):
        """Pickles carry the dense `inv_cov` the reference's `pdf` reads (hmm_state.py:42), so a model written here can
        be loaded AND evaluated by the reference; it is derived data and is dropped again on load."""
        state = dict(self.__dict__)
        state["inv_cov"] = self.inv_cov
        return state

    def __setstate__(self, state):
        state = dict(state)
        state.pop("inv_cov", None)
        self.__dict__.update(state)

    def pdf(self, x):
        """N(x; mean, diag cov), hmm_state.py:36-45.  x: [D] (or [N, D] -> [N])."""
        x = np.asarray(x)
        mean = np.asarray(self.mean)
        if x.shape[-1] != mean.shape[0]:
            raise NameError("The dimensions of the input don't match")
        pm = engine.PackedModel(mean[None, None, :], self.diag_cov()[None, None, :], np.ones((1, 1)))
        out = engine.gmm_score(np.atleast_2d(x), pm, want_E=True)["E"]
        p = torch.exp(-out[:, 0].double()).cpu().numpy()
        return float(p[0]) if x.ndim == 1 else p


def mahalanobis(v1, v2, variance):
    """0.5*log((2 pi)^D prod var) + 0.5*sum (v1-v2)^2/var (hmm_state.py:48-58); v1 may be [N, D] -> [N]."""
    v1 = np.asarray(v1, dtype=np.float64)
    out, _ = engine.local_dist(np.atleast_2d(v1), np.atleast_2d(np.asarray(v2, np.float64)),
                               var=np.atleast_2d(np.asarray(variance, np.float64)))
    r = out[:, 0].double().cpu().numpy()
    return float(r[0]) if v1.ndim == 1 else r


class HMMState:
    """Base class of all HMM states (hmm_state.py:61-78): identity is the random `id`."""

    def __init__(self):
        self.id = uuid.uuid4().int
        self.parent = None

    def evaluate(self, x):
        raise NotImplementedError()

    def __eq__(self, other):
        raise NotImplementedError()

    def __hash__(self):
        return hash(self.id)


class NES(HMMState):
    """Non-emitting state (hmm_state.py:81-97): cost 0, compared by id, recognised by exact type."""

    def __init__(self):
        super().__init__()

    def evaluate(self, x):
        return 0

    def __eq__(self, other):
        return self.id == other.id

    def __hash__(self):
        return hash(self.id)


class GMM(HMMState):
    """Gaussian-mixture state (hmm_state.py:100-179): weights `w[n]`, components `dists[m].mean / .cov`.

    Weights are used as they are (the reference never normalises them: untrained components keep 1/n)."""

    def __init__(self, mu, sigma, n_gaussians):
        super().__init__()
        self.n_gaussians = n_gaussians
        self.w = np.full(n_gaussians, 1 / n_gaussians)
        self.dists = [MultivariateNormal(mean=mu, cov=sigma) for _ in range(n_gaussians)]
        self.mu_old = np.tile(mu, (n_gaussians, 1))
        self.sigma_old = np.tile(sigma, (n_gaussians, 1))
        self.w_old = np.full(n_gaussians, 1 / n_gaussians)

    # ---- scoring (K1)
    def evaluate(self, x, return_neg_log_likelihood=True):
        """x: [D] -> -log sum_m w_m N_m(x), or the vector w_m N_m(x) (hmm_state.py:114-120).
        x: [N, D] scores a batch in one launch -> [N] / [N, n_gaussians]."""
        x = np.asarray(x)
        pm = _ops.pack_gmms([self])
        res = engine.gmm_score(np.atleast_2d(x), pm, want_E=return_neg_log_likelihood,
                               want_comp=not return_neg_log_likelihood)
        if return_neg_log_likelihood:
            c = res["E"][:, 0].double().cpu().numpy()
            return float(c[0]) if x.ndim == 1 else c
        comp = torch.exp(res["comp"][:, 0, :self.n_gaussians].double()).cpu().numpy()
        return comp[0] if x.ndim == 1 else comp

    # ---- EM (K5 + M-step)
    def em(self, data, n_gaussians, max_iteration=10000):
        """EM over the first `n_gaussians` components on the pooled frames `data` [N, D]
        (hmm_state.py:122-159): posteriors, weighted moments and the update run on the device; the host only
        checks np.allclose convergence on the 2*n*D+n updated numbers per iteration."""
        Xd = _ops.to_device(data)
        self._em_device(Xd, n_gaussians, max_iteration)

    def _em_device(self, Xd, n, max_iteration, verbose=True):
        pm = _ops.pack_gmms([self])
        label = torch.zeros(max(Xd.shape[0], 1), dtype=torch.int32, device=Xd.device)
        for it in range(max_iteration):
            stats = engine.gmm_stats(Xd, label, pm, n)
            engine.mstep(stats, pm, n)
            mu = pm.mu[0, :n].cpu().numpy()
            sigma = pm.var[0, :n].cpu().numpy()
            w = pm.w[0, :n].cpu().numpy()
            self.update_models(mu, sigma, w)
            if np.allclose(mu, self.mu_old[:n, :]) and np.allclose(sigma, self.sigma_old[:n, :]) and \
                    np.allclose(w, self.w_old[:n]):
                if verbose:
                    print("EM converged at iteration:", it)
                break
            self.mu_old[:n, :] = mu
            self.sigma_old[:n, :] = sigma
            self.w_old[:n] = w

    def update_models(self, mus, sigmas, weights):
        k = mus.shape[0]
        self.w[:k] = weights
        for m in range(k):
            self.dists[m].mean = mus[m, :]
            self.dists[m].cov = sigmas[m, :]

    def __eq__(self, other):
        if self.n_gaussians != other.n_gaussians or not np.allclose(self.w, other.w):
            return False
        return all(np.allclose(a.mean, b.mean) and np.allclose(a.cov, b.cov) for a, b in zip(self.dists, other.dists))

    def __len__(self):
        return self.n_gaussians

    def __hash__(self):
        return hash(self.id)
