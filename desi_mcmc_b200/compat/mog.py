"""CelestePy/util/dists/mog.py call surface.  Parameter algebra (convolve,
affine maps, convex combinations: a few dozen flops on (K,2,2) arrays) stays
on the host like the reference; every density evaluation runs on the device
through cg_gmm_like_2d (A3 == log of A2)."""
from __future__ import annotations

import numpy as np
import torch

from .. import engine


def _gmm_device(x, ws, mus, sigs):
    dev = engine.device()
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
    return engine.gmm_like_2d_device(d(x), d(ws), d(mus), d(sigs)).cpu().numpy()


def mog_loglike(x, means, icovs, dets, pis):
    """mog.py:5-21: logsumexp_k[-1/2 d^T icov_k d - log 2 pi - 1/2 log det_k + log pi_k] at the rows
    of x, as written (inverse covariances and determinants in, log-density out).  Evaluated by the
    float64 log-space kernel (cg_mog_logpdf), so the far tails stay finite exactly where the
    reference's logsumexp does (a density evaluated as log(sum exp) underflows to -inf there)."""
    x = np.asarray(x, dtype=np.float64)
    xx = np.atleast_2d(x)
    icovs = np.asarray(icovs, dtype=np.float64)
    if xx.shape[1] != 2 or icovs.shape[1:] != (2, 2):
        raise ValueError("must be two-dimensional!")
    lp = engine.mog_logpdf(xx, means, icovs, dets, pis)
    return lp[0] if x.ndim == 1 else lp


class MixtureOfGaussians(object):
    """mog.py:38-112.  dets / icovs / chols are closed-form 2x2 (the reference
    calls numpy.linalg per component, :55-57)."""

    def __init__(self, means, covs, pis):
        means = np.asarray(means, dtype=np.float64)
        self.K, self.D = means.shape
        self.update_params(means, covs, pis)

    def update_params(self, means, covs, pis):
        covs = np.asarray(covs, dtype=np.float64)
        pis = np.asarray(pis, dtype=np.float64)
        assert covs.shape[1] == covs.shape[2] == self.D
        assert self.K == covs.shape[0] == len(pis), "%d != %d != %d" % (self.K, covs.shape[0], len(pis))
        self.means, self.covs, self.pis = np.asarray(means, dtype=np.float64), covs, pis
        if self.D == 2:
            a, b, c, d = covs[:, 0, 0], covs[:, 0, 1], covs[:, 1, 0], covs[:, 1, 1]
            self.dets = a * d - b * c
            self.icovs = np.stack([np.stack([d, -b], -1), np.stack([-c, a], -1)], -2) / self.dets[:, None, None]
            l00 = np.sqrt(a)
            l10 = c / l00
            self.chols = np.zeros_like(covs)
            self.chols[:, 0, 0], self.chols[:, 1, 0] = l00, l10
            self.chols[:, 1, 1] = np.sqrt(d - l10 * l10)
        else:
            self.dets = np.array([np.linalg.det(c) for c in covs])
            self.icovs = np.array([np.linalg.inv(c) for c in covs])
            self.chols = np.array([np.linalg.cholesky(c) for c in covs])

    def logpdf(self, x):
        # :59-61 -> mog_loglike (:5-21): float64 log-space kernel for every D (finite far tails)
        return engine.mog_logpdf(np.asarray(x, dtype=np.float64), self.means, self.icovs, self.dets, self.pis)

    def responsibilities(self, x):
        """logpdf and r_nk = p(k | x_n) (device tensor [N, K]) -- the E-step of
        the prior refit, whose sufficient statistics go into the sweep all-reduce."""
        return engine.mog_logpdf(np.asarray(x, dtype=np.float64), self.means, self.icovs, self.dets,
                                 self.pis, return_resp=True)

    def pdf(self, x):
        x = np.asarray(x, dtype=np.float64)
        p = _gmm_device(np.atleast_2d(x), self.pis, self.means, self.covs)
        return p[0] if x.ndim == 1 else p

    def mean(self, x=None):
        return np.dot(self.pis, self.means)

    def var(self, x=None):
        return np.sum(self.covs * self.pis[:, None, None], axis=0)

    def rvs(self, size=1):
        idx = self.pis.shape[0] - np.sum(np.random.rand(size)[:, None] < np.cumsum(self.pis), axis=1)
        white = np.random.randn(size, self.D)
        return np.einsum('ikj,ij->ik', self.chols[idx], white) + self.means[idx]

    def convolve(self, mog):                                   # mog.py:75-81
        means = np.reshape(self.means[:, None] + mog.means[None, :], (-1, 2))
        weights = np.reshape(self.pis[:, None] * mog.pis[None, :], (-1,))
        covs = np.reshape(self.covs[:, None] + mog.covs[None, :], (-1, 2, 2))
        return MixtureOfGaussians(means, covs, weights)

    def apply_affine(self, A, b):                              # mog.py:84-92
        A = np.asarray(A, dtype=np.float64)
        return MixtureOfGaussians(means=np.dot(self.means, A.T) + b,
                                  covs=np.einsum('ij,kjl,ml->kim', A, self.covs, A),
                                  pis=self.pis)

    @staticmethod
    def convex_combine(mogs, mixing_weights):                  # mog.py:94-100
        return MixtureOfGaussians(means=np.vstack([m.means for m in mogs]),
                                  covs=np.vstack([m.covs for m in mogs]),
                                  pis=np.concatenate([w * m.pis for w, m in zip(mixing_weights, mogs)]))

    def evaluate_grid(self, xlim, ylim, pts=None):             # mog.py:102-112
        assert (ylim[1] > ylim[0]) and (xlim[1] > xlim[0]), "bad limits."
        y_grid = np.arange(ylim[0], ylim[1], dtype=np.float64)
        x_grid = np.arange(xlim[0], xlim[1], dtype=np.float64)
        xx, yy = np.meshgrid(x_grid, y_grid, indexing='xy')
        if pts is None:
            pts = np.column_stack((xx.ravel(order='C'), yy.ravel(order='C')))
        return np.reshape(_gmm_device(pts, self.pis, self.means, self.covs), xx.shape)
