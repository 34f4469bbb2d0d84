"""CelestePy/util/dists/flux_prior.py:4-93: the empirical priors (mixtures of
Gaussians over transformed parameters).  Densities are evaluated by the
device kernel behind compat.mog.MixtureOfGaussians.logpdf (cg_mog_logpdf)."""
from __future__ import annotations

import numpy as np

from .mog import MixtureOfGaussians


class FluxColorMoG(MixtureOfGaussians):
    """:4-40: law on [lu - lr, lg - lr, li - lr, lz - lr, lr] of the log fluxes."""
    A = np.zeros((5, 5))
    A[0, 0] = A[1, 1] = A[3, 2] = A[4, 3] = 1.
    A[2, :] = -1
    A[2, -1] = 1.
    Ainv = np.linalg.inv(A)

    @staticmethod
    def to_colors(x):
        return np.dot(x, FluxColorMoG.A)

    @staticmethod
    def to_fluxes(colors):
        return np.dot(colors, FluxColorMoG.Ainv)

    def logpdf(self, x):
        return super(FluxColorMoG, self).logpdf(np.dot(x, self.A))


class GalShapeMoG(MixtureOfGaussians):
    """:43-59."""

    @staticmethod
    def to_unconstrained(x):
        sigma, ab, phi = x[0], x[1], x[2]
        return np.array([np.log(sigma), np.log(ab) - np.log(1. - ab), np.log(phi) - np.log(np.pi - phi)])

    def logpdf(self, x):
        return super(GalShapeMoG, self).logpdf(self.to_unconstrained(x))


class GalRadiusMoG(MixtureOfGaussians):
    """:62-76."""

    @staticmethod
    def to_unconstrained(x):
        return np.log(x)

    def logpdf(self, x):
        return super(GalRadiusMoG, self).logpdf(np.reshape(self.to_unconstrained(x), (-1, 1))
                                                if np.ndim(x) else np.array([self.to_unconstrained(x)]))


class GalAbMoG(MixtureOfGaussians):
    """:79-93."""

    @staticmethod
    def to_unconstrained(x):
        return np.log(x) - np.log(1. - x)

    def logpdf(self, x):
        return super(GalAbMoG, self).logpdf(np.reshape(self.to_unconstrained(x), (-1, 1))
                                            if np.ndim(x) else np.array([self.to_unconstrained(x)]))
