"""CelestePy/mixture_profiles.py:9-25: the exponential / de Vaucouleurs
mixture-of-Gaussians profiles (amplitudes normalised exactly as the reference
does) and the small container the galaxy code reads (.amp, .mean, .var).
The device kernels carry the same constants (csrc/celeste_b200.cu)."""
from __future__ import annotations

import numpy as np

exp_amp = np.array([2.34853813e-03, 3.07995260e-02, 2.23364214e-01,
                    1.17949102e+00, 4.33873750e+00, 5.99820770e+00])
exp_var = np.array([1.20078965e-03, 8.84526493e-03, 3.91463084e-02,
                    1.39976817e-01, 4.60962500e-01, 1.50159566e+00])
exp_amp /= np.sum(exp_amp)

dev_amp = np.array([4.26347652e-02, 2.40127183e-01, 6.85907632e-01, 1.51937350e+00,
                    2.83627243e+00, 4.46467501e+00, 5.72440830e+00, 5.60989349e+00])
dev_var = np.array([2.23759216e-04, 1.00220099e-03, 4.18731126e-03, 1.69432589e-02,
                    6.84850479e-02, 2.87207080e-01, 1.33320254e+00, 8.40215071e+00])
dev_amp /= np.sum(dev_amp)


class MixtureOfGaussians(object):
    """mixture_profiles.py:27-108 (the parts CelestePy uses)."""

    def __init__(self, amp, mean, var):
        self.amp = np.atleast_1d(amp).astype(float)
        self.mean = np.atleast_2d(np.array(mean)).astype(float)
        self.K, self.D = self.mean.shape
        var = np.asarray(var, dtype=float)
        if var.size == self.K:
            self.var = np.zeros((self.K, self.D, self.D))
            for d in range(self.D):
                self.var[:, d, d] = var
        else:
            self.var = var.copy()
        self.var = 0.5 * (self.var + np.transpose(self.var, (0, 2, 1)))

    def normalize(self):
        self.amp /= np.sum(self.amp)

    def __add__(self, other):
        s = MixtureOfGaussians(np.append(self.amp, other.amp), np.append(self.mean, other.mean, axis=0),
                               np.append(self.var, other.var, axis=0))
        s.normalize()
        return s

    def apply_affine(self, shift, scale):
        newvar = np.einsum('ji,kjl,lm->kim', scale, self.var, scale)
        return MixtureOfGaussians(self.amp, self.mean + shift, newvar)

    def convolve(self, other):
        amp = (other.amp[:, None] * self.amp[None, :]).ravel()
        mean = (other.mean[:, None, :] + self.mean[None, :, :]).reshape(-1, self.D)
        var = (other.var[:, None] + self.var[None, :]).reshape(-1, self.D, self.D)
        return MixtureOfGaussians(amp, mean, var)


def get_exp_mixture():
    return MixtureOfGaussians(exp_amp, np.zeros((exp_amp.size, 2)), exp_var)


def get_dev_mixture():
    return MixtureOfGaussians(dev_amp, np.zeros((dev_amp.size, 2)), dev_var)
