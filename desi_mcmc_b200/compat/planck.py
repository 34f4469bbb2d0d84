"""CelestePy/planck.py:94-158 on the device (cg_planck_fluxes).  The kernel
returns photons_expected_brightness for all five bands of a batch of stars;
the scalar reference calls are thin slices of it."""
from __future__ import annotations

import numpy as np

from .. import engine

# planck.py:13-25
h = 6.6260693e-34
k = 1.3806488e-23
c = 299792458
lens_area = .75 * np.pi * (1.25 ** 2)
exposure_duration = 54
solar_L = 3.846 * 1e26
m_per_ly = c * 31556952.

_BANDS = ['u', 'g', 'r', 'i', 'z']


def photons_expected_brightness(T, b, band):
    """planck.py:155-158."""
    return float(engine.planck_fluxes([T], [b])[0, _BANDS.index(band)])


def photons_expected(T, solar_l, d, band):
    """planck.py:144-152: the lens sees L / (4 pi D^2), i.e. brightness
    b = solar_l / (4 pi d^2) in photons_expected_brightness's units."""
    return photons_expected_brightness(T, solar_l / (4 * np.pi * d * d), band)


def photons_per_joule(T, band):
    """planck.py:94-142, recovered from the device flux at unit brightness."""
    lens_watts = lens_area * 1.0 * solar_L / (m_per_ly ** 2)
    return photons_expected_brightness(T, 1.0, band) / (lens_watts * exposure_duration)


def photons_expected_brightness_batch(T, b):
    """[n] temperatures / brightnesses -> [n, 5] photo-electrons (ugriz)."""
    return engine.planck_fluxes(np.asarray(T, dtype=np.float64), np.asarray(b, dtype=np.float64))
