"""Packs the SDSS ugriz filter response table the reference ships as a text
data file (CelestePy/filter_curves, read by planck.py:31-47) into
desi_mcmc_b200/data/filter_curves.npz: per band, wavelengths in Angstrom and
sensitivities in file order.  Data, not code; run once in the build container
(where /root/reference is mounted)."""
import os
import sys

import numpy as np

ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/CelestePy/filter_curves"
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "desi_mcmc_b200", "data",
                   "filter_curves.npz")
wl = {b: [] for b in "ugriz"}
se = {b: [] for b in "ugriz"}
with open(ref) as f:
    for line in f:
        band, w, s = line.strip().split("\t")
        wl[band].append(float(w))
        se[band].append(float(s))
np.savez(out, **{"wl_" + b: np.array(wl[b]) for b in "ugriz"},
         **{"sens_" + b: np.array(se[b]) for b in "ugriz"})
print({b: len(wl[b]) for b in "ugriz"})
