"""Generates tests/golden/stamp_r_253.npz from the stamp the reference ships
(data/stamps/stamp-r-253.1147-11.6072.fits, 51x51 float64): the pixel array and
the header cards FitsImage reads (fits_image.py:85-138).  Run in the build
container, where /root/reference is mounted; the GPU box only sees the .npz."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from desi_mcmc_b200.compat.fits_image import read_fits_primary  # noqa: E402  (pure numpy, no CUDA)

SRC = "/root/reference/data/stamps/stamp-r-253.1147-11.6072.fits"
img, hdr = read_fits_primary(SRC)
keys = ["CALIB", "SKY", "GAIN", "DARKVAR", "CRPIX1", "CRPIX2", "CRVAL1", "CRVAL2",
        "CD1_1", "CD1_2", "CD2_1", "CD2_2"] + ["PSF_P%d" % i for i in range(18)]
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "stamp_r_253.npz"), img=img,
                    keys=np.array(keys), values=np.array([float(hdr[k]) for k in keys]))
print(img.shape, img.dtype, {k: hdr[k] for k in keys[:12]})
