"""CelestePy/fits_image.py:16-155: the FitsImage constructor (N4)
for real SDSS frames / stamps: header + pixel array -> engine.Image (pixels
resident in HBM).  fitsio / astropy are not needed: `read_fits_primary` parses
the primary HDU of a FITS file (2880-byte blocks of 80-character cards, then
big-endian pixels), which is all the reference's stamps use."""
from __future__ import annotations

import numpy as np

from .. import engine

_BITPIX = {8: ">u1", 16: ">i2", 32: ">i4", 64: ">i8", -32: ">f4", -64: ">f8"}


def _card_value(raw):
    v = raw.split("/")[0].strip() if not raw.strip().startswith("'") else raw.strip()
    if v.startswith("'"):
        return v[1:v.index("'", 1)].rstrip()
    if v in ("T", "F"):
        return v == "T"
    try:
        return int(v)
    except ValueError:
        try:
            return float(v.replace("D", "E"))
        except ValueError:
            return v


def read_fits_primary(path):
    """-> (data [NAXIS2, NAXIS1] in native float64 / int, header dict)."""
    raw = open(path, "rb").read()
    header, pos, done = {}, 0, False
    while not done:
        block = raw[pos:pos + 2880]
        if len(block) < 2880:
            raise ValueError("truncated FITS header")
        for i in range(0, 2880, 80):
            card = block[i:i + 80].decode("ascii", "replace")
            key = card[:8].strip()
            if key == "END":
                done = True
                break
            if card[8:10] == "= ":
                header[key] = _card_value(card[10:])
        pos += 2880
    naxis = int(header.get("NAXIS", 0))
    if naxis == 0:
        return None, header
    shape = [int(header["NAXIS%d" % (k + 1)]) for k in range(naxis)][::-1]
    dt = np.dtype(_BITPIX[int(header["BITPIX"])])
    n = int(np.prod(shape))
    data = np.frombuffer(raw, dtype=dt, count=n, offset=pos).reshape(shape)
    data = data.astype(dt.newbyteorder("="))
    if "BSCALE" in header or "BZERO" in header:
        data = data * float(header.get("BSCALE", 1.)) + float(header.get("BZERO", 0.))
    return data, header


class FitsImage(engine.Image):
    """fits_image.py:48-155.  Either a FITS file (`filename` or
    `fits_file_template % band`), or `img` + `header` (+ calib, gain, sky,
    darkvar, psfvec, invvar for the tractor-image route, :70-78,:118-131).
    Fields as in the reference: img, dn, nelec (= round(dn * gain), :85-90),
    kappa, epsilon = sky * kappa, epsilon0, darkvar, calib, rho_n (0-based),
    phi_n, Ups_n, weights / means / covars, R."""

    def __init__(self, band, filename=None, fits_file_template=None, img=None, header=None,
                 exposure_num=0, calib=None, gain=None, darkvar=None, sky=None, psfvec=None,
                 invvar=None, frame=None, fits_table=None):
        from_file = bool(fits_file_template or filename)
        if fits_file_template:
            self.band_file = fits_file_template % band
            img, header = read_fits_primary(self.band_file)
        elif filename is not None:
            self.band_file = filename
            img, header = read_fits_primary(self.band_file)
        elif img is None or header is None:
            raise ValueError("FitsImage needs a FITS file or img + header")
        else:
            self.band_file = None
        self.img = np.asarray(img, dtype=np.float64)
        self.header, self.frame, self.fits_table = header, frame, fits_table
        if from_file or calib is None:
            calib, sky, gain = header["CALIB"], header["SKY"], header["GAIN"]
            darkvar = header.get("DARKVAR", darkvar)
        if psfvec is None:
            psfvec = [header["PSF_P%d" % i] for i in range(18)]
        self.dn = self.img / calib + sky
        nelec = np.round(self.dn * gain)
        self.darkvar = darkvar
        psfvec = np.asarray(psfvec, dtype=np.float64)
        weights = psfvec[0:3]
        means = psfvec[3:9].reshape(3, 2)
        cv = psfvec[9:].reshape(3, 3)                    # [var_x, var_y, cov_xy] per row (:135)
        covars = np.zeros((3, 2, 2))
        for i in range(3):
            covars[i] = np.array([[cv[i, 0], cv[i, 2]], [cv[i, 2], cv[i, 1]]])
        super(FitsImage, self).__init__(
            band=band, nelec=nelec, calib=float(calib), gain=float(gain), sky=float(sky),
            rho_n=np.array([header["CRPIX1"], header["CRPIX2"]], dtype=np.float64) - 1,
            phi_n=np.array([header["CRVAL1"], header["CRVAL2"]], dtype=np.float64),
            Ups_n=np.array([[header["CD1_1"], header["CD1_2"]], [header["CD2_1"], header["CD2_2"]]],
                           dtype=np.float64),
            psf_weights=weights, psf_means=means, psf_covars=covars, invvar=invvar)
        self.epsilon0 = self.epsilon


# ---- catalogue rows -> SrcParams (util/data/get_data.py:38-42,89-136) -----------
BANDS = ['u', 'g', 'r', 'i', 'z']


def mags2nanomaggies(mags):
    return np.power(10., (np.asarray(mags, dtype=np.float64) - 22.5) / -2.5)


def nanomaggies2mags(nanos):
    return (-2.5) * np.log10(nanos) + 22.5


def photoobj_to_celestepy_src(photoobj_row):
    """SDSS photoObj row (pandas Series or dict) -> engine.SrcParams (:89-136)."""
    row = photoobj_row
    get = lambda k: float(row[k])
    u = np.array([get('ra'), get('dec')])
    fluxes = [float(mags2nanomaggies(get('psfMag_%s' % b))) for b in BANDS]
    if int(row['type']) == 6:                              # photoObj type 6 = star, 3 = galaxy
        return engine.SrcParams(u, a=0, fluxes=fluxes)
    mean = lambda stem: float(np.mean([get('%s_%s' % (stem, b)) for b in BANDS]))
    prob_dev = mean('fracDeV')
    AB = prob_dev * mean('deVAB') + (1. - prob_dev) * mean('expAB')
    Rad = prob_dev * mean('deVRad') + (1. - prob_dev) * mean('expRad')
    Phi = prob_dev * mean('deVPhi') + (1. - prob_dev) * mean('expPhi')
    devFlux = np.array([mags2nanomaggies(get('deVMag_%s' % b)) for b in BANDS])
    expFlux = np.array([mags2nanomaggies(get('expMag_%s' % b)) for b in BANDS])
    fluxes = prob_dev * devFlux + (1. - prob_dev) * expFlux
    return engine.SrcParams(u, a=1, v=u, theta=1.0 - prob_dev, phi=-1. * Phi, sigma=Rad, rho=AB,
                            fluxes=fluxes)
