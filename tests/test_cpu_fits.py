"""N4: the FitsImage input contract from real data -- the
minimal FITS reader, the header -> image fields mapping on the stamp the
reference ships (committed as tests/golden/stamp_r_253.npz), and the
catalogue-row conversion."""
import os

import numpy as np

from desi_mcmc_b200.compat import fits_image as fi
from oracle import ref_cpu as ref

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "stamp_r_253.npz")


def _fixture():
    z = np.load(GOLD)
    return z["img"], {str(k): float(v) for k, v in zip(z["keys"], z["values"])}


def _write_fits(path, data, cards):
    def card(k, v):
        if isinstance(v, bool):
            s = "%-8s= %20s" % (k, "T" if v else "F")
        elif isinstance(v, str):
            s = "%-8s= %-20s" % (k, "'%s'" % v)
        else:
            s = "%-8s= %20s" % (k, repr(v))
        return (s + " / comment").ljust(80)[:80]
    bitpix = {np.dtype(">f4"): -32, np.dtype(">f8"): -64, np.dtype(">i2"): 16}[data.dtype]
    hdr = [card("SIMPLE", True), card("BITPIX", bitpix), card("NAXIS", 2), card("NAXIS1", data.shape[1]),
           card("NAXIS2", data.shape[0])] + [card(k, v) for k, v in cards] + ["END".ljust(80)]
    raw = "".join(hdr).encode("ascii")
    raw += b" " * (-len(raw) % 2880)
    body = data.tobytes()
    body += b"\0" * (-len(body) % 2880)
    open(path, "wb").write(raw + body)


def test_read_fits_primary_round_trip(tmp_path):
    rng = np.random.RandomState(0)
    a = rng.normal(size=(7, 11)).astype(">f4")
    _write_fits(str(tmp_path / "a.fits"), a, [("CALIB", 0.005), ("CTYPE1", "RA---TAN"), ("GAIN", 4.6),
                                              ("NEXP", 3)])
    data, hdr = fi.read_fits_primary(str(tmp_path / "a.fits"))
    assert data.shape == (7, 11) and np.array_equal(data, a.astype("<f4"))
    assert hdr["CALIB"] == 0.005 and hdr["CTYPE1"] == "RA---TAN" and hdr["NEXP"] == 3 and hdr["SIMPLE"] is True
    b = rng.randint(-300, 300, size=(5, 4)).astype(">i2")
    _write_fits(str(tmp_path / "b.fits"), b, [("BZERO", 32768.0), ("BSCALE", 1.0)])
    data, hdr = fi.read_fits_primary(str(tmp_path / "b.fits"))
    assert np.array_equal(data, b.astype(np.float64) + 32768.0)


def test_real_stamp_header_mapping():
    img, hdr = _fixture()
    assert img.shape == (51, 51)
    assert abs(hdr["CALIB"] - 0.00521019566804171) < 1e-18 and abs(hdr["SKY"] - 94.0242767333984) < 1e-12
    o = ref.fits_image_from_header("r", img, hdr)
    assert np.array_equal(o.nelec, np.round((img / hdr["CALIB"] + hdr["SKY"]) * hdr["GAIN"]))   # fits_image.py:85-87
    assert np.array_equal(o.rho_n, [25., 25.]) and np.array_equal(o.phi_n, [253.11475, 11.60716])
    assert abs(o.epsilon - hdr["SKY"] * hdr["GAIN"]) < 1e-12 and o.kappa == hdr["GAIN"]
    assert np.allclose(o.weights, [0.601, 0.324, 0.076], atol=2e-3) and abs(o.weights.sum() - 1) < 2e-3
    assert np.allclose(o.covars[:, 0, 1], o.covars[:, 1, 0])
    assert 21.0 < o.R < 22.0                                  # err = .001 -> R ~ 21.4 px
    # the reference point maps to itself
    assert np.allclose(o.equa2pixel(o.phi_n), o.rho_n) and np.allclose(o.pixel2equa(o.rho_n), o.phi_n)


def test_photoobj_row_conversion():
    star = dict(ra=10.5, dec=-3.25, type=6)
    gal = dict(ra=11.5, dec=-3.5, type=3)
    for i, b in enumerate("ugriz"):
        for r in (star, gal):
            r["psfMag_" + b] = 20. + .1 * i
        gal.update({"fracDeV_" + b: .25, "deVAB_" + b: .5, "deVRad_" + b: 2., "deVPhi_" + b: 30.,
                    "expAB_" + b: .7, "expRad_" + b: 3., "expPhi_" + b: 50., "deVMag_" + b: 19. + .1 * i,
                    "expMag_" + b: 19.5 + .1 * i})
    s = fi.photoobj_to_celestepy_src(star)
    assert s.a == 0 and np.allclose(s.u, [10.5, -3.25])
    assert np.allclose(s.fluxes, ref.mags2nanomaggies(20. + .1 * np.arange(5)))
    g = fi.photoobj_to_celestepy_src(gal)
    assert g.a == 1 and abs(g.theta - .75) < 1e-15 and abs(g.rho - (.25 * .5 + .75 * .7)) < 1e-15
    assert abs(g.sigma - (.25 * 2. + .75 * 3.)) < 1e-15 and abs(g.phi + (.25 * 30. + .75 * 50.)) < 1e-13
    want = .25 * ref.mags2nanomaggies(19. + .1 * np.arange(5)) + .75 * ref.mags2nanomaggies(19.5 + .1 * np.arange(5))
    assert np.allclose(g.fluxes, want, rtol=1e-14)
    assert abs(fi.nanomaggies2mags(fi.mags2nanomaggies(21.3)) - 21.3) < 1e-12
