"""Seeded synthetic inputs of the BASELINE.json configurations.

Pure input generation (NumPy ``RandomState`` so the legacy stream is
reproducible); no rendering or likelihood arithmetic lives here.  Image
parameters follow the header of the r-band stamp the reference ships
(data/stamps/stamp-r-253.1147-11.6072.fits: CALIB .00521, SKY 94, GAIN 4.6, PSF
weights .60/.32/.08 with variances 1.9/7.0/29.8 px^2, CD +-1.1e-4 deg/px); the
galaxy-shape draws follow experiments/galaxy/sample_stamp.py:119-125.
"""
from __future__ import annotations

import numpy as np

BANDS = ["u", "g", "r", "i", "z"]
PIX_DEG = 1.1e-4


def make_psf(rng):
    w = np.array([.60, .32, .08]) * rng.uniform(.9, 1.1, size=3)
    w = w / w.sum()
    means = rng.normal(0., .2, size=(3, 2))
    covs = np.zeros((3, 2, 2))
    for k, v in enumerate((1.9, 7.0, 30.0)):
        vx, vy = v * rng.uniform(.8, 1.25, size=2)
        c = rng.uniform(-.15, .15) * np.sqrt(vx * vy)
        covs[k] = [[vx, c], [c, vy]]
    return w, means, covs


def make_image_meta(rng, band, H, W, center=(253.1147, 11.6072), rotate=True):
    """What FitsImage would read from a frame header (fits_image.py:85-147)."""
    w, means, covs = make_psf(rng)
    ang = rng.uniform(-.02, .02) if rotate else 0.
    ca, sa = np.cos(ang), np.sin(ang)
    ups = PIX_DEG * np.array([[-ca, sa], [sa, ca]])
    return dict(band=band,
                calib=.005 * rng.uniform(.9, 1.1),
                gain=4.6 * rng.uniform(.95, 1.05),
                sky=94. * rng.uniform(.8, 1.2),
                rho_n=np.array([W / 2., H / 2.]) + rng.uniform(-.5, .5, size=2),
                phi_n=np.array(center, dtype=np.float64),
                Ups_n=ups,
                psf_weights=w, psf_means=means, psf_covars=covs,
                shape=(H, W))


def pixel2equa(meta, p):
    """fits_image.py:176-181 on a meta dict (for placing synthetic sources)."""
    phi1rad = meta["phi_n"][1] / 180. * np.pi
    iwc = np.dot(meta["Ups_n"], np.asarray(p, dtype=np.float64) - meta["rho_n"])
    return np.array([iwc[0] / np.cos(phi1rad) + meta["phi_n"][0], iwc[1] + meta["phi_n"][1]])


def make_source(rng, meta, is_star, pix=None, planck=False, margin=0.):
    H, W = meta["shape"]
    if pix is None:
        pix = np.array([rng.uniform(margin, W - margin), rng.uniform(margin, H - margin)])
    u = pixel2equa(meta, pix)
    colors = np.exp(rng.normal(0., .3, size=5))
    if is_star:
        flux = 10. * np.exp(rng.normal(0., .7)) * colors
        s = dict(a=0, u=u, fluxes=flux, t=None, b=None,
                 theta=None, sigma=None, phi=None, rho=None)
        if planck:
            s["t"] = rng.uniform(3000., 10000.)
            s["b"] = np.exp(rng.normal(np.log(5e-11), .5))   # ~7e3 photo-electrons
        return s
    flux = 30. * np.exp(rng.normal(0., .7)) * colors
    return dict(a=1, u=u, fluxes=flux, t=None, b=None,
                theta=float(np.clip(rng.beta(.4, .4), .01, .99)),
                sigma=float(np.exp(rng.normal(np.log(1.5), .5))),
                phi=float(rng.uniform(0., 180.)),   # degrees: celeste_galaxy_conditionals.py:97
                rho=float(rng.uniform(.2, 1.)))


def make_field(seed, H=1489, W=2048, n_src=500, star_frac=.6, planck_frac=0., margin=0.):
    """C2 / C5: one 5-band field with n_src mixed sources."""
    rng = np.random.RandomState(seed)
    metas = [make_image_meta(rng, b, H, W) for b in BANDS]
    srcs = []
    for _ in range(n_src):
        star = rng.uniform() < star_frac
        srcs.append(make_source(rng, metas[2], star, planck=star and rng.uniform() < planck_frac,
                                margin=margin))
    return dict(seed=seed, images=metas, sources=srcs)


def make_stamp(seed, size=25, n_star=1, n_gal=1):
    """C1: 1 star + 1 galaxy on a 5-band 25x25 stamp."""
    rng = np.random.RandomState(seed)
    metas = [make_image_meta(rng, b, size, size) for b in BANDS]
    c = np.array([size / 2., size / 2.])
    srcs = []
    for i in range(n_star):
        srcs.append(make_source(rng, metas[2], True, pix=c + rng.uniform(-.5, .5, size=2) + 3. * i))
    for i in range(n_gal):
        ang = rng.uniform(0, 2 * np.pi)
        srcs.append(make_source(rng, metas[2], False,
                                pix=c + 4. * np.array([np.cos(ang), np.sin(ang)]) * (i + 1)))
    return dict(seed=seed, images=metas, sources=srcs)


def make_proposals(seed, cur, P):
    """C3: P Gaussian perturbations of a galaxy's state (u in pixels .05, flux
    2 %, theta/rho .02, phi 3 deg, sigma 3 %) + log-uniforms from the same
    legacy stream."""
    rng = np.random.RandomState(seed)
    props = []
    for _ in range(P):
        p = dict(cur)
        p["u"] = cur["u"] + rng.normal(0., .05, size=2) * PIX_DEG
        p["fluxes"] = cur["fluxes"] * (1. + rng.normal(0., .02, size=5))
        if cur["a"] == 1:
            p["theta"] = float(np.clip(cur["theta"] + rng.normal(0., .02), .01, .99))
            p["rho"] = float(np.clip(cur["rho"] + rng.normal(0., .02), .05, 1.))
            p["phi"] = float(cur["phi"] + rng.normal(0., 3.))
            p["sigma"] = float(cur["sigma"] * (1. + rng.normal(0., .03)))
        props.append(p)
    log_u = np.log(rng.uniform(size=P))
    return props, log_u


def image_kwargs(meta):
    return {k: meta[k] for k in ("band", "calib", "gain", "sky", "rho_n", "phi_n", "Ups_n",
                                 "psf_weights", "psf_means", "psf_covars")}


def make_eigen_basis(seed=1004, K=4, V=1364):
    """A17 inputs of the shape of the reference's cache/basis_locked/*K-4_V-1364* fits: a positive
    rest-frame basis B (K, V) with rows normalised to 1 (quasar_infer_photometry.py:223-224) on a
    log-spaced wavelength grid lam0 (Angstrom)."""
    rng = np.random.RandomState(seed)
    lam0 = np.exp(np.linspace(np.log(700.), np.log(9000.), V))
    B = np.zeros((K, V))
    for k in range(K):
        cont = (lam0 / 2500.) ** rng.uniform(-2., 0.)
        lines = sum(a * np.exp(-.5 * ((lam0 - c) / w) ** 2)
                    for a, c, w in zip(rng.uniform(0., 3., 4), rng.uniform(1000., 7000., 4), rng.uniform(20., 200., 4)))
        B[k] = cont + lines
    B = B / B.sum(axis=1, keepdims=True)
    return B, lam0


def make_crowded_field(seed=1004, H=1489, W=2048, n_src=5000, star_frac=.1, eigen_frac=.05, K=4):
    """C4: a galaxy-heavy crowded field.  10 % stars whose fluxes come from the black-body route
    (T ~ U(3000, 10000) K, planck.py), 5 % point sources whose fluxes come from the eigen-spectrum
    route (redshift z, simplex weights w, scale m), the rest galaxies with free fluxes.
    -> dict(seed, images, sources); eigen sources carry s["eigen"] = dict(z, w, m, row)."""
    rng = np.random.RandomState(seed)
    metas = [make_image_meta(rng, b, H, W) for b in BANDS]
    srcs, row = [], 0
    for _ in range(n_src):
        r = rng.uniform()
        if r < star_frac:
            s = make_source(rng, metas[2], True, planck=True)
        elif r < star_frac + eigen_frac:
            s = make_source(rng, metas[2], True)
            z = rng.uniform(.3, 4.)
            w = np.exp(rng.normal(size=K))
            s["eigen"] = dict(z=z, w=w / w.sum(), m=float(np.exp(rng.normal(np.log(5000.), .5)) * (1. + z)), row=row)
            s["fluxes"] = None                        # a function of (z, w, m): filled by the consumer
            row += 1
        else:
            s = make_source(rng, metas[2], False)
        s.setdefault("eigen", None)
        srcs.append(s)
    return dict(seed=seed, images=metas, sources=srcs, n_eigen=row)


def instantiate_crowded(engine, field, B, lam0, bound_cap=25., simulate_seed=None, **sweep_kw):
    """Device objects of a make_crowded_field dict: images (pixels drawn ON THE DEVICE as
    Poisson(model image of the field) when simulate_seed is given, else zeros), packed sources,
    the eigen-spectrum state and the engine.GibbsSweep over them."""
    import torch
    dev = engine.device()
    H, W = field["images"][0]["shape"]
    imgs = [engine.Image(nelec=torch.zeros((H, W), dtype=torch.float32, device=dev), **image_kwargs(m))
            for m in field["images"]]
    srcs = [engine.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"], phi=s["phi"],
                             sigma=s["sigma"], rho=s["rho"], fluxes=s["fluxes"],
                             flux_row=s["eigen"]["row"] if s.get("eigen") else None) for s in field["sources"]]
    packed = engine.pack_sources(srcs)
    es = [s["eigen"] for s in field["sources"] if s.get("eigen")]
    eig = None
    if es:
        eig = engine.EigenSpectra([e["z"] for e in es], [e["w"] for e in es], [e["m"] for e in es], B, lam0)
        eig.project()
    side = int(2 * bound_cap + 3)
    if simulate_seed is not None:
        _, models = engine.field_loglik(imgs, packed, bound_cap=bound_cap, max_patch_px=side * side,
                                        return_model=True)
        gen = torch.Generator(device=dev)
        gen.manual_seed(int(simulate_seed))
        for im, lam in zip(imgs, models):
            im.set_nelec(torch.poisson(lam, generator=gen))
    sweep = engine.GibbsSweep(imgs, packed, bound_cap=bound_cap, eigen=eig, **sweep_kw)
    return imgs, packed, eig, sweep
