"""CelestePy/models.py call surface: ``CelesteBase`` / ``Celeste`` and ``Field``
(the live Gibbs sweep of SURVEY.md 3.1) on the CUDA path.

  * ``Field.resample_photons`` (:123-160): ONE render + ONE allocation launch
    for the five bands of the field (the reference loops bands in Python and
    scans H*W*S boxes per band), then the gamma draw of each epsilon.
  * ``CelesteBase.render_model_image`` / ``img_log_likelihood`` (:88-108): the
    fused full-frame kernel (cg_field_loglik).
  * ``CelesteBase.resample_model`` / ``resample_sources`` (:75-83): per-source
    conditionals through ``sources.Source`` (device likelihood evaluations).

The images are ``engine.Image`` objects (the FitsImage data contract); a field's
``img_dict`` is keyed by band exactly as in the reference.
"""
from __future__ import annotations

import numpy as np

from .. import engine
from . import celeste_mcmc as cel_mcmc
from ._util import BANDS, record_of
from .sources import Source


class CelesteBase(object):
    """:15-108."""
    _source_type = None

    def __init__(self, gal_flux_prior_distn=None, star_flux_prior_distn=None):
        self.field_list = []
        self.bands = list(BANDS)
        self.star_flux_prior_distn = star_flux_prior_distn
        self.gal_flux_prior_distn = gal_flux_prior_distn

    def initialize_sources(self, init_srcs=None, init_src_params=None, photoobj_df=None):
        """:30-40 (the photoobj route needs the reference's data utilities and
        is out of scope: SURVEY.md 8f N4)."""
        if init_srcs is not None:
            self.srcs = init_srcs
        elif init_src_params is not None:
            self.srcs = [self._source_type(s, self) for s in init_src_params]
        else:
            raise NotImplementedError

    def intialize_from_fields(self):
        raise NotImplementedError

    def add_field(self, img_dict):
        """:46-53."""
        for k in img_dict.keys():
            assert k in self.bands, "Celeste model doesn't support band %s" % k
        self.field_list.append(Field(img_dict))

    @property
    def source_types(self):
        return np.array([s.object_type for s in self.srcs])

    def get_brightest(self, object_type='star', num_srcs=1, band='r', return_idx=False):
        """:59-69."""
        fluxes = np.array([s.params.flux_dict[band] for s in self.srcs])
        type_idx = np.where(self.source_types == object_type)[0]
        type_fluxes = fluxes[type_idx]
        type_idx = type_idx[np.argsort(type_fluxes)[::-1]][:num_srcs]
        blist = [self.srcs[i] for i in type_idx]
        if return_idx:
            return blist, type_idx
        return blist

    # ---- resample methods (:75-83) -------------------------------------------
    def resample_model(self):
        for field in self.field_list:
            field.resample_photons(self.srcs)
        self.resample_sources()

    def resample_sources(self):
        for src in self.srcs:
            src.resample()

    def resample_sources_batched(self, seed=0):
        """resample_sources with the likelihood evaluations of ALL sources in
        one launch per slice-sampler round (sources.resample_sources_batched)."""
        from .sources import resample_sources_batched
        return resample_sources_batched(self.srcs, seed=seed)

    def resample_sources_on_device(self, seed=0, rngs=None):
        """resample_sources with every source's location slice sampler run on the device in ONE
        launch (sources.resample_sources_on_device; bit-identical to the loops above)."""
        from .sources import resample_sources_on_device
        return resample_sources_on_device(self.srcs, seed=seed, rngs=rngs)

    # ---- model image / likelihood (:88-108) -----------------------------------
    def render_model_image(self, fimg, xlim=None, ylim=None, exclude=None):
        """epsilon + sum of the sources' model patches on fimg (full frame, or
        cropped to xlim/ylim when both are given)."""
        source_list = [s for s in self.srcs if s is not exclude]
        packed = np.concatenate([record_of(s.params) for s in source_list]) if source_list else \
            np.zeros(0, dtype=engine.SRC_DTYPE)
        _, models = engine.field_loglik([fimg], packed, return_model=True)
        mod_img = models[0].cpu().numpy().astype(np.float64)
        if xlim is not None and ylim is not None:
            mod_img = mod_img[ylim[0]:ylim[1], xlim[0]:xlim[1]]
        return mod_img

    def img_log_likelihood(self, fimg, mod_img=None):
        if mod_img is None:
            packed = np.concatenate([record_of(s.params) for s in self.srcs]) if self.srcs else \
                np.zeros(0, dtype=engine.SRC_DTYPE)
            return float(engine.field_loglik([fimg], packed).cpu().numpy()[0])
        from .sources import poisson_loglike
        return poisson_loglike(np.asarray(fimg.nelec.cpu().numpy(), dtype=np.float64), mod_img,
                               np.ones(np.asarray(mod_img).shape))

    # priors: the base model has flat priors (model_sources.py adds the empirical ones)
    def logprior(self, params):
        return 0.

    def prior_sample(self, src_type, u=None):
        raise NotImplementedError("the empirical priors live in model_sources (SURVEY.md 8f N3)")


class Field(object):
    """:110-160: the image data of one field."""

    def __init__(self, img_dict):
        self.img_dict = img_dict
        for k, img in self.img_dict.items():
            img.set_epsilon(float(np.median(img.nelec.cpu().numpy())))
        self.a_0 = 5
        self.b_0 = .005

    def resample_photons(self, srcs, verbose=False, random_state=None):
        for src in srcs:
            src.clear_sample_images()
        bands = list(self.img_dict.keys())
        imgs = [self.img_dict[b] for b in bands]
        samp, noise = cel_mcmc.sample_source_photons_field(imgs, [s.params for s in srcs], random_state)
        noise_sums = {}
        for band, img, samp_imgs, noise_sum in zip(bands, imgs, samp, noise):
            for src, samp_img in zip(srcs, samp_imgs):
                if samp_img is not None:
                    y_grid = np.arange(samp_img.y0, samp_img.y1, dtype=np.float64)
                    x_grid = np.arange(samp_img.x0, samp_img.x1, dtype=np.float64)
                    xx, yy = np.meshgrid(x_grid, y_grid, indexing='xy')
                    pixel_grid = np.column_stack((xx.ravel(order='C'), yy.ravel(order='C')))
                    src.sample_image_list.append((samp_img, img, pixel_grid))
            noise_sums[band] = noise_sum
        rng = random_state or np.random
        for band, img in self.img_dict.items():
            a_n = self.a_0 + noise_sums[band]
            b_n = self.b_0 + img.shape[0] * img.shape[1]
            img.set_epsilon(float(rng.gamma(a_n, 1. / b_n)))
        return noise_sums


class Celeste(CelesteBase):
    _source_type = Source
