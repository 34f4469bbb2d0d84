"""CelestePy/sources.py call surface on the CUDA path: the Poisson
likelihoods a Source evaluates inside the Gibbs / MH loop."""
from __future__ import annotations

import numpy as np
import torch

from .. import _ffi, engine
from . import celeste, celeste_galaxy_conditionals as gal_funs
from ._util import BANDS, fluxes_of, record_of, render_one
from .slicesample import slicesample


def poisson_loglike(data, model_img, mask):
    """sources.py:6-12 (float64 arrays in, float64 arithmetic on the device)."""
    data, model_img, mask = np.asarray(data), np.asarray(model_img), np.asarray(mask)
    assert model_img.shape == mask.shape
    assert data.shape == model_img.shape
    dev = engine.device()
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
    d_d, d_m, d_k = d(data), d(model_img), d(mask)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    _ffi.check(engine.lib().cg_poisson_loglike(d_d.data_ptr(), d_m.data_ptr(), d_k.data_ptr(),
                                               d_d.numel(), out.data_ptr(), engine.stream_ptr()))
    return float(out.item())


class Source(object):
    """sources.py:15-396: holds the chain state of one source and evaluates
    its conditional / isolated / MH likelihoods on the device.
    sample_image_list holds (samp_img, fits_img, pixel_grid) tuples where
    samp_img has NativePatch's fields (data, x0, x1, y0, y1)."""

    def __init__(self, params, model):
        self.params = params
        self.model = model
        self.sample_image_list = []
        self.u_lower = self.params.u - .0025
        self.u_upper = self.params.u + .0025
        self.du = self.u_upper - self.u_lower
        self.loc_samps, self.flux_samps, self.shape_samps, self.ll_samps = [], [], [], []
        self._cache = None

    def clear_sample_images(self):
        self.sample_image_list = []
        self._cache = None

    def __str__(self):
        label = "%s at Ra, Dec = (%2.5f, %2.5f) " % (self.object_type, self.params.u[0], self.params.u[1])
        mags = " ".join(["%s = %2.4f" % (b, self.nanomaggies2mags(f))
                         for b, f in zip(BANDS, fluxes_of(self.params))])
        if self.is_star():
            return label + " with Mags " + mags
        shape = "re=%2.3f, ab=%2.3f, phi=%2.3f" % (self.params.sigma, self.params.rho, self.params.phi)
        return label + " with Mags " + mags + " and Galaxy Shape: " + shape

    def mags2nanomaggies(self, mags):
        return np.power(10., (mags - 22.5) / -2.5)

    def nanomaggies2mags(self, nanos):
        return (-2.5) * np.log10(nanos) + 22.5

    @property
    def object_type(self):
        return "star" if self.is_star() else ("galaxy" if self.is_galaxy() else "none")

    def is_star(self):
        return self.params.a == 0

    def is_galaxy(self):
        return self.params.a == 1

    @property
    def id(self):
        return "in_%s_to_%s" % (str(self.u_lower), str(self.u_upper))

    @staticmethod
    def get_bounding_box(params, img):
        """:81-94."""
        if params.is_star():
            bound = img.R
        elif params.is_galaxy():
            bound = gal_funs.gen_galaxy_psf_image_bound(params, img)
        else:
            raise ValueError("source type unknown")
        px, py = img.equa2pixel(params.u)
        xlim = (np.max([0, np.floor(px - bound)]), np.min([img.shape[1], np.ceil(px + bound)]))
        ylim = (np.max([0, np.floor(py - bound)]), np.min([img.shape[0], np.ceil(py + bound)]))
        return xlim, ylim

    location_samples = property(lambda self: np.array(self.loc_samps))
    flux_samples = property(lambda self: np.array(self.flux_samps))
    shape_samples = property(lambda self: np.array(self.shape_samps))
    loglike_samples = property(lambda self: np.array(self.ll_samps))

    def store_sample(self):
        self.loc_samps.append(self.params.u.copy())
        self.flux_samps.append(np.array(fluxes_of(self.params)))
        self.shape_samps.append(self.params.shape.copy())

    def store_loglike(self):
        self.ll_samps.append(self.log_likelihood())

    def flux_in_image(self, fits_image, fluxes=None):
        """:120-129."""
        f = fluxes_of(self.params) if fluxes is None else np.asarray(fluxes)
        return (f[BANDS.index(fits_image.band)] / fits_image.calib) * fits_image.kappa

    # ---- device state of the sampled photon images --------------------------
    def _sample_cache(self):
        key = (id(self.sample_image_list), len(self.sample_image_list))
        if self._cache is not None and self._cache["key"] == key:
            return self._cache
        imgs, zs, boxes = [], [], []
        for samp_img, fits_img, _ in self.sample_image_list:
            imgs.append(fits_img)
            zs.append(np.ascontiguousarray(np.asarray(samp_img.data), dtype=np.float32).ravel())
            boxes.append((int(samp_img.x0), int(samp_img.x1), int(samp_img.y0), int(samp_img.y1)))
        offs = np.zeros(len(zs), dtype=np.int64)
        if zs:
            offs[1:] = np.cumsum([z.size for z in zs])[:-1]
        aux = torch.from_numpy(np.concatenate(zs)).to(engine.device()) if zs else None
        n = len(imgs)
        self._cache = dict(key=key, imgs=imgs, aux=aux, offs=offs, boxes=boxes,
                           items_cond=engine.make_items([0] * n, np.arange(n), box=boxes, aux=np.arange(n)),
                           items_iso=engine.make_items([0] * n, np.arange(n), flags=_ffi.CG_ITEM_NO_MASK,
                                                       box=boxes))
        return self._cache

    def _record(self, u=None, fluxes=None, shape=None):
        p = self.params
        th = p.shape if shape is None else shape
        return engine.pack_sources([engine.SrcParams(
            u=p.u if u is None else u, a=p.a, theta=th[0], sigma=th[1], phi=th[2], rho=th[3],
            fluxes=fluxes_of(p) if fluxes is None else np.asarray(fluxes, dtype=np.float64))])

    def log_likelihood(self, u=None, fluxes=None, shape=None, fp64=False):
        """:134-183, all sampled images of the source in one launch.  fp64: float64 tables,
        exponentials and logarithms on the device (the guard-band evaluation of the samplers)."""
        f = fluxes_of(self.params) if fluxes is None else np.asarray(fluxes)
        assert np.all(~np.isnan(f)), 'passing in NAN fluxes.'
        c = self._sample_cache()
        if not c["imgs"]:
            return 0
        ll, _ = engine.items_loglik(c["imgs"], self._record(u, fluxes, shape), c["items_cond"],
                                    mode=_ffi.CG_LL_CONDITIONAL, aux=c["aux"], aux_offsets=c["offs"],
                                    use_fp64=fp64)
        return float(ll.sum().item())

    def location_likelihood(self, u):
        return self.log_likelihood(u=u)

    #: |likelihood - slice level| below which the slice sampler decides in float64 (slicesample_gen)
    slice_guard = 1e-2

    def log_likelihood_isolated(self, u=None, fluxes=None, shape=None):
        """:188-237."""
        c = self._sample_cache()
        if not c["imgs"]:
            return 0
        ll, _ = engine.items_loglik(c["imgs"], self._record(u, fluxes, shape), c["items_iso"],
                                    mode=_ffi.CG_LL_POISSON_BG)
        return float(ll.sum().item())

    # ---- resampling ----------------------------------------------------------
    def resample(self, rng=None):
        """:242-245.  rng: optional numpy RandomState for this source's draws
        (default: the global numpy.random stream, as in the reference)."""
        assert len(self.sample_image_list) != 0, "resample source needs sampled images"
        self.resample_fluxes(rng=rng)
        self.resample_location(rng=rng)

    def resample_type(self, proposal_fun=None):
        proposal_fun = self.propose_other_type_prior if proposal_fun is None else proposal_fun
        proposal, logpdf, logreverse, logdet = proposal_fun()
        fimgs = [self.model.field_list[0].img_dict[b] for b in self.model.bands]
        accept_logprob = self.calculate_acceptance_logprob(proposal, logpdf, logreverse, logdet, fimgs)
        if np.log(np.random.rand()) < accept_logprob:
            self.params = proposal

    def propose_other_type_prior(self):
        if self.is_star():
            params, logprob = self.model.prior_sample('galaxy', u=self.params.u)
        else:
            params, logprob = self.model.prior_sample('star', u=self.params.u)
        return params, logprob, self.model.logprior(self.params), 0.

    def calculate_acceptance_logprob(self, proposal, logprob_proposal, logprob_reverse, logdet, images):
        """:275-306: current and proposed state scored in one launch over the
        cached bounding boxes / background patches (self.bounding_boxes,
        self.background_image_dict)."""
        boxes, bgs = [], []
        for img in images:
            xlim, ylim = self.bounding_boxes[img]
            boxes.append((int(xlim[0]), int(xlim[1]), int(ylim[0]), int(ylim[1])))
            bgs.append(self.background_image_dict[img])
        scorer = engine.ProposalScorer(images, boxes, bgs, max_proposals=1, guard=0.0)
        packed = np.concatenate([record_of(self.params), record_of(proposal)])
        delta, _, _ = scorer.score(packed, np.zeros(1))
        curr_logprior = self.model.logprior(self.params)
        prop_logprior = self.model.logprior(proposal)
        return float(delta[0]) + (prop_logprior - curr_logprior) + \
            (logprob_reverse - logprob_proposal) + logdet

    def _location_slice_args(self, rng=None):
        self.n_refined = [0]
        return dict(step=self.du / 5, step_out=False, upper_bound=self.u_upper,
                    lower_bound=self.u_lower, rng=rng, guard=self.slice_guard, n_refined=self.n_refined,
                    refine=lambda u: self.log_likelihood(u=u, fp64=True))

    def resample_location(self, u=None, rng=None):
        """:308-319."""
        if u is None:
            u = self.params.u.copy()
        u, ll = slicesample(init_x=u, logprob=lambda u: self.location_likelihood(u),
                            **self._location_slice_args(rng))
        self.params.u = u
        return u

    def resample_shape(self):
        return

    def resample_fluxes(self, rng=None, psf_sums_by_image=None):
        """:327-350: gamma conditional of the five fluxes.  psf_sums_by_image:
        optional precomputed sum of the unit-flux patch on every sample image
        (resample_sources_batched renders them for all sources in one launch)."""
        a_0, b_0 = 5., .005
        band_counts = {b: 0 for b in BANDS}
        psf_sums = {b: 0 for b in BANDS}
        for k, (src_img, fits_img, pixel_grid) in enumerate(self.sample_image_list):
            band_counts[fits_img.band] += np.sum(np.array(src_img.data))
            if psf_sums_by_image is None:
                psf_ns, ylim, xlim = self.compute_scatter_on_pixels(fits_img)
                tot = np.sum(psf_ns)
            else:
                tot = psf_sums_by_image[k]
            psf_sums[fits_img.band] += tot * fits_img.kappa / fits_img.calib
        a_n = a_0 + np.array([band_counts[b] for b in BANDS])
        b_n = b_0 + np.array([psf_sums[b] for b in BANDS])
        self.params.fluxes = (rng or np.random).gamma(a_n, 1. / b_n)

    def compute_scatter_on_pixels(self, fits_image, u=None, shape=None, xlim=None, ylim=None,
                                  pixel_grid=None, force_type=None):
        """:351-388."""
        u = self.params.u if u is None else u
        render_star = self.is_star() if force_type is None else (force_type == 'star')
        render_gal = self.is_galaxy() if force_type is None else (force_type == 'galaxy')
        if render_star:
            return celeste.gen_point_source_psf_image(u, fits_image, xlim=xlim, ylim=ylim,
                                                      pixel_grid=pixel_grid)
        elif render_gal:
            if shape is None:
                shape = self.params.shape
            return gal_funs.gen_galaxy_psf_image(shape, u, fits_image, xlim=xlim, ylim=ylim,
                                                 check_overlap=True, unconstrained=False,
                                                 return_patch=True)
        raise NotImplementedError("only stars and galaxies have photon scattering images")

    def compute_model_patch(self, fits_image, u=None, xlim=None, ylim=None):
        """:390-395."""
        patch, ylim, xlim = self.compute_scatter_on_pixels(fits_image, u=u, xlim=xlim, ylim=ylim)
        band_flux = (fluxes_of(self.params)[BANDS.index(fits_image.band)] / fits_image.calib) * fits_image.kappa
        return band_flux * patch, ylim, xlim


def make_bbox_dict(params, images, pixel_radius=None):
    """:434-455."""
    if pixel_radius is None:
        raise NotImplementedError

    def image_bbox(params, img):
        img_ymax, img_xmax = img.shape
        px, py = img.equa2pixel(params.u)
        xlim = (np.max([0, int(np.floor(px - pixel_radius))]), np.min([img_xmax, int(np.ceil(px + pixel_radius))]))
        ylim = (np.max([0, int(np.floor(py - pixel_radius))]), np.min([img_ymax, int(np.ceil(py + pixel_radius))]))
        return xlim, ylim
    return {img: image_bbox(params, img) for img in images}


def get_active_sources(source, source_list, image):
    """:457-472."""
    def intersect(sa, sb, image):
        xlima, ylima = sa.bounding_boxes[image]
        xlimb, ylimb = sb.bounding_boxes[image]
        widtha, heighta = xlima[1] - xlima[0], ylima[1] - ylima[0]
        widthb, heightb = xlimb[1] - xlimb[0], ylimb[1] - ylimb[0]
        return (np.abs(xlima[0] - xlimb[0]) * 2 < (widtha + widthb)) and \
               (np.abs(ylima[0] - ylimb[0]) * 2 < (heighta + heightb))
    return [s for s in source_list if intersect(s, source, image) and s is not source]


def generate_background_patch(source, source_list, image):
    """:476-483: the active sources rendered onto the source's box in one
    launch + epsilon."""
    active = get_active_sources(source, source_list, image)
    xlim, ylim = source.bounding_boxes[image]
    box = (int(xlim[0]), int(xlim[1]), int(ylim[0]), int(ylim[1]))
    h, w = box[3] - box[2], box[1] - box[0]
    if not active:
        return np.zeros((h, w)) + image.epsilon
    recs = np.concatenate([record_of(s.params) for s in active])
    items = engine.make_items(np.arange(len(active)), [0] * len(active), box=[box] * len(active))
    patches, _ = engine.render_patches([image], recs, items=items)
    bg = patches[:, :h * w].sum(dim=0).cpu().numpy().astype(np.float64).reshape(h, w)
    return bg + image.epsilon


# ---------------------------------------------------------------------------
# all sources of a model at once (no reference analogue: models.py:81-83 loops
# `for src in srcs: src.resample()`, one likelihood launch per slice evaluation
# per source; here every round of the lock-step slice samplers is ONE launch)
# ---------------------------------------------------------------------------
class BatchedConditional(object):
    """Source.log_likelihood (:134-183) of MANY sources in one cg_items_loglik
    launch.  Built once per sweep from the sources' sample_image_list."""

    def __init__(self, srcs):
        self.srcs = list(srcs)
        imgs, index = [], {}
        src_idx, img_idx, boxes, zs = [], [], [], []
        for s, src in enumerate(self.srcs):
            for samp_img, fits_img, _ in src.sample_image_list:
                if id(fits_img) not in index:
                    index[id(fits_img)] = len(imgs)
                    imgs.append(fits_img)
                src_idx.append(s)
                img_idx.append(index[id(fits_img)])
                boxes.append((int(samp_img.x0), int(samp_img.x1), int(samp_img.y0), int(samp_img.y1)))
                zs.append(np.ascontiguousarray(np.asarray(samp_img.data), dtype=np.float32).ravel())
        if len(imgs) > _ffi.CG_MAX_IMAGES:
            raise ValueError("more than %d distinct images: batch field by field" % _ffi.CG_MAX_IMAGES)
        self.imgs = imgs
        self.src_idx = np.asarray(src_idx, dtype=np.int64)
        n = len(zs)
        self.offs = np.zeros(n, dtype=np.int64)
        if n:
            self.offs[1:] = np.cumsum([z.size for z in zs])[:-1]
        self.aux = torch.from_numpy(np.concatenate(zs)).to(engine.device()) if n else None
        self.items = engine.make_items(self.src_idx, np.asarray(img_idx), box=boxes, aux=np.arange(n)) \
            if n else None
        self.max_patch_px = max([z.size for z in zs] + [1])

    def refresh(self):
        """Re-pack the sources' current parameters (call after they changed)."""
        self._base = np.concatenate([src._record() for src in self.srcs])
        return self._base

    def records(self, u=None):
        """CgSource records of all sources; u: optional {source index: location}.
        Only the location column differs from the packed base, as in
        Source._record(u=...)."""
        base = getattr(self, "_base", None)
        if base is None:
            base = self.refresh()
        if not u:
            return base
        recs = base.copy()
        idx = np.fromiter(u.keys(), dtype=np.int64, count=len(u))
        recs["u"][idx] = np.asarray(list(u.values()), dtype=np.float64)
        return recs

    def loglik(self, live, us):
        """log_likelihood(u=us[k]) of source live[k], all in one launch."""
        if self.items is None:
            return np.zeros(len(live))
        over = {s: np.asarray(x, dtype=np.float64) for s, x in zip(live, us)}
        recs = self.records(over)
        assert np.all(~np.isnan(recs["flux"])), 'passing in NAN fluxes.'   # :140
        sel = np.isin(self.src_idx, np.asarray(live))
        ll, _ = engine.items_loglik(self.imgs, recs, self.items[sel], mode=_ffi.CG_LL_CONDITIONAL,
                                    aux=self.aux, aux_offsets=self.offs, max_patch_px=self.max_patch_px)
        per_src = np.bincount(self.src_idx[sel], weights=ll.cpu().numpy(), minlength=len(self.srcs))
        return per_src[np.asarray(live)]

    def unit_patch_sums(self):
        """sum of the unit-flux patch of every (source, sample image) pair on
        its natural footprint (compute_scatter_on_pixels with no limits), one
        render launch; returned per source in sample_image_list order."""
        if self.items is None:
            return [[] for _ in self.srcs]
        items = self.items.copy()
        items["flags"] = _ffi.CG_ITEM_FLUX_UNIT
        items["aux"] = -1
        cap = min(max(im.shape[0] * im.shape[1] for im in self.imgs), 203 * 203)
        patches, boxes = engine.render_patches(self.imgs, self.records(), items=items, max_patch_px=cap)
        tot = patches.double().sum(dim=1).cpu().numpy()
        out = [[] for _ in self.srcs]
        for k, s in enumerate(self.src_idx):
            out[s].append(float(tot[k]))
        return out


def resample_sources_batched(srcs, seed=0):
    """Source.resample() for every source, with the likelihood evaluations of
    all sources batched: one render launch for the flux conditionals, then the
    location slice samplers advance in lock-step (slicesample.run_lockstep), one
    cg_items_loglik launch per round.  Source s draws from RandomState(seed + s);
    the result equals `src.resample(rng=RandomState(seed + s))` source by source."""
    from .slicesample import run_lockstep, slicesample_gen
    srcs = list(srcs)
    for src in srcs:
        assert len(src.sample_image_list) != 0, "resample source needs sampled images"
    rngs = [np.random.RandomState(seed + s) for s in range(len(srcs))]
    bc = BatchedConditional(srcs)
    sums = bc.unit_patch_sums()
    for src, rng, ps in zip(srcs, rngs, sums):
        src.resample_fluxes(rng=rng, psf_sums_by_image=ps)
    bc.refresh()
    gens = [slicesample_gen(src.params.u.copy(), **src._location_slice_args(rng))
            for src, rng in zip(srcs, rngs)]
    res = run_lockstep(gens, bc.loglik)
    for src, (u, ll) in zip(srcs, res):
        src.params.u = u
    return run_lockstep.last_rounds


def resample_sources_on_device(srcs, seed=0, rngs=None, n_uniforms=160):
    """Source.resample() for every source with the location slice samplers run ON THE DEVICE
    (cg_slice_sample_locations: one launch for all sources, however many shrink rounds they take;
    SURVEY.md 8f N2).  Source s draws from rngs[s] (default RandomState(seed + s)) in the reference's
    order: gamma conditionals of the fluxes, the coordinate shuffle, then the sampler's uniforms -- a
    block of n_uniforms is handed to the device and the generator is advanced by exactly the number
    consumed, so the stream continues as if the loop had run on the host.  Comparisons inside the
    float64 guard band are decided in the kernel; a source the device flags (block exhausted, NaN) is
    redone by the lock-step host sampler from the same generator state.  The result equals resample_sources_batched / the
    per-source loop bit for bit.  -> dict(device=..., fallback=..., evals=...)."""
    from .slicesample import run_lockstep, slicesample_gen
    srcs = list(srcs)
    for src in srcs:
        assert len(src.sample_image_list) != 0, "resample source needs sampled images"
    if rngs is None:
        rngs = [np.random.RandomState(seed + s) for s in range(len(srcs))]
    bc = BatchedConditional(srcs)
    sums = bc.unit_patch_sums()
    for src, rng, ps in zip(srcs, rngs, sums):
        src.resample_fluxes(rng=rng, psf_sums_by_image=ps)
    bc.refresh()
    n_img = len(bc.imgs)
    S = len(srcs)
    uniform_layout = all(len(src.sample_image_list) == n_img and
                         all(fi is bi for (_, fi, _), bi in zip(src.sample_image_list, bc.imgs)) for src in srcs)
    if not uniform_layout or n_img > 8:
        gens = [slicesample_gen(src.params.u.copy(), **src._location_slice_args(rng)) for src, rng in zip(srcs, rngs)]
        for src, (u, ll) in zip(srcs, run_lockstep(gens, bc.loglik)):
            src.params.u = u
        return dict(device=0, fallback=S, evals=0)
    # device layout: photon patches and boxes item-major with a common stride
    stride = int(bc.max_patch_px)
    photons = torch.zeros((S * n_img, stride), dtype=torch.float32, device=engine.device())
    boxes = np.zeros(S * n_img, dtype=_ffi.BOX_DTYPE)
    flat = bc.aux.cpu().numpy()
    host = np.zeros((S * n_img, stride), dtype=np.float32)
    for k in range(S * n_img):
        bx = bc.items["box"][k]
        n = int((bx[1] - bx[0]) * (bx[3] - bx[2]))
        host[k, :n] = flat[bc.offs[k]:bc.offs[k] + n]
        boxes[k] = tuple(int(v) for v in bx)
    photons.copy_(torch.from_numpy(host))
    before, after, order, uni = [], [], np.zeros((S, 2), dtype=np.int32), np.zeros((S, n_uniforms))
    for s, rng in enumerate(rngs):
        before.append(rng.get_state())
        o = [0, 1]
        rng.shuffle(o)                               # slicesample.py:228-230
        order[s] = o
        after.append(rng.get_state())
        uni[s] = rng.random_sample(n_uniforms)
    recs = bc.records()
    d_srcs = engine.to_device(recs)
    d_boxes = engine.to_device(boxes)
    ll, used, status, evals = engine.slice_sample_locations(bc.imgs, d_srcs, S, d_boxes, photons, stride, uni, order,
                                                            sigma=1.0, guard=Source.slice_guard)
    new = d_srcs.cpu().numpy().view(_ffi.SRC_DTYPE)
    redo = []
    for s, (src, rng) in enumerate(zip(srcs, rngs)):
        if status[s] == 0:
            src.params.u = new["u"][s].copy()
            rng.set_state(after[s])
            rng.random_sample(int(used[s]))          # the draws the device consumed
        else:
            rng.set_state(before[s])
            redo.append(s)
    if redo:
        sub = BatchedConditional([srcs[s] for s in redo])
        sub.refresh()
        gens = [slicesample_gen(srcs[s].params.u.copy(), **srcs[s]._location_slice_args(rngs[s])) for s in redo]
        for s, (u, l) in zip(redo, run_lockstep(gens, sub.loglik)):
            srcs[s].params.u = u
    return dict(device=S - len(redo), fallback=len(redo), evals=int(evals.sum()), status=status)
