"""CPU BASELINE / CHECKER -- TEST INFRASTRUCTURE ONLY (bench.py's cpu_baseline
and ``--impl reference`` legs, tests).  Never imported by the product package.

The reference's fastest CPU route for the hot path ("B-fast"):
the compiled Cython ``gmm_like_2d`` (OpenMP, float64; oracle/_ref, built from
/root/reference by oracle/build_ref.py) evaluated on every (source, band)
footprint with mixture tables from the compiled ``celeste_fast``, patches
added into the model image, then the NumPy Poisson log-likelihood
(celeste.py:237-240).  When oracle/_ref is absent the NumPy restatement
(oracle/ref_cpu.py) is used instead and the baseline is labelled "port".
"""
from __future__ import annotations

import os
import time

import numpy as np

from . import build_ref, ref_cpu as ref


def backend():
    """-> (kind, gmm_like_2d(probs,x,ws,mus,sigs), prof_params, joint_params)"""
    gm, cf = build_ref.load()
    if gm is not None:
        return ("reference", gm.gmm_like_2d, cf.gen_galaxy_prof_psf_mixture_params,
                cf.gen_galaxy_psf_mixture_params)

    def _g(probs, x, ws, mus, sigs):
        ref.gmm_like_2d(x, ws, mus, sigs, probs=probs)
    return ("port", _g, ref.gen_galaxy_prof_psf_mixture_params, ref.gen_galaxy_psf_mixture_params)


_omp = {"lib": None, "n": None}


def set_threads(n):
    """OpenMP team size of the compiled reference (gmm_like_2d's prange).  torchrun exports
    OMP_NUM_THREADS=1 to its ranks; the baseline must not inherit that silently, so the count is set
    explicitly through the runtime (libgomp) the compiled module links.  Returns the count in force."""
    import ctypes
    n = max(1, int(n))
    if _omp["lib"] is None:
        for name in ("libgomp.so.1", "libgomp.so"):
            try:
                _omp["lib"] = ctypes.CDLL(name)
                break
            except OSError:
                continue
    if _omp["lib"] is None:
        env = os.environ.get("OMP_NUM_THREADS")
        _omp["n"] = int(env) if env else (os.cpu_count() or 1)
        return _omp["n"]
    _omp["lib"].omp_set_num_threads(n)
    _omp["lib"].omp_get_max_threads.restype = ctypes.c_int
    _omp["n"] = int(_omp["lib"].omp_get_max_threads())
    return _omp["n"]


def host_threads():
    """threads this process may use (cgroup / affinity aware)"""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def threads_used(kind):
    if kind != "reference":
        return 1
    if _omp["n"] is not None:
        return _omp["n"]
    env = os.environ.get("OMP_NUM_THREADS")
    return int(env) if env else (os.cpu_count() or 1)


def _box(px, py, bound, H, W, trunc):
    if trunc:      # celeste.py:139-140
        return (max(0, int(px - bound)), min(int(px + bound + 1), W),
                max(0, int(py - bound)), min(int(py + bound + 1), H))
    return (int(max(0, np.floor(px - bound))), int(min(W, np.ceil(px + bound))),    # :208-211
            int(max(0, np.floor(py - bound))), int(min(H, np.ceil(py + bound))))


def render_source(src, img, bound_cap, be):
    """One (source, band) rate patch through the compiled route.
    -> (patch or None, (x0,x1,y0,y1), evals)"""
    _, gmm, _, joint = be
    H, W = img.nelec.shape
    v = img.equa2pixel(src.u)
    if src.a == 0:
        ws, mus, sigs = img.weights, img.means + v, img.covars
        bound = img.R
        trunc = True
    else:
        Tinv = ref.gen_galaxy_transformation(src.sigma, src.rho, src.phi, img.cd_at_pixel(v[0], v[1]))
        Wm = np.dot(Tinv, Tinv.T)
        ws, mus, sigs = joint(np.array([src.theta, 1. - src.theta]), Wm, v, img.weights, img.means,
                              img.covars, ref.exp_amp, ref.exp_var, ref.dev_amp, ref.dev_var)
        bound = ref.calc_bounding_radius(ws, mus, sigs, 1e-5, center=v)
        trunc = False
    if bound_cap > 0:
        bound = min(bound, bound_cap)
    x0, x1, y0, y1 = _box(v[0], v[1], bound, H, W, trunc)
    if x1 <= x0 or y1 <= y0:
        return None, (x0, x0, y0, y0), 0
    xx, yy = np.meshgrid(np.arange(x0, x1, dtype=np.float64), np.arange(y0, y1, dtype=np.float64))
    pts = np.column_stack((xx.ravel(), yy.ravel()))
    probs = np.zeros(pts.shape[0])
    gmm(probs, pts, np.ascontiguousarray(ws), np.ascontiguousarray(mus), np.ascontiguousarray(sigs))
    flux = img.nmgy2counts(src.flux_dict[img.band])
    return (probs * flux).reshape(xx.shape), (x0, x1, y0, y1), pts.shape[0] * ws.shape[0]


def field_loglik(images, srcs, bound_cap=0.0, be=None):
    """A7/A8 on the CPU: per-image log-likelihood + component evaluations."""
    be = be or backend()
    lls, evals = [], 0
    for img in images:
        lam = np.full(img.nelec.shape, img.epsilon)
        for s in srcs:
            p, (x0, x1, y0, y1), n = render_source(s, img, bound_cap, be)
            evals += n
            if p is not None:
                lam[y0:y1, x0:x1] += p
        lls.append(float(np.sum(img.nelec * np.log(lam) - lam)))
    return lls, evals


def timed_sample(images, srcs, bound_cap, budget_s=15.0, be=None):
    """Runs field_loglik over growing source subsets until ~budget_s of CPU
    work is spent.  -> dict(evals, seconds, n_src, n_img)"""
    be = be or backend()
    n = min(len(srcs), 16)
    field_loglik(images[:1], srcs[:2], bound_cap, be)      # warm-up
    t0 = time.perf_counter()
    _, ev = field_loglik(images, srcs[:n], bound_cap, be)
    dt = time.perf_counter() - t0
    if dt < budget_s / 3 and n < len(srcs):
        n2 = int(min(len(srcs), max(n + 1, n * (budget_s * .8) / max(dt, 1e-3))))
        t0 = time.perf_counter()
        _, ev = field_loglik(images, srcs[:n2], bound_cap, be)
        dt = time.perf_counter() - t0
        n = n2
    # every source fits the budget several times over (many host threads): repeat whole passes
    while n == len(srcs) and dt < budget_s * .6:
        t0 = time.perf_counter()
        _, e2 = field_loglik(images, srcs, bound_cap, be)
        dt += time.perf_counter() - t0
        ev += e2
    return dict(evals=ev, seconds=dt, n_src=n, n_img=len(images))


def live_field_loglik(images, srcs, bound_cap=0.0):
    """B-live: the reference's LIVE numpy route -- util/dists/mog.py:5-21,75-112 through
    celeste.py:114-176 / celeste_galaxy_conditionals.py:185-214 -- as restated in oracle/ref_cpu.py
    (pinned bit for bit on the reference's own Python, tests/test_cpu_oracle_l3.py), single thread.
    -> (per-image log-likelihoods, component evaluations)"""
    lls, evals = [], 0
    for img in images:
        H, W = img.nelec.shape
        lam = np.full((H, W), img.epsilon)
        for s in srcs:
            v = img.equa2pixel(s.u)
            if s.a == 0:
                bound, trunc, K = img.R, True, img.weights.shape[0]
            else:
                bound, trunc, K = ref.gen_galaxy_psf_image_bound(s.shape, s.u, img), False, 14 * img.weights.shape[0]
            if bound_cap > 0:
                bound = min(bound, bound_cap)
            x0, x1, y0, y1 = _box(v[0], v[1], bound, H, W, trunc)
            if x1 <= x0 or y1 <= y0:
                continue
            p, _, _ = ref.compute_model_patch(s, img, xlim=(x0, x1), ylim=(y0, y1))
            lam[y0:y1, x0:x1] += p
            evals += (x1 - x0) * (y1 - y0) * K
        lls.append(float(np.sum(img.nelec * np.log(lam) - lam)))
    return lls, evals


def proposal_score(cur, prop, images, boxes, backgrounds, be):
    """Source.calculate_acceptance_logprob (sources.py:275-306) around the compiled route: for the
    current and the proposed state, per image, render on the cached box (compiled mixture table +
    gmm_like_2d), add the cached background, masked Poisson log-likelihood (sources.py:6-12).
    -> (delta, component evaluations)"""
    _, gmm, _, joint = be
    tot, evals = [0., 0.], 0
    for k, src in enumerate((cur, prop)):
        for img, (x0, x1, y0, y1), bg in zip(images, boxes, backgrounds):
            v = img.equa2pixel(src.u)
            if src.a == 0:
                ws, mus, sigs = img.weights, img.means + v, img.covars
            else:
                Tinv = ref.gen_galaxy_transformation(src.sigma, src.rho, src.phi, img.cd_at_pixel(v[0], v[1]))
                ws, mus, sigs = joint(np.array([src.theta, 1. - src.theta]), np.dot(Tinv, Tinv.T), v, img.weights,
                                      img.means, img.covars, ref.exp_amp, ref.exp_var, ref.dev_amp, ref.dev_var)
            xx, yy = np.meshgrid(np.arange(x0, x1, dtype=np.float64), np.arange(y0, y1, dtype=np.float64))
            pts = np.column_stack((xx.ravel(), yy.ravel()))
            probs = np.zeros(pts.shape[0])
            gmm(probs, pts, np.ascontiguousarray(ws), np.ascontiguousarray(mus), np.ascontiguousarray(sigs))
            model = bg + (probs * img.nmgy2counts(src.flux_dict[img.band])).reshape(xx.shape)
            tot[k] += ref.poisson_loglike(img.nelec[y0:y1, x0:x1], model, img.invvar[y0:y1, x0:x1])
            evals += pts.shape[0] * ws.shape[0]
    return tot[1] - tot[0], evals


def timed_proposals(cur, props, images, boxes, backgrounds, budget_s=8.0, be=None):
    """proposals/s of the reference's MH score on the host: proposals are scored one after the
    other (as sources.py:247-258 does) until ~budget_s is spent."""
    be = be or backend()
    proposal_score(cur, props[0], images, boxes, backgrounds, be)          # warm-up
    n, ev = 0, 0
    t0 = time.perf_counter()
    while n < len(props) and time.perf_counter() - t0 < budget_s:
        _, e = proposal_score(cur, props[n], images, boxes, backgrounds, be)
        ev += e
        n += 1
    dt = time.perf_counter() - t0
    return dict(proposals=n, seconds=dt, evals=ev)
