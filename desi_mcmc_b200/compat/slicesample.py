"""CelestePy/util/infer/slicesample.py:89-247: the slice sampler's control
flow (host), drawing from numpy.random in the reference's order so that a
seeded chain reproduces; `logprob` is a device likelihood.  The sampler is
written once as a generator (slicesample_gen) so that many chains can be
advanced in lock-step with one batched launch per round (run_lockstep)."""
from __future__ import annotations

import numpy as np
import numpy.random as npr


def slicesample_gen(init_x, **slice_sample_args):
    """The sampler of slicesample() as a generator: it YIELDS every point whose
    log-probability it needs and is SENT the value, and returns (new_x, new_llh).
    One code path serves the sequential driver (slicesample) and the lock-step
    driver (run_lockstep: many chains, one batched device launch per round).
    `rng` (keyword) supplies rand / randn / shuffle; default numpy.random, the
    reference's global stream.

    Guard band (keywords `refine`, `guard`; no reference analogue -- the reference's logprob is
    float64): the device likelihoods are fp32-rate sums, good to ~1e-3 absolute.  Every comparison of
    a likelihood with the slice level that falls within `guard` of a tie is decided again with
    refine(x), a float64 evaluation of the same likelihood (level and value both), so that a chain
    takes the float64 reference's decisions by construction, not by margin.  `n_refined` (a one-element
    list, keyword) counts the refinements."""
    rng = slice_sample_args.get('rng', None) or npr
    refine = slice_sample_args.get('refine', None)
    guard = slice_sample_args.get('guard', 0.0) if refine is not None else 0.0
    n_refined = slice_sample_args.get('n_refined', None)
    sigma = slice_sample_args.get('sigma', 1.0)
    step_out = slice_sample_args.get('step_out', True)
    max_steps_out = slice_sample_args.get('max_steps_out', 1000)
    compwise = slice_sample_args.get('compwise', True)
    numdir = slice_sample_args.get('numdir', 2)
    doubling_step = slice_sample_args.get('doubling_step', True)
    upper_bound = slice_sample_args.get('upper_bound', np.inf)
    lower_bound = slice_sample_args.get('lower_bound', -np.inf)

    def direction_slice(direction, init_x):
        def at(z):
            return direction * z + init_x

        level = {}      # float64 slice level, computed on the first near-tie of this direction

        def above(val, z, llh_s):
            """val > llh_s, decided in float64 when the two are within the guard band"""
            if guard > 0. and abs(val - llh_s) < guard:
                if 'exact' not in level:
                    level['exact'] = level['log_u'] + refine(at(0.0))
                if n_refined is not None:
                    n_refined[0] += 1
                return refine(at(z)) > level['exact']
            return val > llh_s

        def acceptable(z, llh_s, L, U):
            while (U - L) > 1.1 * sigma:
                middle = 0.5 * (L + U)
                splits = (middle > 0 and z >= middle) or (middle <= 0 and z < middle)
                if z < middle:
                    U = middle
                else:
                    L = middle
                if splits:                       # `and` chain of the reference, short-circuit kept
                    if not above((yield at(U)), U, llh_s):
                        if not above((yield at(L)), L, llh_s):
                            return False
            return True

        upper = sigma * rng.rand()
        lower = upper - sigma
        level['log_u'] = np.log(rng.rand())
        llh_s = level['log_u'] + (yield at(0.0))
        l_steps_out = 0
        u_steps_out = 0
        if step_out:
            if doubling_step:
                while True:
                    out = above((yield at(lower)), lower, llh_s)
                    if not out:
                        out = above((yield at(upper)), upper, llh_s)
                    if not (out and (l_steps_out + u_steps_out) < max_steps_out):
                        break
                    if rng.rand() < 0.5:
                        l_steps_out += 1
                        lower -= (upper - lower)
                    else:
                        u_steps_out += 1
                        upper += (upper - lower)
            else:
                while above((yield at(lower)), lower, llh_s) and l_steps_out < max_steps_out:
                    l_steps_out += 1
                    lower -= sigma
                while above((yield at(upper)), upper, llh_s) and u_steps_out < max_steps_out:
                    u_steps_out += 1
                    upper += sigma
        start_upper, start_lower = upper, lower
        while True:
            new_z = (upper - lower) * rng.rand() + lower
            new_llh = yield at(new_z)
            if np.isnan(new_llh):
                raise Exception("Slice sampler got a NaN")
            if above(new_llh, new_z, llh_s) and (yield from acceptable(new_z, llh_s, start_lower, start_upper)):
                break
            elif new_z < 0:
                lower = new_z
            elif new_z > 0:
                upper = new_z
            else:
                raise Exception("Slice sampler shrank to zero!")
        return new_z * direction + init_x, new_llh

    if type(init_x) == float or isinstance(init_x, np.number):
        init_x = np.array([init_x])
        scalar = True
    else:
        scalar = False
    if type(upper_bound) == float or isinstance(upper_bound, np.number):
        upper_bound = np.tile(upper_bound, init_x.shape)
    if type(lower_bound) == float or isinstance(lower_bound, np.number):
        lower_bound = np.tile(lower_bound, init_x.shape)
    assert np.all(init_x < upper_bound), "init_x (%s) >= ub (%s)" % \
        (np.array_str(init_x), np.array_str(upper_bound))
    assert np.all(init_x > lower_bound), "init_x (%s) <= lb (%s)" % \
        (np.array_str(init_x), np.array_str(lower_bound))
    dims = init_x.shape[0]
    if compwise:
        ordering = list(range(dims))
        rng.shuffle(ordering)
        new_x = init_x.copy()
        for d in ordering:
            direction = np.zeros((dims))
            direction[d] = 1.0
            new_x, new_llh = yield from direction_slice(direction, new_x)
    else:
        new_x = init_x
        for d in range(numdir):
            direction = rng.randn(dims)
            direction = direction / np.sqrt(np.sum(direction ** 2))
            new_x, new_llh = yield from direction_slice(direction, new_x)
    if scalar:
        return float(new_x[0]), new_llh
    return new_x, new_llh


def slicesample(init_x, logprob, *logprob_args, **slice_sample_args):
    """util/infer/slicesample.py:89-247, same signature and keyword set."""
    gen = slicesample_gen(init_x, **slice_sample_args)
    try:
        x = next(gen)
        while True:
            x = gen.send(logprob(x, *logprob_args))
    except StopIteration as done:
        return done.value


def run_lockstep(gens, batch_logprob):
    """Advance many sampler generators together: each round collects the point
    every live chain is waiting on and evaluates them in ONE call
    batch_logprob(indices, points) -> log-probabilities (a single device
    launch for all chains of a field instead of one per chain).  Every chain
    sees exactly the values, and draws exactly the random numbers, it would
    see when driven alone.  Returns the generators' return values."""
    results = [None] * len(gens)
    pending = {}
    for i, g in enumerate(gens):
        try:
            pending[i] = next(g)
        except StopIteration as done:
            results[i] = done.value
    rounds = 0
    while pending:
        idx = sorted(pending)
        lls = batch_logprob(idx, [pending[i] for i in idx])
        rounds += 1
        for i, ll in zip(idx, lls):
            try:
                pending[i] = gens[i].send(float(ll))
            except StopIteration as done:
                results[i] = done.value
                del pending[i]
    run_lockstep.last_rounds = rounds
    return results


def slicesample_step_out(xx, llh_func, last_llh=None, step=1, step_out=True, x_l=None, x_r=None,
                         lb=-np.inf, ub=np.inf):
    """util/infer/slicesample.py:4-86: coordinate-wise slice sampler with
    stepping out, dimensions in a random permutation.  Same draws from
    numpy.random in the same order as the reference."""
    xx = np.atleast_1d(np.asarray(xx, dtype=np.float64)).copy()
    dims = xx.shape[0]
    perm = list(range(dims))
    np.random.shuffle(perm)
    lb = np.broadcast_to(np.asarray(lb, dtype=np.float64), xx.shape)
    ub = np.broadcast_to(np.asarray(ub, dtype=np.float64), xx.shape)
    assert np.all(xx < ub), "xx (%s) >= ub (%s)" % (np.array_str(xx), np.array_str(ub))
    assert np.all(xx > lb), "xx (%s) <= lb (%s)" % (np.array_str(xx), np.array_str(lb))
    if np.isscalar(step):
        step = np.tile(step, dims)
    else:
        step = np.asarray(step, dtype=np.float64)
    if last_llh is None:
        last_llh = llh_func(xx)
    for d in perm:
        llh0 = last_llh + np.log(np.random.rand())
        rr = np.random.rand()
        if x_l is None:
            x_l = xx.copy()
            x_l[d] = max(x_l[d] - rr * step[d], lb[d])
        else:
            x_l = np.atleast_1d(np.asarray(x_l, dtype=np.float64)).copy()
            assert x_l.shape == xx.shape
            assert np.all(x_l <= xx)
        if x_r is None:
            x_r = xx.copy()
            x_r[d] = min(x_r[d] + (1 - rr) * step[d], ub[d])
        else:
            x_r = np.atleast_1d(np.asarray(x_r, dtype=np.float64)).copy()
            assert x_r.shape == xx.shape
            assert np.all(x_r >= xx)
        if step_out:
            llh_l = llh_func(x_l)
            while llh_l > llh0 and x_l[d] > lb[d]:
                x_l[d] = max(x_l[d] - step[d], lb[d])
                llh_l = llh_func(x_l)
            llh_r = llh_func(x_r)
            while llh_r > llh0 and x_r[d] < ub[d]:
                x_r[d] = min(x_r[d] + step[d], ub[d])
                llh_r = llh_func(x_r)
        assert np.isfinite(llh0)
        x_cur = xx.copy()
        while True:
            xd = np.random.rand() * (x_r[d] - x_l[d]) + x_l[d]
            x_cur[d] = xd
            last_llh = llh_func(x_cur)
            if last_llh > llh0:
                xx[d] = xd
                break
            elif xd > xx[d]:
                x_r[d] = xd
            elif xd < xx[d]:
                x_l[d] = xd
            else:
                raise Exception("Slice sampler shrank too far.")
    if not np.isfinite(last_llh):
        raise Exception("Likelihood is not finite at sampled point")
    return xx, last_llh
