"""CelestePy/util/infer/slicesample.py:89-247: the slice sampler's control
flow (host), drawing from numpy.random in the reference's order so that a
seeded chain reproduces; `logprob` is a device likelihood."""
from __future__ import annotations

import numpy as np
import numpy.random as npr


def slicesample(init_x, logprob, *logprob_args, **slice_sample_args):
    sigma = slice_sample_args.get('sigma', 1.0)
    step_out = slice_sample_args.get('step_out', True)
    max_steps_out = slice_sample_args.get('max_steps_out', 1000)
    compwise = slice_sample_args.get('compwise', True)
    numdir = slice_sample_args.get('numdir', 2)
    doubling_step = slice_sample_args.get('doubling_step', True)
    upper_bound = slice_sample_args.get('upper_bound', np.inf)
    lower_bound = slice_sample_args.get('lower_bound', -np.inf)

    def direction_slice(direction, init_x):
        def dir_logprob(z):
            return logprob(direction * z + init_x, *logprob_args)

        def acceptable(z, llh_s, L, U):
            while (U - L) > 1.1 * sigma:
                middle = 0.5 * (L + U)
                splits = (middle > 0 and z >= middle) or (middle <= 0 and z < middle)
                if z < middle:
                    U = middle
                else:
                    L = middle
                if splits and llh_s >= dir_logprob(U) and llh_s >= dir_logprob(L):
                    return False
            return True

        upper = sigma * npr.rand()
        lower = upper - sigma
        llh_s = np.log(npr.rand()) + dir_logprob(0.0)
        l_steps_out = 0
        u_steps_out = 0
        if step_out:
            if doubling_step:
                while (dir_logprob(lower) > llh_s or dir_logprob(upper) > llh_s) and \
                        (l_steps_out + u_steps_out) < max_steps_out:
                    if npr.rand() < 0.5:
                        l_steps_out += 1
                        lower -= (upper - lower)
                    else:
                        u_steps_out += 1
                        upper += (upper - lower)
            else:
                while dir_logprob(lower) > llh_s and l_steps_out < max_steps_out:
                    l_steps_out += 1
                    lower -= sigma
                while dir_logprob(upper) > llh_s and u_steps_out < max_steps_out:
                    u_steps_out += 1
                    upper += sigma
        start_upper, start_lower = upper, lower
        while True:
            new_z = (upper - lower) * npr.rand() + lower
            new_llh = dir_logprob(new_z)
            if np.isnan(new_llh):
                raise Exception("Slice sampler got a NaN")
            if new_llh > llh_s and acceptable(new_z, llh_s, start_lower, start_upper):
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
        npr.shuffle(ordering)
        new_x = init_x.copy()
        for d in ordering:
            direction = np.zeros((dims))
            direction[d] = 1.0
            new_x, new_llh = direction_slice(direction, new_x)
    else:
        new_x = init_x
        for d in range(numdir):
            direction = npr.randn(dims)
            direction = direction / np.sqrt(np.sum(direction ** 2))
            new_x, new_llh = direction_slice(direction, new_x)
    if scalar:
        return float(new_x[0]), new_llh
    return new_x, new_llh


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
