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
