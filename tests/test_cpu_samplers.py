"""Host control flow of the samplers (util/infer/hmc.py, slicesample.py): the
loops run on the CPU around device likelihoods; here they are driven by
closed-form densities, seeded, and checked for the stationary distribution."""
import numpy as np

from desi_mcmc_b200.compat.hmc import hmc
from desi_mcmc_b200.compat.slicesample import slicesample, slicesample_step_out


def test_slicesample_step_out_stationary_gaussian():
    np.random.seed(0)
    mu, sd = np.array([1., -2.]), np.array([.5, 2.])
    llh = lambda x: float(-.5 * np.sum(((x - mu) / sd) ** 2))
    x, xs = np.array([0., 0.]), []
    for _ in range(1500):
        x, l = slicesample_step_out(x, llh, step=[1., 3.], step_out=True)
        assert l == llh(x)
        xs.append(x.copy())
    xs = np.array(xs[200:])
    assert np.all(np.abs(xs.mean(0) - mu) < 4 * sd / np.sqrt(len(xs) / 4))
    assert np.all(np.abs(xs.std(0) / sd - 1) < .15)


def test_slicesample_step_out_respects_bounds_and_seed():
    llh = lambda x: float(-np.sum(x))                 # exponential on [0, 5]
    np.random.seed(4)
    a = slicesample_step_out(np.array([1.]), llh, step=2., lb=np.array([0.]), ub=np.array([5.]))
    np.random.seed(4)
    b = slicesample_step_out(np.array([1.]), llh, step=2., lb=np.array([0.]), ub=np.array([5.]))
    assert a[0][0] == b[0][0] and 0. <= a[0][0] <= 5.
    try:
        slicesample_step_out(np.array([6.]), llh, lb=np.array([0.]), ub=np.array([5.]))
        assert False
    except AssertionError:
        pass


def test_hmc_stationary_and_reversible():
    np.random.seed(1)
    prec = np.array([4., .25])
    llh = lambda x: float(-.5 * np.sum(prec * x * x))
    grad = lambda x: -prec * x
    x, p, xs = np.zeros(2), None, []
    for _ in range(800):
        x, p, eps, _ = hmc(x, llh, grad, eps=.15, num_steps=12, mass=np.ones(2), p_curr=p, refresh_alpha=.3)
        xs.append(x.copy())
    xs = np.array(xs[100:])
    assert np.all(np.abs(xs.std(0) * np.sqrt(prec) - 1) < .2)
    # an exactly energy-conserving problem is always accepted; momentum is negated twice (:66-80)
    x1, p1, _, _ = hmc(np.array([.3, -.2]), lambda x: 0., lambda x: np.zeros(2), eps=.1, num_steps=5,
                       mass=np.ones(2), p_curr=np.array([1., 2.]), refresh_alpha=1.0)
    assert np.allclose(x1, np.array([.3, -.2]) + .5 * np.array([1., 2.])) and np.allclose(p1, [1., 2.])


def test_slicesample_direction_sampler_seeded():
    llh = lambda x: float(-.5 * np.sum(x * x))
    np.random.seed(7)
    a, la = slicesample(np.array([.1, .2]), llh)
    np.random.seed(7)
    b, lb = slicesample(np.array([.1, .2]), llh)
    assert np.array_equal(a, b) and la == lb == llh(a)


def test_lockstep_equals_sequential_chains():
    """run_lockstep (one batched evaluation per round for all chains) gives
    every chain exactly the trace it has when driven alone with the same RNG."""
    from desi_mcmc_b200.compat.slicesample import run_lockstep, slicesample_gen
    mu, sd = np.array([.3, -1., 2.]), np.array([.2, 1.5, .7])
    f = lambda x: float(-.5 * np.sum(((x - mu) / sd) ** 2))
    calls = []

    def batch(idx, xs):
        calls.append(len(idx))
        return [f(x) for x in xs]

    for kw in (dict(), dict(compwise=False, numdir=3), dict(step_out=False, upper_bound=np.full(3, 9.),
                                                          lower_bound=np.full(3, -9.))):
        gens = [slicesample_gen(np.zeros(3), rng=np.random.RandomState(10 + i), **kw) for i in range(9)]
        res = run_lockstep(gens, batch)
        for i in range(9):
            x, l = slicesample(np.zeros(3), f, rng=np.random.RandomState(10 + i), **kw)
            assert np.array_equal(x, res[i][0]) and l == res[i][1]
    assert max(calls) == 9 and len(calls) < sum(calls) / 3      # rounds are shared by the chains
