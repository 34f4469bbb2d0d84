"""CelestePy/util/infer/hmc.py:16-98: Hamiltonian Monte Carlo (leapfrog)
control flow on the host, drawing from numpy.random in the reference's order so
a seeded chain reproduces.  `llhfunc` / `grad_llhfunc` are device evaluations
(e.g. celeste_galaxy_conditionals.galaxy_source_like / _grad, whose ten
finite-difference renders go to the GPU as one batch)."""
from __future__ import annotations

import numpy as np
import numpy.random as npr


def hmc(x_curr, llhfunc, grad_llhfunc, eps, num_steps, mass=None, num_iter=1, p_curr=None,
        refresh_alpha=0.0, adaptive_step_sz=False, tgt_accept_rate=0.9, avg_accept_time_const=0.95,
        avg_accept_rate=0.9, min_step_sz=0.00005, max_step_sz=1.0, negative_log_prob=True):
    new_accept_rate = 0.
    if mass is None:
        mass = np.ones(np.asarray(x_curr).shape[0])
    imass = 1. / mass

    def energy(X):
        return -llhfunc(X)

    def grad_energy(X):
        return -grad_llhfunc(X)

    def hamiltonian(X, P):
        return energy(X) + .5 * np.sum(imass * P * P)

    def leapstep(xx0, pp0):                                   # :52-59
        xx, pp = xx0.copy(), pp0.copy()
        pph = pp - .5 * eps * grad_energy(xx)
        for l in range(num_steps):
            xx = xx + eps * imass * pph
            eps_mom = .5 * eps if l == num_steps - 1 else eps
            pph = pph - eps_mom * grad_energy(xx)
        return xx, pph

    X = x_curr.copy()
    if p_curr is None:
        P = np.sqrt(mass) * npr.randn(X.shape[0])
    else:
        P = p_curr.copy()
    ll_curr = -hamiltonian(X, P)
    for i in range(num_iter):
        P = refresh_alpha * P + np.sqrt(1.0 - refresh_alpha ** 2) * np.sqrt(mass) * npr.randn(X.shape[0])
        Xp, Pp = leapstep(X, P)
        Pp = -Pp
        ll_prop = -hamiltonian(Xp, Pp)
        accept = np.log(npr.rand()) < ll_prop - ll_curr
        if accept:
            X, P, ll_curr = Xp, Pp, ll_prop
        P = -P
        if adaptive_step_sz:
            new_accept_rate = avg_accept_time_const * avg_accept_rate + (1.0 - avg_accept_time_const) * accept
            if avg_accept_rate > tgt_accept_rate:
                eps = eps * 1.02
            else:
                eps = eps * 0.98
            eps = np.clip(eps, min_step_sz, max_step_sz)
    return X, P, eps, new_accept_rate
