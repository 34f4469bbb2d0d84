"""Generates tests/golden/reference_vectors.npz from the COMPILED reference
(oracle/_ref: gmm_like_fast / celeste_fast built from /root/reference by
oracle/build_ref.py, and the vendored randomkit).  Run in the build container:
    python tests/golden/make_golden.py
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import build_ref, ref_cpu as ref           # noqa: E402
from tests.test_gpu_parity import kat_inputs          # noqa: E402

assert build_ref.build(), "needs /root/reference"
gm, cf = build_ref.load()

X, ws, means, covs = kat_inputs()
probs = np.zeros(X.shape[0])
gm.gmm_like_2d(probs, X, ws, means, covs)
idx = np.arange(0, X.shape[0], 7)

X3, ws3, means3, covs3 = kat_inputs(K=3, n=100)
ci = np.linalg.inv(covs3[2])
_, logdet = np.linalg.slogdet(covs3[2])
lls = np.zeros(X3.shape[0])
gm.multivariate_normal_2d_covinv_logdet_logpdf(lls, X3, means3[2], ci, logdet)

rng = np.random.RandomState(3)
W = rng.normal(size=(2, 2)); W = W.dot(W.T)
v = rng.uniform(0, 50, size=2)
iw = rng.dirichlet(np.ones(3)); im = rng.normal(size=(3, 2))
ic = np.array([np.eye(2) * s + .1 for s in (2., 7., 30.)])
pw, pm, pc = cf.gen_galaxy_prof_psf_mixture_params(W, v, iw, im, ic, ref.dev_amp, ref.dev_var)
th = np.array([.3, .7])
jw, jm, jc = cf.gen_galaxy_psf_mixture_params(th, W, v, iw, im, ic, ref.exp_amp, ref.exp_var,
                                              ref.dev_amp, ref.dev_var)

lib = C.CDLL(os.path.join(build_ref.OUT, "librandomkit.so"))
state = C.create_string_buffer(8192)
lib.rk_seed.argtypes = [C.c_ulong, C.c_void_p]
lib.rk_double.restype = C.c_double; lib.rk_double.argtypes = [C.c_void_p]
lib.rk_binomial.restype = C.c_long; lib.rk_binomial.argtypes = [C.c_void_p, C.c_long, C.c_double]
lib.rk_seed(89, state)
rk_doubles = np.array([lib.rk_double(state) for _ in range(16)])
lib.rk_seed(0, state)
rk_binom = np.array([lib.rk_binomial(state, 914, 6.29379e-16)] +
                    [lib.rk_binomial(state, 1000, .3) for _ in range(7)])

np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"),
                    kat_idx=idx, kat_probs=probs[idx], mvn_lls=lls[::5],
                    mp_W=W, mp_v=v, mp_iw=iw, mp_im=im, mp_ic=ic, mp_th=th,
                    prof_w=np.asarray(pw), prof_m=np.asarray(pm), prof_c=np.asarray(pc),
                    joint_w=np.asarray(jw), joint_m=np.asarray(jm), joint_c=np.asarray(jc),
                    rk_doubles=rk_doubles, rk_binom=rk_binom)
print("wrote reference_vectors.npz", probs[idx].size, "KAT points")
