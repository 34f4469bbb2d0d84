"""Worst relative error of the fused field kernel's model image and of render_kernel's
patches against the float64 oracle on a few seeded small fields (run on the GPU box)."""
import sys, numpy as np
sys.path.insert(0, ".")
from desi_mcmc_b200 import synth, engine, _ffi
from tests import helpers as hp
from tests.test_gpu_parity import _check_patches
from oracle import ref_cpu as oracle
engine.init(0)
for rec in (1, 0):
    engine.set_option(_ffi.CG_OPT_ROW_RECURRENCE, rec)
    wm, wp = 0.0, 0.0
    for seed in (78, 301, 302, 303):
        field = synth.make_field(seed, H=150, W=170, n_src=10, star_frac=.5)
        zeros = [np.zeros((150, 170))] * 5
        gimgs = hp.gpu_images(engine, field, zeros)
        oimgs = hp.oracle_images(oracle, field, zeros)
        ll, models = engine.field_loglik(gimgs, hp.gpu_sources(engine, field), return_model=True)
        osrcs = hp.oracle_sources(oracle, field)
        for i, img in enumerate(oimgs):
            lam = oracle.gen_model_image(osrcs, img)
            wm = max(wm, hp.rel_err(models[i].cpu().numpy(), lam, 1.0))
        wp = max(wp, _check_patches(engine, oracle, field, _ffi.CG_ITEM_FLUX_BAND, False))
    print("recurrence", rec, "worst model-image error %.3g, worst patch error %.3g" % (wm, wp))
