"""The reference's Python call surface (same names, argument meaning and error
behaviour), routed through libceleste_b200.so.  Host NumPy arrays in and out,
like the reference; each call uploads its inputs, runs the CUDA path and
downloads the result.  Batched engine calls (desi_mcmc_b200.engine) are the
fast path; these exist so reference-side callers run unchanged.

  gmm_like_fast   <- CelestePy/util/like/gmm_like_fast.pyx, util/like/__init__.py
  celeste_fast    <- CelestePy/celeste_fast.pyx
  mixture_profiles<- CelestePy/mixture_profiles.py (constants + accessors)
  bounding_box    <- CelestePy/util/bound/bounding_box.py
  celeste         <- CelestePy/celeste.py
  celeste_galaxy_conditionals <- CelestePy/celeste_galaxy_conditionals.py
  sources         <- CelestePy/sources.py (likelihood closures)
  planck          <- CelestePy/planck.py
  celeste_sample_sources <- CelestePy/celeste_sample_sources.pyx (device photon allocation)
  celeste_mcmc    <- CelestePy/celeste_mcmc.py (photon allocation entry point, star / galaxy conditionals)
  models          <- CelestePy/models.py (CelesteBase / Celeste / Field: the live Gibbs sweep)
  slicesample, hmc<- CelestePy/util/infer/{slicesample,hmc}.py (host control flow, device likelihoods)
"""
