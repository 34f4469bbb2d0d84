"""desi_mcmc_b200 -- B200-native (sm_100a) drop-in for CelestePy's
per-proposal image log-likelihood path (HIPS/DESI-MCMC).

Layout
  csrc/celeste_b200.cu   CUDA kernels + the C ABI of include/celeste_b200.h
  _ffi.py                ctypes binding of libceleste_b200.so
  engine.py              device buffers (torch) + engine calls
  compat/                the reference's Python call surface
                         (celeste_fast, gmm_like_fast, celeste, sources, ...)
  synth.py               seeded synthetic inputs of the BASELINE configs
  dist.py                field sharding + the per-sweep all-reduce

Importing the package does not load CUDA; the first engine call does, and it
raises if the shared library or a GPU is missing (no CPU fallback).
"""
__version__ = "0.1.0"
