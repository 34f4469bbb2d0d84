"""ctypes binding of libceleste_b200.so (include/celeste_b200.h).

The library is the product; there is no CPU fallback.  Importing this module
without the built shared object raises, and every call checks the status code
and re-raises the library's message (ValueError for CG_EINVAL, as the
reference's Cython raises ValueError on shape mismatches,
CelestePy/util/like/gmm_like_fast.pyx:146-149; RuntimeError otherwise).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libceleste_b200_checked.so" if os.environ.get("CG_CHECKED") == "1"
                        else "libceleste_b200.so")     # CG_CHECKED=1: the build with bounds asserts in the kernels
if os.environ.get("CG_LIB_VARIANT"):                   # kernel experiments: libceleste_b200_<variant>.so (tools/gpu_variant.sh)
    LIB_PATH = os.path.join(PKG, "libceleste_b200_%s.so" % os.environ["CG_LIB_VARIANT"])

CG_ABI_VERSION = 2
CG_MAX_PSF = 4
CG_NUM_BANDS = 5
CG_MAX_COMP = 56
CG_MAX_IMAGES = 64

CG_OK, CG_EINVAL, CG_ECUDA, CG_ENOMEM, CG_ESTATE = 0, -1, -2, -3, -4

CG_SRC_PLANCK = 1
CG_SRC_EIGEN = 2

CG_ITEM_FLUX_BAND = 0x00
CG_ITEM_FLUX_UNIT = 0x01
CG_ITEM_FIXED_BOX = 0x04
CG_ITEM_TRUNC_BOX = 0x08
CG_ITEM_PROF_EXP = 0x10
CG_ITEM_PROF_DEV = 0x20
CG_ITEM_HEADER_CD = 0x40
CG_ITEM_NO_MASK = 0x80

CG_WS_RENDER, CG_WS_FIELD, CG_WS_ITEMS, CG_WS_SCORE = 0, 1, 2, 3
CG_WS_SWEEP_ALLOC, CG_WS_SWEEP_SCORE, CG_WS_SWEEP_BACKGROUND = 4, 5, 6

CG_LL_POISSON_BG = 0
CG_LL_CONDITIONAL = 1
CG_LL_CONDITIONAL_SUM = 2


class CgImageDesc(C.Structure):
    _fields_ = [
        ("H", C.c_int32), ("W", C.c_int32), ("band", C.c_int32), ("n_psf", C.c_int32),
        ("epsilon", C.c_double), ("kappa", C.c_double), ("calib", C.c_double),
        ("psf_w", C.c_double * CG_MAX_PSF),
        ("psf_mu", (C.c_double * 2) * CG_MAX_PSF),
        ("psf_cov", (C.c_double * 3) * CG_MAX_PSF),
        ("rho_n", C.c_double * 2), ("phi_n", C.c_double * 2),
        ("ups", C.c_double * 4),
        ("d_nelec", C.c_void_p), ("d_mask", C.c_void_p),
    ]


class CgBox(C.Structure):
    _fields_ = [("x0", C.c_int32), ("x1", C.c_int32), ("y0", C.c_int32), ("y1", C.c_int32)]


CG_OPT_CULL = 1
CG_OPT_ROW_RECURRENCE = 2
CG_MAX_FIELDS = 16

# numpy mirrors of the POD arrays that cross the boundary in bulk
SRC_DTYPE = np.dtype([("type", "<i4"), ("flags", "<i4"), ("u", "<f8", (2,)),
                      ("flux", "<f8", (CG_NUM_BANDS,)), ("theta", "<f8"), ("sigma", "<f8"),
                      ("phi", "<f8"), ("rho", "<f8"), ("t", "<f8"), ("b", "<f8"), ("flux_row", "<i8")])
ITEM_DTYPE = np.dtype([("src", "<i4"), ("img", "<i4"), ("flags", "<i4"),
                       ("box", "<i4", (4,)), ("aux", "<i4")])
BOX_DTYPE = np.dtype([("x0", "<i4"), ("x1", "<i4"), ("y0", "<i4"), ("y1", "<i4")])
assert SRC_DTYPE.itemsize == 120 and ITEM_DTYPE.itemsize == 32 and BOX_DTYPE.itemsize == 16

_vp, _i32, _i64, _dbl, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_double, C.c_size_t
_dp = C.POINTER(C.c_double)

# name -> (restype, argtypes); every symbol include/celeste_b200.h declares
PROTOTYPES = {
    "cg_abi_version": (C.c_int, []),
    "cg_init": (C.c_int, [C.c_int]),
    "cg_shutdown": (C.c_int, []),
    "cg_last_error": (C.c_int, [C.c_char_p, _sz]),
    "cg_sm_count": (C.c_int, []),
    "cg_image_create": (C.c_int, [C.POINTER(CgImageDesc), C.POINTER(_vp)]),
    "cg_image_destroy": (C.c_int, [_vp]),
    "cg_image_set_epsilon": (C.c_int, [_vp, _dbl]),
    "cg_image_radius": (_dbl, [_vp]),
    "cg_set_filter_curves": (C.c_int, [_vp, _vp, _vp, _vp]),
    "cg_set_flux_table": (C.c_int, [_vp, _i64]),
    "cg_gmm_like_2d": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i32, _vp]),
    "cg_mvn2d_logpdf": (C.c_int, [_vp, _vp, _vp, _vp, _dbl, _i64, _vp]),
    "cg_galaxy_prof_psf_mixture_params": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32,
                                                    _vp, _vp, _vp]),
    "cg_galaxy_psf_mixture_params": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32,
                                               _vp, _vp, _i32, _vp, _vp, _vp]),
    "cg_calc_bounding_radius": (_dbl, [_vp, _vp, _i32, _dbl, _vp]),
    "cg_workspace_bytes": (_sz, [_i32, _i64, _i32, _i64, _i64]),
    "cg_field_tiles": (_i32, [_i32, _i32]),
    "cg_render_patches": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _i64, _i32, _dbl, _i64, _vp, _vp,
                                    _vp, _sz, _vp]),
    "cg_field_loglik": (C.c_int, [_vp, _i32, _vp, _i32, _dbl, _i64, _vp, _vp, _vp, _sz, _vp]),
    "cg_fields_loglik": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _dbl, _i64, _vp, _vp, _vp, _sz, _vp]),
    "cg_set_option": (C.c_int, [_i32, _i32]),
    "cg_items_loglik": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _i64, _i32, _dbl, _i64, _vp, _vp,
                                  _vp, _vp, _i32, _vp, _sz, _vp]),
    "cg_score_proposals": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _vp, _vp, _vp, _vp, _dbl,
                                     _vp, _vp, _vp, _vp, _sz, _vp]),
    "cg_poisson_loglike": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "cg_poisson_loglike_static": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "cg_alloc_workspace_bytes": (_sz, [_i32, _i64]),
    "cg_sample_source_counts": (C.c_int, [_vp, _i32, _i32, _vp, _vp, _i64, C.c_uint64, _vp, _vp, _vp, _sz, _vp]),
    "cg_binomial_draws": (C.c_int, [_i64, _dbl, C.c_uint64, _i64, _vp, _vp]),
    "cg_sweep_workspace_bytes": (_sz, [_i32, _i32, _i32, _i32, _i64, _i64]),
    "cg_field_sample_photons": (C.c_int, [_vp, _i32, _vp, _i32, _dbl, _i64, C.c_uint64, _vp, _vp, _vp, _vp, _vp,
                                          _vp, _sz, _vp]),
    "cg_resample_fluxes": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _vp, _dbl, _dbl, C.c_uint64, _vp]),
    "cg_resample_epsilon": (C.c_int, [_vp, _i32, _vp, _dbl, _dbl, C.c_uint64, _vp, _vp]),
    "cg_image_sync_epsilon": (C.c_int, [_vp, _vp, _dp]),
    "cg_make_proposals": (C.c_int, [_vp, _i32, _i32, _vp, C.c_uint64, _vp, _vp]),
    "cg_sweep_guard_cap": (_i32, [_i32, _i32]),
    "cg_score_proposals_multi": (C.c_int, [_vp, _i32, _vp, _i32, _i32, _vp, _i32, _vp, _i64, _vp, _vp, _dbl,
                                           _vp, _vp, _vp, _vp, _sz, _vp]),
    "cg_select_proposals": (C.c_int, [_vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    "cg_background_patches": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _vp, _vp, _i64, _i32, _vp, _vp, _sz, _vp]),
    "cg_planck_fluxes": (C.c_int, [_vp, _vp, _i64, _vp, _vp]),
    "cg_eigen_fluxes": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "cg_mog_logpdf": (C.c_int, [_vp, _i64, _i32, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp]),
    "cg_prior_stats": (C.c_int, [_vp, _vp, _i64, _i32, _i32, _vp, _vp]),
    "cg_slice_sample_locations": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _vp, _i64, _vp, _vp, _vp, _i32, _dbl, _dbl,
                                            _vp, _vp, _vp, _vp, _vp, _vp]),
    "cg_hmc_galaxy_params": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _i32, _vp, _vp, _i64, _vp, _vp, _vp, _dbl, _vp, _i32,
                                       _vp, _vp, _vp, _vp, _vp]),
    "cg_flux_prior_scratch_bytes": (_sz, [_i64, _i32]),
    "cg_flux_prior_stats": (C.c_int, [_vp, _i64, _i32, _vp, _vp, _vp, _vp, _i32, _vp, _sz, _vp, _vp, _vp]),
    "cg_sweep_stats": (C.c_int, [_vp, _i32, _vp, _vp, _i32, _vp, _i64, _vp, _vp, _i32, _vp, _vp]),
    "cg_launch_count": (_i64, [C.c_int]),
    "cg_last_eval_count": (_i64, [_vp]),
    "cg_last_issued_count": (_i64, [_vp]),
    "cg_profile_render": (C.c_int, [C.c_int]),
    "cg_profile_render_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "cg_pipe_probe": (C.c_int, [C.c_int, C.POINTER(C.c_double), _vp]),
}

_lib = None


def load(build_if_missing: bool = False):
    """dlopen the library and bind every prototype.  Does not touch CUDA."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if build_if_missing:
            from . import build as _b
            _b.build()
        else:
            raise RuntimeError(
                "%s is missing: build it with `python -m desi_mcmc_b200.build` "
                "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)      # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.cg_abi_version() != CG_ABI_VERSION:
        raise RuntimeError("libceleste_b200.so ABI %d != binding ABI %d"
                           % (lib.cg_abi_version(), CG_ABI_VERSION))
    _lib = lib
    return lib


def last_error() -> str:
    buf = C.create_string_buffer(1024)
    load().cg_last_error(buf, 1024)
    return buf.value.decode("utf-8", "replace")


def check(rc: int) -> None:
    if rc == CG_OK:
        return
    msg = last_error()
    if rc == CG_EINVAL:
        raise ValueError(msg)
    if rc == CG_ENOMEM:
        raise MemoryError(msg)
    raise RuntimeError("libceleste_b200 error %d: %s" % (rc, msg))


def np_ptr(a: np.ndarray) -> int:
    return a.ctypes.data
