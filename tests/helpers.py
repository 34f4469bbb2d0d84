"""Builds oracle-side and product-side objects from the same synthetic dicts."""
import numpy as np

from desi_mcmc_b200 import synth


def oracle_images(ref, field, nelecs, invvars=None):
    out = []
    for i, m in enumerate(field["images"]):
        out.append(ref.SyntheticImage(nelec=nelecs[i], invvar=None if invvars is None else invvars[i],
                                      **synth.image_kwargs(m)))
    return out


def oracle_sources(ref, field_or_list):
    lst = field_or_list["sources"] if isinstance(field_or_list, dict) else field_or_list
    return [ref.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"], phi=s["phi"],
                          sigma=s["sigma"], rho=s["rho"], fluxes=s["fluxes"]) for s in lst]


def gpu_images(eng, field, nelecs, invvars=None):
    out = []
    for i, m in enumerate(field["images"]):
        out.append(eng.Image(nelec=nelecs[i], invvar=None if invvars is None else invvars[i],
                             **synth.image_kwargs(m)))
    return out


def gpu_sources(eng, field_or_list):
    lst = field_or_list["sources"] if isinstance(field_or_list, dict) else field_or_list
    return [eng.SrcParams(u=s["u"], a=s["a"], b=s["b"], t=s["t"], theta=s["theta"], phi=s["phi"],
                          sigma=s["sigma"], rho=s["rho"], fluxes=s["fluxes"]) for s in lst]


def rel_err(a, b, floor):
    """max |a-b| / max(|b|, floor)"""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor))) if a.size else 0.0


def simulate_nelec(ref, field, seed):
    """nelec ~ Poisson(model) with the oracle's model image (small fields only)."""
    rng = np.random.RandomState(seed)
    srcs = oracle_sources(ref, field)
    out = []
    for m in field["images"]:
        H, W = m["shape"]
        img = ref.SyntheticImage(nelec=np.zeros((H, W)), **synth.image_kwargs(m))
        lam = ref.gen_model_image(srcs, img)
        out.append(rng.poisson(lam).astype(np.float64))
    return out
