"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Compiles the reference's own native code for the hot path, from the sources
where they lie under /root/reference, into ``oracle/_ref/`` (git-ignored,
travels to the GPU box with the snapshot).  Nothing is copied into the repo:
the Cython sources are streamed through a type-alias shim into a scratch
directory under /tmp, cythonized there, and only the resulting ``.so`` files
land in ``oracle/_ref/``.

What is built
  * ``gmm_like_fast``  <- CelestePy/util/like/gmm_like_fast.pyx   (A2, OpenMP)
  * ``celeste_fast``   <- CelestePy/celeste_fast.pyx              (A1)
  * ``librandomkit.so``<- deps/randomkit/{randomkit,distributions}.c (N1 RNG)

The shim changes no arithmetic.  NumPy 2 removed ``np.float``/``np.int`` and
the ``np.int_t``/``np.float_t``/``np.ulong_t`` Cython typedef names the
reference uses (celeste_fast.pyx:20-24, gmm_like_fast.pyx:22-26); the shim maps
them to the C types they always meant (long / double / unsigned long).
"""
from __future__ import annotations

import os
import re
import shutil
import subprocess
import sys
import sysconfig
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("CELESTE_REFERENCE", "/root/reference")

PYX = {
    "gmm_like_fast": "CelestePy/util/like/gmm_like_fast.pyx",
    "celeste_fast": "CelestePy/celeste_fast.pyx",
}
RANDOMKIT = ["deps/randomkit/randomkit.c", "deps/randomkit/distributions.c"]


def _shim(text: str) -> str:
    text = re.sub(r"^FLOAT\s*=\s*np\.float\s*$", "FLOAT = np.float64", text, flags=re.M)
    text = re.sub(r"^INT\s*=\s*np\.int\s*$", "INT = np.int64", text, flags=re.M)
    text = text.replace("ctypedef np.int_t INT_t", "ctypedef long INT_t")
    text = text.replace("ctypedef np.float_t FLOAT_t", "ctypedef double FLOAT_t")
    text = text.replace("ctypedef np.ulong_t INDEX_t", "ctypedef unsigned long INDEX_t")
    text = text.replace("dtype=np.float)", "dtype=np.float64)")
    return text


def reference_present() -> bool:
    return all(os.path.exists(os.path.join(REF, p)) for p in PYX.values())


def built() -> bool:
    if not os.path.isdir(OUT):
        return False
    names = os.listdir(OUT)
    return (any(n.startswith("gmm_like_fast") and n.endswith(".so") for n in names)
            and any(n.startswith("celeste_fast") and n.endswith(".so") for n in names))


def build(force: bool = False, verbose: bool = False) -> bool:
    """Build oracle/_ref from /root/reference.  Returns True when the compiled
    reference is available afterwards."""
    if built() and not force:
        return True
    if not reference_present():
        return built()
    import numpy as np
    os.makedirs(OUT, exist_ok=True)
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    pyinc = sysconfig.get_paths()["include"]
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    quiet = None if verbose else subprocess.DEVNULL
    with tempfile.TemporaryDirectory(prefix="celeste_ref_") as tmp:
        for mod, rel in PYX.items():
            with open(os.path.join(REF, rel)) as f:
                src = _shim(f.read())
            pyx = os.path.join(tmp, mod + ".pyx")
            with open(pyx, "w") as f:
                f.write(src)
            subprocess.check_call([sys.executable, "-m", "cython", "--cplus", "-3",
                                   "-o", os.path.join(tmp, mod + ".cpp"), pyx],
                                  stdout=quiet, stderr=quiet)
            flags = ["-O3", "-w", "-DNDEBUG", "-std=c++11", "-fPIC", "-shared",
                     "-DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION"]
            if mod == "gmm_like_fast":
                flags.append("-fopenmp")
            subprocess.check_call([cxx, *flags, "-I", pyinc, "-I", np.get_include(),
                                   os.path.join(tmp, mod + ".cpp"),
                                   "-o", os.path.join(OUT, mod + ext)],
                                  stdout=quiet, stderr=quiet)
        rk = [os.path.join(REF, p) for p in RANDOMKIT]
        if all(os.path.exists(p) for p in rk):
            subprocess.check_call([cc, "-O2", "-fPIC", "-shared", *rk,
                                   "-I", os.path.join(REF, "deps/randomkit"),
                                   "-o", os.path.join(OUT, "librandomkit.so"), "-lm"],
                                  stdout=quiet, stderr=quiet)
    return built()


def load():
    """Import the compiled reference modules (gmm_like_fast, celeste_fast) or
    return (None, None) when oracle/_ref is absent."""
    if not built():
        return None, None
    if OUT not in sys.path:
        sys.path.insert(0, OUT)
    import importlib
    return importlib.import_module("gmm_like_fast"), importlib.import_module("celeste_fast")


if __name__ == "__main__":
    ok = build(force="--force" in sys.argv, verbose=True)
    print("oracle/_ref built:", ok)
    sys.exit(0 if ok else 1)
