"""TEST INFRASTRUCTURE ONLY -- runs the reference's OWN Python modules in place.

The L3 layer of the reference (CelestePy/celeste.py, celeste_galaxy_conditionals.py,
sources.py, fits_image.py, planck.py, util/...) is Python-2 source with imports of
packages that are absent here (autograd, fitsio, tractor, astropy, pyprind).  This
module imports those files FROM /root/reference, unmodified on disk, through a
meta-path finder that

  * applies a mechanical, syntax-only Python-2 -> 3 rewrite to the text it read
    (``print x`` statements, ``lambda(th):``, ``has_key``, ``xrange``, ``iteritems``,
    string / comma ``raise``, tabs, and the NumPy-1 aliases ``np.float`` / ``np.int`` / ``np.Inf``)
    -- no arithmetic is touched, nothing is written anywhere;
  * resolves the package's Python-2 implicit relative imports (``from util.dists import
    mog`` inside CelestePy) and the absolute ``CelestePy.x`` spelling to the same module;
  * supplies stand-ins for the absent third-party packages: ``autograd.numpy`` is NumPy,
    ``autograd.scipy.misc.logsumexp`` is ``scipy.special.logsumexp``, ``autograd.grad``
    raises; fitsio / tractor / astropy / pyprind / matplotlib are empty modules (the
    functions exercised never touch them);
  * binds ``util.like.gmm_like_fast`` and ``celeste_fast`` to the reference's own Cython
    modules compiled into oracle/_ref (oracle/build_ref.py).

It exists to PIN oracle/ref_cpu.py (the restatement the GPU tests check against) on
outputs of the reference itself: tests/golden/make_golden_l3.py generates
tests/golden/reference_l3.npz with it, tests/test_cpu_oracle_l3.py compares.  It only
works where /root/reference is mounted (the build container); the GPU box uses the
committed vectors.  Never imported by the product package.
"""
from __future__ import annotations

import importlib.abc
import importlib.machinery
import importlib.util
import os
import re
import sys
import types

REF_ROOT = os.environ.get("CELESTE_REFERENCE", "/root/reference")
PKG_DIR = os.path.join(REF_ROOT, "CelestePy")
PREFIX = "CelestePy"


def available() -> bool:
    return os.path.isdir(PKG_DIR)


_PRINT = re.compile(r"^(\s*)print\s+(?!\()(.*?)\s*$")
_PRINT_PAREN_PCT = re.compile(r"^(\s*)print\s+(\(.*)$")


def _comment_start(line: str) -> int:
    """index of the '#' that starts a comment (not one inside a string literal), or -1"""
    q = None
    i = 0
    while i < len(line):
        c = line[i]
        if q:
            if c == "\\":
                i += 1
            elif c == q:
                q = None
        elif c in "\"'":
            q = c
        elif c == "#":
            return i
        i += 1
    return -1


def py2to3(text: str) -> str:
    """Syntax-only rewrite, line by line (see the module docstring)."""
    out = []
    for line in text.expandtabs(8).split("\n"):
        cut = _comment_start(line)
        code, sep, comment = (line, "", "") if cut < 0 else (line[:cut], "", line[cut:])
        m = _PRINT.match(code)
        if m and not code.strip().startswith("print("):
            body = m.group(2)
            if body.endswith(","):
                body = body[:-1]
            code = "%sprint(%s)" % (m.group(1), body)
        elif code.strip() == "print":
            code = code.replace("print", "print()")
        code = re.sub(r"lambda\s*\(\s*(\w+)\s*\)\s*:", r"lambda \1:", code)
        code = re.sub(r"\.has_key\(", ".__contains__(", code)
        code = re.sub(r"^(\s*\w+\s*=\s*)range\((.*)\)\s*$", r"\1list(range(\2))", code)   # a list in Python 2 (2to3 does the same)
        code = re.sub(r"\bxrange\b", "range", code)
        code = re.sub(r"\.iteritems\(\)", ".items()", code)
        code = re.sub(r"\.itervalues\(\)", ".values()", code)
        code = re.sub(r"^(\s*)raise\s+(\w+)\s*,\s*(.+)$", r"\1raise \2(\3)", code)
        code = re.sub(r"^(\s*)raise\s+(\"[^\"]*\")\s*$", r"\1raise Exception(\2)", code)
        code = re.sub(r"\bnp\.float\b(?![\w])", "float", code)
        code = re.sub(r"\bnp\.int\b(?![\w])", "int", code)
        code = re.sub(r"\bnp\.Inf\b", "np.inf", code)
        out.append(code + sep + comment)
    return "\n".join(out)


def _stub(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []          # behaves as a package: any submodule import finds another stub
    return m


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    """Empty modules for the third-party packages the exercised code never calls."""
    ROOTS = ("fitsio", "tractor", "astropy", "astrometry", "pyprind", "matplotlib", "seaborn",
             "pyfits", "rtree", "statsmodels", "line_profiler", "mpl_toolkits", "sklearn_stub")

    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] in self.ROOTS:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        return _stub(spec.name)

    def exec_module(self, module):
        def _missing(name):
            if name.startswith("__"):
                raise AttributeError(name)
            return _stub(module.__name__ + "." + name)
        module.__getattr__ = _missing


class _RefFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    """CelestePy.* and the package's Python-2 implicit relative spellings."""

    _siblings = None

    def _sibling(self, name):
        """Python-2 sibling import inside a sub-package (``from mog import ...`` in
        util/dists/flux_prior.py): the unique file of that name below CelestePy/."""
        if _RefFinder._siblings is None:
            idx = {}
            for d, _, files in os.walk(PKG_DIR):
                for f in files:
                    if f.endswith(".py") and f != "__init__.py":
                        idx.setdefault(f[:-3], []).append(os.path.join(d, f))
            _RefFinder._siblings = idx
        hits = _RefFinder._siblings.get(name, [])
        return hits[0] if len(hits) == 1 else None

    def _locate(self, fullname):
        parts = fullname.split(".")
        if parts[0] == PREFIX:
            parts = parts[1:]
        if not parts:
            return PKG_DIR, True
        if ".".join(parts) in _OVERRIDES:
            return os.path.join(PKG_DIR, *parts) + ".<compiled in oracle/_ref>", False
        base = os.path.join(PKG_DIR, *parts)
        if os.path.isdir(base) and os.path.exists(os.path.join(base, "__init__.py")):
            return base, True
        if os.path.exists(base + ".py"):
            return base + ".py", False
        return None, False

    def find_spec(self, fullname, path=None, target=None):
        top = fullname.split(".")[0]
        if top in ("test",):
            return None
        if top != PREFIX and not os.path.exists(os.path.join(PKG_DIR, top)) \
                and not os.path.exists(os.path.join(PKG_DIR, top + ".py")) and top not in _OVERRIDES:
            sib = self._sibling(fullname) if "." not in fullname and fullname not in sys.builtin_module_names else None
            if sib is None:
                return None
            rel = os.path.relpath(sib, PKG_DIR)[:-3].replace(os.sep, ".")
            return importlib.machinery.ModuleSpec(fullname, _Alias(PREFIX + "." + rel))
        loc, is_pkg = self._locate(fullname)
        if loc is None:
            return None
        spec = importlib.machinery.ModuleSpec(fullname, self, is_package=is_pkg,
                                              origin=loc if not is_pkg else os.path.join(loc, "__init__.py"))
        if is_pkg:
            spec.submodule_search_locations = [loc]
        return spec

    def create_module(self, spec):
        # one module object per file, whatever it is imported as
        canon = spec.name if spec.name.startswith(PREFIX) else PREFIX + "." + spec.name
        other = spec.name[len(PREFIX) + 1:] if spec.name.startswith(PREFIX + ".") else canon
        for alias in (canon, other):
            if alias in sys.modules and alias != spec.name:
                return sys.modules[alias]
        return None

    def exec_module(self, module):
        if getattr(module, "_ref_loaded", False):
            return
        module._ref_loaded = True
        spec = module.__spec__
        name = spec.name[len(PREFIX) + 1:] if spec.name.startswith(PREFIX + ".") else spec.name
        override = _OVERRIDES.get(name)
        if override is not None:
            module.__dict__.update(override())
            return
        with open(spec.origin) as f:
            src = py2to3(f.read())
        module.__file__ = spec.origin
        exec(compile(src, spec.origin, "exec"), module.__dict__)


class _Alias(importlib.abc.Loader):
    """bare sibling name -> the canonical CelestePy.x.y module object"""

    def __init__(self, canon):
        self.canon = canon

    def create_module(self, spec):
        return importlib.import_module(self.canon)

    def exec_module(self, module):
        pass


def _compiled(which):
    def load():
        from . import build_ref
        gm, cf = build_ref.load()
        mod = gm if which == "gmm" else cf
        if mod is None:
            raise ImportError("oracle/_ref is not built")
        return {k: getattr(mod, k) for k in dir(mod) if not k.startswith("__")}
    return load


def _like_pkg():
    # util/like/__init__.py star-imports two sibling modules that need autograd internals;
    # only its gmm_like_2d wrapper (:7-11) is on the path
    from . import build_ref
    gm, _ = build_ref.load()
    import numpy as np

    def gmm_like_2d(x, ws, mus, sigs, probs=None):           # util/like/__init__.py:7-11
        if probs is None:
            probs = np.zeros(x.shape[0], dtype=float)
        gm.gmm_like_2d(probs, x, ws, mus, sigs)
        return probs
    return {"gmm_like_2d": gmm_like_2d, "__path__": [os.path.join(PKG_DIR, "util", "like")]}


def _empty():
    # plotting / timing helpers (multi-line Python-2 print statements, matplotlib): not on the path
    def _unused(*a, **k):
        raise NotImplementedError("plot / timer utility stubbed out by oracle/ref_import.py")
    return {"compare_to_model": _unused, "subplot_imshow_colorbar": _unused, "add_colorbar_to_axis": _unused}


_OVERRIDES = {
    "util.misc.plot_util": _empty,
    "util.misc.timer_util": _empty,
    "util.like": _like_pkg,
    "util.like.gmm_like_fast": _compiled("gmm"),
    "celeste_fast": _compiled("cf"),
}

_installed = False


def install():
    """Idempotent: register the finders and the autograd stand-ins."""
    global _installed
    if _installed:
        return
    if not available():
        raise ImportError("%s is not mounted" % PKG_DIR)
    import numpy
    import numpy.linalg
    import scipy.special

    def _no_grad(*a, **k):
        raise NotImplementedError("autograd is not installed; gradients are outside the pinned path")

    anp = _stub("autograd.numpy")
    anp.__dict__.update({k: getattr(numpy, k) for k in dir(numpy) if not k.startswith("__")})
    anp.float, anp.int = float, int
    anp.linalg = numpy.linalg
    asm = _stub("autograd.scipy.misc", logsumexp=scipy.special.logsumexp)
    asp = _stub("autograd.scipy", misc=asm)
    ag = _stub("autograd", numpy=anp, scipy=asp, grad=_no_grad, elementwise_grad=_no_grad,
               jacobian=_no_grad, hessian=_no_grad)
    import numpy.random
    anp.random = numpy.random
    sys.modules.update({"autograd": ag, "autograd.numpy": anp, "autograd.numpy.linalg": numpy.linalg,
                        "autograd.numpy.random": numpy.random,
                        "autograd.scipy": asp, "autograd.scipy.misc": asm,
                        "autograd.util": _stub("autograd.util", quick_grad_check=_no_grad, check_grads=_no_grad)})
    from . import build_ref
    build_ref.load()             # the compiled Cython modules enter sys.modules before the finder can shadow them
    sys.meta_path.insert(0, _RefFinder())
    sys.meta_path.append(_StubFinder())
    _installed = True


def load(name: str):
    """load("celeste") -> the reference's CelestePy/celeste.py as a module."""
    install()
    return importlib.import_module(PREFIX + "." + name)


def load_path(relpath: str, name: str, preset: dict | None = None):
    """A reference file outside the CelestePy package (experiments/...), same rewrite.
    preset: module globals the file expects from its author's machine (site paths)."""
    install()
    path = os.path.join(REF_ROOT, relpath)
    mod = types.ModuleType(name)
    mod.__file__ = path
    mod.__dict__.update(preset or {})
    with open(path) as f:
        src = py2to3(f.read())
    exec(compile(src, path, "exec"), mod.__dict__)
    return mod
