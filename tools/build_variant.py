"""Builds libceleste_b200_<name>.so with extra -D flags (kernel experiments; loaded with CG_LIB_VARIANT=<name>).
usage: python tools/build_variant.py <name> [-DFOO=1 ...] [--ptxas]"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from desi_mcmc_b200 import build
name = sys.argv[1]
flags = [f for f in sys.argv[2:] if f.startswith("-D")]
out = os.path.join(build.PKG, "libceleste_b200_%s.so" % name)
cmd = [build._nvcc(), *build.NVCC_FLAGS, "-Xptxas", "-v", *flags, "-I", os.path.join(build.ROOT, "include"), build.SRC, "-o", out]
r = subprocess.run(cmd, capture_output=True, text=True)
if r.returncode:
    sys.exit(r.stdout + r.stderr)
L = (r.stdout + r.stderr).splitlines()
for i, l in enumerate(L):
    if "Compiling entry function '_Z17field_tile_kernel" in l or "Compiling entry function '_Z17hmc_galaxy_kernel" in l or "Compiling entry function '_Z14prepare_kernelIf" in l:
        print(L[i + 1].strip()); print(L[i + 2].strip()); print(L[i + 3].strip())
print(out)
