"""Generates tests/golden/reference_l3.npz by RUNNING THE REFERENCE'S OWN PYTHON
(/root/reference/CelestePy/*.py, experiments/redshift/redshift_utils.py) through
oracle/ref_import.py on the seeded cases of tests/golden/l3_cases.py.  Run in the build
container (the reference is not on the GPU box):
    python tests/golden/make_golden_l3.py
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
warnings.simplefilter("ignore")

from oracle import build_ref, ref_import as ri           # noqa: E402
from tests.golden import l3_cases as L                    # noqa: E402

assert build_ref.build(), "needs /root/reference"
out = L.run(L.reference_backend())
ru = ri.load_path("experiments/redshift/redshift_utils.py", "redshift_utils",
                  dict(SPECTRA_LOC="", DR10QSO_LOC="", TRAIN_TEST_SPLIT_LOC="", EIGEN_FILE_LOC=""))
out.update(L.run_a17(ru.project_to_bands, ru.softmax))
out.update(L.sampler_traces(ri.load("util.infer.slicesample").slicesample, ri.load("util.infer.hmc").hmc))
np.savez_compressed(os.path.join(HERE, "reference_l3.npz"), **out)
print("wrote reference_l3.npz: %d arrays, %d values" % (len(out), sum(v.size for v in out.values())))
