#!/usr/bin/env python
"""Generate tests/golden/ref_vectors.npz from the REFERENCE's own header-only arithmetic.

Run in the development container (needs /root/reference):

    make -C oracle ref && python tests/golden/make_ref_vectors.py

`oracle/_ref/libmts_ref.so` (oracle/ref_harness.cpp compiled against the reference headers where they lie) evaluates the
case table `tests/helpers.py:math_cases`; the outputs are stored and later compared with the plain-C oracle
(tests/test_oracle_kat.py::test_oracle_matches_reference_vectors), also on machines without the reference tree.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H   # noqa: E402

ref = H.ref_math()
if ref is None:
    sys.exit("oracle/_ref/libmts_ref.so missing: run `make -C oracle ref` where /root/reference exists")
vec = H.math_cases(ref)
np.savez_compressed(os.path.join(HERE, "ref_vectors.npz"), **vec)
print("wrote", len(vec), "arrays,", sum(v.size for v in vec.values()), "values")
