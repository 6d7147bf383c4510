#!/usr/bin/env python
"""Generate tests/golden/rgb2spec_ref16.npz from the REFERENCE's own sRGB->spectrum code.

    make -C oracle rgb2spec && python tests/golden/make_rgb2spec_vectors.py

`oracle/_ref/rgb2spec_opt` (ext/rgb2spec/rgb2spec_opt.cpp compiled where it lies) wrote a resolution-16 model table,
`oracle/_ref/librgb2spec.so` (ext/rgb2spec/rgb2spec.c) looks colours up in it: these are the two halves
`mitsuba2_b200/rgb2spec.py` restates.  Stored: the colours, the coefficients the reference fetches for them, and the
spectra those coefficients evaluate to (rgb2spec_eval_precise) at a fixed set of wavelengths.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle", "_ref")
lib = C.CDLL(os.path.join(REF, "librgb2spec.so"))
lib.rgb2spec_load.restype = C.c_void_p
lib.rgb2spec_fetch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
lib.rgb2spec_eval_precise.restype = C.c_float
lib.rgb2spec_eval_precise.argtypes = [C.c_void_p, C.c_float]
model = lib.rgb2spec_load(os.path.join(REF, "srgb_16.coeff").encode())
assert model
rng = np.random.default_rng(5)
rgb = np.concatenate([rng.random((200, 3)), 0.2 + 0.6 * rng.random((40, 3)),
                      [[0.5, 0.5, 0.5], [1, 0, 0], [0, 1, 0], [0, 0, 1], [0.9, 0.9, 0.1], [0.05, 0.02, 0.01], [0.5, 0.25, 0.125]]]).astype(np.float32)
lam = np.linspace(360, 830, 48).astype(np.float32)
coeff = np.zeros_like(rgb)
spec = np.zeros((rgb.shape[0], lam.size), np.float32)
for i in range(rgb.shape[0]):
    c = rgb[i].copy()
    lib.rgb2spec_fetch(model, c.ctypes.data, coeff[i].ctypes.data)
    for j, l in enumerate(lam):
        spec[i, j] = lib.rgb2spec_eval_precise(coeff[i].ctypes.data, float(l))
np.savez_compressed(os.path.join(HERE, "rgb2spec_ref16.npz"), rgb=rgb, coeff=coeff, wavelengths=lam, spectra=spec)
print("wrote", rgb.shape[0], "colours")
