#!/usr/bin/env python
"""Copies run-time DATA files the benchmark workload names from the reference checkout into mitsuba2_b200/data/
(git-ignored; travels to the GPU box with the snapshot): resources/data/envmap/museum.exr -> museum_envmap.npy
(float32 RGB, decoded with OpenCV), the environment map SURVEY 8(d) specifies for C3.  Where the file is absent the
scenes fall back to a synthetic sky and say so in the bench line's `config.envmap`."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("REF", "/root/reference")

if __name__ == "__main__":
    src = os.path.join(REF, "resources", "data", "envmap", "museum.exr")
    if not os.path.exists(src):
        sys.exit(f"{src} not present")
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    import cv2
    im = cv2.imread(src, cv2.IMREAD_UNCHANGED)
    rgb = np.ascontiguousarray(im[:, :, 2::-1], np.float32)
    out = os.path.join(ROOT, "mitsuba2_b200", "data", "museum_envmap.npy")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    np.save(out, rgb)
    print(out, rgb.shape, float(rgb.mean()))
