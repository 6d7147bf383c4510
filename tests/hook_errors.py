#!/usr/bin/env python
"""Measured float error of every stage hook of the C ABI against the oracle on shared inputs (the numbers behind the per-
function tolerances of tests/test_gpu_parity.py::test_stage_hooks_match_oracle).  GPU only; prints one line per function."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))      # tests/ -> repo root
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np                      # noqa: E402
import helpers as H                     # noqa: E402
import mitsuba2_b200 as mi              # noqa: E402


def err(g, c, floor):
    g, c = np.asarray(g, np.float64), np.asarray(c, np.float64)
    e = np.abs(g - c) / np.maximum(np.abs(c), floor)
    e = e[np.isfinite(e)]
    return float(e.max()), float(np.quantile(e, 0.999))


def main():
    mi.set_variant("cuda_b200_rgb")
    for name, sd in (("material", H.material_scene()), ("n4", H.n4_scene()), ("n1", H.n1_scene())):
        scene, orc = mi.Scene(sd), H.Oracle(sd)
        rng = np.random.default_rng(5)
        n = 4096
        pos = rng.random((n, 2)).astype(np.float32)
        for lab, a, b in zip(("o_mint", "d_maxt"), scene.sensor_sample_ray(pos), orc.sensor_sample_ray(pos)):
            print(name, "sensor", lab, "max %.2e p99.9 %.2e" % err(a[:, :3], b[:, :3], 1e-3))
        wi = rng.normal(size=(n, 3)); wi[:, 2] = np.abs(wi[:, 2]) + 0.05; wi /= np.linalg.norm(wi, axis=1, keepdims=True)
        wi[: n // 8, 2] *= -1
        wo = rng.normal(size=(n, 3)); wo /= np.linalg.norm(wo, axis=1, keepdims=True)
        uv, s1, s2 = rng.random((n, 2)), rng.random(n), rng.random((n, 2))
        for bid in range(len(sd.bsdfs)):
            g, c = scene.bsdf_sample(bid, wi, uv, s1, s2), orc.bsdf_sample(bid, wi, uv, s1, s2)
            print(name, f"bsdf {bid} type {sd.bsdfs[bid]['type']} sample", "max %.2e p99.9 %.2e" % err(g, c, 1e-2))
            (eg, pg), (ec, pc) = scene.bsdf_eval_pdf(bid, wi, uv, wo), orc.bsdf_eval_pdf(bid, wi, uv, wo)
            print(name, f"bsdf {bid} eval", "max %.2e p99.9 %.2e" % err(eg, ec, 1e-3), "pdf", "max %.2e p99.9 %.2e" % err(pg, pc, 1e-3))
        ref_p = rng.normal(size=(n, 3)).astype(np.float32)
        g, c = scene.sample_emitter_direction(ref_p, s2), orc.sample_emitter_direction(ref_p, s2)
        print(name, "emitter d,dist,pdf", "max %.2e p99.9 %.2e" % err(g[:, :5], c[:, :5], 1e-2), "value", "max %.2e p99.9 %.2e" % err(g[:, 5:8], c[:, 5:8], 1e-2))


if __name__ == "__main__":
    main()
