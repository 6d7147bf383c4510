"""Generates tests/golden/c1_cbox_reference.npz: the C1 configuration of BASELINE.json (Cornell box, 256 x 256,
max_depth 6, box filter) rendered by the UNMODIFIED reference CLI (`oracle/_ref/mitsuba/dist/mitsuba -m scalar_rgb`,
built by oracle/build_reference.sh) at 4096 spp.  Run in the development container (needs the reference build):
    python tests/golden/make_reference_renders.py
The image is what `hdrfilm` develops (linear RGB, float32)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import helpers as H  # noqa: E402

if __name__ == "__main__":
    spp = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    sd = H.cbox_scene(res=256, spp=64)
    img, secs = H.RefBuild(sd).render(spp=spp)
    np.savez_compressed(os.path.join(HERE, "c1_cbox_reference.npz"), image=img.astype(np.float32), spp=spp,
                        seconds=secs, threads=os.cpu_count(), variant="scalar_rgb")
    print(f"c1 reference: {spp} spp in {secs:.1f} s, mean {img.mean(axis=(0, 1))}")
