"""Generate the input fixtures the GPU box needs (it has no /root/reference).

Run in the build container, where the reference is mounted read-only:

    python tests/golden/make_fixtures.py

Reads the reference's scene data (src/scenes/cornell.json, src/scenes/CIE.json), runs it through
this repo's packer (computeraytracer_b200/scene_loader.py, the restatement of src/main.js:114-393)
and stores the resulting *binding buffers* -- the shader's input contract -- as one .npz.  No
reference source is copied; the .npz holds packed f32/u32 buffers only.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from computeraytracer_b200.scene_loader import load_scene  # noqa: E402

REF = os.environ.get("CRT_REFERENCE", "/root/reference")


def main() -> None:
    with open(os.path.join(REF, "src/scenes/cornell.json")) as f:
        scene = json.load(f)
    with open(os.path.join(REF, "src/scenes/CIE.json")) as f:
        cie = json.load(f)
    ps = load_scene(scene, cie)
    out = os.path.join(ROOT, "computeraytracer_b200", "data", "default_scene.npz")
    np.savez_compressed(
        out,
        planar_patches=np.frombuffer(ps.planar_patches.tobytes(), dtype=np.uint8),
        camera=ps.camera,
        cie=ps.cie,
        spectra=ps.spectra,
        lights=np.frombuffer(ps.lights.tobytes(), dtype=np.uint8),
        primitives=np.frombuffer(ps.primitives.tobytes(), dtype=np.uint8),
        spectrum_names=np.array(list(ps.spectrum_index.keys())),
    )
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
