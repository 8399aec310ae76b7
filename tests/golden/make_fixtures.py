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
    make_golden_outputs()


def make_golden_outputs() -> None:
    """Golden OUTPUT vectors of the CPU oracle (oracle/crt_oracle.c) on the reference's default scene.

    The reference ships no golden vectors and its WGSL cannot run here (no graphics stack), so these are the
    oracle's outputs, frozen after the oracle was checked against SURVEY.md Appendix B (tests/test_oracle_kat.py).
    They pin the oracle against regressions (CPU tests) and give the CUDA path a second, file-based target.
    """
    from computeraytracer_b200 import scenes
    from tests import orc
    o = orc.load()
    here = os.path.dirname(os.path.abspath(__file__))
    # (1) full frames: 64x64, 4 frames, reference defaults (MAXDEPTH 100)
    ps = scenes.default_scene(64, 64)
    acc, fb, cnt = o.scene(ps).render(4)
    # (2) same with max_depth 4 (BASELINE config 1's bounce limit), 48x32 (ragged: not a multiple of 8x4 tiles... 48x30)
    ps2 = scenes.default_scene(50, 30)
    acc2, fb2, cnt2 = o.scene(ps2, orc.default_params(max_depth=4)).render(3)
    # (3) closest hits of a fixed ray set, with and without excludes
    rays = scenes.camera_rays(ps, 4096, seed=11)
    rng = np.random.default_rng(12)
    ex = rng.integers(0, 19, size=rays.shape[0]).astype(np.uint32)
    ex[ex == 18] = 0xFFFFFFFF
    sc = o.scene(ps)
    i0, t0, _ = sc.trace_rays(rays, None)
    i1, t1, _ = sc.trace_rays(rays, ex)
    out = os.path.join(here, "cornell_oracle_golden.npz")
    np.savez_compressed(out, acc_64x64_s4=acc, fb_64x64_s4=fb, cnt_64x64_s4=cnt,
                        acc_50x30_s3_d4=acc2, fb_50x30_s3_d4=fb2, cnt_50x30_s3_d4=cnt2,
                        rays=rays, exclude=ex, hit_idx=i0, hit_t=t0, hit_idx_ex=i1, hit_t_ex=t1)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
