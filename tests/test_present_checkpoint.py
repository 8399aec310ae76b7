"""Headless present (PNG/PPM of binding 0) and accumulator checkpoint/resume -- SURVEY.md section 8f items 2 and 3.
CPU tests use the oracle's framebuffer and a stand-in renderer; the GPU tests go through the C ABI."""
import os

import numpy as np
import pytest

from computeraytracer_b200 import checkpoint, present, scenes
from computeraytracer_b200.renderer import Params


# ------------------------------------------------------------------------------------------------ present
def test_png_round_trip_of_oracle_framebuffer(oracle, tmp_path):
    ps = scenes.default_scene(96, 64)
    _, fb, _ = oracle.scene(ps).render(2)
    p = str(tmp_path / "frame.png")
    out = present.present(fb, p)
    assert out is not None and np.array_equal(present.read_png(p), fb)
    assert fb[..., 3].min() == 255                        # CS:115 writes alpha 1.0
    present.write_png(p, fb, keep_alpha=False)
    assert np.array_equal(present.read_png(p), fb[..., :3])


def test_png_is_readable_by_an_independent_decoder(oracle, tmp_path):
    Image = pytest.importorskip("PIL.Image")
    ps = scenes.default_scene(40, 30)
    _, fb, _ = oracle.scene(ps).render(1)
    p = str(tmp_path / "frame.png")
    present.write_png(p, fb)
    with Image.open(p) as im:
        assert im.mode == "RGBA" and im.size == (40, 30)
        assert np.array_equal(np.asarray(im), fb)          # row 0 = top row, as the blit shows it (TextureRenderShader.wgsl:27-34)


def test_ppm_layout(tmp_path):
    rng = np.random.default_rng(3)
    fb = rng.integers(0, 256, size=(5, 7, 4), dtype=np.uint8)
    p = str(tmp_path / "f.ppm")
    present.present(fb, p)
    raw = open(p, "rb").read()
    assert raw.startswith(b"P6\n7 5\n255\n")
    assert raw[len(b"P6\n7 5\n255\n"):] == fb[..., :3].tobytes()


def test_present_rejects_wrong_input(tmp_path):
    with pytest.raises(ValueError):
        present.write_png(str(tmp_path / "x.png"), np.zeros((4, 4, 3), dtype=np.uint8))
    with pytest.raises(ValueError):
        present.write_png(str(tmp_path / "x.png"), np.zeros((4, 4, 4), dtype=np.float32))
    with pytest.raises(ValueError):
        present.present(np.zeros((4, 4, 4), dtype=np.uint8), str(tmp_path / "x.jpg"))


def test_read_png_handles_all_filters(tmp_path):
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(4)
    img = (rng.integers(0, 256, size=(9, 11, 4)) // 32 * 32).astype(np.uint8)
    p = str(tmp_path / "pil.png")
    Image.fromarray(img, "RGBA").save(p, optimize=True)   # PIL picks Sub/Up/Average/Paeth adaptively
    assert np.array_equal(present.read_png(p), img)


# ------------------------------------------------------------------------------------------------ checkpoint
class StandIn:
    """What checkpoint.save/load need of a Renderer, without a GPU."""

    def __init__(self, w, h, acc=None, sample=0, max_depth=100):
        self.width, self.height = w, h
        self.acc = np.zeros((h, w, 4), dtype=np.float32) if acc is None else acc
        self._sample = sample
        self.p = Params(); self.p.max_depth = max_depth; self.p.grid_size = 1; self.p.t_min = 0.001; self.p.exposure = 1.0
        self.resolved = 0

    sample = property(lambda self: self._sample)

    def accumulator(self):
        return self.acc.copy()

    def get_params(self):
        return self.p

    def set_binding(self, b, data):
        if b == 1:
            self.acc = np.frombuffer(data if isinstance(data, (bytes, bytearray)) else np.ascontiguousarray(data).tobytes(),
                                     dtype=np.float32).reshape(self.height, self.width, 4).copy()
        elif b == 2:
            self._sample = int(np.asarray(data).reshape(-1)[0])

    def resolve_framebuffer(self):
        self.resolved += 1


def test_checkpoint_round_trip_and_refusals(oracle, tmp_path):
    ps = scenes.default_scene(48, 32)
    acc, _, _ = oracle.scene(ps).render(3)
    src = StandIn(48, 32, acc, sample=3)
    p = str(tmp_path / "acc.crt")
    hdr = checkpoint.save(p, src, ps)
    assert hdr["sample"] == 3 and checkpoint.read_header(p) == hdr
    assert os.path.getsize(p) == 12 + len(__import__("json").dumps(hdr, sort_keys=True)) + 48 * 32 * 16
    dst = StandIn(48, 32)
    assert checkpoint.load(p, dst, ps) == 3
    assert dst.sample == 3 and dst.resolved == 1 and np.array_equal(dst.acc.view(np.uint32), acc.view(np.uint32))
    with pytest.raises(checkpoint.CheckpointError, match="64x32"):
        checkpoint.load(p, StandIn(64, 32), ps)
    with pytest.raises(checkpoint.CheckpointError, match="rendered with"):
        checkpoint.load(p, StandIn(48, 32, max_depth=8), ps)
    with pytest.raises(checkpoint.CheckpointError, match="different scene"):
        checkpoint.load(p, StandIn(48, 32), scenes.mesh_scene(48, 32, n_lon=8, n_lat=4))
    raw = bytearray(open(p, "rb").read()); raw[-5] ^= 0x40
    open(p, "wb").write(raw)
    with pytest.raises(checkpoint.CheckpointError, match="checksum"):
        checkpoint.load(p, StandIn(48, 32), ps)
    open(p, "wb").write(raw[:-16])
    with pytest.raises(checkpoint.CheckpointError, match="truncated"):
        checkpoint.load(p, StandIn(48, 32), ps)
    open(p, "wb").write(b"not a checkpoint at all")
    with pytest.raises(checkpoint.CheckpointError, match="not a libcrt"):
        checkpoint.read_header(p)


def test_scene_digest_sees_every_binding():
    a = scenes.default_scene(32, 32)
    b = scenes.default_scene(32, 32)
    assert checkpoint.scene_digest(a) == checkpoint.scene_digest(b)
    b.primitives = b.primitives.copy(); b.primitives.view(np.uint8).reshape(-1)[20] ^= 1
    assert checkpoint.scene_digest(a) != checkpoint.scene_digest(b)
    assert checkpoint.scene_digest(a) != checkpoint.scene_digest(scenes.default_scene(32, 33))


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_gpu_present_matches_oracle_image(oracle, tmp_path):
    from computeraytracer_b200.renderer import Renderer
    ps = scenes.default_scene(120, 80)
    with Renderer(0) as r:
        r.set_scene(ps); r.reset(); r.frame(4)
        p = str(tmp_path / "gpu.png")
        shown = present.present(r, p)
    _, fb, _ = oracle.scene(ps).render(4)
    assert np.array_equal(shown, fb) and np.array_equal(present.read_png(p), fb)


@pytest.mark.gpu
@pytest.mark.parametrize("maker,kw", [(scenes.default_scene, {}), (scenes.mesh_scene, dict(n_lon=64, n_lat=32))])
def test_gpu_checkpoint_resume_is_exact(tmp_path, maker, kw):
    from computeraytracer_b200.renderer import Renderer, default_params
    ps = maker(96, 64, **kw)
    p = str(tmp_path / "acc.crt")
    with Renderer(0) as r:
        r.set_scene(ps); r.set_params(default_params(max_depth=8)); r.reset()
        r.frame(5)
        checkpoint.save(p, r, ps)
        r.frame(4)
        want_acc, want_fb = r.accumulator(), r.framebuffer()
    with Renderer(0) as r2:                                  # a fresh context, as after a restart
        r2.set_scene(ps); r2.set_params(default_params(max_depth=8))
        assert checkpoint.load(p, r2, ps) == 5
        shown = r2.framebuffer()                             # binding 0 is rebuilt by load
        r2.frame(4)
        assert r2.sample == 9
        assert np.array_equal(r2.accumulator().view(np.uint32), want_acc.view(np.uint32))
        assert np.array_equal(r2.framebuffer(), want_fb)
        assert shown.shape == want_fb.shape and shown[..., 3].min() == 255
        r2.set_params(default_params(max_depth=9))
        with pytest.raises(checkpoint.CheckpointError):
            checkpoint.load(p, r2, ps)
