"""Accumulator checkpoint / resume.

The reference has no checkpointing; its accumulator buffer only carries a `COPY_SRC` usage hint
(`src/main.js:301`).  Resuming is exact by construction: a pixel-sample's random stream is a pure function
of (x, y, sample) (`ComputeShader.wgsl:98`), so a render is fully described by binding 1 (the XYZ sums),
binding 2 (the frame counter) and the inputs that shaped the samples.  A checkpoint stores the two bindings
plus a fingerprint of those inputs, and `load` refuses a file whose fingerprint does not match the
renderer it is loaded into.

File layout (little endian):

    8 bytes   magic  b"CRTACC01"
    4 bytes   u32 n  length of the JSON header
    n bytes   JSON   {"width", "height", "sample", "params": {"max_depth", "grid_size", "t_min"},
                      "exposure", "scene_sha256", "acc_sha256"}
    W*H*16    f32    binding 1, row-major, 16-byte pixels (x, y, z, pad)

`exposure` is recorded but not checked: it only enters the resolve (CS:112), not the samples.
"""
from __future__ import annotations

import hashlib
import json
import struct
from typing import Optional

import numpy as np

from .renderer import Renderer
from .scene_loader import PackedScene

MAGIC = b"CRTACC01"

__all__ = ["CheckpointError", "scene_digest", "save", "load", "read_header"]


class CheckpointError(ValueError):
    pass


def scene_digest(ps: PackedScene) -> str:
    """SHA-256 over bindings 3..8 as uploaded (the camera, and so W x H, included)."""
    h = hashlib.sha256()
    for b, data in sorted(ps.bindings().items()):
        h.update(struct.pack("<IQ", b, len(data)))
        h.update(data)
    return h.hexdigest()


def _params_dict(r: Renderer) -> dict:
    p = r.get_params()
    return {"max_depth": int(p.max_depth), "grid_size": int(p.grid_size), "t_min": float(np.float32(p.t_min))}


def save(path: str, r: Renderer, ps: Optional[PackedScene] = None) -> dict:
    """Write bindings 1 and 2 of `r` (synchronises).  `ps` is the scene that was uploaded; without it the
    scene is not fingerprinted and `load` can only check the image size and parameters."""
    acc = r.accumulator()
    hdr = {
        "width": r.width, "height": r.height, "sample": r.sample, "params": _params_dict(r),
        "exposure": float(np.float32(r.get_params().exposure)),
        "scene_sha256": scene_digest(ps) if ps is not None else None,
        "acc_sha256": hashlib.sha256(acc.tobytes()).hexdigest(),
    }
    blob = json.dumps(hdr, sort_keys=True).encode()
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<I", len(blob)))
        f.write(blob)
        f.write(acc.tobytes())
    return hdr


def read_header(path: str) -> dict:
    with open(path, "rb") as f:
        if f.read(8) != MAGIC:
            raise CheckpointError(f"{path}: not a libcrt accumulator checkpoint")
        (n,) = struct.unpack("<I", f.read(4))
        if n > (1 << 20):
            raise CheckpointError(f"{path}: corrupt header length {n}")
        return json.loads(f.read(n).decode())


def load(path: str, r: Renderer, ps: Optional[PackedScene] = None) -> int:
    """Restore bindings 1 and 2 into `r`, whose scene and parameters are already set, and rebuild binding 0.
    Returns the frame counter.  Raises CheckpointError when the file belongs to a different render."""
    hdr = read_header(path)
    with open(path, "rb") as f:
        f.seek(8)
        (n,) = struct.unpack("<I", f.read(4))
        f.seek(12 + n)
        raw = f.read()
    w, h = int(hdr["width"]), int(hdr["height"])
    if (w, h) != (r.width, r.height):
        raise CheckpointError(f"{path}: image is {w}x{h}, the renderer's camera is {r.width}x{r.height}")
    if len(raw) != w * h * 16:
        raise CheckpointError(f"{path}: accumulator has {len(raw)} bytes, expected {w * h * 16} (truncated file?)")
    if hashlib.sha256(raw).hexdigest() != hdr["acc_sha256"]:
        raise CheckpointError(f"{path}: accumulator checksum mismatch (corrupt file)")
    if hdr["params"] != _params_dict(r):
        raise CheckpointError(f"{path}: rendered with {hdr['params']}, the renderer has {_params_dict(r)}")
    if ps is not None and hdr.get("scene_sha256") not in (None, scene_digest(ps)):
        raise CheckpointError(f"{path}: rendered from a different scene")
    r.set_binding(1, raw)
    r.set_binding(2, np.array([int(hdr["sample"])], dtype="<u4"))
    r.resolve_framebuffer()
    return int(hdr["sample"])
