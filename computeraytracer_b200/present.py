"""Headless present: binding 0 (the RGBA8 framebuffer) written to an image file.

The reference presents the framebuffer by drawing a screen-filling quad that samples it 1:1
(`src/shaders/TextureRenderShader.wgsl:16-48`, `src/main.js:42-112, 612-617`): texture coordinate (0, 0)
sits at clip-space (-1, 1), so row 0 of the framebuffer is the top row of the canvas, the canvas has the
framebuffer's size (`main.js:25`), and the `rgba8unorm` swapchain receives the texel values unchanged.
That needs a graphics API (unavailable on this box); the headless equivalent is the same bytes in a file.
Tone mapping, the sRGB curve (with the reference's green-channel quirk, `ComputeShader.wgsl:453`) and the
unorm rounding all happen in the compute pass (`CS:109-115`) and are part of binding 0 already, so nothing
is recomputed here.

No imaging library is needed: PNG is zlib + CRC32 (both in the standard library), PPM is a header.
"""
from __future__ import annotations

import struct
import zlib
from typing import Union

import numpy as np

__all__ = ["write_png", "write_ppm", "read_png", "present"]


def _as_rgba8(rgba) -> np.ndarray:
    a = np.asarray(rgba)
    if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] != 4:
        raise ValueError(f"expected uint8[H, W, 4] (binding 0), got {a.dtype}{list(a.shape)}")
    return np.ascontiguousarray(a)


def _chunk(tag: bytes, data: bytes) -> bytes:
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def write_png(path: str, rgba, keep_alpha: bool = True, level: int = 6) -> None:
    """8-bit RGBA (or RGB) PNG, rows top to bottom as the blit shows them.  No gamma / sRGB chunk is written
    beyond `sRGB` (rendering intent 0): the values already are display-referred sRGB (CS:441-467)."""
    a = _as_rgba8(rgba)
    h, w, _ = a.shape
    if not keep_alpha:
        a = np.ascontiguousarray(a[:, :, :3])
    color_type = 6 if keep_alpha else 2
    rows = np.empty((h, 1 + a.shape[1] * a.shape[2]), dtype=np.uint8)
    rows[:, 0] = 0                                        # filter type 0 (None) on every scanline
    rows[:, 1:] = a.reshape(h, -1)
    png = b"\x89PNG\r\n\x1a\n"
    png += _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, color_type, 0, 0, 0))
    png += _chunk(b"sRGB", b"\x00")
    png += _chunk(b"IDAT", zlib.compress(rows.tobytes(), level))
    png += _chunk(b"IEND", b"")
    with open(path, "wb") as f:
        f.write(png)


def read_png(path: str) -> np.ndarray:
    """Reader for the files write_png produces (8-bit RGB/RGBA, non-interlaced, any of the five filters);
    used by the round-trip tests so that they do not depend on an imaging library."""
    with open(path, "rb") as f:
        data = f.read()
    if data[:8] != b"\x89PNG\r\n\x1a\n":
        raise ValueError("not a PNG file")
    pos, idat, w = 8, b"", None
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
        if zlib.crc32(tag + body) & 0xFFFFFFFF != crc:
            raise ValueError(f"bad CRC in chunk {tag!r}")
        if tag == b"IHDR":
            w, h, depth, ctype, _, _, interlace = struct.unpack(">IIBBBBB", body)
            if depth != 8 or ctype not in (2, 6) or interlace:
                raise ValueError("only 8-bit non-interlaced RGB/RGBA is supported")
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    if w is None:
        raise ValueError("no IHDR chunk")
    bpp = 4 if ctype == 6 else 3
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + w * bpp)
    out = np.zeros((h, w * bpp), dtype=np.uint8)
    prev = np.zeros(w * bpp, dtype=np.int32)
    for y in range(h):
        ft, line = int(raw[y, 0]), raw[y, 1:].astype(np.int32)
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 0xFF
        else:                                             # Sub / Average / Paeth depend on the pixel to the left
            cur = np.zeros_like(line)
            for i in range(w * bpp):
                a_ = cur[i - bpp] if i >= bpp else 0
                b_ = prev[i]
                c_ = prev[i - bpp] if i >= bpp else 0
                if ft == 1:
                    p = a_
                elif ft == 3:
                    p = (a_ + b_) >> 1
                else:
                    pa, pb, pc = abs(b_ - c_), abs(a_ - c_), abs(a_ + b_ - 2 * c_)
                    p = a_ if (pa <= pb and pa <= pc) else (b_ if pb <= pc else c_)
                cur[i] = (line[i] + p) & 0xFF
        out[y] = cur
        prev = cur
    return out.reshape(h, w, bpp)


def write_ppm(path: str, rgba) -> None:
    """Binary PPM (P6): RGB, alpha dropped, rows top to bottom."""
    a = _as_rgba8(rgba)
    h, w, _ = a.shape
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(np.ascontiguousarray(a[:, :, :3]).tobytes())


def present(source: Union[np.ndarray, "object"], path: str) -> np.ndarray:
    """The headless stand-in for the blit pass + canvas (`main.js:612-617`): writes binding 0 to `path`
    (.png or .ppm) and returns the uint8[H, W, 4] image written.  `source` is a Renderer or the array."""
    rgba = source.framebuffer() if hasattr(source, "framebuffer") else source
    rgba = _as_rgba8(rgba)
    if path.lower().endswith(".ppm"):
        write_ppm(path, rgba)
    elif path.lower().endswith(".png"):
        write_png(path, rgba)
    else:
        raise ValueError("present(): path must end in .png or .ppm")
    return rgba
