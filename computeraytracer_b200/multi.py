"""Multi-GPU rendering, one process per GPU: a thin layer over libcrt's communicator calls.

New capability (the reference is single-device).  Every pixel-sample is independent and its RNG
stream depends only on (x, y, sample) (ComputeShader.wgsl:98), so any rank can compute any
(pixel, sample) and get the single-GPU value.  Two shardings (include/crt.h ``crt_shard_mode``):

* ``SHARD_TILES``      rank r owns the 8x4-pixel tiles with ``tile_id % world == r``; the gathered
  image is *bit-identical* to one GPU.
* ``SHARD_SPLIT_SPP``  rank r renders the frames with ``(sample - 1) % world == r`` over the full
  image; the sum over ranks equals the single-GPU sum up to float summation order.

Nothing is exchanged while rendering.  The one exchange step, ``crt_combine``, lives behind the C ABI
(csrc/crt_api.cu): ``ncclReduce`` of the W*H*4 f32 partial sums into a separate combined buffer on
rank ``dst`` for split-spp, a grouped ``ncclSend``/``ncclRecv`` of the *owned tiles only*
(W*H*16/world bytes per rank) for tiles, then the tone-map epilogue on ``dst``.  It is repeatable:
partial sums stay partial, so frames / combine / frames / combine is the same image as one render.

What is left for the host layer is to carry NCCL's 128-byte unique id from rank 0 to the other ranks
(:func:`init` does it over ``torch.distributed``, any backend; a Node host would use its own channel),
and the pure bookkeeping helpers below, which the CPU tests run under gloo.  The single-process form
(one host thread, n GPUs: ``Renderer(devices=[...])`` = ``crt_create_multi``) needs none of this.
"""
from __future__ import annotations

from typing import Optional

import numpy as np

from .renderer import COMM_ID_BYTES, SHARD_NONE, SHARD_SPLIT_SPP, SHARD_TILES, Renderer  # noqa: F401


def broadcast_comm_id(comm_id: Optional[bytes], src: int = 0, group=None) -> bytes:
    """Rank ``src`` passes the id it got from ``Renderer.comm_unique_id()``, the others ``None``; all get the id."""
    import torch
    import torch.distributed as dist
    dev = "cpu"
    if dist.get_backend(group) == "nccl":
        dev = f"cuda:{torch.cuda.current_device()}"
    t = torch.zeros(COMM_ID_BYTES, dtype=torch.uint8, device=dev)
    if dist.get_rank(group) == src:
        assert comm_id is not None and len(comm_id) == COMM_ID_BYTES
        t.copy_(torch.frombuffer(bytearray(comm_id), dtype=torch.uint8))
    dist.broadcast(t, src=src, group=group)
    return bytes(t.cpu().numpy().tobytes())


def init(r: Renderer, mode: int, group=None) -> None:
    """Join this process' renderer to a libcrt communicator spanning the torch.distributed group."""
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    cid = broadcast_comm_id(r.comm_unique_id() if rank == 0 else None, 0, group)
    r.comm_init(cid, rank, world, mode)


def combine(r: Renderer, dst: int = 0) -> None:
    """``crt_combine``: collective, asynchronous on the renderer's stream; rank ``dst`` ends with the combined image."""
    r.combine(dst)


# ------------------------------------------------------------------------------------------------ bookkeeping
def owned_tiles(width: int, height: int, rank: int, world: int) -> int:
    """Number of 8x4 tiles rank ``rank`` owns under SHARD_TILES."""
    n = ((width + 7) // 8) * ((height + 3) // 4)
    return (n - rank + world - 1) // world if n > rank else 0


def owned_samples(sample_start: int, n_frames: int, rank: int, world: int) -> int:
    """Number of the frames sample_start+1 .. sample_start+n_frames that rank ``rank`` renders under SHARD_SPLIT_SPP."""
    return sum(1 for s in range(sample_start + 1, sample_start + n_frames + 1) if (s - 1) % world == rank)


def tile_units(width: int, height: int, rank: int, world: int) -> np.ndarray:
    """Pixel index (x + y*W, or -1 outside the image) of every unit of rank ``rank``'s tiles, in unit order -- the order
    of libcrt's packed tile buffer (csrc/crt_kernels.cu pack_tiles_kernel / unit_pixel_of)."""
    tiles_x = (width + 7) // 8
    n = owned_tiles(width, height, rank, world)
    unit = np.arange(n * 32, dtype=np.int64)
    tile = (unit >> 5) * world + rank
    lane = unit & 31
    px = (tile % tiles_x) * 8 + (lane & 7)
    py = (tile // tiles_x) * 4 + (lane >> 3)
    return np.where((px < width) & (py < height), px + py * width, -1)


def pack_tiles(acc: np.ndarray, rank: int, world: int) -> np.ndarray:
    """numpy mirror of pack_tiles_kernel: acc[H, W, 4] -> [units, 4] (zeros for units outside the image)."""
    h, w = acc.shape[:2]
    idx = tile_units(w, h, rank, world)
    out = np.zeros((idx.size, 4), dtype=acc.dtype)
    ok = idx >= 0
    out[ok] = acc.reshape(-1, 4)[idx[ok]]
    return out


def unpack_tiles(packed: np.ndarray, acc: np.ndarray, rank: int, world: int) -> None:
    """numpy mirror of unpack_tiles_kernel: overwrite rank ``rank``'s tiles of acc[H, W, 4] (never adds)."""
    h, w = acc.shape[:2]
    idx = tile_units(w, h, rank, world)
    ok = idx >= 0
    acc.reshape(-1, 4)[idx[ok]] = packed[ok]
