"""Multi-GPU rendering: one process per GPU, scene replicated, pixels or samples sharded.

New capability (the reference is single-device).  Every pixel-sample is independent and its RNG
stream depends only on (x, y, sample) (ComputeShader.wgsl:98), so any rank can compute any
(pixel, sample) and get the single-GPU value.  Two shardings (include/crt.h ``crt_shard_mode``):

* ``SHARD_TILES``      rank r owns the 8x4-pixel tiles with ``tile_id % world == r``; its
  accumulator stays zero elsewhere, so a sum over ranks is *bit-identical* to one GPU.
* ``SHARD_SPLIT_SPP``  rank r renders the frames with ``(sample - 1) % world == r`` over the full
  image; the sum over ranks equals the single-GPU sum up to float summation order.

Neither needs communication while rendering.  The only exchange step is the combine at readback:
one ``reduce(SUM)`` of the W*H*4 f32 accumulator to rank 0 (NCCL over NVLink on GPUs; gloo in the
CPU tests), after which rank 0 re-runs the tone-map epilogue on the combined sums.
``torch.distributed`` is plumbing only; the tensors alias libcrt's device buffers.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from .renderer import SHARD_NONE, SHARD_SPLIT_SPP, SHARD_TILES, Renderer  # noqa: F401


class _DeviceBuffer:
    """Exposes a raw device pointer through __cuda_array_interface__ (zero-copy torch view)."""

    def __init__(self, ptr: int, n_floats: int):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 3}


def accumulator_tensor(r: Renderer) -> torch.Tensor:
    """float32 view of binding 1 (W*H*4) in libcrt's device memory."""
    ptr, nbytes = r.device_ptr(1)
    return torch.as_tensor(_DeviceBuffer(ptr, nbytes // 4), device=f"cuda:{r.device}")


def owned_tiles(width: int, height: int, rank: int, world: int) -> int:
    """Number of 8x4 tiles rank ``rank`` owns under SHARD_TILES."""
    n = ((width + 7) // 8) * ((height + 3) // 4)
    return (n - rank + world - 1) // world if n > rank else 0


def owned_samples(sample_start: int, n_frames: int, rank: int, world: int) -> int:
    """Number of the frames sample_start+1 .. sample_start+n_frames that rank ``rank`` renders under SHARD_SPLIT_SPP."""
    return sum(1 for s in range(sample_start + 1, sample_start + n_frames + 1) if (s - 1) % world == rank)


def combine_tensor(acc: torch.Tensor, dst: int = 0, group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """Sum the ranks' accumulators into rank ``dst`` (in place).  Works for both shardings."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(acc, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return acc


def combine(r: Renderer, stream: Optional[torch.cuda.Stream] = None, dst: int = 0) -> None:
    """Reduce binding 1 over NCCL to rank ``dst`` and rebuild binding 0 there.

    ``stream`` must be the stream the renderer launches on (``Renderer.set_stream``) so the reduce is
    ordered after the frames already queued.
    """
    acc = accumulator_tensor(r)
    if stream is not None:
        with torch.cuda.stream(stream):
            combine_tensor(acc, dst)
    else:
        r.synchronize()
        combine_tensor(acc, dst)
        torch.cuda.synchronize(r.device)
    if not dist.is_initialized() or dist.get_rank() == dst:
        r.resolve_framebuffer()
