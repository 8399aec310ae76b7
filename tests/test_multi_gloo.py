"""World-size-2 (and 3) gloo tests of the multi-GPU host logic on CPU.

libcrt's combine (csrc/crt_api.cu crt_combine) is: split-spp -- reduce(SUM) of the ranks' partial sums into a SEPARATE
combined buffer on the destination, the partials stay partial; tiles -- every rank packs the tiles it owns in unit order
and sends only those, the destination overwrites (never adds).  Here the same protocol runs over gloo with the oracle
standing in for the per-rank renderer (it implements the same two shardings) and the numpy mirrors of the pack / unpack
kernels (computeraytracer_b200.multi), twice in a row with more frames in between: what the GPU path does over NCCL,
proven repeatable (an in-place reduce added the other ranks' old partials again on the second combine).
Also: the 128-byte communicator id reaches every rank."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H = 52, 30


def _combine(part: np.ndarray, rank: int, world: int, mode: int, multi) -> np.ndarray:
    """The protocol of crt_combine over gloo; returns the combined accumulator on rank 0 (None elsewhere)."""
    if mode == multi.SHARD_SPLIT_SPP:
        t = torch.from_numpy(part.reshape(-1).copy())       # a scratch copy: `part` itself is never written
        dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
        return t.numpy().reshape(H, W, 4) if rank == 0 else None
    if rank != 0:
        packed = multi.pack_tiles(part, rank, world)
        assert packed.shape[0] == multi.owned_tiles(W, H, rank, world) * 32          # W*H*16/world bytes, not W*H*16
        dist.send(torch.from_numpy(packed.copy()), dst=0)
        return None
    comb = part.copy()
    for r in range(1, world):
        buf = torch.zeros((multi.owned_tiles(W, H, r, world) * 32, 4), dtype=torch.float32)
        dist.recv(buf, src=r)
        multi.unpack_tiles(buf.numpy(), comb, r, world)
    return comb


def _worker(rank, world, mode, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from computeraytracer_b200 import multi, scenes
    from tests import orc
    # the communicator id travels from rank 0 to everyone (on the GPU box it comes from crt_comm_unique_id)
    want_id = bytes((7 * i + 3) % 256 for i in range(multi.COMM_ID_BYTES))
    got_id = multi.broadcast_comm_id(want_id if rank == 0 else None)
    assert got_id == want_id
    ps = scenes.default_scene(W, H)
    sc = orc.load().scene(ps)
    part, _, cnt = sc.render(4, rank=rank, world=world, mode=mode, threads=1, want_fb=False)
    if mode == multi.SHARD_TILES:
        tiles = multi.owned_tiles(W, H, rank, world)
        assert int(np.count_nonzero(part[..., :3].any(axis=-1))) <= tiles * 32
        assert int(cnt[0]) <= tiles * 32 * 4
    else:
        assert int(cnt[0]) == W * H * multi.owned_samples(0, 4, rank, world)
    first = _combine(part, rank, world, mode, multi)
    # progressive preview: three more frames, combine again -- must equal one 7-frame render, not 7 frames + stale partials
    part2, _, _ = sc.render(3, sample_start=4, acc=part, rank=rank, world=world, mode=mode, threads=1, want_fb=False)
    second = _combine(part2, rank, world, mode, multi)
    again = _combine(part2, rank, world, mode, multi)
    if rank == 0:
        assert np.array_equal(second.view(np.uint32), again.view(np.uint32))      # combining twice changes nothing
        np.save(os.path.join(out_dir, f"combined4_{mode}_{world}.npy"), first)
        np.save(os.path.join(out_dir, f"combined7_{mode}_{world}.npy"), second)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,mode", [(2, 1), (3, 1), (2, 2), (3, 2)])
def test_sharded_render_combines_to_single_rank_result(tmp_path, world, mode, oracle):
    from computeraytracer_b200 import scenes
    port = 29500 + os.getpid() % 2000 + world * 7 + mode
    mp.spawn(_worker, args=(world, mode, port, str(tmp_path)), nprocs=world, join=True)
    sc = oracle.scene(scenes.default_scene(W, H))
    for frames in (4, 7):
        got = np.load(tmp_path / f"combined{frames}_{mode}_{world}.npy")
        want, _, _ = sc.render(frames, threads=1, want_fb=False)
        if mode == 1:   # tiles: bit-identical
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        else:           # split-spp: float summation order only
            d = (got[..., :3].astype(np.float64) - want[..., :3].astype(np.float64)) / frames
            assert np.abs(d).max() <= 1e-3 and np.sqrt((d * d).mean()) <= 1e-4


def test_ownership_partitions_everything():
    from computeraytracer_b200 import multi
    for w, h in ((1920, 1080), (52, 30), (7, 3)):
        for world in (1, 2, 3, 4, 8):
            assert sum(multi.owned_tiles(w, h, r, world) for r in range(world)) == ((w + 7) // 8) * ((h + 3) // 4)
            assert sum(multi.owned_samples(5, 13, r, world) for r in range(world)) == 13
            # the packed tile buffers of all ranks hold every pixel exactly once
            px = np.concatenate([multi.tile_units(w, h, r, world) for r in range(world)])
            px = px[px >= 0]
            assert px.size == w * h and np.array_equal(np.sort(px), np.arange(w * h))


def test_pack_unpack_round_trip():
    from computeraytracer_b200 import multi
    rng = np.random.default_rng(0)
    acc = rng.random((30, 52, 4), dtype=np.float32)
    out = np.zeros_like(acc)
    for r in range(3):
        multi.unpack_tiles(multi.pack_tiles(acc, r, 3), out, r, 3)
    assert np.array_equal(acc, out)
