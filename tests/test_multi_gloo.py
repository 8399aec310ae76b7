"""World-size-2 (and 3) gloo tests of the multi-GPU host logic on CPU: shard ownership, the
reduce-to-rank-0 combine and the frame-counter bookkeeping, with the oracle standing in for the
per-rank renderer (it implements the same two shardings)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, mode, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from computeraytracer_b200 import multi, scenes
    from tests import orc
    ps = scenes.default_scene(52, 30)
    sc = orc.load().scene(ps)
    part, _, cnt = sc.render(4, rank=rank, world=world, mode=mode, threads=1, want_fb=False)
    if mode == multi.SHARD_TILES:
        tiles = multi.owned_tiles(52, 30, rank, world)
        assert int(np.count_nonzero(part[..., :3].any(axis=-1))) <= tiles * 32
        assert int(cnt[0]) <= tiles * 32 * 4
    else:
        assert int(cnt[0]) == 52 * 30 * multi.owned_samples(0, 4, rank, world)
    t = torch.from_numpy(part.reshape(-1).copy())
    multi.combine_tensor(t, dst=0)
    if rank == 0:
        np.save(os.path.join(out_dir, f"combined_{mode}_{world}.npy"), t.numpy().reshape(30, 52, 4))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,mode", [(2, 1), (3, 1), (2, 2)])
def test_sharded_render_combines_to_single_rank_result(tmp_path, world, mode, oracle):
    from computeraytracer_b200 import scenes
    port = 29500 + os.getpid() % 2000 + world * 7 + mode
    mp.spawn(_worker, args=(world, mode, port, str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / f"combined_{mode}_{world}.npy")
    want, _, _ = oracle.scene(scenes.default_scene(52, 30)).render(4, threads=1, want_fb=False)
    if mode == 1:   # tiles: bit-identical
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    else:           # split-spp: float summation order only
        d = (got[..., :3].astype(np.float64) - want[..., :3].astype(np.float64)) / 4
        assert np.abs(d).max() <= 1e-3 and np.sqrt((d * d).mean()) <= 1e-4


def test_ownership_partitions_everything():
    from computeraytracer_b200 import multi
    for w, h in ((1920, 1080), (52, 30), (7, 3)):
        for world in (1, 2, 3, 4, 8):
            assert sum(multi.owned_tiles(w, h, r, world) for r in range(world)) == ((w + 7) // 8) * ((h + 3) // 4)
            assert sum(multi.owned_samples(5, 13, r, world) for r in range(world)) == 13
