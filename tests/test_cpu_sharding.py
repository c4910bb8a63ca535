"""CPU suite: the host logic of the multi-GPU path, world_size 2 over gloo.

bench.py shards a frame by row-cyclic bands (rtb_set_shard), all-gathers the packed bands and
scatters them back (rtb_shard_pack / rtb_shard_unpack).  The band index arithmetic is exported by
librtb.so as plain host functions (rtb_shard_row_of, rtb_shard_row_count); here two gloo ranks run
the same pack -> all_gather -> unpack sequence on host tensors with that arithmetic, with a
per-pixel function standing in for the render (pixels are independent functions of their index and
the frame seed, src/raytracer.cu:127).  Sample slicing (partitioning B) is covered the same way with
an all_reduce.  No compute call of the product runs here."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import raytracer_b200 as rtb


def _rows(height, rank, world, band):
    return [y for y in range(height) if (y // band) % world == rank]


@pytest.mark.parametrize("height", [1, 7, 8, 150, 1080, 2160])
@pytest.mark.parametrize("world,band", [(1, 8), (2, 8), (3, 5), (4, 16), (8, 8), (8, 1)])
def test_band_arithmetic(height, world, band):
    lib = rtb.lib
    seen = []
    for rank in range(world):
        want = _rows(height, rank, world, band)
        assert lib.rtb_shard_row_count(height, rank, world, band) == len(want)
        got = [lib.rtb_shard_row_of(lr, rank, world, band) for lr in range(len(want))]
        assert got == want
        assert rtb.shard_rows_host(height, rank, world, band) == want
        seen += want
    assert sorted(seen) == list(range(height))  # every row has exactly one owner


def _pixel_function(W, H, seed):
    """stand-in for the render: any function of (pixel index, seed)"""
    idx = np.arange(W * H * 3, dtype=np.int64).reshape(H, W, 3)
    return (((idx * 3145739 + seed * 6291469) % 1000003) / 1000003.0).astype(np.float32)


def _worker(rank, world, port, W, H, band, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lib = rtb.lib
    try:
        # --- partitioning A: tiles.  render my bands, pack, all-gather, unpack
        full = _pixel_function(W, H, 12345)
        mine = lib.rtb_shard_row_count(H, rank, world, band)
        pad = max(lib.rtb_shard_row_count(H, r, world, band) for r in range(world))
        image = np.zeros((H, W, 3), np.float32)
        for lr in range(mine):
            y = lib.rtb_shard_row_of(lr, rank, world, band)
            image[y] = full[y]  # "rendered" rows of this rank only
        dense = torch.zeros(pad, W, 3)
        for lr in range(mine):
            dense[lr] = torch.from_numpy(image[lib.rtb_shard_row_of(lr, rank, world, band)])
        gathered = [torch.empty_like(dense) for _ in range(world)]
        dist.all_gather(gathered, dense)
        result = image.copy()
        for r in range(world):
            for lr in range(pad):
                y = lib.rtb_shard_row_of(lr, r, world, band)
                if y < H and lr < lib.rtb_shard_row_count(H, r, world, band):
                    result[y] = gathered[r][lr].numpy()
        ok_tiles = bool((result == full).all())
        # --- partitioning B: sample slices.  frames f with f % world == rank, all-reduce, divide
        frames = 6
        acc = torch.zeros(H, W, 3, dtype=torch.float64)
        for f in range(rank, frames, world):
            acc += torch.from_numpy(_pixel_function(W, H, 1000 + f)).double()
        dist.all_reduce(acc)
        mean = (acc / frames).numpy()
        want = sum(_pixel_function(W, H, 1000 + f).astype(np.float64) for f in range(frames)) / frames
        ok_slices = bool(np.allclose(mean, want, rtol=0, atol=1e-12))
        # --- timing reduction used by bench.py: max over ranks
        t = torch.tensor([float(rank + 1)])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ok_max = t.item() == float(world)
        open(os.path.join(out_dir, f"rank{rank}.txt"), "w").write(f"{int(ok_tiles)} {int(ok_slices)} {int(ok_max)}")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("W,H,band", [(64, 50, 8), (33, 37, 5)])
def test_two_rank_gather_and_reduce_over_gloo(tmp_path, W, H, band):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    world = 2
    mp.spawn(_worker, args=(world, port, W, H, band, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert open(tmp_path / f"rank{r}.txt").read() == "1 1 1"
