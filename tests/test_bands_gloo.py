"""CPU, world_size 2 over gloo: the N>1 host path -- band partition + gather-v to rank 0 -- with the
oracle standing in for the per-rank renderer.  The gathered frame must equal the full frame bit for bit."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, height, width, out_path):
    import torch
    import torch.distributed as dist

    from oracle import pyoracle as O
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.bands import band_pixels, band_rows, gather_bands
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = S.scene_model("Gun", width, height)
        tiles_y = (height + 15) // 16
        r0, r1 = band_rows(tiles_y, world, rank)
        c, d, _ = O.render_scene(sc, threads=1, tile_rows=(r0, r1))
        fb = torch.from_numpy(np.concatenate([c, d[..., None]], axis=-1).copy())      # colour + depth, [H, W, 5]
        y0, y1 = band_pixels(height, world, rank)
        if rank != 0:
            fb[:y0] = -1.0                      # poison everything outside the own band
            fb[y1:] = -1.0
        gather_bands(fb, height, rank, world, dist)
        if rank == 0:
            np.save(out_path, fb.numpy())
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,height", [(2, 200), (3, 180)])
def test_band_gather_reproduces_the_full_frame(tmp_path, oracle, world, height):
    import torch.multiprocessing as mp

    from softwarerenderer_b200 import scene as S
    width = 320
    out = str(tmp_path / "gathered.npy")
    mp.spawn(_worker, args=(world, _free_port(), height, width, out), nprocs=world, join=True)
    got = np.load(out)
    c, d, _ = oracle.render_scene(S.scene_model("Gun", width, height), threads=0)
    assert np.array_equal(got[..., :4], c)
    assert np.array_equal(got[..., 4], d)
