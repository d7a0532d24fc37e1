"""-m gpu: the fused band gather.  Two processes share cuda:0 (host-level barriers only, no kernel waits
on another process): rank 1 renders the lower band and stores its final colour straight into rank 0's
colour buffer through a CUDA IPC mapping; rank 0's buffer must then hold the full frame."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_path, rgba8=False):
    import torch.distributed as dist

    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.bands import balanced_cuts
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = S.scene_city(1280, 720, n_instances=24)
        win = srb.MainWindow(sc.width, sc.height, device=0)
        S.submit_scene(win, sc)                                   # full frame once: per-row work for the cuts
        cuts = balanced_cuts(win.TileRowPairs().astype(np.float64) + 0.85 * 80, world)
        win.SetBand(cuts[rank], cuts[rank + 1])
        box = [(win.ExportGatherRgba8() if rgba8 else win.ExportColorHandle()) if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        if rank != 0:
            (win.SetGatherRgba8Target if rgba8 else win.SetColorTarget)(box[0])
        dist.barrier()
        S.submit_scene(win, sc)
        win.Sync()
        dist.barrier()                                            # every band has landed in rank 0's buffer
        if rank == 0:
            np.savez(out_path, color=win.GatheredRgba8() if rgba8 else win.ColorBuffer, cuts=np.array(cuts))
        if rgba8:                                                     # float4 colour stays on the rank that rendered it
            y0, y1 = cuts[rank] * 16, min(cuts[rank + 1] * 16, sc.height)
            np.save(out_path + f".band{rank}.npy", win.ColorBuffer[y0:y1])
        dist.barrier()
        (win.SetGatherRgba8Target if rgba8 else win.SetColorTarget)(None)
        win.Close()
    finally:
        dist.destroy_process_group()


def test_fused_band_gather_reproduces_the_full_frame(tmp_path, oracle):
    import torch.multiprocessing as mp

    from softwarerenderer_b200 import scene as S
    from tests.util import to_u8
    out = str(tmp_path / "fused.npz")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    z = np.load(out)
    oc, od, _ = oracle.render_scene(S.scene_city(1280, 720, n_instances=24), threads=8)
    cuts = z["cuts"].tolist()
    assert cuts[0] == 0 and cuts[-1] == 45 and 0 < cuts[1] < 45
    d8 = np.abs(to_u8(z["color"]) - to_u8(oc)).max(axis=-1)
    assert (d8 <= 1).mean() >= 0.999 and d8.max() <= 1


def test_rgba8_band_gather_assembles_the_quantised_frame(tmp_path, oracle):
    """The same gather at display precision (srb_gather_rgba8_export / _target): 4-byte peer stores; rank 0's gather
    frame must be the whole frame, equal to each rank's own float4 band quantised, and within the gate of the oracle."""
    import torch.multiprocessing as mp

    from softwarerenderer_b200 import scene as S
    from tests.util import to_u8
    out = str(tmp_path / "fused8.npz")
    mp.spawn(_worker, args=(2, _free_port(), out, True), nprocs=2, join=True)
    z = np.load(out)
    cuts = z["cuts"].tolist()
    frame = z["color"]
    assert frame.dtype == np.uint8 and frame.shape == (720, 1280, 4)
    for r in range(2):
        band = np.load(out + f".band{r}.npy")
        y0 = cuts[r] * 16
        want = np.rint(np.clip(band, 0.0, 1.0) * np.float32(255.0)).astype(np.uint8)
        assert np.array_equal(frame[y0:y0 + band.shape[0]], want)
    oc, _, _ = oracle.render_scene(S.scene_city(1280, 720, n_instances=24), threads=8)
    d8 = np.abs(frame.astype(np.int32) - to_u8(oc)).max(axis=-1)
    assert (d8 <= 1).mean() >= 0.999 and d8.max() <= 1


# ---------------------------------------------------------------------------
# Real multi-GPU: one rank per DEVICE (skipped on a one-GPU box), NCCL frame barrier on the raster stream, two
# DIFFERENT frames back to back with no host synchronisation in between -- a band rank's peer stores of frame B must
# not reach rank 0 before rank 0 has taken frame A, and the assembled frames are compared with the ORACLE.
# ---------------------------------------------------------------------------
def _worker_multi(rank, world, port, out_path):
    import torch
    import torch.distributed as dist

    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import _native as N
    from softwarerenderer_b200 import scene as S
    from softwarerenderer_b200.bands import balanced_cuts
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        a = S.scene_city(1280, 720, n_instances=24)
        b = S.scene_model("Gun", 1280, 720, view=S.gun_camera(3))
        win = srb.MainWindow(1280, 720, device=rank)
        win.SetBinSplit(64, 64, 32)                               # split bins on every band rank as well
        S.submit_scene(win, a)
        cuts = balanced_cuts(win.TileRowPairs().astype(np.float64) + 0.85 * 80, world)
        win.SetBand(cuts[rank], cuts[rank + 1])
        box = [win.ExportColorHandle() if rank == 0 else None]
        box8 = [win.ExportGatherRgba8() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        dist.broadcast_object_list(box8, src=0)
        ext = torch.cuda.ExternalStream(N.lib().srb_stream(win._ctx), device=dev)
        token = torch.zeros(1, device=dev)
        bar_rows = [None] * world                                 # the library's own barrier (flags in peer memory)
        dist.all_gather_object(bar_rows, win.BarrierExport())
        dist.barrier()
        win.BarrierConnect(rank, world, bar_rows)

        def barrier(fmt):                                         # float4 frames: NCCL all-reduce; rgba8 frames: srb_frame_barrier
            if fmt == "rgba8":
                win.FrameBarrier(0)
            else:
                dist.all_reduce(token)
        got = {}
        for fmt in ("float4", "rgba8"):
            if rank != 0:
                win.SetGatherRgba8Target(box8[0] if fmt == "rgba8" else None)
                win.SetColorTarget(box[0] if fmt == "float4" else None)
            dist.barrier()
            snaps = []
            with torch.cuda.stream(ext):
                for sc in (a, b, a, b):
                    S.submit_scene(win, sc)                       # no host sync: frames are pipelined on every rank
                    barrier(fmt)                                  # every band of this frame has landed ...
                    if rank == 0:                                 # ... rank 0 takes the frame on the same stream
                        src = torch.as_tensor(_DevArray(N.lib().srb_device_gathered_rgba8_ptr(win._ctx), (720, 1280), "<u4") if fmt == "rgba8"
                                              else _DevArray(N.lib().srb_device_color_ptr(win._ctx), (720, 1280, 4)), device=dev)
                        snaps.append(src.clone())
                    barrier(fmt)                                  # ... before any rank stores the next one
            torch.cuda.synchronize(dev)
            if rank == 0:
                got[fmt] = [s.cpu().numpy() for s in snaps]
        if rank == 0:
            np.savez(out_path, cuts=np.array(cuts), **{f"{f}_{i}": x for f, xs in got.items() for i, x in enumerate(xs)})
        dist.barrier()
        if rank != 0:
            win.SetGatherRgba8Target(None)
            win.SetColorTarget(None)
        win.BarrierConnect(rank, world, None)
        dist.barrier()
        win.Close()
    finally:
        dist.destroy_process_group()


class _DevArray:
    def __init__(self, ptr, shape, typestr="<f4"):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (int(ptr), False), "version": 2}


def test_bands_on_different_gpus_assemble_the_oracle_frame(tmp_path, oracle):
    import torch
    import torch.multiprocessing as mp

    from softwarerenderer_b200 import scene as S
    from tests.util import to_u8
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least two GPUs")
    out = str(tmp_path / "multi.npz")
    mp.spawn(_worker_multi, args=(world, _free_port(), out), nprocs=world, join=True)
    z = np.load(out)
    want = [oracle.render_scene(S.scene_city(1280, 720, n_instances=24), threads=8)[0],
            oracle.render_scene(S.scene_model("Gun", 1280, 720, view=S.gun_camera(3)), threads=8)[0]]
    for i in range(4):
        o8 = to_u8(want[i & 1])
        d8 = np.abs(to_u8(z[f"float4_{i}"]) - o8).max(axis=-1)
        assert (d8 <= 1).mean() >= 0.999 and d8.max() <= 1, (i, "float4")
        g8 = z[f"rgba8_{i}"].view(np.uint8).reshape(720, 1280, 4).astype(np.int32)
        d8 = np.abs(g8 - o8).max(axis=-1)
        assert (d8 <= 1).mean() >= 0.999 and d8.max() <= 1, (i, "rgba8")
