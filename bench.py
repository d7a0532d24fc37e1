#!/usr/bin/env python
"""bench.py -- frames/s and Mpixels/s of the rasterization hot path on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload city4k|gun1080|stress8k]
  python bench.py --impl reference ...      # the CPU path (oracle port) on the host cores

A step = one frame of the workload.  `value` = frames/s with the draw list and
all meshes/textures resident in HBM (srb_replay_frame, frames pipelined); `e2e`
= the same frame through the public API with host buffers, every step:
srb_begin_frame + srb_submit_draws (host draw records -> pinned -> H2D) +
srb_end_frame + srb_present_rgb_async (the packed RGB float frame
MainWindow.OnRender uploads, D2H on the copy stream while the next frame
renders).  N>1: sort-first bands on 16-row tile boundaries, one rank per GPU;
the bands reach rank 0's frame as peer stores over NVLink from inside the tile
kernel (`--gather nccl`: grouped NCCL send/recv instead); for `e2e` every rank
delivers its own rows to one shared pinned host frame over its own PCIe link.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[2] / [3]: the 4K roofline config (and its band-partitioned form)
    "city4k": dict(kind="city", width=3840, height=2160, desc="74x{dust2,Gun,gordon}=1,007,140 tris, 1332 draws, FrustumCuller active, 3840x2160"),
    # configs[1]
    "gun1080": dict(kind="gun", width=1920, height=1080, desc="bundled Gun (3,910 tris, 2048^2 texture), 1920x1080"),
    # configs[4]
    "stress8k": dict(kind="stress", width=7680, height=4320, desc="155 grid meshes x 64,800 = 10,044,000 small tris, 1024^2 texture, 7680x4320"),
    # configs[4] AS STATED (bilinear + 4 lights incl. point lights): extensions beyond the reference, oracle = our restatement only
    "stress8k_ext": dict(kind="stress_ext", width=7680, height=4320, desc="stress8k with the extensions of BASELINE config 5: bilinear texture filter, "
                         "4 lights (stock directional + 1 directional + 2 point lights with Light.cs attenuation); oracle = our restatement only"),
    # small variants for quick checks
    "city720": dict(kind="city", width=1280, height=720, desc="city scene at 1280x720"),
}


def build_scene(name):
    from softwarerenderer_b200 import scene as S
    w = WORKLOADS[name]
    if w["kind"] == "city":
        return S.scene_city(w["width"], w["height"])
    if w["kind"] == "gun":
        return S.scene_model("Gun", w["width"], w["height"])
    return S.scene_stress(w["width"], w["height"], extensions=w["kind"] == "stress_ext")


from softwarerenderer_b200.bands import (balanced_cuts, band_rows, gather_bands, rebalance_from_periods,  # noqa: E402
                                        relieve_worst_band)


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 8 for i in range(4) if r[4 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def base_config(workload, scene):
    """The workload description both arms print (identical dictionaries: the driver compares them); everything
    that only the GPU arm knows goes under `config_detail`."""
    W, H = scene.width, scene.height
    l2 = ("no explicit flush: per-frame working set (framebuffer %.0f MB + geometry/records) exceeds the 126 MB L2" % (20 * W * H / 1e6)
          if 20 * W * H > 126e6 else "working set fits L2 (launch-latency-bound config); no flush")
    return {"workload": workload, "scene": WORKLOADS[workload]["desc"], "width": W, "height": H,
            "triangles": scene.triangle_count, "draws": len(scene.draws), "l2": l2}


def bind_to_gpu_numa(gpu_index):
    """Pin this process to the CPUs next to its GPU (NVML's ideal affinity), so that host memory it touches first --
    its rows of the shared host frame, its pinned staging -- lands on the NUMA node its PCIe link is attached to.
    Returns the number of CPUs bound to, or None when NVML or the call is unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * i + b for i, w in enumerate(mask) for b in range(64) if (int(w) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return None


def algorithmic_bytes(scene, visible, p_cov):
    """compulsory bytes per frame."""
    geo, texs = 0, {}
    for d, vis in zip(scene.draws, visible):
        if not vis:
            continue
        geo += 48 * d.mesh.Vertices.shape[0] + 2 * d.mesh.Indices.size + 192
        if d.mesh.Texture is not None:
            texs[id(d.mesh.Texture)] = 4 * d.mesh.Texture.Width * d.mesh.Texture.Height
    tex = min(sum(texs.values()), 4 * p_cov)
    fb = 20 * scene.width * scene.height
    return dict(geometry=geo, texels=tex, framebuffer=fb, total=geo + tex + fb)


def run_reference(args):
    """--impl reference: the reference's CPU path.  The C# itself cannot run here (no .NET), so this
    is the C++ restatement (oracle 'port'), tile-row-parallel on all host threads."""
    from oracle import pyoracle as O
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    O.build()
    scene = build_scene(args.workload)
    cores = os.cpu_count() or 1
    L = O.lib()
    u = O.make_uniforms(scene)
    arr, n, keep = O.make_draws(scene, O.frustum_flags(scene))
    win = L.orc_create(scene.width, scene.height)
    cc = (C.c_float * 4)(*[float(x) for x in scene.clear_color])

    def frame():
        L.orc_clear_frame_mt(win, cc, cores)             # Parallel.For clears, as in the reference (MainWindow.cs:400-436)
        L.orc_render_frame_mt(win, arr, n, C.byref(u), cores, 0, 1 << 30)

    budget = float(os.environ.get("SRB_REF_BUDGET_S", "150"))
    t_start = time.perf_counter()
    # Warm-up frames as asked (the driver compares both arms' counts), within a quarter of the time budget.
    warm = 0
    while warm < args.warmup and (warm == 0 or time.perf_counter() - t_start < budget / 4):
        frame()
        warm += 1
    t0 = time.perf_counter()
    steps = 0
    while steps < args.steps:
        frame()
        steps += 1
        if time.perf_counter() - t_start > budget and steps >= 1:
            break
    dt = time.perf_counter() - t0
    fps = steps / dt
    W, H = scene.width, scene.height
    line = {
        "impl": "reference", "metric": "frames_per_sec", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": 1000.0 * dt / steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "mpixels_per_sec": W * H * fps / 1e6,
        "config": base_config(args.workload, scene),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
                         "sample": f"{steps} full frame(s) of the workload (requested {args.steps}; bounded to ~{budget:.0f}s); "
                                   "C++ restatement of the C# path, parallel over triangles, then over runs of 8 tiles (longest list first); the C# reference cannot run here (no .NET)"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


class _DevArray:
    def __init__(self, ptr, shape, typestr="<f4"):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (int(ptr), False), "version": 2}


STAGES = ["frustum", "vertex", "setup_count", "scan", "fill", "sort", "tile_raster"]
DRAW_REC_BYTES = 384


def run_gpu(args):
    import torch
    import torch.distributed as dist

    import softwarerenderer_b200 as srb
    from softwarerenderer_b200 import _native as N
    from softwarerenderer_b200 import scene as S

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback; use --impl reference for the CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = N.lib()
    scene = build_scene(args.workload)
    W, H = scene.width, scene.height
    tiles_y = (H + 15) // 16
    win = srb.MainWindow(W, H, device=local)
    ctx = win._ctx
    if args.pipeline:
        win.SetPipeline(args.pipeline)

    # ---- first frame through the mirror API: uploads meshes/textures, settles capacities ----
    S.submit_scene(win, scene)
    st = win.Stats()
    # ---- N>1: sort-first bands on tile-row boundaries.  Every rank rendered the whole frame once above,
    # so all ranks derive the same cost-balanced cuts from the same per-row bin counts (no exchange).
    cuts = [band_rows(tiles_y, world, r)[0] for r in range(world)] + [tiles_y]
    full_frame_ms = float(st.gpu_ms)
    w_rows = win.TileRowPairs().astype(np.float64) + 0.85 * ((W + 15) // 16)         # an empty tile costs about 0.85 pair
    if world > 1 and args.bands == "balanced":
        # fused gather: the pixels of bands 1.. go to rank 0 as peer stores over NVLink, which costs their
        # tile kernel about --remote-cost pairs per tile; band 0 takes correspondingly more rows
        extra = args.remote_cost * ((W + 15) // 16) if args.gather == "fused" else 0.0
        cuts = balanced_cuts(w_rows, world, remote_extra=extra)
    r0, r1 = cuts[rank], cuts[rank + 1]
    fused = world > 1 and args.gather == "fused"
    rgba8_gather = fused and args.gather_format == "rgba8"
    box = box8 = [None]
    refine_log = None

    def use_gather(fmt):
        """Band ranks: where the tile kernel stores final colour.  'rgba8': 4-byte peer stores into rank 0's RGBA8
        frame (float4 colour stays local); 'float4': 16-byte peer stores into rank 0's ColorBuffer; None: local."""
        if rank == 0:
            return
        win.SetGatherRgba8Target(box8[0] if fmt == "rgba8" else None)
        win.SetColorTarget(box[0] if fmt == "float4" else None)

    if world > 1:
        win.SetBand(r0, r1)
        if fused and args.frame_barrier == "peer":
            bar_rows = [None] * world
            dist.all_gather_object(bar_rows, win.BarrierExport())
            dist.barrier()                      # every row is zeroed before anyone posts into it
            win.BarrierConnect(rank, world, bar_rows)
        if fused:
            # rank 0 exports its ColorBuffer and an RGBA8 frame (its own tile kernel fills its band of that one too)
            box = [win.ExportColorHandle() if rank == 0 else None]
            box8 = [win.ExportGatherRgba8() if rank == 0 else None]
            dist.broadcast_object_list(box, src=0)
            dist.broadcast_object_list(box8, src=0)
            use_gather(args.gather_format)
        S.submit_scene(win, scene)              # settle band-sized capacities
        if args.bands == "balanced" and args.refine_cuts > 0:
            # The model behind the first cuts (bin pairs + a constant per tile) knows nothing of a band's geometry
            # work or of what its stores cost over NVLink.  Measure instead: every rank times its own band, uncoupled
            # (no barrier), the periods are exchanged, each band's period is spread over its rows like the model's
            # weights, and the cuts are balanced again on that; the best cuts seen (smallest worst period) are kept.
            # Every rank computes the same cuts from the same gathered numbers.
            def band_period(k=48):
                for _ in range(16):
                    N.check(ctx, L.srb_replay_frame(ctx))
                win.Sync()
                t0 = time.perf_counter()
                for _ in range(k):
                    N.check(ctx, L.srb_replay_frame(ctx))
                win.Sync()
                return (time.perf_counter() - t0) / k * 1e3

            best_worst, best_cuts, best_ts, tried = None, list(cuts), None, []
            for it in range(args.refine_cuts + 1):
                ts = [None] * world
                dist.all_gather_object(ts, float(band_period()))
                tried.append({"cuts": [int(x) for x in cuts], "period_ms": [round(float(x), 4) for x in ts]})
                if best_worst is None or max(ts) < best_worst:
                    best_worst, best_cuts, best_ts = max(ts), list(cuts), [float(x) for x in ts]
                if it == args.refine_cuts:
                    break
                new = rebalance_from_periods(cuts, ts, w_rows)
                if list(new) == list(cuts):
                    break
                cuts = list(new)
                win.SetBand(cuts[rank], cuts[rank + 1])
                S.submit_scene(win, scene)
            # Local search from the best cuts: take one tile row off the slowest band at a time (faster neighbour
            # first) for as long as the measured worst period falls.  What the proportional model cannot see -- the
            # part of a band's period that does not move with its rows -- is found this way.
            for _ in range(2 * args.refine_cuts):
                improved = False
                for cand in relieve_worst_band(best_cuts, best_ts):
                    cuts = list(cand)
                    win.SetBand(cuts[rank], cuts[rank + 1])
                    S.submit_scene(win, scene)
                    ts = [None] * world
                    dist.all_gather_object(ts, float(band_period()))
                    tried.append({"cuts": [int(x) for x in cuts], "period_ms": [round(float(x), 4) for x in ts]})
                    if max(ts) < 0.995 * best_worst:
                        best_worst, best_cuts, best_ts, improved = max(ts), list(cuts), [float(x) for x in ts], True
                        break
                if not improved:
                    break
            if list(cuts) != best_cuts:
                cuts = best_cuts
                win.SetBand(cuts[rank], cuts[rank + 1])
                S.submit_scene(win, scene)
            r0, r1 = cuts[rank], cuts[rank + 1]
            refine_log = tried
        st = win.Stats()
    y0, y1 = r0 * 16, min(r1 * 16, H)
    # host-side command array for the e2e loop (what the C# shim would pin and pass)
    rows = []
    from softwarerenderer_b200.rasterizer import Rasterizer
    from softwarerenderer_b200.shaders import Shaders
    S.submit_scene.__doc__  # noqa: B018
    win.ClearDepthBuffer(); win.ClearColorBuffer(scene.clear_color)
    for i, d in enumerate(scene.draws):
        Rasterizer.RenderMesh(win, d.mesh.Vertices, d.mesh.Indices, d.model, d.view, d.proj, Shaders.StockVertexShader,
                              Shaders.StockFragmentShader(d.mesh.Texture), d.cull, d.depth, d.blend, submissionIndex=i,
                              sphereBounds=d.mesh.SphereBounds if d.frustum_cull else None)
    cmds = np.concatenate(win._draws)
    # the draw records are what end_frame uploads (the vertex/triangle block tables are built on the device)
    win._draws = []
    win._clear_flags = 0
    clear = (C.c_float * 4)(*[float(x) for x in scene.clear_color])

    ext = torch.cuda.ExternalStream(L.srb_stream(ctx), device=dev)
    fb = torch.as_tensor(_DevArray(L.srb_device_color_ptr(ctx), (H, W, 4)), device=dev)

    token = torch.zeros(1, device=dev)
    bar_stream = torch.cuda.Stream(device=dev) if world > 1 else None
    bar_pending = []

    def frame_barrier():
        """Fused gather: the pixels already went over NVLink as peer stores; what is left of the gather is the
        "every band of frame k has landed" barrier.  'peer' (default): srb_frame_barrier -- every rank posts its frame
        counter into every rank's flag row over NVLink and waits for the others' previous frame; no collective, no host
        work beyond one launch.  The NCCL forms remain for comparison, a 4-byte all-reduce.  'side': the all-reduce of
        frame k runs on its own stream behind tile kernel k, and tile kernel k+2 waits for it -- the ranks stay
        within one frame of each other, and the collective's launch and latency are off the raster stream.
        'inline': on the raster stream, between tile kernels k and k+1 (round 1)."""
        if args.frame_barrier == "peer":        # the library's own barrier: flags in peer memory, one 32-thread kernel
            N.check(ctx, L.srb_frame_barrier(ctx, 1))
            return
        if args.frame_barrier == "inline":
            dist.all_reduce(token)
            return
        if args.frame_barrier == "none":        # diagnostic only: ranks run free (not a result)
            return
        e = torch.cuda.Event()
        e.record(ext)
        bar_stream.wait_event(e)
        with torch.cuda.stream(bar_stream):
            dist.all_reduce(token)
            eb = torch.cuda.Event()
            eb.record(bar_stream)
        bar_pending.append(eb)
        if len(bar_pending) > 1:
            ext.wait_event(bar_pending.pop(0))

    def hard_barrier():
        """Every rank has reached this point of its raster stream (device-side; the host does not wait)."""
        if args.frame_barrier == "peer" and fused:
            N.check(ctx, L.srb_frame_barrier(ctx, 0))
        else:
            dist.all_reduce(token)

    def gather(landed=False):
        """landed=True: the caller reads the assembled frame next, so every band of THIS frame must be in place
        (barrier on the raster stream); otherwise the ranks only have to stay within a frame of each other."""
        if world == 1:
            return
        if fused and landed:
            hard_barrier()
        elif fused:
            frame_barrier()
        else:
            gather_bands(fb, H, rank, world, dist, cuts=cuts)

    def step_resident():
        N.check(ctx, L.srb_replay_frame(ctx))
        gather()

    host = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True)
    host_rgb = [torch.empty((H, W, 3), dtype=torch.float32, pin_memory=True) for _ in range(2)]
    e2e_count = [0]

    def step_e2e():
        """Host draw list in, presentable frame out (MainWindow.OnRender's packed RGB float array,
        MainWindow.cs:232-246), every step.  The D2H copy of frame k runs on the library's copy stream while
        frame k+1 is submitted and rendered; two pinned host frames alternate."""
        N.check(ctx, L.srb_begin_frame(ctx, clear, N.SRB_CLEAR_COLOR | N.SRB_CLEAR_DEPTH))
        N.check(ctx, L.srb_submit_draws(ctx, N.ptr(cmds), cmds.shape[0]))
        N.check(ctx, L.srb_end_frame(ctx))
        gather(landed=True)
        if rank == 0:
            N.check(ctx, L.srb_present_rgb_async(ctx, C.c_void_p(host_rgb[e2e_count[0] & 1].data_ptr())))
            e2e_count[0] += 1
        if fused:
            hard_barrier()                      # band ranks must not start the next frame before rank 0 has flattened this one

    host_rgba8 = [torch.empty((H, W, 4), dtype=torch.uint8, pin_memory=True) for _ in range(2)]

    def step_e2e_rgba8():
        """Same loop, the frame delivered at display precision (srb_present_rgba8_async: what the reference's blit
        leaves in the window's 8-bit framebuffer, MainWindow.cs:251-253)."""
        N.check(ctx, L.srb_begin_frame(ctx, clear, N.SRB_CLEAR_COLOR | N.SRB_CLEAR_DEPTH))
        N.check(ctx, L.srb_submit_draws(ctx, N.ptr(cmds), cmds.shape[0]))
        N.check(ctx, L.srb_end_frame(ctx))
        gather(landed=True)
        if rank == 0:
            if rgba8_gather:        # the assembled 8-bit frame is already in rank 0's gather frame: copy it out
                N.check(ctx, L.srb_present_gathered_rgba8_async(ctx, C.c_void_p(host_rgba8[e2e_count[0] & 1].data_ptr())))
            else:
                N.check(ctx, L.srb_present_rgba8_async(ctx, C.c_void_p(host_rgba8[e2e_count[0] & 1].data_ptr())))
            e2e_count[0] += 1
        if fused:
            hard_barrier()

    def step_e2e_colorbuffer():
        """Same, but the result is the float4 ColorBuffer itself, read synchronously (srb_read_color)."""
        N.check(ctx, L.srb_begin_frame(ctx, clear, N.SRB_CLEAR_COLOR | N.SRB_CLEAR_DEPTH))
        N.check(ctx, L.srb_submit_draws(ctx, N.ptr(cmds), cmds.shape[0]))
        N.check(ctx, L.srb_end_frame(ctx))
        gather(landed=True)
        if rank == 0:
            N.check(ctx, L.srb_read_color(ctx, C.c_void_p(host.data_ptr())))
        if fused:
            hard_barrier()

    def timed(step, K, Wm, sampler=None):
        with torch.cuda.stream(ext):
            for _ in range(Wm):
                step()
            torch.cuda.synchronize(dev)
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)
            if sampler:
                sampler.start()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record(ext)
            for _ in range(K):
                step()
            e1.record(ext)
            torch.cuda.synchronize(dev)
            wall = time.perf_counter() - t0
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)
            clocks = sampler.stop() if sampler else None
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms, wall * 1000.0], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1]) / 1000.0
        return ms, wall, clocks

    K, Wm = args.steps, max(args.warmup, 3)
    sampler = ClockSampler(local) if rank == 0 else None
    # Initialisation, not measurement: let the context settle on this frame before anything is timed -- grids sized from
    # the frame before, the extra CTAs of split bins, and the captured geometry chain of every buffer set (a chain is
    # captured the second time a set sees the same shape: 2 x 4 sets).  The W warm-up steps of every timed loop follow.
    with torch.cuda.stream(ext):
        for _ in range(24):
            step_resident()
    torch.cuda.synchronize(dev)
    ms, wall, clocks = timed(step_resident, K, Wm, sampler)
    fps = K / (ms / 1000.0)
    gather_verified = None
    if rgba8_gather:
        # the frame assembled on rank 0 must be every rank's own band, quantised
        full8 = torch.from_numpy(win.GatheredRgba8()).to(dev) if rank == 0 else torch.empty((H, W, 4), dtype=torch.uint8, device=dev)
        dist.broadcast(full8, src=0)
        mine = np.rint(np.clip(win.ColorBuffer[y0:y1], 0.0, 1.0) * np.float32(255.0)).astype(np.uint8)
        oks = [None] * world
        dist.all_gather_object(oks, bool(np.array_equal(full8[y0:y1].cpu().numpy(), mine)))
        gather_verified = all(oks)
        del full8
    # e2e: host buffers in, host frame out, every step (wall clock: it includes host work by design)
    Ke = max(5, min(K, args.e2e_steps))
    if rgba8_gather:      # rank 0 holds only its own band as float4: the float forms of the via-GPU-0 loop do not apply
        ms_e = wall_e = ms_c = wall_c = float("nan")
        fps_e = fps_c = None
    else:
        ms_e, wall_e, _ = timed(step_e2e, Ke, 3)
        fps_e = Ke / max(wall_e, ms_e / 1000.0)          # timed() ends with a device-wide synchronize: the last copies are inside
        N.check(ctx, L.srb_present_wait(ctx))
        ms_c, wall_c, _ = timed(step_e2e_colorbuffer, Ke, 3)
        fps_c = Ke / max(wall_c, ms_c / 1000.0)
    K8 = max(Ke, min(K, 4 * args.e2e_steps))
    ms_8, wall_8, _ = timed(step_e2e_rgba8, K8, 3)
    fps_8 = K8 / max(wall_8, ms_8 / 1000.0)
    N.check(ctx, L.srb_present_wait(ctx))
    rgba8_ok = None
    if rank == 0 and not rgba8_gather:
        cb = win.ColorBuffer
        rgba8_ok = bool(np.array_equal(host_rgba8[(e2e_count[0] - 1) & 1].numpy(),
                                       np.rint(np.clip(cb, 0.0, 1.0) * np.float32(255.0)).astype(np.uint8)))

    # ---- N>1 e2e: every band rank delivers its own rows to ONE shared pinned host frame over its own PCIe
    # link (srb_present_rgb_rows_async); no NVLink gather on this path, the frame is assembled in host memory.
    # A 4-byte all-reduce behind a device-side fence on the previous copies is the "frame delivered" barrier.
    e2e_parallel = None
    if world > 1:
        if fused:
            use_gather(None)                    # final colour into the local buffer only
        # Band cuts for THIS loop: a rank's frame time is its rendering plus the copy of its rows over its own PCIe
        # link, and at 12 B per pixel the copy is ten times the rendering -- the raster-balanced cuts of `value`
        # (rank 0 owns half the rows, all of them sky) would leave one link with half the frame.  Minimise the
        # largest (render + copy) instead: per tile row, the measured full-frame time in proportion to the row's
        # bin load, plus its bytes at the link rate.
        t_render = full_frame_ms * w_rows / max(w_rows.sum(), 1e-9)
        t_copy = np.full(tiles_y, 16 * W * 12 / (args.pcie_gbs * 1e6))
        t_copy[-1] = (H - 16 * (tiles_y - 1)) * W * 12 / (args.pcie_gbs * 1e6)
        cuts_e2e = balanced_cuts(t_render + t_copy, world)
        e0r, e1r = cuts_e2e[rank], cuts_e2e[rank + 1]
        win.SetBand(e0r, e1r)
        S.submit_scene(win, scene)
        win._draws = []
        ye0, ye1 = e0r * 16, min(e1r * 16, H)
        nbytes = 2 * H * W * 12
        shm_path = "/dev/shm/srb200_frame_%s" % os.environ.get("MASTER_PORT", "0")
        if rank == 0:
            with open(shm_path, "wb") as f:
                f.truncate(nbytes)                       # sparse: no page exists yet
        dist.barrier()
        shared = np.memmap(shm_path, dtype=np.float32, mode="r+", shape=(2, H, W, 3))
        # ONE frame in shared host memory; every rank touches ITS rows first, from a CPU next to its GPU, so that on a
        # multi-socket host the pages live on the node its PCIe link is attached to.  (The 8-GPU boxes of this pool
        # are single-node guests -- profiles/r02o_nvidia_smi_topo.txt -- and take 65-100 GB/s from all GPUs together,
        # whatever the placement: that, not the links, bounds the 12 B/pixel frame at N >= 4.)
        numa_cpus = None if args.no_numa_bind else bind_to_gpu_numa(local)
        for b in range(2):
            shared[b, ye0:ye1] = 0.0                     # first touch of this rank's rows
        dist.barrier()
        rc = torch.cuda.cudart().cudaHostRegister(shared.ctypes.data, nbytes, 0)
        assert int(rc) == 0, "cudaHostRegister failed: %s" % rc
        par_count = [0]

        trace = {}

        def tick(name, t0):
            t1 = time.perf_counter()
            trace[name] = trace.get(name, 0.0) + (t1 - t0)
            return t1

        def step_e2e_parallel():
            t = time.perf_counter()
            N.check(ctx, L.srb_begin_frame(ctx, clear, N.SRB_CLEAR_COLOR | N.SRB_CLEAR_DEPTH))
            N.check(ctx, L.srb_submit_draws(ctx, N.ptr(cmds), cmds.shape[0]))
            t = tick("begin+submit", t)
            N.check(ctx, L.srb_end_frame(ctx))
            t = tick("end_frame", t)
            N.check(ctx, L.srb_present_fence(ctx))   # stream waits for the copies of the frames before this one
            hard_barrier()                           # => on every rank: all bands of the previous frame have landed
            t = tick("fence+barrier", t)
            N.check(ctx, L.srb_present_rgb_rows_async(ctx, C.c_void_p(shared[par_count[0] & 1].ctypes.data), ye0, ye1))
            tick("present_rows", t)
            par_count[0] += 1

        # the same frame at display precision, 4 B per pixel: the first third of each float buffer's rows serves as
        # the RGBA8 frame (same registered pages)
        flat8 = [shared[b].reshape(-1).view(np.uint8)[: H * W * 4].reshape(H, W, 4) for b in range(2)]
        par8_count = [0]

        def step_e2e_parallel_rgba8():
            N.check(ctx, L.srb_begin_frame(ctx, clear, N.SRB_CLEAR_COLOR | N.SRB_CLEAR_DEPTH))
            N.check(ctx, L.srb_submit_draws(ctx, N.ptr(cmds), cmds.shape[0]))
            N.check(ctx, L.srb_end_frame(ctx))
            N.check(ctx, L.srb_present_fence(ctx))
            hard_barrier()
            N.check(ctx, L.srb_present_rgba8_rows_async(ctx, C.c_void_p(flat8[par8_count[0] & 1].ctypes.data), ye0, ye1))
            par8_count[0] += 1

        W_par = 10          # the bands changed (D2H-aware cuts): the captured geometry chains of the buffer sets settle again
        ms_p, wall_p, _ = timed(step_e2e_parallel, Ke, W_par)
        N.check(ctx, L.srb_present_wait(ctx))
        dist.barrier()
        fps_p = Ke / max(wall_p, ms_p / 1000.0)
        # the assembled frame must be the frame: compare with this rank's own rows read the slow way
        ok = bool(np.array_equal(shared[(par_count[0] - 1) & 1][ye0:ye1], win.ColorBuffer[ye0:ye1, :, :3]))
        oks = [None] * world
        dist.all_gather_object(oks, ok)
        e2e_parallel_rgba8 = None
        if True:
            K8p = max(Ke, min(K, 4 * args.e2e_steps))
            ms_q, wall_q, _ = timed(step_e2e_parallel_rgba8, K8p, 3)
            N.check(ctx, L.srb_present_wait(ctx))
            dist.barrier()
            want8 = np.rint(np.clip(win.ColorBuffer[ye0:ye1], 0.0, 1.0) * np.float32(255.0)).astype(np.uint8)
            ok8 = bool(np.array_equal(flat8[(par8_count[0] - 1) & 1][ye0:ye1], want8))
            oks8 = [None] * world
            dist.all_gather_object(oks8, ok8)
            e2e_parallel_rgba8 = {"value": K8p / max(wall_q, ms_q / 1000.0), "unit": "frames/s", "steps": K8p,
                                  "d2h_bytes_per_step": int(W * H * 4), "rows_verified": all(oks8),
                                  "call": "srb_present_rgba8_rows_async, every rank its own rows into one shared pinned host frame"}
        e2e_parallel = {"value": fps_p, "unit": "frames/s", "ms_per_step": 1000.0 * max(wall_p, ms_p / 1000.0) / Ke,
                        "rgba8_present_parallel": e2e_parallel_rgba8,
                        "d2h_bytes_per_step": int(W * H * 12), "rows_verified": all(oks),
                        "cuts_tile_rows": [int(x) for x in cuts_e2e], "numa_bound_cpus": numa_cpus,
                        "host_us_per_call_rank0": {k: round(1e6 * v / (Ke + W_par), 1) for k, v in trace.items()},
                        "cuts_policy": "minimise max over ranks of (render + rows x 12 B / %.0f GB/s)" % args.pcie_gbs,
                        "result": "every band rank copies its own rows of the packed RGB float frame into one shared "
                                  "pinned host frame over its own PCIe link (srb_present_rgb_rows_async); no NVLink gather"}
        torch.cuda.cudart().cudaHostUnregister(shared.ctypes.data)
        del flat8
        del shared
        dist.barrier()
        if rank == 0:
            os.unlink(shm_path)
        win.SetBand(r0, r1)                     # back to the cuts and the gather of `value`
        if fused:
            use_gather(args.gather_format)
        S.submit_scene(win, scene)
        win._draws = []

    # ---- single-frame latency (median and p5/p95 over >= 200 frames): one frame at a time,
    # synchronised, device time of begin..end of the pipeline from the library's own CUDA events (no overlap
    # between frames here, unlike the throughput loop above) ----
    lat = []
    for i in range(203):
        N.check(ctx, L.srb_replay_frame(ctx))
        s_ = win.Stats()                         # synchronises; gpu_ms = CUDA events around the whole frame
        if i >= 3:
            lat.append(float(s_.gpu_ms))
    lat = np.array(lat)
    frame_latency = {"median_ms": float(np.median(lat)), "p5_ms": float(np.percentile(lat, 5)),
                     "p95_ms": float(np.percentile(lat, 95)), "frames": int(lat.size),
                     "note": "one synchronised frame at a time (no frame pipelining); this rank's band only"}

    # ---- per-stage device times (CUDA events on the pipeline stream), one synced frame at a time ----
    N.check(ctx, L.srb_set_profiling(ctx, 1))
    Kp = min(K, 50)
    acc = np.zeros(len(STAGES))
    buf = (C.c_float * len(STAGES))()
    for i in range(Kp + 3):
        N.check(ctx, L.srb_replay_frame(ctx))
        N.check(ctx, L.srb_get_stage_ms(ctx, buf, len(STAGES)))
        if i >= 3:
            acc += np.array(list(buf))
    N.check(ctx, L.srb_set_profiling(ctx, 0))
    stage_ms = {n_: float(a / Kp) for n_, a in zip(STAGES, acc)}
    stage_ms_ranks = None
    if world > 1:
        stage_ms_ranks = [None] * world
        dist.all_gather_object(stage_ms_ranks, stage_ms)

    # ---- N>1: the same resident loop once more with the OTHER gather format (reported beside `value`, never in
    # its place): band ranks switch between 4-byte stores into rank 0's RGBA8 frame and 16-byte stores into its
    # float4 ColorBuffer ----
    other_gather = None
    if fused and not args.no_other_gather:
        other = "float4" if rgba8_gather else "rgba8"
        use_gather(other)
        dist.barrier()
        Kx = min(K, 300)
        ms_x, _, _ = timed(step_resident, Kx, 3)
        if other == "rgba8":
            full8 = torch.from_numpy(win.GatheredRgba8()).to(dev) if rank == 0 else torch.empty((H, W, 4), dtype=torch.uint8, device=dev)
            dist.broadcast(full8, src=0)
            mine = np.rint(np.clip(win.ColorBuffer[y0:y1], 0.0, 1.0) * np.float32(255.0)).astype(np.uint8)
            ok_x = bool(np.array_equal(full8[y0:y1].cpu().numpy(), mine))
            del full8
        else:
            # rank 0's ColorBuffer now holds every band: each rank compares its rows of it with a local re-render
            fullc = torch.from_numpy(win.ColorBuffer).to(dev) if rank == 0 else torch.empty((H, W, 4), dtype=torch.float32, device=dev)
            dist.broadcast(fullc, src=0)
            got = fullc[y0:y1].cpu().numpy()
            del fullc
            use_gather(None)
            N.check(ctx, L.srb_replay_frame(ctx))
            ok_x = bool(np.array_equal(got, win.ColorBuffer[y0:y1]))
        oks_x = [None] * world
        dist.all_gather_object(oks_x, ok_x)
        other_gather = {"format": other, "value": Kx / (ms_x / 1000.0), "unit": "frames/s", "steps": Kx, "verified": all(oks_x),
                        "note": "same loop, bands gathered as " + ("display-precision RGBA8 (4 B per pixel into one frame on GPU 0; float4 "
                                "bands stay on the rendering GPU)" if other == "rgba8" else
                                "float4 (16 B per pixel into GPU 0's ColorBuffer: exact API parity; bound by GPU 0's NVLink ingress, "
                                "%.0f MB per frame)" % (W * H * 16 * (world - 1) / world / 1e6))}
        use_gather(args.gather_format)
        dist.barrier()

    # ---- roofline bookkeeping ----
    depth = win.DepthBuffer
    p_cov = int((depth[y0:y1] != np.float32(-3.4028235e38)).sum())
    from softwarerenderer_b200.rasterizer import FrustumCuller
    cullable = [d for d in scene.draws if d.frustum_cull]
    visible = [True] * len(scene.draws)
    if cullable:
        flags = FrustumCuller.TestSpheres(win, [d.mesh.SphereBounds for d in cullable], [d.model for d in cullable],
                                          scene.draws[0].view, scene.draws[0].proj)
        it = iter(flags)
        visible = [bool(next(it)) if d.frustum_cull else True for d in scene.draws]
    balg = algorithmic_bytes(scene, visible, p_cov)
    balg["framebuffer"] = 20 * W * (y1 - y0)
    balg["total"] = balg["geometry"] + balg["texels"] + balg["framebuffer"]
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
    dom = max(stage_ms, key=stage_ms.get)
    dom_bytes = balg["framebuffer"] + balg["texels"] if dom == "tile_raster" else balg["geometry"]
    achieved = dom_bytes / (stage_ms[dom] / 1000.0) / 1e9 if stage_ms[dom] > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath)).get(args.workload, {}).get(dom)
    frame_ms = ms / K
    st = win.Stats()

    line = None
    if rank == 0:
        line = {
            "metric": "frames_per_sec", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": frame_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "mpixels_per_sec": W * H * fps / 1e6,
            "config": base_config(args.workload, scene),
            "config_detail": {"draws_visible": int(sum(visible)), "triangles_setup": int(st.triangles_setup),
                              "triangles_no_coverage": int(st.triangles_no_coverage), "split_parts": int(st.split_parts),
                              "tile_pairs": int(st.tile_pairs), "pipeline": "visibility" if st.pipeline_used == 2 else "ordered",
                              "parallelism": f"bands{world}" if world > 1 else "single",
                              "bands": {"cuts_tile_rows": [int(x) for x in cuts], "policy": args.bands,
                                        "gather": (("RGBA8 (4 B/pixel) " if rgba8_gather else "float4 (16 B/pixel) ") +
                                                   "peer stores from the tile kernel into rank 0's frame over NVLink; frame barrier: "
                                                   + args.frame_barrier) if fused else "NCCL grouped send/recv of float4 bands",
                                        "gather_format": args.gather_format if fused else "float4", "gather_verified": gather_verified,
                                        "cuts_refinement": refine_log}
                              if world > 1 else None},
            "e2e": {"value": fps_e, "unit": "frames/s", "h2d_bytes_per_step": int(cmds.shape[0] * DRAW_REC_BYTES),
                    "d2h_bytes_per_step": int(W * H * 12), "steps": Ke,
                    "ms_per_step": 1000.0 * max(wall_e, ms_e / 1000.0) / Ke if fps_e else None,
                    "mpixels_per_sec": W * H * fps_e / 1e6 if fps_e else None,
                    "result": "packed RGB float frame, the array MainWindow.OnRender uploads (MainWindow.cs:232-246), "
                              "srb_present_rgb_async into two alternating pinned host frames; D2H of frame k overlaps frame k+1",
                    "colorbuffer_float4_sync": {"value": fps_c, "unit": "frames/s", "d2h_bytes_per_step": int(W * H * 16),
                                                "call": "srb_read_color"},
                    "rgba8_present": {"value": fps_8, "unit": "frames/s", "d2h_bytes_per_step": int(W * H * 4), "steps": K8,
                                      "call": "srb_present_rgba8_async", "verified": rgba8_ok,
                                      "result": "the same frame at display precision (clamp, x255, round to nearest: what the "
                                                "reference's blit leaves in the window's 8-bit framebuffer, MainWindow.cs:251-253)"}},
            "e2e_parallel_present": e2e_parallel,     # N>1: promoted to `e2e` below; the gather-to-GPU-0 loop is kept beside it
            "gpu_launches": int(st.kernel_launches) * K,          # timed region of `value`; each e2e step adds the flatten kernel
            "clocks": clocks,
            "value_other_gather": other_gather,
            "frame_latency": frame_latency,
            "stage_ms": stage_ms,
            "stage_ms_ranks": stage_ms_ranks,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "algorithmic_bytes": dom_bytes, "peak_source": peak_src},
            "roofline_frame": {"algorithmic_bytes": balg, "achieved": balg["total"] / (frame_ms / 1000.0) / 1e9,
                               "frac": balg["total"] / (frame_ms / 1000.0) / 1e9 / peak, "unit": "GB/s"},
        }
        if e2e_parallel is not None and e2e_parallel["rows_verified"]:
            via_gpu0 = line["e2e"]
            line["e2e"] = dict(e2e_parallel, h2d_bytes_per_step=via_gpu0["h2d_bytes_per_step"] * world, steps=Ke,
                               mpixels_per_sec=W * H * e2e_parallel["value"] / 1e6)
            line["e2e_via_gpu0_gather"] = via_gpu0
            del line["e2e_parallel_present"]
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(scene)
            fr = line["cpu_baseline"].pop("fragments", None)
            if fr:
                # informational: what an immediate-mode rasterizer moves for the same frame --
                # 4 B per depth-tested fragment + 36 B per written fragment -- against the tiled path's bytes
                imm = 4 * fr["tested"] + 36 * fr["shaded"]
                line["roofline_frame"]["immediate_mode_equivalent_bytes"] = int(imm)
                line["roofline_frame"]["fragments_tested"] = int(fr["tested"])
                line["roofline_frame"]["fragments_shaded_in_submission_order"] = int(fr["shaded"])
    if world > 1:
        if fused:                                # unmap rank 0's frames before rank 0 frees them
            use_gather(None)
            if args.frame_barrier == "peer":
                win.Sync()
                dist.barrier()
                win.BarrierConnect(rank, world, None)
        torch.cuda.synchronize(dev)
        dist.barrier()
    win.Close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line))


def cpu_baseline(scene, budget_s=25.0):
    """Oracle port timed on the box's host cores on a bounded sample (whole frames)."""
    from oracle import pyoracle as O
    O.build()
    L = O.lib()
    cores = os.cpu_count() or 1
    u = O.make_uniforms(scene)
    arr, n, keep = O.make_draws(scene, O.frustum_flags(scene))
    win = L.orc_create(scene.width, scene.height)
    cc = (C.c_float * 4)(*[float(x) for x in scene.clear_color])
    times = []
    t_all = time.perf_counter()
    for i in range(6):
        t0 = time.perf_counter()
        L.orc_clear_frame_mt(win, cc, cores)
        L.orc_render_frame_mt(win, arr, n, C.byref(u), cores, 0, 1 << 30)
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_all > budget_s and len(times) >= 2:
            break
    st = O.OrcStats()
    L.orc_get_stats(win, C.byref(st))            # counters restart at every orc_clear_depth: the last frame's
    # (i) of SURVEY.md 8(d): the literal single-thread walk (reference loop order), one frame, if the budget allows
    single = None
    if float(np.median(times)) * cores < 20.0:
        t0 = time.perf_counter()
        L.orc_clear_depth(win)
        L.orc_clear_color(win, cc)
        for k in range(n):
            L.orc_render_mesh(win, C.byref(arr[k]), C.byref(u))
        single = 1.0 / (time.perf_counter() - t0)
    L.orc_destroy(win)
    meas = times[1:] if len(times) > 1 else times
    med = float(np.median(meas))
    return {"fragments": {"tested": int(st.fragments_tested), "shaded": int(st.fragments_shaded)},
            "value": 1.0 / med, "unit": "frames/s", "cores": cores, "kind": "port",
            "single_thread_literal_value": single,
            "sample": f"{len(meas)} full frame(s) after 1 warm-up, median; C++ restatement of the C# path "
                      "(parallel over triangles, then over runs of 8 tiles, longest list first; bit-identical to the literal oracle); the C# reference cannot run here"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="srb200", choices=["srb200", "reference"])
    ap.add_argument("--workload", default="city4k", choices=sorted(WORKLOADS))
    ap.add_argument("--pipeline", type=int, default=0, help="0 auto, 1 ordered, 2 visibility")
    ap.add_argument("--e2e-steps", type=int, default=100)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gather", default="fused", choices=["fused", "nccl"], help="N>1: how bands reach rank 0")
    ap.add_argument("--gather-format", default="rgba8", choices=["float4", "rgba8"],
                    help="N>1, fused gather: what the band ranks store into rank 0's frame -- the frame at display precision "
                         "(RGBA8, the format of the parity gate; default) or the float4 ColorBuffer (four times the NVLink "
                         "traffic: at 4K and N=8, 116 MB per frame into GPU 0, which caps the rate near 7,000 frames/s)")
    ap.add_argument("--no-other-gather", action="store_true", help="N>1: skip the additional measurement with the other gather format")
    ap.add_argument("--frame-barrier", default="peer", choices=["peer", "side", "inline", "none"],
                    help="N>1, fused gather: where the per-frame 4-byte all-reduce runs (see frame_barrier)")
    ap.add_argument("--no-numa-bind", action="store_true", help="N>1 e2e: do not pin the rank to the CPUs next to its GPU")
    ap.add_argument("--pcie-gbs", type=float, default=50.0, help="N>1 e2e: host link rate assumed by the band cuts of the e2e loop")
    ap.add_argument("--refine-cuts", type=int, default=4,
                    help="N>1, balanced bands: rounds of re-balancing the cuts on the ranks' measured band periods (0: model only)")
    ap.add_argument("--remote-cost", type=float, default=12.0,
                    help="N>1, fused gather: extra cost of a tile whose pixels are stored to rank 0 over NVLink, in bin pairs")
    ap.add_argument("--bands", default="balanced", choices=["balanced", "equal"], help="N>1: band cut policy")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
