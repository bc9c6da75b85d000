#!/usr/bin/env python
"""bench.py — headline benchmark of the voxel_volume render pass (BASELINE.json: Mrays/s and ms/frame at 4K on the
synthetic 1024^3 volume; 1/2/4/8 B200).

    python bench.py --gpus 1 --steps K --warmup W              # the CUDA path through the C ABI (libvrt.so)
    torchrun --nproc-per-node N ... bench.py --gpus N ...      # image tiles sharded over N GPUs, frame gathered on rank 0
    python bench.py --impl reference --steps K --warmup W      # the reference algorithm on the host cores (oracle port)

A step = one frame of the camera path of the config (camera and noise frame index change every step).
Prints ONE JSON line (rank 0). The product path never touches oracle/: only the cpu_baseline leg and the
--impl reference arm do."""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "voxel-raytracing_b200"))

METRIC, UNIT = "Mrays/s", "Mrays/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="vrt", choices=["vrt", "reference"])
    ap.add_argument("--config", default="cfg4")
    ap.add_argument("--whd", type=int, nargs=3, default=None, help="override the synthetic volume size (debug)")
    ap.add_argument("--size", type=int, nargs=2, default=None, help="override the resolution (debug)")
    ap.add_argument("--traversal", type=int, default=0, help="0 default, 1 dense u8, 2 brick")
    ap.add_argument("--pipeline", type=int, default=0, help="0 default (wavefront), 1 megakernel, 2 wavefront")
    ap.add_argument("--passes", type=int, default=None, help="override the pass mask (measurement of pass subsets)")
    ap.add_argument("--ao", type=int, default=None, help="override aoSamples")
    ap.add_argument("--no-gbuffer", action="store_true")
    ap.add_argument("--shard", type=int, nargs=2, default=None, metavar=("RANK", "WORLD"),
                    help="debug: render only the tile shard RANK of WORLD on one GPU (what one rank of an N-GPU run does), no gather")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--frames-in-flight", type=int, default=3, help="frame objects used round-robin (>= 2)")
    ap.add_argument("--exchange", choices=["gather", "push"], default="gather",
                    help="N>1: NCCL gather + de-tile on rank 0, or every rank stores its tiles straight into rank 0's frame over NVLink (symmetric memory)")
    ap.add_argument("--serial-frames", action="store_true", help="one stream: frame k+1 starts when frame k has finished (default: two frames in flight on their own streams)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(device)],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 8:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for n, v in zip(names, c[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # median over the upper half: samples taken between frames (idle) would drag a plain median down
        busy = sorted(sm)[len(sm) // 2:]
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def l2_block(primary_alg_bytes, primary_ms):
    """BASELINE's target is quoted against the L2 roofline for PRIMARY rays: algorithmic bytes of the primary kernel (1 B per reference
    getVoxel call + its per-pixel outputs) / its CUDA-event time, against the L2 read bandwidth measured by tools/l2peak.cu."""
    p = os.path.join(ROOT, "profiles", "r1_l2_peak.json")
    try:
        peak = float(json.load(open(p))["l2_read_gbs"])
    except Exception:
        return None
    ach = primary_alg_bytes / (primary_ms * 1e-3) / 1e9 if primary_ms > 0 else 0.0
    return {"kernel": "wave_primary_kernel", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "peak_source": "measured (tools/l2peak.cu, 64 MiB read-only stream, profiles/r1_l2_peak.json)"}


def ncu_traffic(kernel="render_kernel"):
    """Per-launch DRAM traffic of the dominant kernel from the committed ncu summary, if one exists."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(kernel.split(" ")[0] + "_dram_bytes_per_launch")
        except Exception:
            return None
    return None


# ----------------------------------------------------------------------------------------------- oracle legs
def oracle_scene_for(cfg, whd):
    """CPU-side inputs of the same workload (oracle leg only)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    import synth
    import assets
    kind, what, kw = cfg["scene"]
    if kind == "gen":
        w, h, d = whd or kw["whd"]
        vol = O.generate(what, kw["seed"], w, h, d)
        pal = synth.generator_palette(what, kw["seed"])
    else:
        import configs
        rc, vol, pal = O.load_vox(open(configs.standin_vox_path(what, **kw), "rb").read())
        assert rc == 0
    sc, keep = O.make_scene(vol, pal, assets.noise_standin(), assets.sky_standin(2048, 1024))
    return O, sc, keep, (vol.shape[2], vol.shape[1], vol.shape[0])


def oracle_camera(O, name, bounds, size, k):
    import configs
    pos, yaw, pitch, frame = configs.camera_pose(name, bounds, k)
    focal = float(np.float32(1.0 / np.tan(np.radians(55.0 / 2))))
    d, r, u = O.camera_vectors(yaw, pitch, focal)
    return O.make_camera(pos, d, r, u, bounds, size, frame)


def oracle_timed(O, sc, cam, st, row_step, nthreads=0):
    H = cam.screenSize[1]
    t0 = time.perf_counter()
    out = O.render(sc, cam, st, rows=(0, H, row_step), nthreads=nthreads, planes=["color", "rays", "fetches"])
    dt = time.perf_counter() - t0
    return dt, int(out["rays"].sum()), int(out["fetches"].sum())


def cpu_baseline(cfg, name, whd, size, target_s):
    O, sc, keep, bounds = oracle_scene_for(cfg, whd)
    st = O.make_settings(pass_mask=cfg["passes"], ao=cfg["ao"])
    cam = oracle_camera(O, name, bounds, size, 0)
    cores = O.lib().orc_num_threads()
    dt, rays, _ = oracle_timed(O, sc, cam, st, 32)                       # calibration: every 32nd row
    step = int(min(32, max(1, round(32 * dt / max(target_s, 1e-3)))))
    if step < 32:
        dt, rays, _ = oracle_timed(O, sc, cam, st, step)
    return {"value": rays / dt / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"frame 0 of the workload, every {step}th row ({size[1] // step} of {size[1]} rows), {dt:.1f} s, oracle port with OpenMP over rows"}


def config_blob(cfg, name, whd, size, extra):
    kind, what, kw = cfg["scene"]
    blob = {"workload": f"{name}: {cfg['label']}", "volume": list(whd) if whd else list(kw.get("whd", [])) or what,
            "resolution": list(size), "pass_mask": cfg["passes"], "aoSamples": cfg["ao"], "denoiser": bool(cfg["denoise"]),
            "l2": "inputs larger than L2 (1 GiB volume + 144 MiB brick levels vs 126 MB); camera pose and noise frame index change every step"}
    blob.update(extra)
    return blob


# ----------------------------------------------------------------------------------------------- reference arm
def run_reference(args):
    """The reference algorithm (CPU restatement of voxel_volume.frag: oracle/) on all host cores. Each step renders a
    bounded row sample of that step's frame. Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import configs
    cfg = configs.CONFIGS[args.config]
    size = tuple(args.size) if args.size else cfg["size"]
    O, sc, keep, bounds = oracle_scene_for(cfg, args.whd)
    st = O.make_settings(pass_mask=cfg["passes"], ao=cfg["ao"])
    cores = O.lib().orc_num_threads()
    # calibrate the per-step sample to ~2 s of host work
    dt, _, _ = oracle_timed(O, sc, oracle_camera(O, args.config, bounds, size, 0), st, 64)
    step = int(min(64, max(1, round(64 * dt / 2.0))))
    for k in range(args.warmup):
        oracle_timed(O, sc, oracle_camera(O, args.config, bounds, size, k), st, step)
    rays = fetches = 0
    t = 0.0
    for k in range(args.steps):
        dt, r, f = oracle_timed(O, sc, oracle_camera(O, args.config, bounds, size, k), st, step)
        t += dt; rays += r; fetches += f
    val = rays / t / 1e6
    sample = f"every {step}th row of each frame ({size[1] // step} of {size[1]} rows per step), {t:.1f} s total"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_blob(cfg, args.config, args.whd, size, {"note": "reference has no CPU implementation and its Vulkan path is unavailable on this box; "
                                                                               "this is the CPU restatement of shader/voxel_volume.frag (oracle port), OpenMP over rows"}),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "algorithmic_bytes_per_s": fetches / t}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------- the CUDA arm
class _DevBuf:
    """Wraps a device pointer owned by libvrt so torch.distributed can move it (no copy)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def run_vrt(args):
    import torch
    import torch.distributed as dist
    import pyvrt
    import configs

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torch.distributed.run")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg = dict(configs.CONFIGS[args.config])
    if args.passes is not None:
        cfg["passes"] = args.passes
    if args.ao is not None:
        cfg["ao"] = args.ao
    size = tuple(args.size) if args.size else cfg["size"]
    W, H = size

    if args.shard and world == 1:
        ctx = pyvrt.Context(local, rank=args.shard[0], world=args.shard[1])
    else:
        ctx = pyvrt.Context(local, rank=rank, world=world)   # raises without a GPU: there is no CPU fallback
    scene = configs.build_scene(ctx, cfg, args.whd)
    bounds = scene.bounds
    st = configs.settings_for(cfg, gbuffer=not args.no_gbuffer)
    st.traversal = args.traversal
    st.pipeline = args.pipeline
    # Two frames in flight (as in the reference, source/engine/engine.hpp:19): two render passes used alternately, each frame on
    # a stream of its own so that the kernels of frame k+1 fill the tails of frame k's (--serial-frames: one stream, back to back).
    if not args.serial_frames:
        ctx.set_concurrent_frames(True)
    F = max(2, args.frames_in_flight)
    passes = tuple(pyvrt.VoxelRenderPass(ctx, scene, st) for _ in range(F))
    dev = torch.device("cuda", local)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    final = "denoised" if cfg["denoise"] and world == 1 else "color"
    frame_bytes = W * H * 4
    host2 = [torch.empty(frame_bytes, dtype=torch.uint8, pin_memory=True) for _ in range(F)] if rank == 0 else None
    # N>1: the colour shards of frame k travel to rank 0 (NCCL gather) on a communication stream while the context stream already
    # traces frame k+1; rank 0 assembles frame k (detile kernel, same communication stream) once its shards are in.
    gathered, asm2, comm, copy_stream = [None] * F, None, None, None
    sym = hdl = None
    if world > 1:
        comm = torch.cuda.Stream(device=dev, priority=-1)
        if rank == 0:
            copy_stream = torch.cuda.Stream(device=dev)
        if args.exchange == "push":
            # rank 0's buffer holds the F assembled frames; every rank maps it (peer memory over NVLink / NVSwitch)
            import torch.distributed._symmetric_memory as symm_mem
            sym = symm_mem.empty(F * frame_bytes, dtype=torch.uint8, device=dev)
            hdl = symm_mem.rendezvous(sym, dist.group.WORLD)
            frame0_ptr = int(hdl.buffer_ptrs[0])
            if rank == 0:
                asm2 = [sym[k * frame_bytes:(k + 1) * frame_bytes] for k in range(F)]
        elif rank == 0:
            asm2 = [torch.empty(frame_bytes, dtype=torch.uint8, device="cuda") for _ in range(F)]
    ev_rendered = [torch.cuda.Event() for _ in range(F)]
    ev_gathered = [torch.cuda.Event() for _ in range(F)]
    ev_ready = [torch.cuda.Event() for _ in range(F)]
    ev_copied = [torch.cuda.Event() for _ in range(F)]
    sent, copied_pending = [False] * F, [False] * F
    fstream = [None] * F                                   # the stream each frame object renders on
    last_slot = None

    def submit(k, readback=False):
        """Queue frame k: render this rank's tiles (+ N>1: gather and assemble; + readback: copy the finished frame to pinned host
        memory). Returns the frame and the kernels launched outside vrt_render (the library counts its own)."""
        nonlocal last_slot
        slot = k % F
        cam = configs.make_camera(args.config, bounds, size, k)
        if world == 1:
            fr = passes[slot].record(cam)
            if readback:
                fr.download_async(final, host2[slot].numpy())
            return fr, 0
        if sent[slot]:
            fstream[slot].wait_event(ev_gathered[slot])      # this slot's shard buffer is still being sent
        fr = passes[slot].record(cam)
        if fstream[slot] is None:
            fstream[slot] = torch.cuda.ExternalStream(fr.stream, device=dev)
        ev_rendered[slot].record(fstream[slot])
        ptr, n = fr.plane_ptr("color")
        shard = torch.as_tensor(_DevBuf(ptr, n), device=dev)
        launches = 0
        comm.wait_event(ev_rendered[slot])
        with torch.cuda.stream(comm):
            if rank == 0 and copied_pending[slot]:
                comm.wait_event(ev_copied[slot])             # asm2[slot] is still being read back
            if args.exchange == "push":
                if readback:
                    hdl.barrier(channel=0)                   # rank 0 has finished with what this slot held
                fr.push_tiles("color", frame0_ptr + slot * frame_bytes, stream=comm.cuda_stream)
                hdl.barrier(channel=1)                       # every rank's tiles of frame k are in rank 0's buffer
                launches = 1
            else:
                if rank == 0 and gathered[slot] is None:
                    gathered[slot] = torch.empty((world, n), dtype=torch.uint8, device="cuda")
                dist.gather(shard, list(gathered[slot].unbind(0)) if rank == 0 else None, dst=0)
                if rank == 0:
                    pyvrt.assemble(ctx, gathered[slot].data_ptr(), world, size, "color", asm2[slot].data_ptr(), stream=comm.cuda_stream)
                    launches = 1
            if rank == 0 and readback:
                ev_ready[slot].record(comm)
                copy_stream.wait_event(ev_ready[slot])
                with torch.cuda.stream(copy_stream):
                    host2[slot].copy_(asm2[slot], non_blocking=True)
                    ev_copied[slot].record(copy_stream)
                copied_pending[slot] = True
            ev_gathered[slot].record(comm)                   # shards sent / frame assembled
        sent[slot] = True
        last_slot = slot
        return fr, launches

    def drain():
        """Join the frames' streams and the communication stream into the context stream (the timed region ends there)."""
        ctx.join()
        if world > 1 and last_slot is not None:
            ext.wait_event(ev_gathered[last_slot])
        return 0

    def wait_readback(k):
        slot = k % F
        if rank != 0:
            return
        if world == 1:
            passes[slot].frame.download_wait()
        elif copied_pending[slot]:
            ev_copied[slot].synchronize()

    def barrier():
        ctx.sync(); torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    if world > 1 and args.exchange == "push":
        # self-check of the peer-memory plumbing (outside every timed region): frame 0 assembled by push must equal the same frame
        # assembled by NCCL gather + de-tile
        fr0, _ = submit(0)
        drain(); barrier()
        ptr, n = fr0.plane_ptr("color")
        g = torch.empty((world, n), dtype=torch.uint8, device="cuda") if rank == 0 else None
        dist.gather(torch.as_tensor(_DevBuf(ptr, n), device=dev), list(g.unbind(0)) if rank == 0 else None, dst=0)
        if rank == 0:
            ref = torch.empty(frame_bytes, dtype=torch.uint8, device="cuda")
            pyvrt.assemble(ctx, g.data_ptr(), world, size, "color", ref.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            if not torch.equal(ref, asm2[0]):
                raise SystemExit("push exchange: assembled frame differs from the gathered one")
        barrier()

    # ---- warm-up, then an untimed counting pass over the K timed poses (rays / voxel visits / kernel time per frame)
    for k in range(max(args.warmup, F)):
        submit(k)
    drain()
    barrier()
    rays = fetches = pixels = lib_launches = 0
    ms_kernel, ms_stage = [], []
    stage_rays, stage_fetches = np.zeros(3), np.zeros(3)
    for k in range(args.steps):
        fr, _ = submit(k)
        s = fr.stats()
        rays += s["rays"]; fetches += s["fetches"]; pixels += s["pixels"]; ms_kernel.append(s["ms_trace"] + s["ms_denoise"])
        ms_stage.append(s["ms_stage"]); stage_rays += s["stage_rays"]; stage_fetches += s["stage_fetches"]
        lib_launches += s["kernel_launches"]
    drain()
    barrier()

    # ---- timed region 1: K frames, inputs resident, device time on the launching stream, max over ranks. The region ends when the
    # last frame is assembled on rank 0 (every gather and detile kernel is inside it).
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = lib_launches    # the timed region replays exactly the frames counted above
    barrier()
    with torch.cuda.stream(ext):
        e0.record()
    t_host = time.perf_counter()
    for k in range(args.steps):
        _, n = submit(k)
        launches += n
    launches += drain()
    t_host = time.perf_counter() - t_host        # host time to QUEUE the K frames (must stay below the device time, else the GPU starves)
    with torch.cuda.stream(ext):
        e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None

    # ---- timed region 2 (e2e): the same K frames through the public API with the finished frame read back into pinned host
    # memory EVERY step; per-step inputs (camera block + settings) travel as kernel parameters. The read-back of frame k runs on
    # the copy stream beside the kernels of the following frames, and the host waits for frame k-(F-1) before it queues frame
    # k+1 (F frame objects in flight). All K read-backs complete inside the timed region.
    if args.shard and world == 1:                           # debug mode: a shard is not a frame, there is nothing to read back
        t_e2e = ms * 1e-3
    else:
        for k in range(F):
            submit(k, True)
        drain()
        for k in range(F):
            wait_readback(k)
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            submit(k, True)
            if k >= F - 1:
                wait_readback(k - (F - 1))            # the host runs at most F-1 frames ahead of the finished read-backs
        drain()
        for k in range(max(0, args.steps - (F - 1)), args.steps):
            wait_readback(k)
        barrier()
        t_e2e = time.perf_counter() - t0

    tot = torch.tensor([float(rays), float(fetches), float(pixels)], dtype=torch.float64, device="cuda")
    mx = torch.tensor([ms, t_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tot); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    rays_all, fetches_all, pixels_all = [float(x) for x in tot.tolist()]
    ms, t_e2e = [float(x) for x in mx.tolist()]

    if rank == 0:
        peak, peak_src = measured_peaks()
        b_out = 37 if (st.out_mask & pyvrt.OUT_GBUFFER) else 4
        frame_bytes_alg = (fetches + pixels * b_out) / args.steps           # this rank, whole pass, per frame
        k_ms = float(np.mean(ms_kernel))
        # dominant kernel of the pass: per-stage CUDA-event times and per-stage getVoxel counts from the library
        stage_ms = np.mean(np.array(ms_stage), axis=0)
        engine = args.traversal in (0, 4, 5) and os.environ.get("VRT_ENGINE_T", "8") != "0"
        names = ["wave_primary_kernel", "wave_bounce_kernel (x5)", "wave_anyhit_engine_kernel" if engine else "wave_anyhit_kernel",
                 "wave_resolve_kernel"] if args.pipeline != 1 else ["render_kernel"] * 4
        dom = int(np.argmax(stage_ms[:3]))
        # algorithmic bytes of that kernel: 1 B per reference getVoxel call + its outputs
        # (primary: colour + G-buffer per pixel; bounce: nothing the reference stores; any-hit: the 1-byte hit flag per ray)
        out_bytes = [pixels * b_out, 0.0, float(stage_rays[2])][dom] if args.pipeline != 1 else pixels * b_out
        alg_bytes = (float(stage_fetches[dom]) + out_bytes) / args.steps
        dom_ms = float(stage_ms[dom])
        achieved = alg_bytes / (dom_ms * 1e-3) / 1e9
        line = {"metric": METRIC, "value": rays_all / (ms * 1e-3) / 1e6, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config_blob(cfg, args.config, args.whd, size, {"parallelism": f"tiles{world}" if world > 1 else "single", "traversal": args.traversal, "pipeline": args.pipeline,
                                                                         "gathered_plane": "color RGBA8" if world > 1 else None, "frames_in_flight": F, "frame_streams": "one" if args.serial_frames else "per frame",
                                                                         "exchange": ("NCCL gather of frame k on a communication stream beside the tracing of the next frames, detile on rank 0" if args.exchange == "gather" else "every rank stores its tiles of frame k straight into rank 0's frame (symmetric memory over NVLink) on a communication stream; device-side barriers") if world > 1 else None}),
                "rays_per_frame": rays_all / args.steps, "voxel_visits_per_frame": fetches_all / args.steps,
                "Mrays_per_s_primary": pixels_all / (ms * 1e-3) / 1e6,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic(names[dom]),
                             "kernel": names[dom], "kernel_ms": dom_ms, "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
                             "frame": {"achieved": frame_bytes_alg / (k_ms * 1e-3) / 1e9, "frac": frame_bytes_alg / (k_ms * 1e-3) / 1e9 / peak,
                                       "algorithmic_bytes": frame_bytes_alg, "ms": k_ms},
                             "l2": l2_block(float(stage_fetches[0]) / args.steps + pixels / args.steps * b_out, float(stage_ms[0])),
                             "stages_ms": {"primary": float(stage_ms[0]), "bounce": float(stage_ms[1]), "anyhit": float(stage_ms[2]), "resolve": float(stage_ms[3])},
                             "stages_Mrays": {"primary": float(stage_rays[0] / args.steps / 1e6), "bounce": float(stage_rays[1] / args.steps / 1e6), "anyhit": float(stage_rays[2] / args.steps / 1e6)},
                             "stages_Gvisits": {"primary": float(stage_fetches[0] / args.steps / 1e9), "bounce": float(stage_fetches[1] / args.steps / 1e9), "anyhit": float(stage_fetches[2] / args.steps / 1e9)},
                             "note": "algorithmic bytes = getVoxel calls of the reference loop x 1 B + the kernel's outputs (primary: %d B/pixel; any-hit: 1 B/ray); the loop is issue-bound, not bandwidth-bound (DESIGN.md)" % b_out},
                "e2e": {"value": rays_all / t_e2e / 1e6, "unit": UNIT, "h2d_bytes_per_step": C.sizeof(pyvrt.Camera) + C.sizeof(pyvrt.RenderSettings),
                        "d2h_bytes_per_step": frame_bytes, "ms_per_step": 1e3 * t_e2e / args.steps},
                "gpu_launches": launches, "clocks": clocks, "host_submit_ms_per_step": 1e3 * t_host / args.steps}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(cfg, args.config, args.whd, size, args.cpu_seconds)
        print(json.dumps(line), flush=True)
    barrier()
    # tensors that were used on the context's stream must be returned to torch's allocator while that stream still exists
    gathered = host2 = asm2 = sym = hdl = None
    for p_ in passes[1:]:
        p_.frame.release()
    del submit
    import gc
    gc.collect()
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    scene.destroy(); ctx.close()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_vrt(a)
