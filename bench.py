#!/usr/bin/env python
"""bench.py — headline benchmark of the voxel_volume render pass (BASELINE.json: Mrays/s and ms/frame at 4K on the
synthetic 1024^3 volume; 1/2/4/8 B200).

    python bench.py --gpus 1 --steps K --warmup W              # the CUDA path through the C ABI (libvrt.so)
    torchrun --nproc-per-node N ... bench.py --gpus N ...      # image tiles sharded over N GPUs, frame assembled on rank 0
    python bench.py --impl reference --steps K --warmup W      # the reference algorithm on the host cores (oracle port)

A step = one frame of the camera path of the config (camera and noise frame index change every step).
Prints ONE JSON line (rank 0). The product path never touches oracle/: only the cpu_baseline / parity leg and the
--impl reference arm do.

N > 1: the exchange is the library's own (vrt_comm_* / vrt_frame_exchange in csrc/vrt_comm.cu): every rank stores its tiles straight
into rank 0's frame over NVLink peer memory and, for the end-to-end number, into a pinned host frame shared by all ranks (every GPU's
PCIe link carries its own tiles). torch.distributed (NCCL) is only used here for the barriers around the timed regions and for
reducing the per-rank counters."""
from __future__ import annotations

import argparse
import ctypes as C
import json
import math
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "voxel-raytracing_b200"))

METRIC, UNIT = "Mrays/s", "Mrays/s"
STEP_INSTRUCTIONS = 7          # SASS instructions of one replayed DDA step (profiles/r2_sass_step.txt): FMNMX3, 3 FSETP, 3 predicated FADD


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="vrt", choices=["vrt", "reference"])
    ap.add_argument("--config", default="cfg4")
    ap.add_argument("--whd", type=int, nargs=3, default=None, help="override the synthetic volume size (debug)")
    ap.add_argument("--size", type=int, nargs=2, default=None, help="override the resolution (debug)")
    ap.add_argument("--traversal", type=int, default=0, help="0 default (octant fields), 1 dense u8, 2 brick, 3 fine, 4 linear, 5 octant")
    ap.add_argument("--pipeline", type=int, default=0, help="0 default (wavefront), 1 megakernel, 2 wavefront")
    ap.add_argument("--passes", type=int, default=None, help="override the pass mask (measurement of pass subsets)")
    ap.add_argument("--ao", type=int, default=None, help="override aoSamples")
    ap.add_argument("--no-gbuffer", action="store_true")
    ap.add_argument("--shard", type=int, nargs=2, default=None, metavar=("RANK", "WORLD"),
                    help="debug: render only the tile shard RANK of WORLD on one GPU (what one rank of an N-GPU run does), no exchange")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the cpu_baseline + parity leg (N = 1)")
    ap.add_argument("--frames-in-flight", type=int, default=3, help="frame objects used round-robin (>= 2)")
    ap.add_argument("--serial-frames", action="store_true", help="one stream: frame k+1 starts when frame k has finished (default: frames in flight on their own streams)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--min-timed-ms", type=float, default=1500.0, help="the K-step region is repeated until this much GPU time has been measured; the median region is reported")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50", "-i", str(device)],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        time.sleep(0.1)
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
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"], "samples": 0}
        # median over the upper half: samples taken between regions (idle) would drag a plain median down
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


def ncu_traffic(kernel, config, world, default_path):
    """Per-launch DRAM traffic (dram__bytes_read + write) of the dominant kernel from the committed `ncu --set full` capture of the
    N = 1 launch of THIS config; a rank of an N-GPU run traces 1/N of the rays over the same fields, so its launch is booked at 1/N of
    it (stated in traffic_source). None when no capture exists for the config / code path."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not default_path or not os.path.exists(p):
        return None, None
    try:
        t = json.load(open(p))
    except Exception:
        return None, None
    v = t.get(f"{config}:{kernel.split(' ')[0]}_dram_bytes_per_launch")
    if v is None:
        return None, None
    src = t.get("source", "profiles/traffic.json")
    return (float(v) / world, src + (f"; N = 1 capture scaled by 1/{world} for the tile shard" if world > 1 else ""))


# ----------------------------------------------------------------------------------------------- oracle legs
def host_threads():
    """torch.distributed.run exports OMP_NUM_THREADS=1: the oracle is told how many threads to use explicitly."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


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


def oracle_timed(O, sc, cam, st, row_step, nthreads, planes=("color", "rays", "fetches")):
    H = cam.screenSize[1]
    t0 = time.perf_counter()
    out = O.render(sc, cam, st, rows=(0, H, row_step), nthreads=nthreads, planes=list(planes) if planes else None)
    dt = time.perf_counter() - t0
    return dt, int(out["rays"].sum()), int(out["fetches"].sum()), out


def cpu_baseline_and_parity(cfg, name, whd, size, target_s, gpu_frame0):
    """The oracle port on all host cores over a bounded row sample of frame 0 of the workload; the same rows of the same frame as
    rendered by the CUDA path (instrumented, outside every timed region) are then compared plane by plane: the parity block."""
    O, sc, keep, bounds = oracle_scene_for(cfg, whd)
    st = O.make_settings(pass_mask=cfg["passes"], ao=cfg["ao"])
    cam = oracle_camera(O, name, bounds, size, 0)
    cores = host_threads()
    dt, rays, _, _ = oracle_timed(O, sc, cam, st, 32, cores)                       # calibration: every 32nd row
    step = int(min(32, max(1, round(32 * dt / max(target_s, 1e-3)))))
    dt, rays, _, want = oracle_timed(O, sc, cam, st, step, cores, planes=None)     # the timed sample, all planes kept
    base = {"value": rays / dt / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"frame 0 of the workload, every {step}th row ({len(range(0, size[1], step))} of {size[1]} rows), {dt:.1f} s, oracle port with OpenMP over rows"}
    rows = slice(0, size[1], step)
    got = gpu_frame0()

    def diff_bits(n):
        a, b = np.ascontiguousarray(got[n][rows]), np.ascontiguousarray(want[n][rows])
        if a.dtype.kind == "f":
            same = (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))
        else:
            same = a == b
        return int((~same).sum())

    def max_rel(n):
        a, b = got[n][rows].astype(np.float64), want[n][rows].astype(np.float64)
        ok = np.isfinite(a) & np.isfinite(b)
        if not ok.any():
            return 0.0
        return float((np.abs(a - b)[ok] / np.maximum(np.abs(b[ok]), 1e-30)).max())

    hit_bad = diff_bits("hit")
    par = {"frame": 0, "rows_compared": len(range(0, size[1], step)), "pixels_compared": len(range(0, size[1], step)) * size[0],
           "hit_xyz_steps_mismatch": hit_bad, "palette_mismatch": diff_bits("hit_palette"),
           "rays_mismatch": diff_bits("rays"), "fetches_mismatch": diff_bits("fetches"), "mask_mismatch": diff_bits("mask"),
           "depth_bits_mismatch": diff_bits("depth"), "position_bits_mismatch": diff_bits("position"), "normal_mismatch": diff_bits("normal"),
           "depth_normal_max_rel": max(max_rel("depth"), max_rel("position")),
           "color_max_abs_255": int(np.abs(got["color"][rows].astype(int) - want["color"][rows].astype(int)).max()),
           "color_f32_bits_mismatch_pixels": int((~(((got["color_f32"][rows].view(np.uint32) == want["color_f32"][rows].view(np.uint32))
                                                     | (np.isnan(got["color_f32"][rows]) & np.isnan(want["color_f32"][rows]))).all(axis=-1))).sum()),
           "oracle": "oracle/orc_render.cpp (CPU restatement of shader/voxel_volume.frag; parity unpinned: the reference holds no vectors for this path)"}
    par["ok"] = all(par[k] == 0 for k in ("hit_xyz_steps_mismatch", "palette_mismatch", "rays_mismatch", "fetches_mismatch", "mask_mismatch", "normal_mismatch")) \
        and par["depth_normal_max_rel"] <= 1e-5 and par["color_max_abs_255"] <= 1
    return base, par


def config_blob(cfg, name, whd, size, world, bounds):
    """Identical in both arms (the driver compares the two `config` objects)."""
    kind, what, kw = cfg["scene"]
    vol_mib = bounds[0] * bounds[1] * bounds[2] / 2 ** 20
    fld_mib = 8 * (bounds[0] + 2) * (bounds[1] + 2) * (bounds[2] + 2) / 2 ** 20
    big = vol_mib + fld_mib > 126
    l2 = (f"inputs {'larger' if big else 'SMALLER'} than the 126 MB L2: volume {vol_mib:.0f} MiB + 8 directional safe-step fields {fld_mib:.0f} MiB"
          + ("" if big else f"; the scene is L2-resident by construction, the {size[0] * size[1] * 37 / 2 ** 20:.0f} MiB of frame planes written per step are not")
          + "; camera pose and noise frame index change every step")
    return {"workload": f"{name}: {cfg['label']}", "volume": list(whd) if whd else list(kw.get("whd", [])) or what,
            "resolution": list(size), "pass_mask": cfg["passes"], "aoSamples": cfg["ao"], "denoiser": bool(cfg["denoise"]),
            "parallelism": f"tiles{world}" if world > 1 else "single",
            "l2": l2}


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
    cores = host_threads()
    # calibrate the per-step sample to ~2 s of host work
    dt, _, _, _ = oracle_timed(O, sc, oracle_camera(O, args.config, bounds, size, 0), st, 64, cores)
    step = int(min(64, max(1, round(64 * dt / 2.0))))
    for k in range(args.warmup):
        oracle_timed(O, sc, oracle_camera(O, args.config, bounds, size, k), st, step, cores)
    rays = fetches = 0
    t = 0.0
    for k in range(args.steps):
        dt, r, f, _ = oracle_timed(O, sc, oracle_camera(O, args.config, bounds, size, k), st, step, cores)
        t += dt; rays += r; fetches += f
    val = rays / t / 1e6
    sample = f"every {step}th row of each frame ({len(range(0, size[1], step))} of {size[1]} rows per step), {t:.1f} s total"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_blob(cfg, args.config, args.whd, size, args.gpus, bounds),
            "note": "the reference has no CPU implementation and its Vulkan path is unavailable on this box; this is the CPU restatement of "
                    "shader/voxel_volume.frag (oracle port), OpenMP over rows, on all host cores whatever --gpus says",
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "algorithmic_bytes_per_s": fetches / t}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------- the CUDA arm
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
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = dict(configs.CONFIGS[args.config])
    if args.passes is not None:
        cfg["passes"] = args.passes
    if args.ao is not None:
        cfg["ao"] = args.ao
    size = tuple(args.size) if args.size else cfg["size"]
    W, H = size
    denoise = bool(cfg["denoise"])

    ctx = pyvrt.Context(local)   # raises without a GPU: there is no CPU fallback
    if world > 1:
        key = [f"bench{os.getpid()}_{int(time.time())}" if rank == 0 else None]
        dist.broadcast_object_list(key, src=0)
        ctx.comm_create(rank, world, key[0])                 # the library's own exchange (csrc/vrt_comm.cu); also sets the tile shard
    elif args.shard:
        ctx.set_shard(args.shard[0], args.shard[1])
    scene = configs.build_scene(ctx, cfg, args.whd)
    scene.prepare(args.traversal)
    mem = scene.memory()
    bounds = scene.bounds
    st = configs.settings_for(cfg, gbuffer=not args.no_gbuffer)
    st.traversal = args.traversal
    st.pipeline = args.pipeline
    if world > 1:
        st.denoise.enable = 0                               # sharded ranks trace; rank 0 denoises the assembled planes
    # Frames in flight (the reference keeps two, source/engine/engine.hpp:19): F render passes used round-robin, each frame on
    # a stream of its own so that the kernels of frame k+1 fill the tails of frame k's (--serial-frames: one stream, back to back).
    if not args.serial_frames:
        ctx.set_concurrent_frames(True)
    F = max(2, args.frames_in_flight)
    S = F if world == 1 else min(8, 2 * F)                  # exchange slots on rank 0: more than frames in flight, so that no rank waits for a slot
    passes = tuple(pyvrt.VoxelRenderPass(ctx, scene, st) for _ in range(F))
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    final = "denoised" if denoise else "color"
    frame_bytes = W * H * 4
    xplanes = ("color", "normal", "position") if denoise else ("color",)
    xhost = () if denoise else ("color",)                    # e2e: every rank stores its tiles into the shared pinned host frame
    host2 = dn_out = dn_scratch = cstream = None
    ds = pyvrt.default_settings().denoise
    # cfg3 on N > 1 GPUs: the denoiser is distributed ("strips": every rank denoises the rows it owns, vrt_render_denoise_exchange);
    # VRT_BENCH_ROUTE_A=1 measures route (a) instead (all three planes to rank 0, denoiser there).
    strips = plain_strips = False
    if world > 1:
        if denoise and not os.environ.get("VRT_BENCH_ROUTE_A"):
            ctx.comm_alloc_frames(size, S, xplanes + ("denoised",), ("denoised",))
            try:
                ctx.comm_alloc_strips(ds)
                strips = True
            except pyvrt.VrtError as e:                       # strips thinner than the denoiser's reach (tiny frames): route (a)
                if e.code != pyvrt.E_UNSUPPORTED:
                    raise
        else:
            ctx.comm_alloc_frames(size, S, xplanes, xhost)
            # read-back path (e2e) of a frame without a denoiser: strips carrying the colour plane alone — tiles cross NVLink into row
            # strips, every rank's copy engine writes whole rows into the shared host frame (VRT_BENCH_PUSH_HOST=1: the push kernels instead)
            if not denoise and not os.environ.get("VRT_BENCH_PUSH_HOST"):
                ctx.comm_alloc_strips(None)
                plain_strips = True
        if rank == 0 and denoise:
            dn_out = [torch.empty(frame_bytes, dtype=torch.uint8, device=dev) for _ in range(S)]
            dn_scratch = torch.empty(frame_bytes, dtype=torch.uint8, device=dev)
        if rank == 0:
            cstream = torch.cuda.ExternalStream(ctx.comm_stream, device=dev)
    if rank == 0 and (world == 1 or denoise):
        host2 = [torch.empty(frame_bytes, dtype=torch.uint8, pin_memory=True) for _ in range(S)]
    ev_done = [torch.cuda.Event() for _ in range(S)]
    pending = [False] * S

    def submit(k, readback=False, route_a=False):
        """Queue frame k: render this rank's tiles; N>1: store them into rank 0's frame (and the shared host frame when the frame is
        read back); rank 0 orders its consumer (denoiser, read-back) behind the arrival of every rank's tiles. Returns the frame and
        the kernels launched outside vrt_render / vrt_frame_exchange (the library counts its own)."""
        slot = k % S                                         # exchange slot (N > 1) / read-back buffer
        cam = configs.make_camera(args.config, bounds, size, k)
        if world == 1 or os.environ.get("VRT_BENCH_NOEXCHANGE"):   # (the env knob: what the ranks would do with a free exchange)
            fr = passes[k % F].record(cam)
            if readback and world == 1:
                fr.download_async(final, host2[slot].numpy())
            return fr, 0
        if strips and not route_a:
            fr = passes[k % F].record_denoise_exchange(cam, slot, to_host=readback)
            if rank == 0:
                ctx.comm_wait(slot)                          # every rank's denoised rows are in the slot's denoised plane
                if readback:
                    pending[slot] = True
                else:
                    ctx.comm_release(slot, on_stream=True)
            return fr, 1 if rank == 0 else 0
        if plain_strips and readback:
            fr = passes[k % F].record_denoise_exchange(cam, slot, to_host=2)
            if rank == 0:
                pending[slot] = True
            return fr, 0
        if os.environ.get("VRT_BENCH_UNFUSED"):               # measurement only: separate push kernels after the pass
            fr = passes[k % F].record(cam)
            fr.exchange(slot, xplanes, xhost if readback else ())
        else:                                                 # the pass's own kernels store into rank 0's frame
            fr = passes[k % F].record_exchange(cam, slot, xplanes, xhost if readback else ())
        extra = 0
        if rank == 0:
            ctx.comm_wait(slot)                              # wait kernel on the communicator's stream
            extra = 1
            if denoise:
                dc, _, _ = ctx.comm_frame_plane(slot, "color"); dn, _, _ = ctx.comm_frame_plane(slot, "normal"); dp, _, _ = ctx.comm_frame_plane(slot, "position")
                pyvrt.denoise_device_on(ctx, dc, dn, dp, size, ds, dn_out[slot].data_ptr(), dn_scratch.data_ptr(), stream=ctx.comm_stream)
                extra += ds.iterations
                if readback:
                    with torch.cuda.stream(cstream):
                        host2[slot].copy_(dn_out[slot], non_blocking=True)
            if readback:
                ev_done[slot].record(cstream)
                pending[slot] = True                         # released by wait_readback once the host has the frame
            else:
                ctx.comm_release(slot, on_stream=True)       # the ranks may overwrite the slot once the consumer got there
        return fr, extra

    def drain():
        """Join the frames' streams and the communicator's stream into the context stream (the timed region ends there)."""
        ctx.join()

    def wait_readback(k):
        slot = k % S
        if rank != 0:
            return
        if world == 1:
            passes[k % F].frame.download_wait()
        elif pending[slot]:
            if not denoise or strips:
                ctx.comm_wait_host(slot)                     # every rank's tiles (strips: denoised rows) are in the shared pinned host frame
            if not strips and not plain_strips:
                ev_done[slot].synchronize()
            ctx.comm_release(slot)
            pending[slot] = False

    def barrier():
        ctx.sync(); torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def checksum(ptr, nbytes):
        t = torch.as_tensor(_DevBuf(ptr, nbytes), device=dev).to(torch.int64)
        return torch.stack([t.sum(), (t * t).sum()])

    # ---- self-check of the exchange (outside every timed region): the bytes of the assembled frame are the bytes of the shards
    if world > 1 and not os.environ.get("VRT_BENCH_NOEXCHANGE"):
        fr0, _ = submit(0, route_a=True)
        drain(); barrier()
        ptr, n = fr0.plane_ptr("color")
        mine = checksum(ptr, n)
        dist.all_reduce(mine)
        if rank == 0:
            dptr, _, nb = ctx.comm_frame_plane(0, "color")
            if not torch.equal(checksum(dptr, nb), mine):
                raise SystemExit("exchange self-check failed: the assembled frame does not hold the ranks' tiles")
        barrier()
        if strips:
            # the distributed denoiser against route (a) on the same frame: rank 0 denoised the assembled planes of slot 0 just now
            # (submit(0, route_a=True) above); the strips deliver the same frame into slot 1; the two denoised frames must be the same bytes
            cam0 = configs.make_camera(args.config, bounds, size, 0)
            passes[1 % F].record_denoise_exchange(cam0, 1, to_host=False)
            if rank == 0:
                ctx.comm_wait(1)
            drain(); barrier()
            if rank == 0:
                dptr, _, nb = ctx.comm_frame_plane(1, "denoised")
                if not torch.equal(torch.as_tensor(_DevBuf(dptr, nb), device=dev), dn_out[0]):
                    raise SystemExit("strips self-check failed: the distributed denoiser's frame differs from rank 0's denoise of the assembled planes")
                ctx.comm_release(1)
            barrier()

    # ---- warm-up, then an untimed counting pass over the K timed poses, each frame ALONE on the GPU (rays / voxel visits / kernel
    # times per frame; this is also the single-frame latency)
    for k in range(max(args.warmup, F)):
        submit(k)
    drain()
    barrier()
    rays = fetches = pixels = lib_launches = extra_launches = 0
    ms_kernel, ms_stage, ms_alone = [], [], []
    stage_rays, stage_fetches = np.zeros(3), np.zeros(3)
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for k in range(args.steps):
        barrier()
        with torch.cuda.stream(ext):
            a0.record()
        fr, n = submit(k)
        drain()
        with torch.cuda.stream(ext):
            a1.record()
        s = fr.stats()
        ctx.sync()
        ms_alone.append(a0.elapsed_time(a1))
        rays += s["rays"]; fetches += s["fetches"]; pixels += s["pixels"]; ms_kernel.append(s["ms_trace"] + s["ms_denoise"])
        ms_stage.append(s["ms_stage"]); stage_rays += s["stage_rays"]; stage_fetches += s["stage_fetches"]
        lib_launches += s["kernel_launches"]; extra_launches += n
    barrier()

    # ---- timed region 1: K frames, inputs resident, device time on the launching stream, max over ranks. The region ends when the
    # last frame is assembled on rank 0 (every push, signal and wait kernel is inside it). The region is repeated until at least
    # --min-timed-ms of GPU time has been measured; the median region is reported and the clocks are sampled over the whole span.
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    region_ms, host_ms = [], []
    repeats = 1
    while len(region_ms) < repeats:
        barrier()
        with torch.cuda.stream(ext):
            e0.record()
        t_host = time.perf_counter()
        for k in range(args.steps):
            submit(k)
        drain()
        host_ms.append(1e3 * (time.perf_counter() - t_host))   # host time to QUEUE the K frames (must stay below the device time, else the GPU starves)
        with torch.cuda.stream(ext):
            e1.record()
        barrier()
        mx = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        region_ms.append(float(mx.item()))
        if len(region_ms) == 1:
            repeats = int(min(400, max(1, math.ceil(args.min_timed_ms / max(region_ms[0], 1e-3)))))
    clocks = sampler.stop() if sampler else None
    ms = float(np.median(region_ms))

    # ---- timed region 2 (e2e): the same K frames through the public API with the finished frame in pinned HOST memory EVERY step;
    # per-step inputs (camera block + settings) travel as kernel parameters. N = 1: the read-back of frame k runs on the copy stream
    # beside the kernels of the following frames. N > 1: every rank stores its own tiles into the shared pinned host frame over its
    # own PCIe link (the denoised frame of cfg3 is produced on rank 0 and copied from there). The host waits for frame k-(F-1)
    # before it queues frame k+1 (F frame objects in flight). All K frames are complete in host memory inside the timed region.
    e2e_probe = None
    if args.shard and world == 1:                           # debug mode: a shard is not a frame, there is nothing to read back
        t_e2e = ms * 1e-3
    else:
        # N = 1: a frame object's planes are the source of its read-back, so the host runs F-1 frames ahead. N > 1: the read-back leaves
        # from the exchange slots (S of them, guarded by the slot hand-shake), so the host may run S-2 frames ahead of the finished ones.
        depth = F - 1 if world == 1 else max(F - 1, S - 2)

        def e2e_region():
            barrier()
            t0 = time.perf_counter()
            for k in range(args.steps):
                submit(k, True)
                if k >= depth:
                    wait_readback(k - depth)              # the host runs at most `depth` frames ahead of the finished read-backs
            drain()
            for k in range(max(0, args.steps - depth), args.steps):
                wait_readback(k)
            barrier()
            return time.perf_counter() - t0
        e2e_region()                                       # warm-up of the read-back path
        if plain_strips and not os.environ.get("VRT_BENCH_STRIPS_HOST"):
            # two read-back paths exist for a frame without a denoiser (row strips + copy engines, or push kernels into the host frame);
            # which one is faster depends on N: both are warmed up and timed once here, outside the timed regions, and the faster one runs
            t_strips = e2e_region()
            plain_strips = False
            e2e_region()
            t_push = e2e_region()
            tt = torch.tensor([t_strips, t_push], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            plain_strips = bool(tt[0] <= tt[1])
            e2e_probe = {"strips_ms_per_step": float(tt[0]) * 1e3 / args.steps, "push_ms_per_step": float(tt[1]) * 1e3 / args.steps}
            e2e_region()
        if world > 1 and not os.environ.get("VRT_BENCH_NOEXCHANGE"):
            # self-check of the read-back path (outside the timed regions): the host frame of the last step holds the ranks' bytes.
            # Without a denoiser: byte sums of the ranks' shards of that frame; with the strips' denoiser: the slot's device plane.
            kl = args.steps - 1
            if denoise:
                if rank == 0 and strips:
                    dptr, hview, nb = ctx.comm_frame_plane(kl % S, "denoised")
                    if not np.array_equal(torch.as_tensor(_DevBuf(dptr, nb), device=dev).cpu().numpy(), hview):
                        raise SystemExit("read-back self-check failed: the shared host frame differs from the assembled denoised frame")
            else:
                ptr, n = passes[kl % F].frame.plane_ptr("color")
                mine = checksum(ptr, n)
                dist.all_reduce(mine)
                if rank == 0:
                    _, hview, _ = ctx.comm_frame_plane(kl % S, "color")
                    hv = hview.astype(np.int64)
                    if [int(hv.sum()), int((hv * hv).sum())] != [int(x) for x in mine.tolist()]:
                        raise SystemExit("read-back self-check failed: the shared host frame does not hold the ranks' tiles")
            barrier()
        reps = int(min(200, max(1, math.ceil(0.5 * repeats))))
        t_e2e = float(np.median([e2e_region() for _ in range(reps)]))

    tot = torch.tensor([float(rays), float(fetches), float(pixels)], dtype=torch.float64, device=dev)
    mx = torch.tensor([t_e2e, float(np.median(ms_alone))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    rays_all, fetches_all, pixels_all = [float(x) for x in tot.tolist()]
    t_e2e, latency_ms = [float(x) for x in mx.tolist()]

    if rank == 0:
        peak, peak_src = measured_peaks()
        b_out = 37 if (st.out_mask & pyvrt.OUT_GBUFFER) else 4
        frame_bytes_alg = (fetches + pixels * b_out) / args.steps           # this rank, whole pass, per frame
        k_ms = float(np.mean(ms_kernel))
        # dominant kernel of the pass: per-stage CUDA-event times and per-stage getVoxel counts from the library
        stage_ms = np.mean(np.array(ms_stage), axis=0)
        default_path = args.traversal in (0, 5) and args.pipeline != 1 and os.environ.get("VRT_ENGINE_T", "8") != "0"
        engine = args.traversal in (0, 4, 5) and os.environ.get("VRT_ENGINE_T", "8") != "0"
        names = ["wave_primary_kernel", "wave_bounce_kernel (x3)", "wave_anyhit_engine_kernel" if engine else "wave_anyhit_kernel",
                 "wave_resolve_kernel"] if args.pipeline != 1 else ["render_kernel"] * 4
        dom = int(np.argmax(stage_ms[:3]))
        # algorithmic bytes of that kernel: 1 B per reference getVoxel call + its outputs
        # (primary: colour + G-buffer per pixel; bounce: nothing the reference stores; any-hit: the 1-byte hit flag per ray)
        out_bytes = [pixels * b_out, 0.0, float(stage_rays[2])][dom] if args.pipeline != 1 else pixels * b_out
        alg_bytes = (float(stage_fetches[dom]) + out_bytes) / args.steps
        dom_ms = float(stage_ms[dom])
        achieved = alg_bytes / (dom_ms * 1e-3) / 1e9
        traffic, traffic_src = ncu_traffic(names[dom], args.config, world, default_path and not args.shard)
        info = ctx.device_info()
        f_sm = (clocks or {}).get("sm_max_mhz") or 1965.0
        issue_peak = info["sm_count"] * 4 * 32 * f_sm * 1e6                  # lane-instructions per second the chip can issue
        issue_ach = float(stage_fetches[dom]) / args.steps * STEP_INSTRUCTIONS / (dom_ms * 1e-3)
        line = {"metric": METRIC, "value": rays_all / (ms * 1e-3) / 1e6, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config_blob(cfg, args.config, args.whd, size, world, bounds),
                "run": {"traversal": args.traversal, "pipeline": args.pipeline, "frames_in_flight": F, "exchange_slots": S if world > 1 else None, "frame_streams": "one" if args.serial_frames else "per frame",
                        "timed_regions": len(region_ms), "timed_ms_total": float(np.sum(region_ms)), "region_ms_min_median_max": [float(np.min(region_ms)), ms, float(np.max(region_ms))],
                        "exchange": ("libvrt vrt_render_exchange: the pass's own kernels store " + "+".join(xplanes) + " of every rank's tiles straight into rank 0's frame (CUDA IPC peer memory over NVLink), "
                                     "arrival flags and rank 0's wait are stream memory operations (no kernel)" + ("; rank 0 denoises the assembled planes" if denoise else "")) if world > 1 and not strips else
                                    ("libvrt vrt_render_denoise_exchange (strips): the pass's own kernels store colour+normal+position of every rank's tiles into the buffer of the rank that owns their rows "
                                     "(CUDA IPC peer memory over NVLink); every rank denoises its rows (+ halo) and copies them into rank 0's denoised plane; flags and waits are stream memory operations") if world > 1 else None,
                        "e2e_readback_probe": e2e_probe,
                        "e2e_readback": (("the colour plane crosses NVLink into row strips; every rank's copy engine writes its rows into a pinned host frame shared by all ranks (own PCIe link each, no push kernel)" if plain_strips else
                                          "every rank stores its tiles into a pinned host frame shared by all ranks (own PCIe link each)") if not denoise else ("every rank copies its denoised rows into a pinned host frame shared by all ranks (own PCIe link each)" if strips else "rank 0 copies the denoised frame to pinned host memory")) if world > 1
                                        else "vrt_frame_download_async on the copy stream, F frames in flight"},
                "latency_ms_single_frame": latency_ms,
                "scene_bytes": mem["volume_bytes"] + mem["field_bytes"], "scene": mem,
                "rays_per_frame": rays_all / args.steps, "voxel_visits_per_frame": fetches_all / args.steps,
                "Mrays_per_s_primary": pixels_all / (ms * 1e-3) / 1e6,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                             "kernel": names[dom], "kernel_ms": dom_ms, "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
                             "limiter": "instruction issue (ALU pipe), not memory: see roofline.issue and DESIGN.md section 4",
                             "issue": {"instr_per_step": STEP_INSTRUCTIONS, "steps_per_launch": float(stage_fetches[dom]) / args.steps, "achieved_lane_instr_per_s": issue_ach,
                                       "peak_lane_instr_per_s": issue_peak, "frac": issue_ach / issue_peak,
                                       "note": "reference getVoxel calls the kernel stands for x 7 SASS instructions per replayed DDA step / kernel time, against SMs x 4 schedulers x 32 lanes x max SM clock"},
                             "frame": {"achieved": frame_bytes_alg / (k_ms * 1e-3) / 1e9, "frac": frame_bytes_alg / (k_ms * 1e-3) / 1e9 / peak,
                                       "algorithmic_bytes": frame_bytes_alg, "ms": k_ms},
                             "l2": l2_block(float(stage_fetches[0]) / args.steps + pixels / args.steps * b_out, float(stage_ms[0])),
                             "stages_ms": {"primary": float(stage_ms[0]), "bounce": float(stage_ms[1]), "anyhit": float(stage_ms[2]), "resolve": float(stage_ms[3])},
                             "stages_Mrays": {"primary": float(stage_rays[0] / args.steps / 1e6), "bounce": float(stage_rays[1] / args.steps / 1e6), "anyhit": float(stage_rays[2] / args.steps / 1e6)},
                             "stages_Gvisits": {"primary": float(stage_fetches[0] / args.steps / 1e9), "bounce": float(stage_fetches[1] / args.steps / 1e9), "anyhit": float(stage_fetches[2] / args.steps / 1e9)},
                             "note": "stage times are of frames running ALONE; algorithmic bytes = getVoxel calls of the reference loop x 1 B + the kernel's outputs (primary: %d B/pixel; any-hit: 1 B/ray)" % b_out},
                "e2e": {"value": rays_all / t_e2e / 1e6, "unit": UNIT, "h2d_bytes_per_step": C.sizeof(pyvrt.Camera) + C.sizeof(pyvrt.RenderSettings),
                        "d2h_bytes_per_step": frame_bytes, "ms_per_step": 1e3 * t_e2e / args.steps},
                "gpu_launches": lib_launches + extra_launches, "clocks": clocks, "host_submit_ms_per_step": float(np.median(host_ms)) / args.steps}
        if world == 1 and not args.no_cpu_baseline and not args.shard:
            def gpu_frame0():
                sti = configs.settings_for(cfg, gbuffer=True, instrument=True)
                sti.traversal, sti.pipeline, sti.denoise.enable = args.traversal, args.pipeline, 0
                rp = pyvrt.VoxelRenderPass(ctx, scene, sti)
                fr = rp.record(configs.make_camera(args.config, bounds, size, 0)); fr.sync()
                out = {n: fr.download(n) for n in ("hit", "hit_palette", "depth", "position", "normal", "mask", "rays", "fetches", "color", "color_f32")}
                fr.release()
                return out
            line["cpu_baseline"], line["parity"] = cpu_baseline_and_parity(cfg, args.config, args.whd, size, args.cpu_seconds, gpu_frame0)
        print(json.dumps(line), flush=True)
        if "parity" in line and not line["parity"]["ok"]:
            raise SystemExit("parity check failed: the CUDA path differs from the oracle (see the parity block)")
    barrier()
    # tensors that were used on the library's streams must be returned to torch's allocator while those streams still exist
    host2 = dn_out = dn_scratch = cstream = None
    for p_ in passes:
        p_.frame.release()
    del submit
    import gc
    gc.collect()
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    scene.destroy(); ctx.close()                             # closes the communicator (collective)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


class _DevBuf:
    """Wraps a device pointer owned by libvrt so torch can read it (no copy)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_vrt(a)
