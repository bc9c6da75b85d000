"""Scene ingestion on the GPU box: vrt_scene_load_vox_memory (host parse + device flatten + field build) against the previous route
(host parse + host flatten, vrt_vox_decode, then vrt_scene_from_dense upload + field build), on the floatingcolored stand-in
(8 instances of one 128^3 model -> 256^3 scene). Wall-clock of whole calls (load time, not a per-frame path); min of 5."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "voxel-raytracing_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np
import pyvrt
from vrtlib import cached_vox

ctx = pyvrt.Context(0)
for name in ("floatingcolored", "treehouse"):
    data = cached_vox(name)
    t_dev, t_host_flat, t_dense = [], [], []
    for _ in range(5):
        t0 = time.perf_counter(); gs = pyvrt.VoxelScene.from_vox_bytes(ctx, data); ctx.sync(); t_dev.append(time.perf_counter() - t0)
        bounds = gs.bounds; gs.destroy()
        t0 = time.perf_counter(); vol, pal = pyvrt.vox_decode(data); t_host_flat.append(time.perf_counter() - t0)
        t0 = time.perf_counter(); gs = pyvrt.VoxelScene.from_dense(ctx, vol, pal); ctx.sync(); t_dense.append(time.perf_counter() - t0)
        gs.destroy()
    print(json.dumps(dict(scene=name, bounds=bounds, vox_bytes=len(data), load_vox_device_flatten_ms=min(t_dev) * 1e3,
                          host_parse_flatten_ms=min(t_host_flat) * 1e3, from_dense_upload_build_ms=min(t_dense) * 1e3)), flush=True)
ctx.close()
