"""N>1 path on CPU (gloo, world size 2 and 3): every rank renders its interleaved tile shard with the product's own
per-pixel code and layout functions (host-compiled: tests/host_harness), the colour shards are gathered on rank 0 with
torch.distributed, assembled with the product's index functions, and must be byte-identical to the 1-rank frame: tile
ownership, compact tile-major indices and the inverse mapping, on the default traversal back-end and on FINE. (The
GPU-side exchange itself — vrt_comm_* over CUDA IPC — needs a device: tests/test_host_cpp.py, GPU tier.)"""
import ctypes as C
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vrtlib import O, assets, harness, harness_render, look_at_camera, random_volume, random_palette, to_vrt_camera, to_vrt_settings, TRAV_FINE, TRAV_OCTANT

TILE = 32


def _inputs():
    rng = np.random.default_rng(11)
    whd = (40, 36, 44)
    vol, pal = random_volume(rng, whd, 0.03), random_palette(rng, metallic_every=4)
    cam = look_at_camera(whd, (200, 120), (-15, 50, -20), (20, 18, 22), frame=5)   # ragged: 200x120 is not a multiple of 32
    return vol, pal, cam, O.make_settings(ao=2)


def _worker(rank, world, port, outdir, trav):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        vol, pal, cam, st = _inputs()
        noise, sky = assets.noise_standin(), assets.sky_standin(64, 32)
        sc, keep = O.make_scene(vol, pal, noise, sky)
        h = harness()
        hs = h.hh_scene_create(C.c_void_p(sc.volume), sc.w, sc.h, sc.d, C.c_void_p(sc.palette), C.c_void_p(sc.noise), C.c_void_p(sc.sky), sc.skyW, sc.skyH)
        W, H = cam.screenSize[0], cam.screenSize[1]
        tiles = ((W + TILE - 1) // TILE) * ((H + TILE - 1) // TILE)
        shard_px = ((tiles + world - 1) // world) * TILE * TILE
        shard = np.zeros(shard_px * 4, np.uint8)
        vs, vc = to_vrt_settings(st), to_vrt_camera(cam)
        h.hh_render_shard.restype = C.c_int
        rc = h.hh_render_shard(C.c_void_p(hs), C.byref(vc), C.byref(vs), trav, rank, world, shard.ctypes.data_as(C.c_void_p), C.c_size_t(shard_px))
        assert rc == 0
        t = torch.from_numpy(shard)
        if rank == 0:
            parts = [torch.empty_like(t) for _ in range(world)]
            dist.gather(t, parts, dst=0)
            gathered = torch.cat(parts).numpy()
            out = np.zeros((H, W, 4), np.uint8)
            h.hh_assemble(gathered.ctypes.data_as(C.c_void_p), world, W, H, C.c_size_t(shard_px), out.ctypes.data_as(C.c_void_p))
            np.save(os.path.join(outdir, "frame.npy"), out)
        else:
            dist.gather(t, None, dst=0)
        h.hh_scene_destroy(C.c_void_p(hs))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,trav", [(2, TRAV_OCTANT), (3, TRAV_OCTANT), (2, TRAV_FINE)])
def test_gloo_sharded_frame_is_byte_identical(world, trav, tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    mp.spawn(_worker, args=(world, port, str(tmp_path), trav), nprocs=world, join=True)
    got = np.load(tmp_path / "frame.npy")
    vol, pal, cam, st = _inputs()
    sc, keep = O.make_scene(vol, pal, assets.noise_standin(), assets.sky_standin(64, 32))
    want = harness_render(sc, cam, st, traversal=trav)["color"]
    assert np.array_equal(got, want)
    assert np.array_equal(want, O.render(sc, cam, st)["color"]) or np.abs(want.astype(int) - O.render(sc, cam, st)["color"].astype(int)).max() <= 1


def test_strip_layout_of_the_distributed_denoiser():
    """vrt_comm_alloc_strips / vrt_render_denoise_exchange: rank q denoises rows [q*SR, (q+1)*SR) and holds rows [q*SR - halo, (q+1)*SR + halo)
    clipped to the image. The kernels store a row into the buffers found by strip_for_each_holder (csrc/vrt_types.h), which only looks at
    the row's owner and its neighbours: for every halo <= SR that must be exactly the set of ALL ranks whose range holds the row, with the
    right row index, every buffer row must be written exactly once, and no index may leave the buffer (SR + 2*halo rows)."""
    import ctypes as C
    from vrtlib import harness
    h = harness()
    q, r = (C.c_int * 4)(), (C.c_int * 4)()
    for H, world in [(2160, 8), (2160, 2), (217, 3), (40, 8), (360, 4), (1, 1), (7, 16), (4320, 8), (150, 2)]:
        SR = (H + world - 1) // world
        for halo in sorted({0, 1, 6, SR // 2, SR}):
            if halo > SR:
                continue
            seen = {}
            for py in range(H):
                n = h.hh_strip_holders(py, world, SR, halo, H, q, r)
                assert 1 <= n <= 3
                got = {(q[i], r[i]) for i in range(n)}
                want = set()
                for k in range(world):
                    lo, hi = max(0, k * SR - halo), min(H, (k + 1) * SR + halo)
                    if lo <= py < hi:
                        want.add((k, py - lo))
                assert got == want, (H, world, halo, py, got, want)
                for k, row in got:
                    assert 0 <= row < SR + 2 * halo
                    assert (k, row) not in seen
                    seen[(k, row)] = py
            # every rank that owns rows got all the rows of its range
            for k in range(world):
                lo, hi = max(0, k * SR - halo), min(H, (k + 1) * SR + halo)
                assert sum(1 for (kk, _) in seen if kk == k) == max(0, hi - lo)


@pytest.mark.parametrize("iterations,step_width,mode", [(1, 2.0, 0), (2, 2.0, 0), (2, 2.0, 1), (3, 1.5, 1), (3, 0.5, 0), (4, 2.5, 1), (2, 5.0, 0), (3, 0.0, 1)])
def test_strip_halo_is_enough_for_the_denoiser(iterations, step_width, mode):
    """The distributed denoiser (strips) runs the denoiser on rows [y0 - halo, y1 + halo) of the inputs and keeps rows [y0, y1): with
    halo = vrt_denoise_reach_rows(settings) those rows must be the rows of the whole-frame result, exactly — here with the oracle's
    denoiser on random planes, for strips inside the image and at its top and bottom (where the buffer ends with the image)."""
    import pyvrt
    import oracle as O
    ds = pyvrt.DenoiserSettings(1, iterations, 20.4, 1e-2, 1e-1, step_width, mode)
    halo = pyvrt.denoise_reach_rows(ds)
    assert 0 < halo <= 64
    rng = np.random.default_rng(11)
    H, W = 3 * 23 + 2 * 40, 57
    color = rng.integers(0, 256, (H, W, 4), dtype=np.uint8); color[..., 3] = 0
    normal = np.zeros((H, W, 4), np.int8); normal[..., :3] = rng.choice([-127, 0, 127], (H, W, 3))
    pos = (rng.random((H, W, 4), dtype=np.float32) * 3).astype(np.float32); pos[..., 3] = 0
    kw = dict(iterations=iterations, step_width=step_width, mode=mode)
    full = O.denoise(color, normal, pos, **kw)
    for y0, y1 in [(0, 23), (60, 83), (H - 23, H), (40, 41)]:
        lo, hi = max(0, y0 - halo), min(H, y1 + halo)
        part = O.denoise(color[lo:hi], normal[lo:hi], pos[lo:hi], **kw)
        assert np.array_equal(part[y0 - lo:y1 - lo], full[y0:y1]), (y0, y1, halo)
    # and it is not a wild over-estimate: one row less than the taps' plain reach breaks it somewhere (integral step widths)
    if float(step_width).is_integer() and step_width > 0:
        reach = int(sum(i * step_width + 1 for i in range(iterations)))
        assert reach <= halo <= reach + iterations
