"""GPU tier, BASELINE.json's five configs AT THEIR STATED SIZES against the CPU oracle: the same scene, camera
path pose, pass mask and frame index that bench.py times. Hit voxel + step count, palette index, depth, position, normal, mask and
the per-pixel ray / getVoxel counts are compared bit for bit, colour within 1/255 (float colour bit for bit up to 1e-4 of the pixels).
cfg5 follows SURVEY.md 8(d): frames k in {0, 85, 170} of the orbit, every 8th row of the 7680x4320 frame, at world = 1 and with the
eight tile shards of an 8-GPU run rendered one after another on this GPU and assembled by direct tile stores."""
import ctypes as C
import os
import numpy as np
import pytest

from vrtlib import O, assets, bits_equal, cached_vox
import pyvrt
import configs
import synth
from test_gpu import vrt_settings, cam_to_py, check, EXACT

pytestmark = pytest.mark.gpu

FOCAL = float(np.float32(1.0 / np.tan(np.radians(55.0 / 2))))


def oracle_cam(name, bounds, size, k):
    pos, yaw, pitch, frame = configs.camera_pose(name, bounds, k)
    d, r, u = O.camera_vectors(yaw, pitch, FOCAL)
    return O.make_camera(pos, d, r, u, bounds, size, frame)


def scene_pair(gpu_ctx, name):
    """The scene of BASELINE config `name`, on the GPU exactly as bench.py builds it and on the host for the oracle."""
    cfg = configs.CONFIGS[name]
    kind, what, kw = cfg["scene"]
    gs = configs.build_scene(gpu_ctx, cfg)
    if kind == "gen":
        w, h, d = kw["whd"]
        vol = O.generate(what, kw["seed"], w, h, d)
        pal = synth.generator_palette(what, kw["seed"])
    else:
        rc, vol, pal = O.load_vox(open(configs.standin_vox_path(what, **kw), "rb").read())
        assert rc == 0
    assert gs.bounds == (vol.shape[2], vol.shape[1], vol.shape[0])
    assert np.array_equal(gs.palette().view(np.uint32), np.ascontiguousarray(pal, np.float32).view(np.uint32))
    sc, keep = O.make_scene(vol, pal, assets.noise_standin(), assets.sky_standin(2048, 1024))
    return cfg, gs, sc, keep, vol


def mem_available_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable"):
                return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


def gpu_frame(gpu_ctx, gs, cam, st, denoise=None):
    rp = pyvrt.VoxelRenderPass(gpu_ctx, gs, vrt_settings(st, denoise=denoise))
    fr = rp.record(cam_to_py(cam)); fr.sync()
    out = {n: fr.download(n) for n, _, _ in O.PLANE_SPECS}
    if denoise is not None:
        out["denoised"] = fr.download("denoised")
    out["_stats"] = fr.stats()
    fr.release()
    return out


def test_cfg2_mandelbulb_256_at_1080p(gpu_ctx):
    cfg, gs, sc, keep, vol = scene_pair(gpu_ctx, "cfg2")
    assert gs.bounds == (256, 256, 256) and np.array_equal(gs.download_volume(), vol)
    st = O.make_settings(pass_mask=cfg["passes"], ao=cfg["ao"])
    for k in (0, 37):
        cam = oracle_cam("cfg2", gs.bounds, cfg["size"], k)
        want = O.render(sc, cam, st)
        assert 0.1 < (want["hit_palette"] != 0).mean() < 1.0
        check(gpu_frame(gpu_ctx, gs, cam, st), want)
    gs.destroy()


@pytest.mark.parametrize("mode", [0, 1])
def test_cfg3_floating_4k_with_denoiser(gpu_ctx, mode):
    cfg, gs, sc, keep, vol = scene_pair(gpu_ctx, "cfg3")
    st = O.make_settings(pass_mask=cfg["passes"], ao=cfg["ao"])
    cam = oracle_cam("cfg3", gs.bounds, cfg["size"], 0)
    assert tuple(cam.screenSize) == (3840, 2160)
    want = O.render(sc, cam, st)
    got = gpu_frame(gpu_ctx, gs, cam, st, denoise=dict(iterations=2, mode=mode))
    check(got, want)
    assert want["rays"].max() > 4                                          # reflection stacks at 4K too
    # the filter itself: the oracle's denoiser over the GPU's own G-buffer (expf is the only non-shared transcendental)
    dn = O.denoise(got["color"], got["normal"], got["position"], iterations=2, mode=mode)
    d = np.abs(dn.astype(int) - got["denoised"].astype(int))
    assert d.max() <= 1 and (d != 0).mean() < 1e-3
    # and end to end: oracle traversal + oracle denoiser vs the denoised plane of the product, within 1/255
    dn2 = O.denoise(want["color"], want["normal"], want["position"], iterations=2, mode=mode)
    assert np.abs(dn2.astype(int) - got["denoised"].astype(int)).max() <= 2
    gs.destroy()


def test_cfg4_terrain_1024_full_frame_4k(gpu_ctx):
    cfg, gs, sc, keep, vol = scene_pair(gpu_ctx, "cfg4")
    assert gs.bounds == (1024, 1024, 1024) and np.array_equal(gs.download_volume(), vol)
    st = O.make_settings(pass_mask=cfg["passes"], ao=cfg["ao"])
    cam = oracle_cam("cfg4", gs.bounds, cfg["size"], 0)
    want = O.render(sc, cam, st)                                            # the full 3840x2160 frame
    got = gpu_frame(gpu_ctx, gs, cam, st)
    check(got, want)
    hit = want["hit_palette"] != 0
    assert 0.3 < hit.mean() and want["hit"][..., 3].max() == 512 and want["hit"][..., :3].max() > 700     # far cells, exhausted rays
    mem = gs.memory()
    assert mem["volume_bytes"] == 1 << 30 and mem["field_bytes"] == 8 * 1026 ** 3                        # nothing but the default back-end's fields
    gs.destroy()


def test_cfg5_sparse_2048_orbit_8k_sampled_rows_world1_and_world8(gpu_ctx):
    import torch
    cfg, gs, sc, keep, vol = scene_pair(gpu_ctx, "cfg5")
    assert gs.bounds == (2048, 2048, 2048)
    if mem_available_gb() > 24:
        assert np.array_equal(gs.download_volume(), vol)                   # 8 GiB: the generator's 64-bit addressing
    else:
        pytest.skip("not enough host memory for the 8 GiB volume comparison")
    st = O.make_settings(pass_mask=cfg["passes"], ao=cfg["ao"])
    size = cfg["size"]
    assert size == (7680, 4320)
    W, H = size
    rows = slice(0, H, 8)
    names = [n for n, _, _ in O.PLANE_SPECS]
    for k in (0, 85, 170):
        cam = oracle_cam("cfg5", gs.bounds, size, k)
        want = O.render(sc, cam, st, rows=(0, H, 8))
        want = {n: v[rows] for n, v in want.items()}
        # ---- world = 1
        rp = pyvrt.VoxelRenderPass(gpu_ctx, gs, vrt_settings(st))
        fr = rp.record(cam_to_py(cam)); fr.sync()
        got = {n: fr.download(n)[rows] for n in names}
        fr.release()
        for n in EXACT:
            ok, bad = bits_equal(got[n], want[n]); assert ok, (k, n, bad)
        assert np.abs(got["color"].astype(int) - want["color"].astype(int)).max() <= 1
        same = (got["color_f32"].view(np.uint32) == want["color_f32"].view(np.uint32)) | (np.isnan(got["color_f32"]) & np.isnan(want["color_f32"]))
        assert 1.0 - same.all(axis=-1).mean() <= 1e-4
        assert want["hit"][..., :3].max() > 1200 and (want["hit_palette"] != 0).mean() > 0.03
        # ---- world = 8: the eight shards of an 8-GPU run, one after another, assembled by direct tile stores (vrt_frame_push_tiles)
        planes = ("color", "depth", "normal", "hit", "hit_palette", "position")
        bpp = {"color": 4, "depth": 4, "normal": 4, "hit": 8, "hit_palette": 1, "position": 16}
        full = {n: torch.zeros(W * H * bpp[n], dtype=torch.uint8, device="cuda") for n in planes}
        torch.cuda.synchronize()
        try:
            for r in range(8):
                gpu_ctx.set_shard(r, 8)
                rp = pyvrt.VoxelRenderPass(gpu_ctx, gs, vrt_settings(st))
                fr = rp.record(cam_to_py(cam))
                for n in planes:
                    fr.push_tiles(n, full[n].data_ptr())
                fr.sync(); fr.release()
        finally:
            gpu_ctx.set_shard(0, 1)
        for n in planes:
            _, dt, ch = pyvrt.PLANES[pyvrt.PLANE_ID[n]]
            a = full[n].cpu().numpy().view(dt).reshape((H, W, ch) if ch > 1 else (H, W))[rows]
            if n == "color":
                assert np.array_equal(a, got[n]), (k, n)                   # N-rank frame == 1-rank frame, byte for byte
            else:
                ok, bad = bits_equal(a, want[n]); assert ok, (k, "world8", n, bad)
        del full
    mem = gs.memory()                                                     # the two-level fields, not 8 x 8.6 GiB of linear ones
    assert mem["field_bytes"] < 8 * 2 ** 30 and mem["volume_bytes"] + mem["field_bytes"] < 40 * 2 ** 30, mem
    gs.destroy()
