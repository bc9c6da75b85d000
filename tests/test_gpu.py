"""GPU tier (B200): the CUDA path, called through the C ABI (libvrt.so via pyvrt), against the CPU oracle on the
same seeded inputs. Bar (BASELINE.json): hit voxel, palette index, step count bit-exact; depth / position within
1e-5 relative (we assert bit-exact); colour within 1/255 per channel (we assert bit-exact float colour except for
at most 1e-4 of pixels, where the binary64 atan2/asin of the sky lookup may round differently)."""
import ctypes as C
import os
import numpy as np
import pytest

from vrtlib import (O, assets, TRAV_BRICK, TRAV_DENSE, TRAV_FINE, ALL_TRAVERSALS, harness, default_camera, look_at_camera, random_volume, random_palette,
                    bits_equal, cached_vox, to_vrt_camera)
import pyvrt
import synth

pytestmark = pytest.mark.gpu

PIPELINES = (pyvrt.PIPELINE_MEGAKERNEL, pyvrt.PIPELINE_WAVEFRONT)
EXACT = ["depth", "motion", "mask", "position", "normal", "hit", "hit_palette", "rays", "fetches"]


def vrt_settings(st: O.Settings, traversal=0, out_mask=3, denoise=None, pipeline=0):
    s = pyvrt.default_settings()
    s.pass_mask, s.out_mask, s.traversal, s.pipeline = st.pass_mask, out_mask, traversal, pipeline
    s.vol.aoSamples, s.vol.ambientIntensity = st.aoSamples, st.ambientIntensity
    s.light.direction[:] = st.lightDir[:]; s.light.intensity = st.lightIntensity; s.light.color[:] = st.lightColor[:]
    s.denoise.enable = 0
    if denoise is not None:
        s.denoise.enable, s.denoise.iterations, s.denoise.mode = 1, denoise.get("iterations", 2), denoise.get("mode", 0)
    return s


def cam_to_py(cam: O.Camera) -> pyvrt.Camera:
    c = pyvrt.Camera(); C.memmove(C.byref(c), C.byref(cam), C.sizeof(cam)); return c


def gpu_render(ctx, scene, cam, st, traversal=0, denoise=None, names=None, pipeline=0):
    rp = pyvrt.VoxelRenderPass(ctx, scene, vrt_settings(st, traversal, denoise=denoise, pipeline=pipeline))
    fr = rp.record(cam_to_py(cam)); fr.sync()
    names = names or [n for n, _, _ in O.PLANE_SPECS] + (["denoised"] if denoise is not None else [])
    out = {n: fr.download(n) for n in names}
    out["_stats"] = fr.stats()
    fr.release()
    return out


def check(got, want, color_outlier_frac=1e-4):
    for n in EXACT:
        ok, bad = bits_equal(got[n], want[n])
        assert ok, f"plane {n}: {bad} of {got[n].size} differ"
    diff = np.abs(got["color"].astype(int) - want["color"].astype(int))
    assert diff.max() <= 1, f"colour differs by {diff.max()}/255"
    same = (got["color_f32"].view(np.uint32) == want["color_f32"].view(np.uint32)) | (np.isnan(got["color_f32"]) & np.isnan(want["color_f32"]))
    frac = 1.0 - same.all(axis=-1).mean()
    assert frac <= color_outlier_frac, f"float colour differs on {frac:.2e} of pixels"
    assert int(want["rays"].sum()) == got["_stats"]["rays"] and int(want["fetches"].sum()) == got["_stats"]["fetches"]


def make_pair(ctx, vol, pal, sky=(256, 128)):
    noise, skyt = assets.noise_standin(), assets.sky_standin(*sky)
    sc, keep = O.make_scene(vol, pal, noise, skyt)
    gs = pyvrt.VoxelScene.from_dense(ctx, vol, pal); gs.set_noise(noise); gs.set_skybox(skyt)
    return sc, keep, gs


# ------------------------------------------------------------------------------------------- BASELINE configs
def test_cfg1_treehouse_standin_primary_720p(gpu_ctx):
    data = cached_vox("treehouse")
    rc, vol, pal = O.load_vox(data)
    gs = pyvrt.VoxelScene.from_vox_bytes(gpu_ctx, data)
    assert gs.bounds == (126, 126, 126) and np.array_equal(gs.download_volume(), vol) and np.array_equal(gs.palette(), pal)
    noise, sky = assets.noise_standin(), assets.sky_standin(256, 128)
    gs.set_noise(noise); gs.set_skybox(sky)
    sc, keep = O.make_scene(vol, pal, noise, sky)
    cam = default_camera((126, 126, 126), (1280, 720))
    st = O.make_settings(pass_mask=O.PASS_PRIMARY)
    want = O.render(sc, cam, st)
    for trav in ALL_TRAVERSALS:
        for pipe in PIPELINES:
            check(gpu_render(gpu_ctx, gs, cam, st, trav, pipeline=pipe), want)
    gs.destroy()


def test_cfg2_mandelbulb_standin_shadow_1080p(gpu_ctx):
    data = cached_vox("mandelbulb", n=128)
    rc, vol, pal = O.load_vox(data)
    sc, keep, gs = make_pair(gpu_ctx, vol, pal)
    cam = look_at_camera(gs.bounds, (1920, 1080), (-40, 150, -90), (64, 64, 64))
    st = O.make_settings(pass_mask=O.PASS_PRIMARY | O.PASS_SHADOW | O.PASS_SHADE, ao=0)
    want = O.render(sc, cam, st)
    for pipe in PIPELINES:
        check(gpu_render(gpu_ctx, gs, cam, st, pipeline=pipe), want)
    gs.destroy()


@pytest.mark.parametrize("mode", [0, 1])
def test_cfg3_floating_standin_ao_shadow_denoise(gpu_ctx, mode):
    rc, vol, pal = O.load_vox(cached_vox("floatingcolored"))
    sc, keep, gs = make_pair(gpu_ctx, vol, pal, sky=(2048, 1024))
    cam = look_at_camera(gs.bounds, (960, 540), (-60, 330, -80), (132, 132, 132), frame=3)
    st = O.make_settings(ao=1)
    want = O.render(sc, cam, st)
    check(gpu_render(gpu_ctx, gs, cam, st, pipeline=pyvrt.PIPELINE_MEGAKERNEL), want)
    got = gpu_render(gpu_ctx, gs, cam, st, denoise=dict(iterations=2, mode=mode))
    check(got, want)
    assert want["rays"].max() > 8                           # metallic palette entries produced reflection stacks
    # denoise the GPU's own G-buffer with the oracle: isolates the filter (expf is within 2 ulp of glibc's)
    dn = O.denoise(got["color"], got["normal"], got["position"], iterations=2, mode=mode)
    d = np.abs(dn.astype(int) - got["denoised"].astype(int))
    assert d.max() <= 1 and (d != 0).mean() < 1e-3
    gs.destroy()


@pytest.mark.parametrize("gen,whd,pos,target", [(1, (256, 192, 256), (20.3, 150.2, 12.1), (128, 90, 128)),
                                                (2, (256, 256, 256), (-30.5, 200.7, -40.2), (128, 128, 128))])
def test_cfg4_cfg5_synthetic_generators_and_full_pass(gpu_ctx, gen, whd, pos, target):
    gs = pyvrt.VoxelScene.generate(gpu_ctx, gen, 42, *whd)
    vol = synth.GENERATORS[gen](42, *whd)
    assert np.array_equal(gs.download_volume(), vol)
    pal = synth.generator_palette(gen, 42)
    assert np.array_equal(gs.palette().view(np.uint32), pal.view(np.uint32))
    noise, sky = assets.noise_standin(), assets.sky_standin(256, 128)
    gs.set_noise(noise); gs.set_skybox(sky)
    sc, keep = O.make_scene(vol, pal, noise, sky)
    cam = look_at_camera(whd, (640, 360), pos, target, frame=9)
    st = O.make_settings()
    want = O.render(sc, cam, st)
    assert 0.05 < (want["hit_palette"] != 0).mean() and want["rays"].max() > 6
    for trav in ALL_TRAVERSALS:
        for pipe in PIPELINES:
            check(gpu_render(gpu_ctx, gs, cam, st, trav, pipeline=pipe), want)
    gs.destroy()


# ------------------------------------------------------------------------------------------- structure + properties
def test_brick_levels_match_cpu_builder(gpu_ctx):
    rng = np.random.default_rng(21)
    for whd in ((50, 37, 61), (64, 64, 64), (9, 130, 5), (128, 20, 36)):
        vol = random_volume(rng, whd, 0.01)
        pal = random_palette(rng)
        gs = pyvrt.VoxelScene.from_dense(gpu_ctx, vol, pal)
        m, d = gs.download_accel()
        h = harness()
        hs = h.hh_scene_create(vol.ctypes.data_as(C.c_void_p), whd[0], whd[1], whd[2], pal.ctypes.data_as(C.c_void_p), None, None, 0, 0)
        m2, d2 = np.zeros_like(m), np.zeros_like(d)
        h.hh_scene_bmask(C.c_void_p(hs), m2.ctypes.data_as(C.c_void_p)); h.hh_scene_bdist(C.c_void_p(hs), d2.ctypes.data_as(C.c_void_p))
        f2 = np.zeros(m.shape + (4, 4, 4), np.uint8)
        h.hh_scene_fdist(C.c_void_p(hs), f2.ctypes.data_as(C.c_void_p))
        assert np.array_equal(m, m2) and np.array_equal(d, d2)
        assert np.array_equal(gs.download_fdist(), f2)
        for o in range(8):                                    # directional fields: bit-volume build == the CPU dynamic programme
            o2 = np.zeros((whd[2] + 2, whd[1] + 2, whd[0] + 2), np.uint8)
            h.hh_scene_odist(C.c_void_p(hs), o, o2.ctypes.data_as(C.c_void_p))
            assert np.array_equal(gs.download_odist(o), o2), f"odist octant {o} differs for {whd}"
            s2 = np.zeros_like(o2)                            # two-level layout: the GPU's classification + fill == the CPU restatement
            h.hh_scene_sdist(C.c_void_p(hs), o, s2.ctypes.data_as(C.c_void_p))
            sd = gs.download_sdist(o)
            assert np.array_equal(sd, s2), f"sparse field octant {o} differs for {whd}"
            dec = lambda a: np.where(a < 128, a.astype(np.int32), a.astype(np.int32) * 4 - 384)
            inside = (o2 != 0) & (o2 != 255)
            assert np.array_equal(sd == 0, o2 == 0) and np.array_equal(sd == 255, o2 == 255) and (dec(sd)[inside] <= dec(o2)[inside]).all() and (dec(sd)[inside] >= 1).all()
        h.hh_scene_destroy(C.c_void_p(hs))
        gs.destroy()


def test_dense_and_brick_traversal_identical_at_4k(gpu_ctx):
    """Size-independent property at BASELINE's full resolution: the acceleration changes no output byte."""
    gs = pyvrt.VoxelScene.generate(gpu_ctx, 1, 42, 512, 256, 512)
    gs.set_noise(assets.noise_standin()); gs.set_skybox(assets.sky_standin(512, 256))
    cam = look_at_camera(gs.bounds, (3840, 2160), (30.3, 190.2, 25.1), (256, 120, 256), frame=1)
    st = O.make_settings()
    a = gpu_render(gpu_ctx, gs, cam, st, TRAV_DENSE, pipeline=pyvrt.PIPELINE_MEGAKERNEL)
    for trav, pipe in ((TRAV_BRICK, pyvrt.PIPELINE_MEGAKERNEL), (TRAV_FINE, pyvrt.PIPELINE_MEGAKERNEL), (TRAV_FINE, pyvrt.PIPELINE_WAVEFRONT),
                       (TRAV_DENSE, pyvrt.PIPELINE_WAVEFRONT), (4, pyvrt.PIPELINE_WAVEFRONT), (5, pyvrt.PIPELINE_WAVEFRONT), (5, pyvrt.PIPELINE_MEGAKERNEL),
                       (6, pyvrt.PIPELINE_WAVEFRONT), (6, pyvrt.PIPELINE_MEGAKERNEL)):
        b = gpu_render(gpu_ctx, gs, cam, st, trav, pipeline=pipe)
        for n, _, _ in O.PLANE_SPECS:
            ok, bad = bits_equal(a[n], b[n]); assert ok, (trav, pipe, n, bad)
        assert a["_stats"]["rays"] == b["_stats"]["rays"] and a["_stats"]["fetches"] == b["_stats"]["fetches"]
    hit = a["hit_palette"] != 0
    assert 0.2 < hit.mean() < 1.0 and a["hit"][..., 3].max() <= 512
    x, y, z = a["hit"][..., 0][hit].astype(int), a["hit"][..., 1][hit].astype(int), a["hit"][..., 2][hit].astype(int)
    sub = slice(None, None, 997)
    vol = gs.download_volume()
    assert np.array_equal(vol[z[sub], y[sub], x[sub]], a["hit_palette"][hit][sub])
    gs.destroy()


def test_production_frames_equal_instrumented_frames(gpu_ctx):
    """Frames without VRT_OUT_INSTRUMENT (what bench.py times): colour and every G-buffer plane equal the oracle's / the instrumented
    frame's on every pipeline, and the ray and getVoxel totals of the frame statistics are the oracle's exactly."""
    gs = pyvrt.VoxelScene.generate(gpu_ctx, 1, 42, 256, 192, 256)
    noise, sky = assets.noise_standin(), assets.sky_standin(512, 256)
    gs.set_noise(noise); gs.set_skybox(sky)
    vol = O.generate(1, 42, 256, 192, 256)
    sc, keep = O.make_scene(vol, synth.generator_palette(1, 42), noise, sky)
    cam = look_at_camera(gs.bounds, (640, 360), (20.3, 150.2, 12.1), (128, 90, 128), frame=7)
    st = O.make_settings()
    want = O.render(sc, cam, st)
    names = ["color", "depth", "motion", "mask", "position", "normal"]
    for trav in (4, 5, 6):
        for pipe in PIPELINES:
            rp = pyvrt.VoxelRenderPass(gpu_ctx, gs, vrt_settings(st, trav, out_mask=pyvrt.OUT_GBUFFER, pipeline=pipe))
            fr = rp.record(cam_to_py(cam)); fr.sync()
            got = {n: fr.download(n) for n in names}
            stats = fr.stats(); fr.release()
            for n in names[1:]:
                ok, bad = bits_equal(got[n], want[n]); assert ok, (trav, pipe, n, bad)
            assert np.array_equal(got["color"], want["color"]) or np.abs(got["color"].astype(int) - want["color"].astype(int)).max() <= 1
            assert stats["rays"] == int(want["rays"].sum())
            assert stats["fetches"] == int(want["fetches"].sum())
    # full resolution, sharp edge: production frame == instrumented frame, byte for byte
    big = pyvrt.VoxelScene.generate(gpu_ctx, 1, 42, 512, 256, 512)
    big.set_noise(noise); big.set_skybox(sky)
    cam = look_at_camera(big.bounds, (3840, 2160), (30.3, 190.2, 25.1), (256, 120, 256), frame=1)
    a = gpu_render(gpu_ctx, big, cam, st, names=names)
    rp = pyvrt.VoxelRenderPass(gpu_ctx, big, vrt_settings(st, 0, out_mask=pyvrt.OUT_GBUFFER))
    fr = rp.record(cam_to_py(cam)); fr.sync()
    for n in names:
        ok, bad = bits_equal(a[n], fr.download(n)); assert ok, (n, bad)
    fr.release(); gs.destroy(); big.destroy()


def test_two_frames_in_flight_readback(gpu_ctx):
    """vrt_frame_download_async: frame k is read back on the copy stream while frame k+1 renders; bytes equal the blocking download."""
    import torch
    gs = pyvrt.VoxelScene.generate(gpu_ctx, 2, 1234, 192, 192, 192)
    gs.set_noise(assets.noise_standin()); gs.set_skybox(assets.sky_standin(256, 128))
    st = pyvrt.default_settings(); st.denoise.enable = 0
    passes = [pyvrt.VoxelRenderPass(gpu_ctx, gs, st) for _ in range(2)]
    host = [torch.empty(480 * 270 * 4, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
    cams = [cam_to_py(look_at_camera(gs.bounds, (480, 270), (-40 + 9 * k, 230, -60), (96, 96, 96), frame=k)) for k in range(6)]
    want = []
    for c in cams:
        fr = passes[0].record(c); want.append(fr.download("color").copy())
    got = []
    for k, c in enumerate(cams):
        fr = passes[k & 1].record(c)
        fr.download_async("color", host[k & 1].numpy())
        if k > 0:
            passes[(k - 1) & 1].frame.download_wait(); got.append(host[(k - 1) & 1].numpy().reshape(270, 480, 4).copy())
    passes[(len(cams) - 1) & 1].frame.download_wait(); got.append(host[(len(cams) - 1) & 1].numpy().reshape(270, 480, 4).copy())
    for a, b in zip(got, want):
        assert np.array_equal(a, b)
    for p in passes:
        p.frame.release()
    gs.destroy()


def test_concurrent_frames_are_byte_identical(gpu_ctx):
    """vrt_context_set_concurrent_frames: two frame objects render on streams of their own and overlap on the GPU; every frame
    must equal the frame rendered alone (per-frame buffers, counters and side streams: nothing is shared between frames)."""
    gs = pyvrt.VoxelScene.generate(gpu_ctx, 1, 42, 160, 160, 160)
    gs.set_noise(assets.noise_standin()); gs.set_skybox(assets.sky_standin(256, 128))
    st = pyvrt.default_settings(); st.denoise.enable = 0
    cams = [cam_to_py(look_at_camera(gs.bounds, (512, 288), (-30 + 7 * k, 200, -50), (80, 80, 80), frame=k)) for k in range(8)]
    solo = pyvrt.VoxelRenderPass(gpu_ctx, gs, st)
    want = [solo.record(c).download("color").copy() for c in cams]
    stats = solo.frame.stats()
    gpu_ctx.set_concurrent_frames(True)
    try:
        passes = [pyvrt.VoxelRenderPass(gpu_ctx, gs, st) for _ in range(2)]
        got = []
        for k in range(0, len(cams), 2):
            frames = [passes[j].record(cams[k + j]) for j in range(2)]           # both queued before either is read
            assert frames[0].stream != frames[1].stream and frames[0].stream != gpu_ctx.stream
            got += [frames[0].download("color").copy(), frames[1].download("color").copy()]
        gpu_ctx.join(); gpu_ctx.sync()
        assert passes[(len(cams) - 1) & 1].frame.stats()["rays"] == stats["rays"]
        for p in passes:
            p.frame.release()
    finally:
        gpu_ctx.set_concurrent_frames(False)
    assert len(got) == len(want)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)
    solo.frame.release(); gs.destroy()


def test_degenerate_rays_on_gpu(gpu_ctx):
    rng = np.random.default_rng(9)
    vol = random_volume(rng, (24, 24, 24), 0.04); pal = random_palette(rng)
    sc, keep, gs = make_pair(gpu_ctx, vol, pal)
    st = O.make_settings(light_dir=np.array([0, 1, 0], np.float32))     # zero components: literal inf/NaN path
    cam = look_at_camera((24, 24, 24), (128, 96), (-8, 30, -9), (12, 12, 12))
    for pipe in PIPELINES:
        check(gpu_render(gpu_ctx, gs, cam, st, pipeline=pipe), O.render(sc, cam, st))
    gs.destroy()
    solid = np.full((12, 12, 12), 5, np.uint8)
    sc, keep, gs = make_pair(gpu_ctx, solid, pal)
    cam = default_camera((12, 12, 12), (64, 40), pos=(6.5, 6.5, 6.5))       # camera inside solid: step-0 hits, NaN normals
    for pipe in PIPELINES:
        check(gpu_render(gpu_ctx, gs, cam, O.make_settings(), pipeline=pipe), O.render(sc, cam, O.make_settings()))
    gs.destroy()


def test_empty_and_ragged_sizes(gpu_ctx):
    pal = random_palette(np.random.default_rng(1))
    sc, keep, gs = make_pair(gpu_ctx, np.zeros((5, 6, 7), np.uint8), pal)
    cam = default_camera((7, 6, 5), (33, 17), pos=(3, 3, -10))          # not a multiple of the 8x4 warp patch or the 32x32 tile
    for pipe in PIPELINES:
        got = gpu_render(gpu_ctx, gs, cam, O.make_settings(), pipeline=pipe)
        check(got, O.render(sc, cam, O.make_settings()))
        assert (got["hit_palette"] == 0).all() and got["_stats"]["pixels"] == 33 * 17
    gs.destroy()


@pytest.mark.parametrize("world", [2, 3, 8])
def test_tile_sharded_frame_is_byte_identical(gpu_ctx, world):
    """N ranks emulated one after another on one GPU (separate contexts); gather = concatenation, then vrt_assemble."""
    import torch
    rng = np.random.default_rng(31)
    vol = random_volume(rng, (60, 40, 60), 0.02); pal = random_palette(rng)
    noise, sky = assets.noise_standin(), assets.sky_standin(256, 128)
    size = (300, 170)
    cam = look_at_camera((60, 40, 60), size, (-20, 55, -25), (30, 20, 30))
    st = O.make_settings()
    gs = pyvrt.VoxelScene.from_dense(gpu_ctx, vol, pal); gs.set_noise(noise); gs.set_skybox(sky)
    full = gpu_render(gpu_ctx, gs, cam, st, names=["color", "position", "hit", "hit_palette"])
    gs.destroy()
    shards = {n: [] for n in ("color", "position", "hit", "hit_palette")}
    total = dict(rays=0, fetches=0, pixels=0)
    pushed = {n: torch.zeros(full[n].nbytes, dtype=torch.uint8, device="cuda") for n in shards}
    torch.cuda.synchronize()
    for r in range(world):
        ctx = pyvrt.Context(0, rank=r, world=world)
        s = pyvrt.VoxelScene.from_dense(ctx, vol, pal); s.set_noise(noise); s.set_skybox(sky)
        rp = pyvrt.VoxelRenderPass(ctx, s, vrt_settings(st)); fr = rp.record(cam_to_py(cam)); fr.sync()
        for n in shards:
            shards[n].append(fr.download(n))
        stt = fr.stats()
        for k in total:
            total[k] += stt[k]
        # the collective-free assembly: every rank stores its tiles straight into the (here: same-GPU) row-major frame
        for n in pushed:
            fr.push_tiles(n, pushed[n].data_ptr())
        fr.sync()
        fr.release(); s.destroy(); ctx.close()
    assert total["pixels"] == size[0] * size[1] and total["rays"] == full["_stats"]["rays"] and total["fetches"] == full["_stats"]["fetches"]
    for n in pushed:
        assert np.array_equal(pushed[n].cpu().numpy(), full[n].view(np.uint8).ravel()), "push " + n
    for n in shards:
        gathered = torch.from_numpy(np.concatenate([x.view(np.uint8).ravel() for x in shards[n]])).cuda()
        _, dt, ch = pyvrt.PLANES[pyvrt.PLANE_ID[n]]
        out = torch.empty(size[0] * size[1] * ch * np.dtype(dt).itemsize, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        pyvrt.assemble(gpu_ctx, gathered.data_ptr(), world, size, n, out.data_ptr())
        gpu_ctx.sync()
        res = out.cpu().numpy()
        assert np.array_equal(res, full[n].view(np.uint8).ravel()), n


def test_standalone_denoise_matches_oracle(gpu_ctx):
    import torch
    rng = np.random.default_rng(7)
    h, w = 270, 480
    color = rng.integers(0, 256, (h, w, 4), dtype=np.uint8); color[..., 3] = 0
    normal = np.zeros((h, w, 4), np.int8); normal[..., :3] = rng.choice([-127, 0, 127], (h, w, 3))
    pos = (rng.random((h, w, 4), dtype=np.float32) * 3); pos[..., 3] = 0
    for mode in (0, 1):
        for iters in (1, 2, 3):
            ds = pyvrt.DenoiserSettings(1, iters, 20.4, 1e-2, 1e-1, 2.0, mode)
            tc, tn, tp = torch.from_numpy(color).cuda(), torch.from_numpy(normal).cuda(), torch.from_numpy(pos).cuda()
            out = torch.empty_like(tc)
            torch.cuda.synchronize()
            pyvrt.denoise_device(gpu_ctx, tc.data_ptr(), tn.data_ptr(), tp.data_ptr(), (w, h), ds, out.data_ptr())
            gpu_ctx.sync()
            want = O.denoise(color, normal, pos, iterations=iters, mode=mode)
            d = np.abs(out.cpu().numpy().astype(int) - want.astype(int))
            assert d.max() <= 1 and (d != 0).mean() < 2e-3, (mode, iters, d.max(), (d != 0).mean())


@pytest.mark.parametrize("sizes", [(480, 270, 480, 270), (1129, 635, 1920, 1080), (1920, 1080, 1280, 720), (333, 217, 1001, 64), (1, 1, 50, 40),
                                   (3840, 2160, 2560, 1440)])
def test_blit_matches_oracle_bit_exact(gpu_ctx, sizes):
    """Output stage (BlitStage::record / blit.frag) through the C ABI against the oracle: every byte equal, UNORM and sRGB stores."""
    import torch
    sw, sh, tw, th = sizes
    rng = np.random.default_rng(sw + 7 * tw)
    src = rng.integers(0, 256, (sh, sw, 4), dtype=np.uint8)
    tsrc = torch.from_numpy(src).cuda()
    for flags in (0, pyvrt.BLIT_SRGB, pyvrt.BLIT_SRGB | pyvrt.BLIT_BGRA):
        out = torch.zeros((th, tw, 4), dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        pyvrt.blit_device(gpu_ctx, tsrc.data_ptr(), (sw, sh), out.data_ptr(), (tw, th), flags)
        gpu_ctx.sync()
        want = O.blit(src, (tw, th), srgb=bool(flags & pyvrt.BLIT_SRGB))
        if flags & pyvrt.BLIT_BGRA:
            want = want[..., [2, 1, 0, 3]]
        got = out.cpu().numpy()
        assert np.array_equal(got, want), (sizes, flags, int((got != want).sum()))


def test_frame_to_window_jittered_sequence(gpu_ctx, tmp_path):
    """The callers either side of the pass, end to end on the device: UpscalerStage.update drives frame / cameraJitter, the pass renders at
    the FSR render resolution, the blit presents at the target resolution; every stage equals the oracle's."""
    import torch
    rng = np.random.default_rng(11)
    whd = (40, 36, 44)
    vol, pal = random_volume(rng, whd, 0.03), random_palette(rng, metallic_every=4)
    noise, sky = assets.noise_standin(), assets.sky_standin(128, 64)
    target = (320, 180)
    rres = pyvrt.render_resolution(True, pyvrt.FSR_BALANCED, target)
    assert rres == O.render_resolution(True, 17, target)
    sc, keep = O.make_scene(vol, pal, noise, sky)
    st = O.make_settings()
    gs = pyvrt.VoxelScene.from_dense(gpu_ctx, vol, pal); gs.set_noise(noise); gs.set_skybox(sky)
    s = pyvrt.default_settings(); s.out_mask = pyvrt.OUT_GBUFFER; s.denoise.enable = 0
    s.light.direction[:] = st.lightDir[:]
    rp = pyvrt.VoxelRenderPass(gpu_ctx, gs, s)
    up = pyvrt.UpscalerStage(rres[0], target[0])
    for k in range(3):
        up.update(0.016)
        cam = look_at_camera(whd, rres, (-15, 45, -20), (20, 18, 22), frame=up.frameCount)
        cam.cameraJitter[:] = [float(up.jitterX), float(up.jitterY)]
        want = O.render(sc, cam, st)
        c = pyvrt.Camera(); C.memmove(C.byref(c), C.byref(cam), C.sizeof(cam))
        fr = rp.record(c); fr.sync()
        color = fr.download("color")
        assert np.abs(color.astype(int) - want["color"].astype(int)).max() <= 1
        out = torch.zeros((target[1], target[0], 4), dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        pyvrt.blit_device(gpu_ctx, fr.plane_ptr("color")[0], rres, out.data_ptr(), target, pyvrt.BLIT_SRGB)
        gpu_ctx.sync()
        assert np.array_equal(out.cpu().numpy(), O.blit(color, target, srgb=True))
    pyvrt.image_write(str(tmp_path / "frame.png"), out.cpu().numpy(), pyvrt.IMAGE_PNG_RGB)
    assert os.path.getsize(str(tmp_path / "frame.png")) > target[0] * target[1] * 3
    fr.release(); gs.destroy()


@pytest.mark.parametrize("name", ["lshape", "twoinst", "nopalette", "overlap9"])
def test_device_flatten_matches_reference_golden(gpu_ctx, name):
    """Scene ingestion on the GPU (csrc/vrt_flatten.cu: every model uploaded once, one kernel per instance in the reference's order)
    against the volumes the REFERENCE's own parser + flatten produced for .vox files written by the reference's own writer."""
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    data = open(os.path.join(gold, name + ".vox"), "rb").read()
    ref = np.load(os.path.join(gold, name + "_ref.npz"))
    gs = pyvrt.VoxelScene.from_vox_bytes(gpu_ctx, data)
    vol = gs.download_volume()
    assert vol.shape == ref["volume"].shape and np.array_equal(vol, ref["volume"])
    assert np.array_equal(gs.palette().view(np.uint32), ref["palette"].view(np.uint32))
    v2, _ = pyvrt.vox_decode(data)                       # the host flatten (vrt_vox_decode) agrees with the device flatten
    assert np.array_equal(v2, vol)
    gs.destroy()


def test_device_flatten_multi_instance_standin(gpu_ctx):
    data = cached_vox("floatingcolored")                  # 2x2x2 instances of one 128^3 model: uploaded once, flattened 8 times
    rc, vol, pal = O.load_vox(data)
    gs = pyvrt.VoxelScene.from_vox_bytes(gpu_ctx, data)
    assert np.array_equal(gs.download_volume(), vol) and np.array_equal(gs.palette(), pal)
    gs.destroy()


def test_sharded_frame_denoised_on_assembling_rank_is_byte_identical(gpu_ctx):
    """SURVEY 8(e), denoiser under sharding, route (a): the ranks trace their tiles, colour / normal / position are assembled on rank 0
    (here by direct tile stores, as over NVLink), and the denoiser runs there on the whole frame: same bytes as the single-GPU frame."""
    import torch
    rng = np.random.default_rng(77)
    whd = (56, 44, 52)
    vol = random_volume(rng, whd, 0.03); pal = random_palette(rng, metallic_every=4)
    noise, sky = assets.noise_standin(), assets.sky_standin(128, 64)
    size = (200, 150)                                    # ragged in 32 x 32 tiles
    cam = look_at_camera(whd, size, (-18, 50, -22), (28, 20, 26), frame=4)
    st = O.make_settings()
    gs = pyvrt.VoxelScene.from_dense(gpu_ctx, vol, pal); gs.set_noise(noise); gs.set_skybox(sky)
    single = gpu_render(gpu_ctx, gs, cam, st, denoise=dict(iterations=2, mode=0), names=["color", "normal", "position", "denoised"])
    gs.destroy()
    world = 4
    names = ("color", "normal", "position")
    frame = {n: torch.zeros(single[n].nbytes, dtype=torch.uint8, device="cuda") for n in names}
    torch.cuda.synchronize()
    for r in range(world):
        ctx = pyvrt.Context(0, rank=r, world=world)
        s = pyvrt.VoxelScene.from_dense(ctx, vol, pal); s.set_noise(noise); s.set_skybox(sky)
        rp = pyvrt.VoxelRenderPass(ctx, s, vrt_settings(st, out_mask=pyvrt.OUT_GBUFFER)); fr = rp.record(cam_to_py(cam))
        for n in names:
            fr.push_tiles(n, frame[n].data_ptr())
        fr.sync(); fr.release(); s.destroy(); ctx.close()
    for n in names:
        assert np.array_equal(frame[n].cpu().numpy(), single[n].view(np.uint8).ravel()), n
    out = torch.zeros(size[0] * size[1] * 4, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    ds = pyvrt.DenoiserSettings(1, 2, 20.4, 1e-2, 1e-1, 2.0, 0)
    pyvrt.denoise_device(gpu_ctx, frame["color"].data_ptr(), frame["normal"].data_ptr(), frame["position"].data_ptr(), size, ds, out.data_ptr())
    gpu_ctx.sync()
    assert np.array_equal(out.cpu().numpy(), single["denoised"].view(np.uint8).ravel())
    # a sharded context refuses to denoise its own shard (taps cross tile borders): the error says where the denoiser belongs
    ctx = pyvrt.Context(0, rank=1, world=2)
    s = pyvrt.VoxelScene.from_dense(ctx, vol, pal); s.set_noise(noise); s.set_skybox(sky)
    with pytest.raises(pyvrt.VrtError) as e:
        pyvrt.VoxelRenderPass(ctx, s, vrt_settings(st, denoise=dict(iterations=2))).record(cam_to_py(cam))
    assert "rank 0" in str(e.value)
    s.destroy(); ctx.close()


def test_denoiser_iteration0_edge_blind_kernel_is_exact(gpu_ctx):
    """Iteration 0 (phi = +inf) runs the edge-blind kernel: no exp, so the result equals the oracle's byte for byte, also where the
    G-buffer holds inf / NaN / huge positions (whose weights are 1 all the same)."""
    import torch
    from vrtlib import blind_blur
    rng = np.random.default_rng(23)
    h, w = 270, 481
    color = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    normal = rng.integers(-127, 128, (h, w, 4)).astype(np.int8)
    pos = (rng.random((h, w, 4), dtype=np.float32) * 1e3).astype(np.float32)
    pos[3, 4, 0] = np.inf; pos[10, 11, 1] = np.nan; pos[20, 30, 2] = 3e38; pos[21, 30, 2] = -3e38; pos[0, 0, :] = -np.inf
    tc, tn, tp = torch.from_numpy(color).cuda(), torch.from_numpy(normal).cuda(), torch.from_numpy(pos).cuda()
    for mode in (0, 1):
        out = torch.zeros_like(tc)
        torch.cuda.synchronize()
        pyvrt.denoise_device(gpu_ctx, tc.data_ptr(), tn.data_ptr(), tp.data_ptr(), (w, h), pyvrt.DenoiserSettings(1, 1, 20.4, 1e-2, 1e-1, 2.0, mode), out.data_ptr())
        gpu_ctx.sync()
        got = out.cpu().numpy()
        assert np.array_equal(got, O.denoise(color, normal, pos, iterations=1, mode=mode)), mode
        assert np.array_equal(got, blind_blur(color, mode)), mode


def test_degenerate_rotation_takes_the_ordered_host_flatten(gpu_ctx):
    from vrtlib import degenerate_rotation_vox
    data = degenerate_rotation_vox()
    rc, vol, pal = O.load_vox(data)
    gs = pyvrt.VoxelScene.from_vox_bytes(gpu_ctx, data)
    assert np.array_equal(gs.download_volume(), vol) and np.array_equal(gs.palette(), pal)
    gs.destroy()


@pytest.mark.parametrize("step_width", [1.5, 0.7, 2.25])
def test_standalone_denoise_fractional_step_width(gpu_ctx, step_width):
    """Non-integral stepWidth (a float slider in the reference GUI): the taps fall between texels and the LINEAR sampler of the
    reference blends four of them (render_image.cpp:60-66). Kernel vs oracle, both modes, within 1/255 (expf is not shared)."""
    import torch
    rng = np.random.default_rng(17)
    h, w = 271, 483
    color = rng.integers(0, 256, (h, w, 4), dtype=np.uint8); color[..., 3] = 0
    normal = np.zeros((h, w, 4), np.int8); normal[..., :3] = rng.choice([-127, 0, 127], (h, w, 3))
    pos = (rng.random((h, w, 4), dtype=np.float32) * 3); pos[..., 3] = 0
    tc, tn, tp = torch.from_numpy(color).cuda(), torch.from_numpy(normal).cuda(), torch.from_numpy(pos).cuda()
    for mode in (0, 1):
        for iters in (2, 3):
            ds = pyvrt.DenoiserSettings(1, iters, 20.4, 1e-2, 1e-1, step_width, mode)
            out = torch.empty_like(tc)
            torch.cuda.synchronize()
            pyvrt.denoise_device(gpu_ctx, tc.data_ptr(), tn.data_ptr(), tp.data_ptr(), (w, h), ds, out.data_ptr())
            gpu_ctx.sync()
            want = O.denoise(color, normal, pos, iterations=iters, step_width=step_width, mode=mode)
            d = np.abs(out.cpu().numpy().astype(int) - want.astype(int))
            assert d.max() <= 1 and (d != 0).mean() < 2e-3, (mode, iters, d.max(), (d != 0).mean())
    assert not np.array_equal(want, O.denoise(color, normal, pos, iterations=3, step_width=float(round(step_width)) or 1.0, mode=1))


def test_temporal_upscaler_own_definition_properties(gpu_ctx):
    """SURVEY 8(f) N4, parity unpinned (FSR2 is closed source: own definition, csrc/vrt_upscale.cu). Tested properties: exact pass-through
    at scale 1 / jitter (0, 0) (first frame and steady state); determinism; with the Halton jitter of the frame driver a static scene
    rendered at half resolution converges towards the full-resolution image (better than a single bilinear frame); reset drops the history."""
    import torch
    rng = np.random.default_rng(5)
    w, h = 321, 187
    img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8); img[..., 3] = 0
    t = torch.from_numpy(img).cuda()
    up = pyvrt.TemporalUpscaler(gpu_ctx, (w, h), (w, h))
    out = torch.zeros_like(t)
    torch.cuda.synchronize()
    for k in range(3):                                     # pass-through, first frame and with history
        up.dispatch(t.data_ptr(), out.data_ptr(), reset=(k == 0)); gpu_ctx.sync()
        assert np.array_equal(out.cpu().numpy(), img), k
    up.destroy()

    # a static analytic scene "rendered" at half resolution with the frame driver's jitter, upscaled 2x
    TW, TH, RW, RH = 512, 288, 256, 144
    def scene(x, y):                                        # continuous position in target pixels -> rgb in [0, 1]
        return np.stack([0.5 + 0.5 * np.sin(x * 0.11) * np.cos(y * 0.07), 0.5 + 0.5 * np.sin((x + y) * 0.05), ((np.floor(x / 16) + np.floor(y / 16)) % 2) * 0.8 + 0.1], axis=-1)
    ty, tx = np.mgrid[0:TH, 0:TW]
    truth = scene(tx + 0.5, ty + 0.5)
    def render(jx, jy):
        ry, rx = np.mgrid[0:RH, 0:RW]
        c = scene((rx + 0.5 - jx) * (TW / RW), (ry + 0.5 + jy) * (TH / RH))          # sample (i, j) sits at (i + 0.5 - jx, j + 0.5 + jy)
        a = np.zeros((RH, RW, 4), np.uint8); a[..., :3] = np.rint(c * 255).astype(np.uint8)
        return a
    def run(frames):
        u = pyvrt.TemporalUpscaler(gpu_ctx, (RW, RH), (TW, TH))
        stage = pyvrt.UpscalerStage(RW, TW)
        o = torch.zeros((TH, TW, 4), dtype=torch.uint8, device="cuda")
        outs = []
        for k in range(frames):
            stage.update(0.016)
            src = torch.from_numpy(render(float(stage.jitterX), float(stage.jitterY))).cuda()
            torch.cuda.synchronize()
            u.dispatch(src.data_ptr(), o.data_ptr(), jitter=(float(stage.jitterX), float(stage.jitterY)), reset=(k == 0)); gpu_ctx.sync()
            outs.append(o.cpu().numpy().copy())
        u.destroy()
        return outs
    a, b = run(24), run(24)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)                          # deterministic
    err = [np.abs(x[..., :3] / 255.0 - truth).mean() for x in a]
    assert err[-1] < 0.6 * err[0] and err[-1] < err[4]        # accumulation of jittered frames beats the single first frame
    # reset: the first frame of a new sequence does not depend on the history
    u = pyvrt.TemporalUpscaler(gpu_ctx, (RW, RH), (TW, TH))
    o = torch.zeros((TH, TW, 4), dtype=torch.uint8, device="cuda")
    src = torch.from_numpy(render(0.25, -0.125)).cuda()
    other = torch.from_numpy(render(-0.3, 0.2)).cuda()
    torch.cuda.synchronize()
    u.dispatch(src.data_ptr(), o.data_ptr(), jitter=(0.25, -0.125), reset=True); gpu_ctx.sync(); first = o.cpu().numpy().copy()
    u.dispatch(other.data_ptr(), o.data_ptr(), jitter=(-0.3, 0.2)); gpu_ctx.sync()
    u.dispatch(src.data_ptr(), o.data_ptr(), jitter=(0.25, -0.125), reset=True); gpu_ctx.sync()
    assert np.array_equal(o.cpu().numpy(), first)
    u.destroy()


def test_two_level_volume_is_byte_identical(gpu_ctx, monkeypatch):
    """scenes over the dense budget keep the palette volume two-level as well (bricks of one value + 512-byte bricks);
    a generated scene is then never materialised densely. Forced here on small scenes: the stored volume expands to the generator's /
    the uploaded bytes, every plane equals the oracle's, and the measurement back-ends that need the dense volume refuse politely."""
    monkeypatch.setenv("VRT_VOLUME_DENSE_MAX_GIB", "0")
    noise, sky = assets.noise_standin(), assets.sky_standin(128, 64)
    whd = (100, 72, 90)
    gs = pyvrt.VoxelScene.generate(gpu_ctx, 2, 1234, *whd)
    vol = synth.GENERATORS[2](1234, *whd)
    assert np.array_equal(gs.download_volume(), vol)
    mem = gs.memory()
    assert mem["volume_bytes"] < vol.size                              # top level + the bricks that hold a shape
    gs.set_noise(noise); gs.set_skybox(sky)
    sc, keep = O.make_scene(vol, synth.generator_palette(2, 1234), noise, sky)
    cam = look_at_camera(whd, (320, 200), (-20.5, 90.7, -30.2), (50, 36, 45), frame=3)
    st = O.make_settings()
    want = O.render(sc, cam, st)
    for pipe in PIPELINES:
        check(gpu_render(gpu_ctx, gs, cam, st, 0, pipeline=pipe), want)                       # default resolves to the two-level fields
    check(gpu_render(gpu_ctx, gs, cam, st, TRAV_DENSE), want)                                  # one fetch per step through the two-level volume
    check(gpu_render(gpu_ctx, gs, cam, st, 5), want)                                           # linear fields built from the two-level volume
    with pytest.raises(pyvrt.VrtError) as e:
        gpu_render(gpu_ctx, gs, cam, st, 4)
    assert e.value.code == pyvrt.E_UNSUPPORTED
    gs.destroy()
    rng = np.random.default_rng(3)                                      # uploaded volumes are compacted after the upload
    v2 = random_volume(rng, (37, 50, 41), 0.02); pal = random_palette(rng)
    g2 = pyvrt.VoxelScene.from_dense(gpu_ctx, v2, pal); g2.set_noise(noise); g2.set_skybox(sky)
    assert np.array_equal(g2.download_volume(), v2)
    sc2, keep2 = O.make_scene(v2, pal, noise, sky)
    cam2 = look_at_camera((37, 50, 41), (160, 120), (-12, 60, -15), (18, 25, 20))
    check(gpu_render(gpu_ctx, g2, cam2, st), O.render(sc2, cam2, st))
    g2.destroy()
    monkeypatch.delenv("VRT_VOLUME_DENSE_MAX_GIB")
    g1 = pyvrt.VoxelScene.generate(gpu_ctx, 1, 42, 96, 64, 80)          # generator 1 through the per-voxel function == the dense kernel
    monkeypatch.setenv("VRT_VOLUME_DENSE_MAX_GIB", "0")
    g1s = pyvrt.VoxelScene.generate(gpu_ctx, 1, 42, 96, 64, 80)
    assert np.array_equal(g1.download_volume(), g1s.download_volume())
    g1.destroy(); g1s.destroy()
