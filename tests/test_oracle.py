"""CPU tier: pins the oracle (oracle/) — against the golden fixtures produced by the reference's own .vox
parser/writer, against closed-form known-answer scenes, and against an independent integer traversal.
The reference itself has no tests of this path (test/test_test.cpp is a factorial placeholder): PARITY UNPINNED."""
import os
import numpy as np
import pytest

from vrtlib import O, assets, default_camera, look_at_camera, random_volume, random_palette, ROOT

GOLD = os.path.join(ROOT, "tests", "golden")


def _gold(name):
    with open(os.path.join(GOLD, name + ".vox"), "rb") as f:
        data = f.read()
    ref = np.load(os.path.join(GOLD, name + "_ref.npz"))
    return data, ref["volume"], ref["palette"]


# ---------------------------------------------------------------- scene loader vs the reference's ogt_vox (golden)
@pytest.mark.parametrize("name", ["lshape", "twoinst", "nopalette", "overlap9"])
def test_loader_matches_reference_golden(name):
    data, vol, pal = _gold(name)
    rc, v, p = O.load_vox(data)
    assert rc == 0
    assert v.shape == vol.shape and np.array_equal(v, vol)
    assert np.array_equal(p.view(np.uint32), pal.view(np.uint32))


def test_axis_swap_lshape():
    # model voxels (x,y,z): (0,0,0)=1 (1,0,0)=2 (0,0,1)=3 in a 4x3x2 model -> scene[x + z*w + y*w*h] (voxel_scene.cpp:99)
    data, _, _ = _gold("lshape")
    rc, v, p = O.load_vox(data)
    assert v.shape == (3, 2, 4)          # depth = dy(vox) = 3, height = dz(vox) = 2, width = 4
    assert v[0, 0, 0] == 1 and v[0, 0, 1] == 2 and v[0, 1, 0] == 3 and (v != 0).sum() == 3
    assert p[2, 4] == np.float32(0.25) and p[1, 4] == 0   # MATL _metal
    assert p[0, 3] == 0                                   # colour 0 alpha forced to 0 (ogt_vox.h:1839)


def test_default_palette_pinned_by_reference():
    data, _, pal = _gold("nopalette")
    rc, v, p = O.load_vox(data)
    # entry 1 of the rotated palette is the first default colour (255,255,255,255) -> pow(1, 2.2) = 1
    assert np.array_equal(p, pal) and np.all(p[1, :4] == 1.0)


@pytest.mark.skipif(O.ref_lib() is None, reason="oracle/_ref not built (no /root/reference on this machine)")
@pytest.mark.parametrize("name", ["treehouse", "floatingcolored"])
def test_loader_matches_reference_live(name):
    from vrtlib import cached_vox
    data = cached_vox(name)
    rc, v, p = O.load_vox(data)
    rc2, v2, p2 = O.ref_load_vox(data)
    assert rc == 0 and rc2 == 0 and np.array_equal(v, v2) and np.array_equal(p, p2)


@pytest.mark.skipif(O.ref_lib() is None, reason="oracle/_ref not built")
def test_loader_random_instances_vs_reference():
    rng = np.random.default_rng(5)
    rots = [np.eye(3), [[0, 1, 0], [-1, 0, 0], [0, 0, 1]], [[-1, 0, 0], [0, -1, 0], [0, 0, 1]], [[0, 0, 1], [0, 1, 0], [-1, 0, 0]], [[1, 0, 0], [0, 0, -1], [0, 1, 0]]]
    for _ in range(6):
        models = [(rng.random(tuple(rng.integers(1, 9, 3))) < 0.5).astype(np.uint8) * rng.integers(1, 255, dtype=np.uint8) for _ in range(3)]
        for m in models:
            m.flat[0] = 9   # never an empty model
        inst = [(int(rng.integers(0, 3)), np.array(rots[rng.integers(0, len(rots))], np.float32), tuple(int(t) for t in rng.integers(-12, 12, 3))) for _ in range(5)]
        data = O.ref_write_vox(models, inst, assets.palette_standin(3), {4: 0.75})
        rc, v, p = O.load_vox(data); rc2, v2, p2 = O.ref_load_vox(data)
        assert rc == rc2 == 0 and v.shape == v2.shape and np.array_equal(v, v2) and np.array_equal(p, p2)


def test_lfs_pointer_and_garbage_rejected():
    stub = b"version https://git-lfs.github.com/spec/v1\noid sha256:0cfd1da0\nsize 493398\n"
    assert O.load_vox(stub)[0] == -2
    assert O.load_vox(b"not a vox file at all")[0] == -1
    assert O.load_vox(b"VOX " + (150).to_bytes(4, "little"))[0] == -3   # parses, but holds no instance


def test_golden_render_unchanged():
    data, _, _ = _gold("twoinst")
    rc, vol, p = O.load_vox(data)
    sc, keep = O.make_scene(vol, p, assets.noise_standin(), assets.sky_standin(64, 32))
    cam = default_camera((vol.shape[2], vol.shape[1], vol.shape[0]), (48, 27), pos=(-6.5, 9.25, -11.0), yaw=58.0, pitch=-27.0, frame=5)
    out = O.render(sc, cam, O.make_settings())
    gold = np.load(os.path.join(GOLD, "render_small.npz"))
    for k in gold.files:
        a, b = out[k], gold[k]
        assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), k
    assert (out["hit_palette"] != 0).sum() > 50


# ---------------------------------------------------------------- closed-form known-answer scenes
def _scene(vol, metallic=None):
    pal = np.zeros((256, 8), np.float32); pal[1:, :4] = 0.5
    if metallic:
        pal[metallic, 4] = 0.5
    return O.make_scene(vol, pal, assets.noise_standin(), assets.sky_standin(64, 32))


def test_single_voxel_known_answer():
    vol = np.zeros((16, 16, 16), np.uint8); vol[8, 5, 3] = 7      # (z,y,x) -> voxel (3,5,8)
    sc, keep = _scene(vol)
    # ray along +z through the voxel centre column, tiny x/y slopes so no component is exactly 0
    r = O.trace_one(sc, (3.5, 5.5, -4.0), (1e-4, 2e-4, 1.0))
    assert r["mapPos"] == (3, 5, 8) and r["material"] == 7
    # entry: tmin = 4, pos.z = -4 + 4.1 = 0.1 -> voxel z=0; 8 steps in z to reach z=8
    assert r["steps"] == 8 and r["fetches"] == 9
    assert np.array_equal(r["normal"], np.array([-0.0, -0.0, -1.0], np.float32))
    assert abs(r["pos"][2] - 8.0) < 1e-4 and abs(r["pos"][0] - 3.5012) < 1e-3


def test_empty_volume_all_sky_and_bounds_exit():
    vol = np.zeros((8, 8, 8), np.uint8)
    sc, keep = _scene(vol)
    r = O.trace_one(sc, (4.2, 4.3, -3.0), (0.01, 0.02, 1.0))
    assert r["material"] == 0 and r["mapPos"][2] == 8 and r["steps"] == r["fetches"] == 8
    cam = default_camera((8, 8, 8), (32, 18), pos=(4, 4, -20))
    out = O.render(sc, cam, O.make_settings())
    assert (out["hit_palette"] == 0).all() and (out["mask"] == 0).all() and (out["depth"] == 0).all()
    assert (out["rays"] == 1).all() and (out["position"] == 0).all()


def test_camera_inside_solid_first_step_hit():
    vol = np.full((8, 8, 8), 3, np.uint8)
    sc, keep = _scene(vol)
    r = O.trace_one(sc, (4.5, 4.5, 4.5), (0.3, 0.2, 0.9))
    # F5: mask is zero-initialised -> normalize(vec3(0)) = NaN normal, d = 0, pos = start
    assert r["steps"] == 0 and r["fetches"] == 1 and r["material"] == 3 and np.isnan(r["normal"]).all()
    assert np.array_equal(r["pos"], np.array([4.5, 4.5, 4.5], np.float32))


def test_max_steps_is_a_miss():
    vol = np.zeros((600, 4, 4), np.uint8); vol[590] = 1            # solid slab 590 voxels deep in z
    sc, keep = _scene(vol)
    r = O.trace_one(sc, (2.2, 2.3, 0.5), (1e-5, 1e-5, 1.0))
    assert r["material"] == 0 and r["steps"] == 512 and r["fetches"] == 512   # F10
    r = O.trace_one(sc, (2.2, 2.3, 100.5), (1e-5, 1e-5, 1.0))
    assert r["material"] == 1 and r["mapPos"][2] == 590 and r["steps"] == 490


def test_mask_store_is_230():
    vol = np.zeros((8, 8, 8), np.uint8); vol[2:6, 2:6, 2:6] = 1
    sc, keep = _scene(vol)
    cam = look_at_camera((8, 8, 8), (16, 16), (4.4, 4.6, -9.0), (4.5, 4.5, 4.5))
    out = O.render(sc, cam, O.make_settings())
    assert set(np.unique(out["mask"])) <= {0, 230} and (out["mask"] == 230).any()


# ---------------------------------------------------------------- independent integer cross-check + invariants
def _supercover(vol, start, dirv, max_steps=512):
    """Exact rational Amanatides-Woo on the same float inputs (python Fractions), tie-free rays only."""
    from fractions import Fraction as F
    D, H, W = vol.shape
    p = [F(float(s)) for s in start]; d = [F(float(x)) for x in dirv]
    cell = [int(np.floor(float(s))) for s in start]
    step = [1 if x > 0 else -1 for x in d]
    tmax = [((cell[k] + (1 if step[k] > 0 else 0)) - p[k]) / d[k] for k in range(3)]
    tdel = [abs(1 / d[k]) for k in range(3)]
    for i in range(max_steps):
        if not (0 <= cell[0] < W and 0 <= cell[1] < H and 0 <= cell[2] < D):
            return tuple(cell), i, 0
        m = vol[cell[2], cell[1], cell[0]]
        if m:
            return tuple(cell), i, int(m)
        k = min(range(3), key=lambda a: tmax[a])
        if sorted(tmax)[0] == sorted(tmax)[1]:
            return None   # tie: skip this ray
        tmax[k] += tdel[k]; cell[k] += step[k]
    return tuple(cell), max_steps, 0


def test_trace_matches_exact_rational_walk():
    rng = np.random.default_rng(2)
    vol = random_volume(rng, (24, 20, 28), 0.03)
    sc, keep = O.make_scene(vol, random_palette(rng))
    checked = 0
    for _ in range(300):
        start = rng.uniform([1, 1, 1], [23, 19, 27]).astype(np.float32)   # inside the box: boxIntersection returns start
        dirv = rng.normal(size=3).astype(np.float32)
        ref = _supercover(vol, start, dirv)
        if ref is None:
            continue
        r = O.trace_one(sc, start, dirv)
        # float accumulation may legitimately differ from exact arithmetic only at near-ties; those are rare
        if (r["mapPos"], r["steps"], r["material"]) == ref:
            checked += 1
    assert checked >= 290


def test_render_invariants():
    rng = np.random.default_rng(3)
    vol = random_volume(rng, (40, 30, 50), 0.02)
    sc, keep = O.make_scene(vol, random_palette(rng), assets.noise_standin(), assets.sky_standin(64, 32))
    cam = look_at_camera((40, 30, 50), (96, 54), (-15, 40, -20), (20, 15, 25))
    out = O.render(sc, cam, O.make_settings())
    hit = out["hit_palette"] != 0
    assert hit.any() and (~hit).any()
    xyz = out["hit"][..., :3].astype(int); steps = out["hit"][..., 3]
    assert steps.max() <= 512
    hx, hy, hz = xyz[hit].T
    assert np.array_equal(vol[hz, hy, hx], out["hit_palette"][hit])      # hit voxel holds the reported material
    assert (out["fetches"] >= steps).all()
    assert (out["rays"][~hit] == 1).all() and (out["rays"][hit] >= 6).all()   # 1 primary + 4 AO + 1 shadow at least
    assert (out["color"][..., 3] == 0).all()


def test_pass_mask_subsets():
    rng = np.random.default_rng(4)
    vol = random_volume(rng, (20, 20, 20), 0.05)
    sc, keep = O.make_scene(vol, random_palette(rng), assets.noise_standin(), assets.sky_standin(64, 32))
    cam = look_at_camera((20, 20, 20), (40, 30), (-10, 25, -12), (10, 10, 10))
    full = O.render(sc, cam, O.make_settings())
    prim = O.render(sc, cam, O.make_settings(pass_mask=O.PASS_PRIMARY))
    assert np.array_equal(full["hit"], prim["hit"]) and (prim["rays"] == 1).all()
    sh = O.render(sc, cam, O.make_settings(pass_mask=O.PASS_PRIMARY | O.PASS_SHADOW | O.PASS_SHADE, ao=0))
    hit = sh["hit_palette"] != 0
    assert (sh["rays"][hit] == 2).all()


# ---------------------------------------------------------------- camera + denoiser
def test_camera_default_vectors():
    d, r, u = O.camera_vectors(90.0, 0.0, 1.0)
    assert np.allclose(d, [0, 0, 1], atol=1e-6) and d[0] != 0          # yaw 90 gives x = -4.37e-8, not 0
    assert np.allclose(r, [1, 0, 0], atol=1e-6) and np.array_equal(u, np.array([0, -1, 0], np.float32))


def test_denoiser_modes_and_iteration0_is_edge_blind():
    rng = np.random.default_rng(7)
    h, w = 24, 32
    color = rng.integers(0, 256, (h, w, 4), dtype=np.uint8); color[..., 3] = 0
    normal = np.zeros((h, w, 4), np.int8); normal[..., 1] = 127
    pos = rng.random((h, w, 4), dtype=np.float32) * 50; pos[..., 3] = 0
    # one iteration: phi = +inf -> all edge weights are exp(-0) = 1 -> plain kernel-weighted blur (F7)
    ex = O.denoise(color, normal, pos, iterations=1, mode=0)
    c = color.astype(np.float32) / np.float32(255)
    k = np.float32(np.exp(np.float32(-0.25)))
    y, x = 5, 7    # interior pixel; AS_EXECUTED taps: (-1,-1) w=k, (1,-1) w=1, (0,0) w=k
    want = (c[y - 1, x - 1] * k + c[y - 1, x + 1] * np.float32(1.0) + c[y, x] * k) / (k + np.float32(1.0) + k)
    assert np.abs(ex[y, x, :3].astype(np.float32) - want[:3] * 255).max() <= 0.51
    it = O.denoise(color, normal, pos, iterations=1, mode=1)
    assert not np.array_equal(ex, it)
    # a constant image is a fixed point in both modes and for several iterations
    flat = np.zeros((h, w, 4), np.uint8); flat[..., :3] = (10, 200, 77)
    for mode in (0, 1):
        assert np.array_equal(O.denoise(flat, normal, pos * 0, iterations=3, mode=mode), flat)


# ---- the steps either side of the pass (oracle/orc_output.cpp) ----
def test_jitter_phase_counts_pinned_by_reference_header():
    # the reference's own table, thirdparty/fsr2/include/ffx_fsr2.h:341-347: scale factor -> phase count
    for scaling, want in ((15, 18), (17, 23), (20, 32), (30, 72)):
        rw, _ = O.render_resolution(True, scaling, (1920, 1080))       # voxel_render_settings.cpp:3-13
        assert O.jitter_phase_count(rw, 1920) == want, (scaling, rw)
    assert O.render_resolution(True, 17, (1920, 1080)) == (1129, 635)   # uint32((10.0f/17)*1920)
    assert O.render_resolution(False, 17, (1920, 1080)) == (1920, 1080)
    assert O.jitter_phase_count(1920, 1920) == 8


def test_jitter_is_halton_2_3_exact_fractions():
    from fractions import Fraction
    def halton(i, b):
        f, r = Fraction(1), Fraction(0)
        while i > 0:
            f /= b; r += f * (i % b); i //= b
        return r
    for phases in (8, 18, 23, 32, 72):
        for idx in range(phases + 3):
            x, y = O.jitter_offset(idx, phases)
            k = idx % phases + 1
            assert abs(Fraction(float(x)) - (halton(k, 2) - Fraction(1, 2))) < Fraction(1, 1 << 22)
            assert abs(Fraction(float(y)) - (halton(k, 3) - Fraction(1, 2))) < Fraction(1, 1 << 22)
            assert -0.5 <= x < 0.5 and -0.5 <= y < 0.5
    # base 2 is exact in binary32
    assert [float(O.jitter_offset(i, 8)[0]) for i in range(4)] == [0.0, -0.25, 0.25, -0.375]


def test_upscaler_update_wraps_one_frame_late():
    fc, seen = 0, []
    for _ in range(20):
        x, y, nfc = O.upscaler_update(1920, 1920, fc)        # 8 phases
        seen.append((fc, float(x), float(y))); fc = nfc
    assert [s[0] for s in seen] == [0, 1, 2, 3, 4, 5, 6, 7, 8, 0, 1, 2, 3, 4, 5, 6, 7, 8, 0, 1]   # upscaler_stage.cpp:67-69: '>' not '>='
    assert seen[8][1:] == seen[0][1:] == seen[9][1:]         # frameCount 8 % 8 == index 0 again


def test_blit_known_answers():
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, (36, 64, 4), dtype=np.uint8)
    assert np.array_equal(O.blit(a, (64, 36), srgb=False), a)                    # same size: texel centres, weights (1,0,0,0)
    flat = np.full((7, 5, 4), 77, np.uint8)
    assert (O.blit(flat, (33, 21), srgb=False) == 77).all()                      # partition of unity
    up = O.blit(a, (128, 72), srgb=False)                                        # 2x: taps at +-1/4 texel -> weights 9/16, 3/16, 3/16, 1/16
    p = a.astype(np.float64)
    want = (9 * p[1, 1] + 3 * p[1, 2] + 3 * p[2, 1] + 1 * p[2, 2]) / 16
    assert np.abs(up[3, 3].astype(np.float64) - want).max() <= 0.5 + 1e-4
    assert np.array_equal(up[0, 0], a[0, 0])                                     # CLAMP_TO_EDGE at the corner
    # fill + centre crop (blit.frag:17-19): a wider target shows the source's middle rows only
    tall = np.zeros((100, 10, 4), np.uint8); tall[:, :, 0] = np.arange(100)[:, None]
    wide = O.blit(tall, (40, 40), srgb=False)                                    # scale = 10/40; rows 45..55 of the source
    assert 44 <= wide[0, 0, 0] <= 46 and 54 <= wide[-1, 0, 0] <= 56
    # sRGB store: closed-form codes of the transfer function; alpha stays linear
    grey = np.zeros((2, 2, 4), np.uint8); grey[...] = (0, 1, 128, 128)
    enc = O.blit(grey, (2, 2), srgb=True)[0, 0]
    assert tuple(enc) == (0, 13, 188, 128)                                       # 1/255 -> 12.71 -> 13; 128/255 -> 187.85 -> 188
    full = np.full((1, 1, 4), 255, np.uint8)
    assert tuple(O.blit(full, (3, 3), srgb=True)[1, 1]) == (255, 255, 255, 255)


def test_iteration0_is_a_plain_blur_whatever_the_gbuffer_holds():
    from vrtlib import blind_blur
    rng = np.random.default_rng(17)
    h, w = 37, 53
    color = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    normal = rng.integers(-127, 128, (h, w, 4)).astype(np.int8)
    pos = (rng.random((h, w, 4), dtype=np.float32) * 1e3).astype(np.float32)
    pos[3, 4, 0] = np.inf; pos[10, 11, 1] = np.nan; pos[20, 30, 2] = 3e38; pos[21, 30, 2] = -3e38; pos[0, 0, :] = -np.inf
    for mode in (0, 1):
        got = O.denoise(color, normal, pos, iterations=1, mode=mode)
        assert np.array_equal(got, blind_blur(color, mode)), mode


def test_denoiser_fractional_step_width_samples_between_texels():
    """stepWidth is a float slider in the reference GUI (voxel_settings_gui.cpp:106) and the denoiser's inputs go through a LINEAR /
    CLAMP_TO_EDGE sampler (render_image.cpp:60-66): for a non-integral offset*stepWidth the taps blend four texels. An independent
    numpy statement of the same equations (AS_INTENDED, 2 iterations, stepWidth0 = 1.5 -> tap distance 2.5 in iteration 1)."""
    f32 = np.float32
    rng = np.random.default_rng(5)
    h, w = 37, 53
    color = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    normal = rng.integers(-127, 128, (h, w, 4)).astype(np.int8)
    pos = (rng.random((h, w, 4), dtype=f32) * 4).astype(f32)
    got = O.denoise(color, normal, pos, iterations=2, step_width=1.5, mode=1)
    assert not np.array_equal(got, O.denoise(color, normal, pos, iterations=2, step_width=2.0, mode=1))

    import ctypes as C
    libm = C.CDLL("libm.so.6"); libm.expf.restype = C.c_float; libm.expf.argtypes = [C.c_float]
    expf = np.vectorize(lambda v: f32(libm.expf(float(v))), otypes=[f32])
    ys, xs = np.mgrid[0:h, 0:w]

    def sample(img, ox, oy):
        ux, uy = (xs.astype(f32) + f32(ox)).astype(f32), (ys.astype(f32) + f32(oy)).astype(f32)
        fi, fj = np.floor(ux), np.floor(uy)
        a, b = (ux - fi).astype(f32)[..., None], (uy - fj).astype(f32)[..., None]
        i0, i1 = np.clip(fi.astype(int), 0, w - 1), np.clip(fi.astype(int) + 1, 0, w - 1)
        j0, j1 = np.clip(fj.astype(int), 0, h - 1), np.clip(fj.astype(int) + 1, 0, h - 1)
        one = f32(1)
        w00, w10, w01, w11 = (one - a) * (one - b), a * (one - b), (one - a) * b, a * b
        mix = (((w00 * img[j0, i0] + w10 * img[j0, i1]).astype(f32) + w01 * img[j1, i0]).astype(f32) + w11 * img[j1, i1]).astype(f32)
        return np.where((a == 0) & (b == 0), img[j0, i0], mix)

    def dot4(t):
        return (((t[..., 0] * t[..., 0] + t[..., 1] * t[..., 1]).astype(f32) + t[..., 2] * t[..., 2]).astype(f32) + t[..., 3] * t[..., 3]).astype(f32)

    n = np.maximum(normal.astype(f32) / f32(127), f32(-1)).astype(f32)
    src = color
    for it in range(2):
        with np.errstate(divide="ignore", invalid="ignore"):
            phi = [f32(f32(1) / f32(it)) * f32(v) for v in (20.4, 1e-2, 1e-1)]
        sw = f32(f32(it) * f32(1.5) + f32(1))
        c = (src.astype(f32) / f32(255)).astype(f32)
        acc = np.zeros((h, w, 4), f32); tot = np.zeros((h, w), f32)
        for oy in (-1, 0, 1):
            for ox in (-1, 0, 1):
                k = f32(libm.expf(float(f32(-(f32(ox * ox) / f32(8) + f32(oy * oy) / f32(8))))))
                oc, on, op = sample(c, f32(ox) * sw, f32(oy) * sw), sample(n, f32(ox) * sw, f32(oy) * sw), sample(pos, f32(ox) * sw, f32(oy) * sw)
                with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
                    cw = np.fmin(expf(-dot4(c - oc) / phi[0]), f32(1))
                    nw = np.fmin(expf(-np.fmax(dot4(n - on) / f32(sw * sw), f32(0)) / phi[1]), f32(1))
                    pw = np.fmin(expf(-dot4(pos - op) / phi[2]), f32(1))
                wgt = ((cw * nw).astype(f32) * pw).astype(f32)
                acc = (acc + ((oc * wgt[..., None]).astype(f32) * k).astype(f32)).astype(f32)
                tot = (tot + (wgt * k).astype(f32)).astype(f32)
        v = (acc / tot[..., None]).astype(f32)
        src = np.rint(np.clip(np.nan_to_num(v, nan=0.0), 0, 1).astype(f32) * f32(255)).astype(np.uint8)
    assert np.array_equal(src, got)
