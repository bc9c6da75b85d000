"""Shared test helpers: paths, ctypes bindings of the host harness, scene/camera fixtures, plane comparison."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "voxel-raytracing_b200")
for p in (PKG, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import assets  # noqa: E402
import oracle as O  # noqa: E402

HARNESS_DIR = os.path.join(ROOT, "tests", "host_harness")
HARNESS_LIB = os.path.join(HARNESS_DIR, "libvrt_hostharness.so")
CACHE = os.environ.get("VRT_ASSET_CACHE", "/tmp/vrt_assets")

TRAV_DENSE, TRAV_BRICK, TRAV_FINE, TRAV_LINEAR, TRAV_OCTANT, TRAV_SPARSE = 1, 2, 3, 4, 5, 6
ALL_TRAVERSALS = (TRAV_DENSE, TRAV_BRICK, TRAV_FINE, TRAV_LINEAR, TRAV_OCTANT, TRAV_SPARSE)


class VrtCamera(C.Structure):   # include/vrt.h vrt_camera
    _fields_ = O.Camera._fields_


class VrtVolParams(C.Structure):
    _fields_ = [("aoSamples", C.c_uint32), ("ambientIntensity", C.c_float)]


class VrtLight(C.Structure):
    _fields_ = [("direction", C.c_float * 3), ("intensity", C.c_float), ("color", C.c_float * 4)]


class VrtDenoise(C.Structure):
    _fields_ = [("enable", C.c_int32), ("iterations", C.c_int32), ("phiColor0", C.c_float), ("phiNormal0", C.c_float),
                ("phiPos0", C.c_float), ("stepWidth", C.c_float), ("mode", C.c_int32)]


class VrtSettings(C.Structure):
    _fields_ = [("pass_mask", C.c_uint32), ("out_mask", C.c_uint32), ("traversal", C.c_uint32), ("pipeline", C.c_uint32),
                ("vol", VrtVolParams), ("light", VrtLight), ("denoise", VrtDenoise)]


def build_harness() -> str:
    src = os.path.join(HARNESS_DIR, "harness.cpp")
    deps = [src] + [os.path.join(PKG, "csrc", f) for f in ("vrt_core.cuh", "vrt_pixel.cuh", "vrt_wave.cuh", "vrt_output.cuh", "vrt_types.h")]
    if not os.path.exists(HARNESS_LIB) or any(os.path.getmtime(d) > os.path.getmtime(HARNESS_LIB) for d in deps):
        gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.check_call([gxx, "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared",
                               "-x", "c++", src, "-o", HARNESS_LIB])
    return HARNESS_LIB


_h = None


def harness():
    global _h
    if _h is None:
        _h = C.CDLL(build_harness())
        _h.hh_scene_create.restype = C.c_void_p
        _h.hh_render.restype = C.c_int
        _h.hh_render_wave.restype = C.c_int
    return _h


def to_vrt_settings(st: O.Settings, out_mask=3, traversal=0, denoise=None) -> VrtSettings:
    s = VrtSettings()
    s.pass_mask, s.out_mask, s.traversal = st.pass_mask, out_mask, traversal
    s.vol.aoSamples, s.vol.ambientIntensity = st.aoSamples, st.ambientIntensity
    s.light.direction[:] = st.lightDir[:]
    s.light.intensity = st.lightIntensity
    s.light.color[:] = st.lightColor[:]
    if denoise:
        s.denoise = VrtDenoise(1, denoise.get("iterations", 2), denoise.get("phi_color", 20.4), denoise.get("phi_normal", 1e-2),
                               denoise.get("phi_pos", 1e-1), denoise.get("step_width", 2.0), denoise.get("mode", 0))
    return s


def to_vrt_camera(cam: O.Camera) -> VrtCamera:
    c = VrtCamera()
    C.memmove(C.byref(c), C.byref(cam), C.sizeof(cam))
    return c


def harness_render(scene: O.Scene, cam: O.Camera, st: O.Settings, traversal=TRAV_BRICK, rows=None, wave=False, out_mask=3):
    h = harness()
    hs = h.hh_scene_create(C.c_void_p(scene.volume), scene.w, scene.h, scene.d, C.c_void_p(scene.palette), C.c_void_p(scene.noise),
                           C.c_void_p(scene.sky), scene.skyW, scene.skyH)
    W, H = cam.screenSize[0], cam.screenSize[1]
    out, p = {}, O.Planes()
    for n, dt, ch in O.PLANE_SPECS:
        out[n] = np.zeros((H, W, ch) if ch > 1 else (H, W), dt)
        setattr(p, n, out[n].ctypes.data_as(C.c_void_p))
    vs, vc = to_vrt_settings(st, out_mask=out_mask), to_vrt_camera(cam)
    rb, re, rs = rows if rows else (0, H, 1)
    fn = h.hh_render_wave if wave else h.hh_render
    rc = fn(C.c_void_p(hs), C.byref(vc), C.byref(vs), traversal, C.byref(p), rb, re, rs)
    h.hh_scene_destroy(C.c_void_p(hs))
    assert rc == 0
    return out


# ------------------------------------------------------------------------------------------------ fixtures
def cached_vox(name: str, **kw) -> bytes:
    os.makedirs(CACHE, exist_ok=True)
    key = name + "".join(f"_{k}{v}" for k, v in sorted(kw.items()))
    path = os.path.join(CACHE, key + ".vox")
    if not os.path.exists(path):
        data = assets.write_vox(**assets.STANDINS[name](**kw))
        with open(path + ".tmp", "wb") as f:
            f.write(data)
        os.replace(path + ".tmp", path)
    with open(path, "rb") as f:
        return f.read()


def default_camera(bounds_whd, size, pos=(8, 8, -50), yaw=90.0, pitch=0.0, frame=0, jitter=(0, 0)):
    """VoxelRenderer defaults: voxel_renderer.cpp:20 (pos (8,8,-50), yaw 90, pitch 0, focal 1/tan(27.5 deg))."""
    focal = float(np.float32(1.0 / np.tan(np.radians(55.0 / 2))))
    d, r, u = O.camera_vectors(yaw, pitch, focal)
    return O.make_camera(pos, d, r, u, bounds_whd, size, frame, jitter)


def look_at_camera(bounds_whd, size, pos, target, frame=0):
    v = np.asarray(target, np.float64) - np.asarray(pos, np.float64)
    yaw = float(np.degrees(np.arctan2(v[2], v[0])))
    pitch = float(np.degrees(np.arctan2(v[1], np.hypot(v[0], v[2]))))
    return default_camera(bounds_whd, size, pos=pos, yaw=yaw, pitch=pitch, frame=frame)


def random_volume(rng, whd, density=0.05, blobs=True):
    w, h, d = whd
    vol = np.zeros((d, h, w), np.uint8)
    m = rng.random((d, h, w)) < density
    vol[m] = rng.integers(1, 256, m.sum(), dtype=np.uint8)
    if blobs:
        for _ in range(6):
            c = rng.integers(0, [d, h, w]); s = rng.integers(2, 9, 3)
            lo = np.maximum(c - s, 0); hi = np.minimum(c + s, [d, h, w])
            vol[lo[0]:hi[0], lo[1]:hi[1], lo[2]:hi[2]] = rng.integers(1, 256)
    return vol


def random_palette(rng, metallic_every=5):
    pal = np.zeros((256, 8), np.float32)
    pal[:, :4] = rng.random((256, 4), dtype=np.float32)
    pal[::metallic_every, 4] = 0.5
    pal[0] = 0
    return pal


# ------------------------------------------------------------------------------------------------ comparison
def bits_equal(a, b):
    """Bit-exact equality for float arrays (NaN == NaN if same payload class; -0 != +0 is reported)."""
    a = np.ascontiguousarray(a); b = np.ascontiguousarray(b)
    if a.dtype.kind == "f":
        ai, bi = a.view(np.uint32), b.view(np.uint32)
        same = (ai == bi) | (np.isnan(a) & np.isnan(b))
        return bool(same.all()), int((~same).sum())
    same = a == b
    return bool(same.all()), int((~same).sum())


def assert_planes_exact(got, want, names=None, rows=None):
    for n, _, _ in O.PLANE_SPECS:
        if names and n not in names:
            continue
        if n not in got or n not in want:
            continue
        g, w = got[n], want[n]
        if rows:
            g, w = g[rows[0]:rows[1]:rows[2]], w[rows[0]:rows[1]:rows[2]]
        ok, bad = bits_equal(g, w)
        assert ok, f"plane {n}: {bad} of {g.size} elements differ"


def blind_blur(color, mode):
    """Iteration 0 of the denoiser as a plain kernel-weighted blur, in the oracle's float operations and order (what the product's
    edge-blind kernel computes): taps clamped to the edge, (c * k) summed left to right, divided by the summed k, UNORM8 store."""
    f32 = np.float32
    h, w = color.shape[:2]
    c = color.astype(f32) / f32(255)
    libm = C.CDLL("libm.so.6"); libm.expf.restype = C.c_float; libm.expf.argtypes = [C.c_float]   # the host tables use glibc's expf (numpy's differs by an ulp)
    g = lambda x, y: f32(libm.expf(float(f32(-(f32(x * x) / f32(8) + f32(y * y) / f32(8))))))
    if mode == 0:      # AS_EXECUTED: the shader's stride-16 view of the tables
        taps = [(-1, -1, g(-1, -1)), (1, -1, g(0, 0)), (0, 0, g(1, 1))]
    else:
        taps = [(x, y, g(x, y)) for y in (-1, 0, 1) for x in (-1, 0, 1)]
    ys, xs = np.mgrid[0:h, 0:w]
    acc = np.zeros((h, w, 4), f32); tot = f32(0)
    for dx, dy, k in taps:
        acc = (acc + c[np.clip(ys + dy, 0, h - 1), np.clip(xs + dx, 0, w - 1)] * k).astype(f32)
        tot = f32(tot + k)
    v = np.clip(acc / tot, 0, 1).astype(f32)
    return np.rint(v * f32(255)).astype(np.uint8)


def degenerate_rotation_vox() -> bytes:
    """overlap9.vox with two instance rotations patched to packed values that name the same axis twice (_r = 0 and 21). The reference's
    parser asserts on them in a debug build (ogt_vox.h:874) and leaves a zero row in a release build; oracle and product model the
    release behaviour. Such an instance maps many voxels onto one cell: the flatten's loop order decides, so the product flattens it on
    the host (vrt_api.cu scene_from_layout)."""
    data = bytearray(open(os.path.join(ROOT, "tests", "golden", "overlap9.vox"), "rb").read())
    assert data[6724:6726] == b"_r" and data[6730:6731] == b"4" and data[6926:6928] == b"_r" and data[6932:6934] == b"17"
    data[6730:6731] = b"0"; data[6932:6934] = b"21"
    return bytes(data)
