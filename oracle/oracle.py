"""ORACLE — TEST INFRASTRUCTURE ONLY. ctypes binding of oracle/libvrt_oracle.so (CPU restatement of the
reference shaders) and of oracle/_ref/libref_vox.so (the reference's vendored .vox parser, compiled unmodified).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import this."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libvrt_oracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libref_vox.so")

PASS_PRIMARY, PASS_SHADOW, PASS_AO, PASS_REFLECT, PASS_SHADE, PASS_ALL = 1, 2, 4, 8, 16, 31


class Camera(C.Structure):
    _fields_ = [("camPos", C.c_float * 4), ("camDir", C.c_float * 4), ("camRight", C.c_float * 4), ("camUp", C.c_float * 4),
                ("volumeBounds", C.c_uint32 * 3), ("frame", C.c_uint32), ("screenSize", C.c_int32 * 2), ("cameraJitter", C.c_float * 2)]


class Settings(C.Structure):
    _fields_ = [("pass_mask", C.c_uint32), ("aoSamples", C.c_uint32), ("ambientIntensity", C.c_float),
                ("lightDir", C.c_float * 3), ("lightIntensity", C.c_float), ("lightColor", C.c_float * 4)]


class Scene(C.Structure):
    _fields_ = [("volume", C.c_void_p), ("w", C.c_uint32), ("h", C.c_uint32), ("d", C.c_uint32), ("palette", C.c_void_p),
                ("noise", C.c_void_p), ("sky", C.c_void_p), ("skyW", C.c_uint32), ("skyH", C.c_uint32)]


PLANE_SPECS = [("color", np.uint8, 4), ("depth", np.float32, 1), ("motion", np.float32, 2), ("mask", np.uint8, 1),
               ("position", np.float32, 4), ("normal", np.int8, 4), ("hit", np.int16, 4), ("hit_palette", np.uint8, 1),
               ("color_f32", np.float32, 3), ("rays", np.uint32, 1), ("fetches", np.uint32, 1)]


class Planes(C.Structure):
    _fields_ = [(n, C.c_void_p) for n, _, _ in PLANE_SPECS]


class DenoiseSettings(C.Structure):
    _fields_ = [("iterations", C.c_int), ("phiColor0", C.c_float), ("phiNormal0", C.c_float), ("phiPos0", C.c_float),
                ("stepWidth", C.c_float), ("mode", C.c_int)]


def build(ref: bool = True) -> None:
    subprocess.check_call(["make", "-s", "-C", HERE, "all"])
    if ref and os.path.isdir("/root/reference"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build(ref=False)
        _lib = C.CDLL(LIB)
        _lib.orc_render.restype = C.c_int
        _lib.orc_denoise.restype = C.c_int
        _lib.orc_load_vox.restype = C.c_int
        _lib.orc_trace_one.restype = C.c_int
        _lib.orc_num_threads.restype = C.c_int
        _lib.orc_jitter_phase_count.restype = C.c_int
        _lib.orc_blit.restype = C.c_int
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_scene(volume, palette, noise=None, sky=None):
    """volume: uint8 [D,H,W] (z,y,x; x fastest). palette: float32 [256,8]. Returns (Scene, keepalive)."""
    volume = np.ascontiguousarray(volume, np.uint8)
    palette = np.ascontiguousarray(palette, np.float32).reshape(256, 8)
    d, h, w = volume.shape
    s = Scene()
    s.volume, s.w, s.h, s.d, s.palette = _ptr(volume), w, h, d, _ptr(palette)
    keep = [volume, palette]
    if noise is not None:
        noise = np.ascontiguousarray(noise, np.uint8); assert noise.shape == (512, 512, 4)
        s.noise = _ptr(noise); keep.append(noise)
    if sky is not None:
        sky = np.ascontiguousarray(sky, np.float32); assert sky.ndim == 3 and sky.shape[2] == 4
        s.sky, s.skyW, s.skyH = _ptr(sky), sky.shape[1], sky.shape[0]; keep.append(sky)
    return s, keep


def make_camera(pos, dir_, right, up, bounds, size, frame=0, jitter=(0.0, 0.0)):
    c = Camera()
    for k in range(3):
        c.camPos[k], c.camDir[k], c.camRight[k], c.camUp[k] = pos[k], dir_[k], right[k], up[k]
    c.camPos[3] = 1.0
    c.volumeBounds[:] = [int(b) for b in bounds]
    c.frame = frame
    c.screenSize[:] = [int(size[0]), int(size[1])]
    c.cameraJitter[:] = [float(jitter[0]), float(jitter[1])]
    return c


def camera_vectors(yaw, pitch, focal):
    d, r, u = (C.c_float * 3)(), (C.c_float * 3)(), (C.c_float * 3)()
    lib().orc_camera_vectors(C.c_float(yaw), C.c_float(pitch), C.c_float(focal), d, r, u)
    return np.array(d[:], np.float32), np.array(r[:], np.float32), np.array(u[:], np.float32)


def make_settings(pass_mask=PASS_ALL, ao=4, ambient=1.0, light_dir=None, light_intensity=1.0, light_color=(1, 1, 1, 1)):
    s = Settings()
    s.pass_mask, s.aoSamples, s.ambientIntensity = pass_mask, ao, ambient
    if light_dir is None:
        light_dir = np.float32(1.0) / np.sqrt(np.float32(3.0)) * np.ones(3, np.float32)  # glm::normalize(vec3(1))
    s.lightDir[:] = [float(np.float32(v)) for v in light_dir]
    s.lightIntensity = light_intensity
    s.lightColor[:] = list(light_color)
    return s


def render(scene, cam, st, rows=None, nthreads=0, planes=None):
    """Returns dict of numpy planes. rows = (begin, end, step) to subsample rows (other rows stay zero)."""
    W, H = cam.screenSize[0], cam.screenSize[1]
    out, p = {}, Planes()
    for n, dt, ch in PLANE_SPECS:
        if planes is not None and n not in planes:
            continue
        out[n] = np.zeros((H, W, ch) if ch > 1 else (H, W), dt)
        setattr(p, n, _ptr(out[n]))
    rb, re, rs = rows if rows else (0, H, 1)
    rc = lib().orc_render(C.byref(scene), C.byref(cam), C.byref(st), C.byref(p), rb, re, rs, nthreads)
    if rc != 0:
        raise RuntimeError(f"orc_render failed: {rc}")
    return out


def trace_one(scene, start, dir_, max_steps=512):
    o = (C.c_int32 * 6)(); f = (C.c_float * 6)()
    s = (C.c_float * 3)(*[float(v) for v in start]); d = (C.c_float * 3)(*[float(v) for v in dir_])
    lib().orc_trace_one(C.byref(scene), s, d, C.c_uint32(max_steps), o, f)
    return dict(mapPos=tuple(o[0:3]), steps=o[3], material=o[4], fetches=o[5], pos=np.array(f[0:3], np.float32), normal=np.array(f[3:6], np.float32))


def denoise(color, normal, position, iterations=2, phi_color=20.4, phi_normal=1e-2, phi_pos=1e-1, step_width=2.0, mode=0, nthreads=0):
    h, w = color.shape[:2]
    ds = DenoiseSettings(iterations, phi_color, phi_normal, phi_pos, step_width, mode)
    color = np.ascontiguousarray(color, np.uint8); normal = np.ascontiguousarray(normal, np.int8)
    position = np.ascontiguousarray(position, np.float32)
    out = np.zeros((h, w, 4), np.uint8)
    rc = lib().orc_denoise(_ptr(color), _ptr(normal), _ptr(position), w, h, C.byref(ds), _ptr(out), nthreads)
    if rc != 0:
        raise RuntimeError(f"orc_denoise failed: {rc}")
    return out


def jitter_phase_count(render_width, display_width):
    return lib().orc_jitter_phase_count(int(render_width), int(display_width))


def jitter_offset(index, phase_count):
    x, y = C.c_float(), C.c_float()
    lib().orc_jitter_offset(C.byref(x), C.byref(y), int(index), int(phase_count))
    return np.float32(x.value), np.float32(y.value)


def upscaler_update(render_width, target_width, frame_count):
    """UpscalerStage::update: returns (jitterX, jitterY, next frameCount)."""
    fc, x, y = C.c_int(frame_count), C.c_float(), C.c_float()
    lib().orc_upscaler_update(int(render_width), int(target_width), C.byref(fc), C.byref(x), C.byref(y))
    return np.float32(x.value), np.float32(y.value), fc.value


def render_resolution(fsr_enable, scaling, target):
    t = (C.c_uint32 * 2)(*target); o = (C.c_uint32 * 2)()
    lib().orc_render_resolution(int(bool(fsr_enable)), int(scaling), t, o)
    return o[0], o[1]


def blit(src, target_size, srgb=True):
    """BlitStage: src (h, w, 4) u8 -> (th, tw, 4) u8; target_size = (tw, th)."""
    src = np.ascontiguousarray(src, np.uint8); sh, sw = src.shape[:2]
    out = np.zeros((target_size[1], target_size[0], 4), np.uint8)
    rc = lib().orc_blit(_ptr(src), sw, sh, _ptr(out), target_size[0], target_size[1], 1 if srgb else 0, 0)
    if rc != 0:
        raise RuntimeError(f"orc_blit failed: {rc}")
    return out


def generate(generator, seed, w, h, d):
    vol = np.empty((d, h, w), np.uint8)
    rc = lib().orc_generate(int(generator), C.c_uint32(seed & 0xFFFFFFFF), _ptr(vol), w, h, d)
    if rc != 0:
        raise RuntimeError(f"orc_generate failed: {rc}")
    return vol


def _load_vox_with(fn, data: bytes):
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    vol = C.c_void_p(); whd = (C.c_uint32 * 3)(); pal = np.zeros((256, 8), np.float32)
    rc = fn(buf, C.c_size_t(len(data)), C.byref(vol), whd, _ptr(pal))
    if rc != 0:
        return rc, None, None
    w, h, d = whd[0], whd[1], whd[2]
    arr = np.ctypeslib.as_array(C.cast(vol, C.POINTER(C.c_uint8)), shape=(d, h, w)).copy()
    return 0, arr, pal


def load_vox(data: bytes):
    rc, vol, pal = _load_vox_with(lib().orc_load_vox, data)
    return rc, vol, pal


_ref = None


def ref_lib():
    """The reference's own ogt_vox parser/writer (oracle/_ref). None if it was never built (no /root/reference)."""
    global _ref
    if _ref is None and os.path.exists(REF_LIB):
        _ref = C.CDLL(REF_LIB)
        _ref.ref_load_vox.restype = C.c_int
        _ref.ref_write_vox.restype = C.c_void_p
    return _ref


def ref_load_vox(data: bytes):
    return _load_vox_with(ref_lib().ref_load_vox, data)


def ref_write_vox(models, instances, palette_rgba, metal=None) -> bytes:
    """models: list of uint8 [x,y,z] arrays; instances: list of (model, rot3x3 (m00..m22 ogt layout), (tx,ty,tz))."""
    r = ref_lib()
    sizes = np.array([m.shape for m in models], np.uint32)
    vox = np.concatenate([np.ascontiguousarray(np.transpose(m, (2, 1, 0))).ravel() for m in models]).astype(np.uint8)
    im = np.array([i[0] for i in instances], np.uint32)
    xf = np.array([list(np.asarray(i[1], np.float32).ravel()) + list(i[2]) for i in instances], np.float32)
    pal = np.ascontiguousarray(palette_rgba, np.uint8)
    mt = np.full(256, -1.0, np.float32)
    for k, v in (metal or {}).items():
        mt[k] = v
    n = C.c_uint32()
    p = r.ref_write_vox(C.c_uint32(len(models)), _ptr(sizes), _ptr(vox), C.c_uint32(len(instances)), _ptr(im), _ptr(xf), _ptr(pal), _ptr(mt), C.byref(n))
    data = C.string_at(p, n.value)
    return data
