"""ctypes binding of libvrt.so (include/vrt.h) for the bench and the tests, plus thin host-side mirrors of the
reference's stage interface: VoxelScene (voxel_scene.hpp:18-34), GeometryStage.record -> GeometryBuffer
(geometry_stage.hpp:19-27,53), DenoiserStage.record (denoiser_stage.hpp:45-46).

The library is the product; this file only marshals arguments. If libvrt.so is missing or no CUDA device is
present every entry point raises: there is no CPU fallback."""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VRT_LIB", os.path.join(HERE, "libvrt.so"))   # VRT_LIB: build variants for measurement only

PASS_PRIMARY, PASS_SHADOW, PASS_AO, PASS_REFLECT, PASS_SHADE, PASS_ALL = 1, 2, 4, 8, 16, 31
OUT_GBUFFER, OUT_INSTRUMENT = 1, 2
TRAVERSAL_DEFAULT, TRAVERSAL_DENSE_U8, TRAVERSAL_BRICK, TRAVERSAL_FINE, TRAVERSAL_LINEAR, TRAVERSAL_OCTANT, TRAVERSAL_SPARSE = 0, 1, 2, 3, 4, 5, 6
PIPELINE_DEFAULT, PIPELINE_MEGAKERNEL, PIPELINE_WAVEFRONT = 0, 1, 2
DENOISE_AS_EXECUTED, DENOISE_AS_INTENDED = 0, 1
E_INVALID = -1
E_LFS_POINTER, E_NO_INSTANCE, E_PARSE, E_IO, E_CUDA, E_NO_DEVICE, E_UNSUPPORTED, E_OOM, E_COMM = -2, -3, -4, -5, -6, -7, -8, -9, -10

# plane id -> (name, dtype, channels)
PLANES = {0: ("color", np.uint8, 4), 1: ("denoised", np.uint8, 4), 2: ("depth", np.float32, 1), 3: ("motion", np.float32, 2),
          4: ("mask", np.uint8, 1), 5: ("position", np.float32, 4), 6: ("normal", np.int8, 4), 7: ("hit", np.int16, 4),
          8: ("hit_palette", np.uint8, 1), 9: ("color_f32", np.float32, 3), 10: ("rays", np.uint32, 1), 11: ("fetches", np.uint32, 1)}
PLANE_ID = {v[0]: k for k, v in PLANES.items()}
TILE = 32


class Camera(C.Structure):
    _fields_ = [("camPos", C.c_float * 4), ("camDir", C.c_float * 4), ("camRight", C.c_float * 4), ("camUp", C.c_float * 4),
                ("volumeBounds", C.c_uint32 * 3), ("frame", C.c_uint32), ("screenSize", C.c_int32 * 2), ("cameraJitter", C.c_float * 2)]


class VolumeParams(C.Structure):
    _fields_ = [("aoSamples", C.c_uint32), ("ambientIntensity", C.c_float)]


class Light(C.Structure):
    _fields_ = [("direction", C.c_float * 3), ("intensity", C.c_float), ("color", C.c_float * 4)]


class DenoiserSettings(C.Structure):
    _fields_ = [("enable", C.c_int32), ("iterations", C.c_int32), ("phiColor0", C.c_float), ("phiNormal0", C.c_float),
                ("phiPos0", C.c_float), ("stepWidth", C.c_float), ("mode", C.c_int32)]


class RenderSettings(C.Structure):
    _fields_ = [("pass_mask", C.c_uint32), ("out_mask", C.c_uint32), ("traversal", C.c_uint32), ("pipeline", C.c_uint32),
                ("vol", VolumeParams), ("light", Light), ("denoise", DenoiserSettings)]


class FrameStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("fetches", C.c_uint64), ("pixels", C.c_uint64), ("ms_trace", C.c_float),
                ("ms_denoise", C.c_float), ("ms_total", C.c_float), ("kernel_launches", C.c_uint32), ("reserved", C.c_uint32),
                ("ms_stage", C.c_float * 4), ("stage_rays", C.c_uint64 * 3), ("stage_fetches", C.c_uint64 * 3)]


class VrtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"vrt error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise VrtError(-100, f"{LIB_PATH} is missing: build it with `make -C {HERE}` (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.vrt_last_error.restype = C.c_char_p
        L.vrt_version.restype = C.c_char_p
        L.vrt_context_stream.restype = C.c_void_p
        L.vrt_frame_stream.restype = C.c_void_p
        L.vrt_comm_stream.restype = C.c_void_p
        for fn in ("vrt_context_destroy", "vrt_scene_destroy", "vrt_frame_release", "vrt_host_free", "vrt_upscaler_destroy", "vrt_camera_vectors", "vrt_default_settings", "vrt_render_resolution"):
            getattr(L, fn).restype = None
        _lib = L
    return _lib


def _ck(rc):
    if rc != 0:
        raise VrtError(rc, lib().vrt_last_error().decode())


def default_settings() -> RenderSettings:
    s = RenderSettings()
    lib().vrt_default_settings(C.byref(s))
    return s


def camera_vectors(yaw, pitch, focal):
    d, r, u = (C.c_float * 3)(), (C.c_float * 3)(), (C.c_float * 3)()
    lib().vrt_camera_vectors(C.c_float(yaw), C.c_float(pitch), C.c_float(focal), d, r, u)
    return np.array(d[:], np.float32), np.array(r[:], np.float32), np.array(u[:], np.float32)


def make_camera(pos, yaw, pitch, bounds_whd, size, frame=0, jitter=(0.0, 0.0), focal=None) -> Camera:
    """Push constants as VoxelRenderer::recordCommands fills them (voxel_renderer.cpp:72-81)."""
    if focal is None:
        focal = float(np.float32(1.0 / np.tan(np.radians(55.0 / 2))))   # voxel_renderer.cpp:20
    d, r, u = camera_vectors(yaw, pitch, focal)
    c = Camera()
    for k in range(3):
        c.camPos[k], c.camDir[k], c.camRight[k], c.camUp[k] = float(pos[k]), d[k], r[k], u[k]
    c.camPos[3] = 1.0
    c.volumeBounds[:] = [int(b) for b in bounds_whd]
    c.frame = int(frame)
    c.screenSize[:] = [int(size[0]), int(size[1])]
    c.cameraJitter[:] = [float(jitter[0]), float(jitter[1])]
    return c


def _read_image(fn, arg, dtype, by_path):
    out = C.c_void_p(); w, h = C.c_uint32(), C.c_uint32()
    if by_path:
        _ck(fn(arg.encode(), C.byref(out), C.byref(w), C.byref(h)))
    else:
        buf = (C.c_uint8 * max(1, len(arg))).from_buffer_copy(arg or b"\0")
        _ck(fn(buf, C.c_size_t(len(arg)), C.byref(out), C.byref(w), C.byref(h)))
    ct = C.c_float if dtype == np.float32 else C.c_uint8
    a = np.ctypeslib.as_array(C.cast(out, C.POINTER(ct)), shape=(h.value, w.value, 4)).copy()
    lib().vrt_host_free(out)
    return a


def read_hdr(path_or_bytes):
    """Radiance RGBE -> float32 [h, w, 4] (alpha 1), host only."""
    by_path = isinstance(path_or_bytes, str)
    return _read_image(lib().vrt_image_read_hdr if by_path else lib().vrt_image_decode_hdr, path_or_bytes, np.float32, by_path)


def read_png(path_or_bytes):
    """PNG -> uint8 [h, w, 4], host only."""
    by_path = isinstance(path_or_bytes, str)
    return _read_image(lib().vrt_image_read_png if by_path else lib().vrt_image_decode_png, path_or_bytes, np.uint8, by_path)


def file_sha256(path) -> str:
    hexbuf = C.create_string_buffer(65)
    _ck(lib().vrt_file_sha256(path.encode(), hexbuf))
    return hexbuf.value.decode()


def lfs_pointer_info(path):
    """(oid, size) if the file is a git-LFS pointer stub, else None."""
    oid = C.create_string_buffer(65); size = C.c_uint64()
    rc = lib().vrt_lfs_pointer_info(path.encode(), oid, C.byref(size))
    return (oid.value.decode(), size.value) if rc == 0 else None


def asset_matches_pointer(candidate, pointer) -> bool:
    """True when `candidate` is the real file the LFS pointer at `pointer` names (sha256 and size equal)."""
    return lib().vrt_asset_matches_pointer(candidate.encode(), pointer.encode()) == 0


def vox_decode(data: bytes):
    """Host-only decode of .vox bytes -> (volume[D,H,W] uint8, palette[256,8] float32). No GPU involved."""
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    vol = C.c_void_p(); whd = (C.c_uint32 * 3)(); pal = np.zeros((256, 8), np.float32)
    _ck(lib().vrt_vox_decode(buf, C.c_size_t(len(data)), C.byref(vol), whd, pal.ctypes.data_as(C.c_void_p)))
    w, h, d = whd[0], whd[1], whd[2]
    arr = np.ctypeslib.as_array(C.cast(vol, C.POINTER(C.c_uint8)), shape=(d, h, w)).copy()
    lib().vrt_host_free(vol)
    return arr, pal


class Context:
    def __init__(self, device=0, rank=0, world=1):
        self.h = C.c_void_p()
        _ck(lib().vrt_context_create(int(device), C.byref(self.h)))
        self.rank, self.world, self.device = rank, world, device
        if world != 1:
            _ck(lib().vrt_context_set_shard(self.h, rank, world))

    def set_shard(self, rank, world):
        """Tile shard of the frames created from now on (vrt_context_set_shard)."""
        _ck(lib().vrt_context_set_shard(self.h, int(rank), int(world)))
        self.rank, self.world = int(rank), int(world)

    def device_info(self):
        sm = C.c_int(); l2 = C.c_size_t(); hbm = C.c_size_t()
        _ck(lib().vrt_device_info(self.h, C.byref(sm), C.byref(l2), C.byref(hbm)))
        return dict(sm_count=sm.value, l2_bytes=l2.value, hbm_bytes=hbm.value)

    def sync(self):
        _ck(lib().vrt_context_sync(self.h))

    @property
    def stream(self):
        return lib().vrt_context_stream(self.h)

    def set_concurrent_frames(self, on=True):
        """Frame objects created afterwards render on streams of their own (frames in flight overlap on the GPU)."""
        _ck(lib().vrt_context_set_concurrent_frames(self.h, 1 if on else 0))

    def join(self):
        """The context stream waits for everything queued on the frames' own streams so far."""
        _ck(lib().vrt_context_join(self.h))

    # ---- multi-GPU exchange (vrt_comm_*): one process per GPU, frame assembled on rank 0 ----
    def comm_create(self, rank, world, key: str):
        _ck(lib().vrt_comm_create(self.h, int(rank), int(world), key.encode()))
        self.rank, self.world = int(rank), int(world)

    def comm_alloc_frames(self, size, slots, planes=("color",), host_planes=()):
        pm = sum(1 << PLANE_ID[n] for n in planes); hm = sum(1 << PLANE_ID[n] for n in host_planes)
        _ck(lib().vrt_comm_alloc_frames(self.h, int(size[0]), int(size[1]), int(slots), C.c_uint32(pm), C.c_uint32(hm)))

    def comm_alloc_strips(self, ds: "DenoiserSettings | None"):
        """vrt_comm_alloc_strips: the denoiser distributed over the ranks (every rank denoises the rows it owns). Raises VrtError with
        code E_UNSUPPORTED when a strip would be thinner than the denoiser's reach. ds=None: no denoiser, the strips carry the colour
        plane alone (the read-back path of a sharded frame)."""
        _ck(lib().vrt_comm_alloc_strips(self.h, C.byref(ds) if ds is not None else None))

    def comm_wait(self, slot, stream=None):
        _ck(lib().vrt_comm_wait(self.h, int(slot), C.c_void_p(stream) if stream else None))

    def comm_wait_host(self, slot):
        _ck(lib().vrt_comm_wait_host(self.h, int(slot)))

    def comm_release(self, slot, stream=None, on_stream=False):
        if on_stream:
            _ck(lib().vrt_comm_release_on(self.h, int(slot), C.c_void_p(stream) if stream else None))
        else:
            _ck(lib().vrt_comm_release(self.h, int(slot)))

    def comm_barrier(self):
        _ck(lib().vrt_comm_barrier(self.h))

    @property
    def comm_stream(self):
        return lib().vrt_comm_stream(self.h)

    def comm_frame_plane(self, slot, name):
        """(device pointer, host numpy view or None, bytes) of an assembled plane of a slot as this process sees it."""
        d, h, n = C.c_void_p(), C.c_void_p(), C.c_size_t()
        _ck(lib().vrt_comm_frame_plane(self.h, int(slot), PLANE_ID[name], C.byref(d), C.byref(h), C.byref(n)))
        host = np.ctypeslib.as_array(C.cast(h, C.POINTER(C.c_uint8)), shape=(n.value,)) if h.value else None
        return d.value, host, n.value

    def comm_destroy(self):
        _ck(lib().vrt_comm_destroy(self.h))

    def close(self):
        if self.h:
            lib().vrt_context_destroy(self.h); self.h = C.c_void_p()


class VoxelScene:
    """Mirror of the reference's VoxelScene: width/height/depth + resident volume, palette, skybox."""

    def __init__(self, ctx: Context, handle):
        self.ctx, self.h = ctx, handle
        whd = (C.c_uint32 * 3)()
        _ck(lib().vrt_scene_bounds(self.h, whd))
        self.width, self.height, self.depth = whd[0], whd[1], whd[2]

    @property
    def bounds(self):
        return (self.width, self.height, self.depth)

    @classmethod
    def from_vox_file(cls, ctx, path):
        h = C.c_void_p(); _ck(lib().vrt_scene_load_vox(ctx.h, path.encode(), C.byref(h))); return cls(ctx, h)

    @classmethod
    def from_vox_bytes(cls, ctx, data: bytes):
        buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
        h = C.c_void_p(); _ck(lib().vrt_scene_load_vox_memory(ctx.h, buf, C.c_size_t(len(data)), C.byref(h))); return cls(ctx, h)

    @classmethod
    def from_dense(cls, ctx, volume, palette):
        volume = np.ascontiguousarray(volume, np.uint8); palette = np.ascontiguousarray(palette, np.float32).reshape(256, 8)
        d, hh, w = volume.shape
        h = C.c_void_p()
        _ck(lib().vrt_scene_from_dense(ctx.h, volume.ctypes.data_as(C.c_void_p), w, hh, d, palette.ctypes.data_as(C.c_void_p), C.byref(h)))
        return cls(ctx, h)

    @classmethod
    def generate(cls, ctx, generator, seed, w, h_, d):
        h = C.c_void_p(); _ck(lib().vrt_scene_generate(ctx.h, int(generator), C.c_uint64(seed), w, h_, d, C.byref(h))); return cls(ctx, h)

    def set_noise(self, rgba8):
        a = np.ascontiguousarray(rgba8, np.uint8); _ck(lib().vrt_scene_set_noise(self.h, a.ctypes.data_as(C.c_void_p), a.shape[1], a.shape[0]))

    def set_skybox(self, rgba32f):
        a = np.ascontiguousarray(rgba32f, np.float32); _ck(lib().vrt_scene_set_skybox(self.h, a.ctypes.data_as(C.c_void_p), a.shape[1], a.shape[0]))

    def load_skybox(self, path):
        """Radiance .hdr -> RGBA32F skybox (stbi_loadf in the reference, voxel_scene.cpp:132)."""
        _ck(lib().vrt_scene_load_skybox(self.h, path.encode()))

    def load_noise(self, path):
        """PNG -> RGBA8 512x512 noise texture (stbi_load in the reference, voxel_renderer.cpp:22)."""
        _ck(lib().vrt_scene_load_noise(self.h, path.encode()))

    def palette(self):
        p = np.zeros((256, 8), np.float32); _ck(lib().vrt_scene_palette(self.h, p.ctypes.data_as(C.c_void_p))); return p

    def download_volume(self):
        v = np.zeros((self.depth, self.height, self.width), np.uint8)
        _ck(lib().vrt_scene_download_volume(self.h, v.ctypes.data_as(C.c_void_p), C.c_size_t(v.size))); return v

    def download_accel(self):
        bx, by, bz = (self.width + 3) // 4, (self.height + 3) // 4, (self.depth + 3) // 4
        m = np.zeros((bz, by, bx), np.uint64); dd = np.zeros((bz, by, bx), np.uint8)
        _ck(lib().vrt_scene_download_accel(self.h, m.ctypes.data_as(C.c_void_p), dd.ctypes.data_as(C.c_void_p))); return m, dd

    def download_fdist(self):
        bx, by, bz = (self.width + 3) // 4, (self.height + 3) // 4, (self.depth + 3) // 4
        f = np.zeros((bz, by, bx, 4, 4, 4), np.uint8)   # [bz][by][bx][z&3][y&3][x&3]
        _ck(lib().vrt_scene_download_fdist(self.h, f.ctypes.data_as(C.c_void_p), C.c_size_t(f.size))); return f

    def download_odist(self, octant):
        f = np.zeros((self.depth + 2, self.height + 2, self.width + 2), np.uint8)   # aproned, [z+1][y+1][x+1]
        _ck(lib().vrt_scene_download_odist(self.h, int(octant), f.ctypes.data_as(C.c_void_p), C.c_size_t(f.size))); return f

    def prepare(self, traversal=0):
        """Build the safe-step fields of a traversal back-end now (the default back-end's are built at load time)."""
        _ck(lib().vrt_scene_prepare(self.h, C.c_uint32(traversal)))

    def memory(self):
        v, f, pk, ms = C.c_size_t(), C.c_size_t(), C.c_size_t(), C.c_double()
        _ck(lib().vrt_scene_memory(self.h, C.byref(v), C.byref(f), C.byref(pk), C.byref(ms)))
        return dict(volume_bytes=v.value, field_bytes=f.value, peak_build_bytes=pk.value, load_ms=ms.value)

    def download_sdist(self, octant):
        """The two-level (sparse) field of an octant, expanded to the aproned linear layout of download_odist."""
        f = np.zeros((self.depth + 2, self.height + 2, self.width + 2), np.uint8)
        _ck(lib().vrt_scene_download_sdist(self.h, int(octant), f.ctypes.data_as(C.c_void_p), C.c_size_t(f.size))); return f

    def accel_info(self):
        a, b, c = C.c_size_t(), C.c_size_t(), C.c_size_t()
        _ck(lib().vrt_scene_accel_info(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(dense_bytes=a.value, brick_mask_bytes=b.value, brick_dist_bytes=c.value)

    def destroy(self):
        if self.h:
            lib().vrt_scene_destroy(self.h); self.h = C.c_void_p()


class Frame:
    """Mirror of GeometryBuffer: named planes living in device memory."""

    def __init__(self, ctx):
        self.ctx, self.h, self.size = ctx, C.c_void_p(), None

    def sync(self):
        _ck(lib().vrt_frame_sync(self.h))

    @property
    def stream(self):
        """cudaStream_t this frame's work is queued on (the context's unless frames are concurrent)."""
        return lib().vrt_frame_stream(self.h)

    def push_tiles(self, name, d_frame_ptr, stream=None):
        """Store this rank's tiles of a plane into a row-major frame that may be peer memory (vrt_frame_push_tiles)."""
        _ck(lib().vrt_frame_push_tiles(self.h, PLANE_ID[name], C.c_void_p(d_frame_ptr), C.c_void_p(stream) if stream else None))

    def exchange(self, slot, planes=("color",), host_planes=(), stream=None):
        """This rank's tiles of the planes into rank 0's assembled frame `slot` (peer memory) and / or the shared host frame."""
        pm = sum(1 << PLANE_ID[n] for n in planes); hm = sum(1 << PLANE_ID[n] for n in host_planes)
        _ck(lib().vrt_frame_exchange(self.h, int(slot), C.c_uint32(pm), C.c_uint32(hm), C.c_void_p(stream) if stream else None))

    def download_async(self, name, out):
        """Queue the read-back behind this frame's kernels on the copy stream (out: pinned uint8 buffer); pair with download_wait()."""
        _, nbytes = self.plane_ptr(name)
        assert out.nbytes == nbytes
        _ck(lib().vrt_frame_download_async(self.h, PLANE_ID[name], out.ctypes.data_as(C.c_void_p), C.c_size_t(nbytes)))

    def download_wait(self):
        if self.h:
            _ck(lib().vrt_frame_download_wait(self.h))

    def plane_ptr(self, name):
        p = C.c_void_p(); n = C.c_size_t()
        _ck(lib().vrt_frame_plane(self.h, PLANE_ID[name], C.byref(p), C.byref(n))); return p.value, n.value

    def download(self, name, out=None):
        _, dt, ch = PLANES[PLANE_ID[name]]
        _, nbytes = self.plane_ptr(name)
        if out is None:
            out = np.empty(nbytes // np.dtype(dt).itemsize, dt)
        _ck(lib().vrt_frame_download(self.h, PLANE_ID[name], out.ctypes.data_as(C.c_void_p), C.c_size_t(nbytes)))
        if self.ctx.world == 1:
            W, H = self.size
            return out.reshape((H, W, ch) if ch > 1 else (H, W))
        return out

    def stats(self) -> dict:
        s = FrameStats(); _ck(lib().vrt_frame_stats_get(self.h, C.byref(s)))
        return {k: (list(getattr(s, k)) if k.startswith(("ms_stage", "stage_")) else getattr(s, k)) for k, _ in FrameStats._fields_ if k != "reserved"}

    def release(self):
        if self.h:
            lib().vrt_frame_release(self.h); self.h = C.c_void_p()


class VoxelRenderPass:
    """GeometryStage + DenoiserStage as one headless pass: record() launches asynchronously on the context stream
    (the reference records a command buffer) and returns the Frame; Frame.sync() is the fence wait."""

    def __init__(self, ctx: Context, scene: VoxelScene, settings: RenderSettings | None = None):
        self.ctx, self.scene = ctx, scene
        self.settings = settings if settings is not None else default_settings()
        self.frame = Frame(ctx)

    def record_exchange(self, camera: Camera, slot, planes=("color",), host_planes=()) -> Frame:
        """record() + Frame.exchange() in one call, the exchange fused into the pass's kernels (vrt_render_exchange)."""
        size = (camera.screenSize[0], camera.screenSize[1])
        if self.frame.size not in (None, size):
            self.frame.release()
        pm = sum(1 << PLANE_ID[n] for n in planes); hm = sum(1 << PLANE_ID[n] for n in host_planes)
        _ck(lib().vrt_render_exchange(self.ctx.h, self.scene.h, C.byref(camera), C.byref(self.settings), int(slot), C.c_uint32(pm), C.c_uint32(hm), C.byref(self.frame.h)))
        self.frame.size = size
        return self.frame

    def record_denoise_exchange(self, camera: Camera, slot, to_host=False) -> Frame:
        """vrt_render_denoise_exchange: the pass + the distributed denoiser; the denoised plane of `slot` on rank 0 is the result
        (strips without a denoiser: the colour plane). to_host: 0 device frame, 1 device + shared host frame, 2 host frame only."""
        size = (camera.screenSize[0], camera.screenSize[1])
        if self.frame.size not in (None, size):
            self.frame.release()
        _ck(lib().vrt_render_denoise_exchange(self.ctx.h, self.scene.h, C.byref(camera), C.byref(self.settings), int(slot), int(to_host), C.byref(self.frame.h)))
        self.frame.size = size
        return self.frame

    def record(self, camera: Camera) -> Frame:
        size = (camera.screenSize[0], camera.screenSize[1])
        if self.frame.size not in (None, size):
            self.frame.release()
        _ck(lib().vrt_render(self.ctx.h, self.scene.h, C.byref(camera), C.byref(self.settings), C.byref(self.frame.h)))
        self.frame.size = size
        return self.frame


def assemble(ctx: Context, d_gathered_ptr: int, world: int, size, plane: str, d_out_ptr: int, stream: int = None):
    if stream is None:
        _ck(lib().vrt_assemble(ctx.h, C.c_void_p(d_gathered_ptr), world, size[0], size[1], PLANE_ID[plane], C.c_void_p(d_out_ptr)))
    else:
        _ck(lib().vrt_assemble_on(ctx.h, C.c_void_p(d_gathered_ptr), world, size[0], size[1], PLANE_ID[plane], C.c_void_p(d_out_ptr), C.c_void_p(stream)))


BLIT_SRGB, BLIT_BGRA = 1, 2
IMAGE_PPM, IMAGE_PNG_RGB, IMAGE_PNG_RGBA = 0, 1, 2
FSR_NONE, FSR_QUALITY, FSR_BALANCED, FSR_PERFORMANCE, FSR_ULTRA_PERFORMANCE = 10, 15, 17, 20, 30   # FsrScaling, voxel_render_settings.hpp:6-13


def blit_device(ctx: Context, d_src, src_size, d_dst, dst_size, flags=BLIT_SRGB, stream: int = None):
    """BlitStage::record: src_size / dst_size are (w, h); device pointers to RGBA8 images. Asynchronous."""
    if stream is None:
        _ck(lib().vrt_blit(ctx.h, C.c_void_p(d_src), src_size[0], src_size[1], C.c_void_p(d_dst), dst_size[0], dst_size[1], C.c_uint32(flags)))
    else:
        _ck(lib().vrt_blit_on(ctx.h, C.c_void_p(d_src), src_size[0], src_size[1], C.c_void_p(d_dst), dst_size[0], dst_size[1], C.c_uint32(flags), C.c_void_p(stream)))


def image_write(path: str, rgba8, fmt=IMAGE_PNG_RGB):
    a = np.ascontiguousarray(rgba8, np.uint8)
    assert a.ndim == 3 and a.shape[2] == 4
    _ck(lib().vrt_image_write(path.encode(), a.ctypes.data_as(C.c_void_p), a.shape[1], a.shape[0], fmt))


def jitter_phase_count(render_width, display_width) -> int:
    return lib().vrt_jitter_phase_count(int(render_width), int(display_width))


def jitter_offset(index, phase_count):
    x, y = C.c_float(), C.c_float()
    _ck(lib().vrt_jitter_offset(int(index), int(phase_count), C.byref(x), C.byref(y)))
    return np.float32(x.value), np.float32(y.value)


def render_resolution(fsr_enable, scaling, target):
    t = (C.c_uint32 * 2)(*target); o = (C.c_uint32 * 2)()
    lib().vrt_render_resolution(int(bool(fsr_enable)), C.c_uint32(scaling), t, o)
    return o[0], o[1]


class UpscalerStage:
    """The frame-driver half of UpscalerStage (upscaler_stage.cpp:59-70): update() once per frame sets jitterX / jitterY for this
    frame and advances frameCount; both feed the camera block (voxel_renderer.cpp:80-82). FSR2's upscaling itself is out of scope."""

    def __init__(self, render_width, target_width):
        self.render_width, self.target_width = int(render_width), int(target_width)
        self.frameCount, self.jitterX, self.jitterY = 0, np.float32(0), np.float32(0)

    def update(self, delta=0.0):
        fc, x, y = C.c_int(self.frameCount), C.c_float(), C.c_float()
        _ck(lib().vrt_upscaler_update(self.render_width, self.target_width, C.byref(fc), C.byref(x), C.byref(y)))
        self.frameCount, self.jitterX, self.jitterY = fc.value, np.float32(x.value), np.float32(y.value)


class TemporalUpscaler:
    """The stage after the denoiser (UpscalerStage::record): an own-definition temporal upscaler (csrc/vrt_upscale.cu); device pointers in."""

    def __init__(self, ctx: Context, render_size, target_size):
        self.ctx, self.h = ctx, C.c_void_p()
        _ck(lib().vrt_upscaler_create(ctx.h, int(render_size[0]), int(render_size[1]), int(target_size[0]), int(target_size[1]), C.byref(self.h)))

    def dispatch(self, d_color, d_out, d_depth=None, d_motion=None, d_mask=None, jitter=(0.0, 0.0), mv_scale=(1.0, 1.0), reset=False, stream=None):
        p = lambda v: C.c_void_p(v) if v else None
        _ck(lib().vrt_upscaler_dispatch(self.h, p(d_color), p(d_depth), p(d_motion), p(d_mask), C.c_float(jitter[0]), C.c_float(jitter[1]),
                                        C.c_float(mv_scale[0]), C.c_float(mv_scale[1]), 1 if reset else 0, p(d_out), p(stream)))

    def destroy(self):
        if self.h:
            lib().vrt_upscaler_destroy(self.h); self.h = C.c_void_p()


def denoise_device_on(ctx: Context, d_color, d_normal, d_position, size, ds: DenoiserSettings, d_out, d_scratch, stream=None):
    """vrt_denoise_on: asynchronous, caller-owned scratch (rank 0 on the assembled planes of a sharded frame)."""
    _ck(lib().vrt_denoise_on(ctx.h, C.c_void_p(d_color), C.c_void_p(d_normal), C.c_void_p(d_position), size[0], size[1], C.byref(ds),
                             C.c_void_p(d_out), C.c_void_p(d_scratch), C.c_void_p(stream) if stream else None))


def denoise_reach_rows(ds: DenoiserSettings) -> int:
    """vrt_denoise_reach_rows: the halo (rows either side) the denoiser's output for a range of rows depends on. No device needed."""
    return int(lib().vrt_denoise_reach_rows(C.byref(ds)))


DENOISE_STATIC_MAX_PIXELS = 3840 * 2160   # include/vrt.h VRT_DENOISE_STATIC_MAX_PIXELS


def denoise_device_static(ctx: Context, d_color, d_normal, d_position, size, ds: DenoiserSettings, d_out, stream=None):
    """vrt_denoise_static: vrt_denoise_on without a scratch argument and without any allocation (the intermediate image is a
    __device__ array of the library, at most DENOISE_STATIC_MAX_PIXELS; calls that overlap on the GPU race on it)."""
    _ck(lib().vrt_denoise_static(ctx.h, C.c_void_p(d_color), C.c_void_p(d_normal), C.c_void_p(d_position), size[0], size[1], C.byref(ds),
                                 C.c_void_p(d_out), C.c_void_p(stream) if stream else None))


def denoise_device(ctx: Context, d_color, d_normal, d_position, size, ds: DenoiserSettings, d_out):
    _ck(lib().vrt_denoise(ctx.h, C.c_void_p(d_color), C.c_void_p(d_normal), C.c_void_p(d_position), size[0], size[1], C.byref(ds), C.c_void_p(d_out)))
