"""The traversal core shared by the kernels and the host harness (csrc/vrt_core.cuh, vrt_pixel.cuh, vrt_wave.cuh, vrt_engine-free paths:
every back-end, both pipelines) executed on the CPU in a harness built with -fsanitize=undefined -fno-sanitize-recover=undefined:
signed overflow, bad shifts, out-of-range float -> int conversions in that code would be undefined behaviour for nvcc as well.
usage: harness_ubsan.py <instrumented libvrt_hostharness .so>   (run by tests/test_sanitize.py in a process of its own)"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import vrtlib  # noqa: E402

vrtlib._h = C.CDLL(sys.argv[1])
vrtlib._h.hh_scene_create.restype = C.c_void_p
vrtlib._h.hh_render.restype = C.c_int
vrtlib._h.hh_render_wave.restype = C.c_int
from vrtlib import O, harness_render, look_at_camera, random_volume, random_palette  # noqa: E402

rng = np.random.default_rng(5)
whd = (70, 40, 55)
vol, pal = random_volume(rng, whd, 0.01), random_palette(rng, metallic_every=3)
vol[:, :6, :] = 3                                             # a floor: long shadow / reflection / AO rays above it
noise = rng.integers(0, 256, (512, 512, 4), dtype=np.uint8)
sky = rng.random((16, 32, 4), dtype=np.float32)
sc, keep = O.make_scene(vol, pal, noise, sky)
st = O.make_settings()
# outside looking in, inside the volume, straight down an axis (zero direction components take the literal loop)
for pos, tgt in [((-20.3, 60.2, -12.1), (35, 10, 27)), ((35.5, 20.5, 27.5), (0, 0, 0)), ((35, 200, 27), (35, 0, 27))]:
    cam = look_at_camera(whd, (96, 54), pos, tgt, frame=3)
    want = O.render(sc, cam, st)
    for trav in vrtlib.ALL_TRAVERSALS:
        for wave in (False, True):
            got = harness_render(sc, cam, st, traversal=trav, wave=wave)
            for n in ("hit", "hit_palette", "rays", "fetches", "normal"):
                assert np.array_equal(got[n], want[n]), (trav, wave, n)
print("harness under UBSan: ok")
