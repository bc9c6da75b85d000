"""C++ host mirror (include/vrt_host.hpp) above the C ABI: the reference is compiled C++, so the drop-in's host side is too.
CPU tier: the check program compiles with g++ against the header, links against every libvrt symbol it uses, and the
no-device error path throws instead of falling back. GPU tier: VoxelRenderer::recordCommands (geometry -> denoiser) on a
.vox file, planes compared byte for byte with the oracle."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from vrtlib import O, ROOT, PKG, cached_vox, CACHE, bits_equal

SRC = os.path.join(ROOT, "tests", "host_cpp", "host_cpp_check.cpp")
EXE = os.path.join(ROOT, "tests", "host_cpp", "host_cpp_check")


def build_check():
    deps = [SRC, os.path.join(ROOT, "include", "vrt_host.hpp"), os.path.join(ROOT, "include", "vrt.h"), os.path.join(PKG, "libvrt.so")]
    if not os.path.exists(EXE) or any(os.path.getmtime(d) > os.path.getmtime(EXE) for d in deps):
        gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.check_call([gxx, "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), SRC, "-o", EXE,
                               "-L" + PKG, "-lvrt", "-Wl,-rpath," + PKG])
    return EXE


def test_cpp_host_compiles_and_links():
    out = subprocess.check_output([build_check(), "--link-only"], text=True)
    assert out.startswith("vrt ")


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="this is the no-GPU behaviour test")
def test_cpp_host_throws_without_gpu():
    out = subprocess.check_output([build_check(), "--no-gpu"], text=True)
    assert "threw" in out and "no CPU fallback" in out


@pytest.mark.gpu
def test_cpp_renderer_matches_oracle(tmp_path):
    data = cached_vox("treehouse")
    vox = os.path.join(CACHE, "treehouse.vox")
    W, H = 640, 360
    prefix = str(tmp_path / "cpp")
    out = subprocess.check_output([build_check(), vox, str(W), str(H), prefix], text=True)
    assert out.startswith("ok ")
    rc, vol, pal = O.load_vox(data)
    noise = np.fromfile(prefix + "_noise.bin", np.uint8).reshape(512, 512, 4)
    sky = np.fromfile(prefix + "_sky.bin", np.float32).reshape(32, 64, 4)
    sc, keep = O.make_scene(vol, pal, noise, sky)
    cam = O.Camera()
    raw = open(prefix + "_camera.bin", "rb").read()
    assert len(raw) == C.sizeof(cam)
    C.memmove(C.byref(cam), raw, len(raw))
    want = O.render(sc, cam, O.make_settings())          # reference defaults: all passes, 4 AO samples
    planes = {0: ("color", np.uint8, 4), 2: ("depth", np.float32, 1), 4: ("mask", np.uint8, 1), 5: ("position", np.float32, 4), 6: ("normal", np.int8, 4)}
    got = {}
    for k, (n, dt, ch) in planes.items():
        a = np.fromfile(f"{prefix}_plane{k}.bin", dt)
        got[n] = a.reshape((H, W, ch) if ch > 1 else (H, W))
    for n in ("depth", "mask", "position", "normal"):
        ok, bad = bits_equal(got[n], want[n])
        assert ok, f"plane {n}: {bad} elements differ"
    assert np.abs(got["color"].astype(int) - want["color"].astype(int)).max() <= 1
    dn = np.fromfile(prefix + "_denoised.bin", np.uint8).reshape(H, W, 4)
    ref = O.denoise(got["color"], got["normal"], got["position"])
    assert np.abs(dn.astype(int) - ref.astype(int)).max() <= 1
    # the whole frame: FSR QUALITY render resolution, two upscaler updates (frameCount 2, Halton index 1), blit into a W x H sRGB window
    cam2 = O.Camera(); raw = open(prefix + "_camera2.bin", "rb").read(); C.memmove(C.byref(cam2), raw, len(raw))
    rw, rh = O.render_resolution(True, 15, (W, H))
    assert (cam2.screenSize[0], cam2.screenSize[1]) == (rw, rh) and cam2.frame == 2
    jx, jy = O.jitter_offset(1, O.jitter_phase_count(rw, W))
    assert bits_equal(np.array(cam2.cameraJitter[:], np.float32), np.array([jx, jy], np.float32))[0]
    final2 = np.fromfile(prefix + "_final2.bin", np.uint8).reshape(rh, rw, 4)
    want2 = O.render(sc, cam2, O.make_settings())
    col2 = np.fromfile(prefix + "_fsr_plane0.bin", np.uint8).reshape(rh, rw, 4)
    pos2 = np.fromfile(prefix + "_fsr_plane5.bin", np.float32).reshape(rh, rw, 4)
    nrm2 = np.fromfile(prefix + "_fsr_plane6.bin", np.int8).reshape(rh, rw, 4)
    assert bits_equal(pos2, want2["position"])[0] and bits_equal(nrm2, want2["normal"])[0]     # the jittered frame, bit for bit
    assert np.abs(col2.astype(int) - want2["color"].astype(int)).max() <= 1
    assert np.abs(final2.astype(int) - O.denoise(col2, nrm2, pos2).astype(int)).max() <= 1
    window = np.fromfile(prefix + "_window.bin", np.uint8).reshape(H, W, 4)
    assert np.array_equal(window, O.blit(final2, (W, H), srgb=True))      # the blit itself is exact on the same input
    assert os.path.getsize(prefix + "_window.png") > W * H * 3
