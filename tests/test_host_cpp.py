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


COMM_SRC = os.path.join(ROOT, "tests", "host_cpp", "comm_check.cpp")
COMM_EXE = os.path.join(ROOT, "tests", "host_cpp", "comm_check")


def build_comm_check():
    deps = [COMM_SRC, os.path.join(ROOT, "include", "vrt.h"), os.path.join(PKG, "libvrt.so")]
    if not os.path.exists(COMM_EXE) or any(os.path.getmtime(d) > os.path.getmtime(COMM_EXE) for d in deps):
        gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.check_call([gxx, "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), COMM_SRC, "-o", COMM_EXE,
                               "-L" + PKG, "-lvrt", "-Wl,-rpath," + PKG])
    return COMM_EXE


def test_cpp_comm_check_compiles_and_links():
    """The multi-rank exchange driver is plain C++ over include/vrt.h (no torch, no NCCL): it must build and link on the CPU tier."""
    assert os.path.exists(build_comm_check())


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="this is the no-GPU behaviour test")
def test_cpp_comm_check_fails_loudly_without_gpu():
    p = subprocess.run([build_comm_check(), "2", "64", "64", "1"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 10 and "no CPU fallback" in p.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("world,size", [(2, (200, 150)), (3, (333, 217)), (4, (640, 360)), (8, (256, 40))])
def test_cpp_ranks_assemble_a_frame_without_torch(world, size):
    """a C++ host linking libvrt.so shards a frame over `world` processes and assembles it on rank 0 through
    vrt_comm_* / vrt_frame_exchange alone (peer stores into rank 0's frame + shared pinned host frame), then denoises it there:
    every plane and the denoised frame equal the single-rank frame byte for byte. Ranks share the GPUs that are present."""
    import torch
    ndev = max(1, torch.cuda.device_count())
    env = dict(os.environ, VRT_COMM_TIMEOUT_S="60")
    p = subprocess.run([build_comm_check(), str(world), str(size[0]), str(size[1]), str(ndev), "5"], capture_output=True, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "identical to the single-rank frames" in p.stdout
    # the denoiser distributed over the ranks (vrt_comm_alloc_strips / vrt_render_denoise_exchange); 8 ranks x 5 rows: refused, loudly
    assert ("strips (denoiser distributed over the ranks): denoised frame identical" in p.stdout) == (world != 8), p.stdout
    assert ("unsupported at this size" in p.stdout) == (world == 8), p.stdout
    assert "strips without a denoiser (colour rows to rank 0 / the host frame): colour frame identical" in p.stdout


def test_cpp_host_compiles_and_links():
    out = subprocess.check_output([build_check(), "--link-only"], text=True)
    assert out.startswith("vrt ")


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="this is the no-GPU behaviour test")
def test_cpp_host_throws_without_gpu():
    out = subprocess.check_output([build_check(), "--no-gpu"], text=True)
    assert "threw" in out and "no CPU fallback" in out


def test_cpp_camera_controller_matches_reference_arithmetic():
    """CameraController::update / mouseCallback (camera_controller.cpp:30-71) in the C++ mirror against a binary32 restatement here,
    direction vectors against the oracle's camera_vectors (camera_controller.cpp:15-28): bit for bit over a scripted input sequence."""
    out = subprocess.check_output([build_check(), "--camera"], text=True).strip().splitlines()
    f32 = np.float32
    pos = np.array([8.0, 8.0, -50.0], f32); yaw, pitch, focal = f32(90.0), f32(0.0), f32(1.920982)
    last = [-1.0, -1.0]
    W, S, A, D = 1, 2, 4, 8
    script = [(0, 0.016, W), (0, 0.016, W | D), (1, 400.0, 300.0, 1), (1, 417.5, 288.25, 1), (0, 0.033, S | A), (1, 10.0, 500.0, 1), (1, 0, 0, 0),
              (1, 90.0, 90.0, 1), (1, 95.0, 700.0, 1), (0, 0.5, A), (0, 0.25, W)]
    assert len(out) == len(script)
    for line, st in zip(out, script):
        nd, right, _ = O.camera_vectors(float(yaw), float(pitch), 1.0)
        if st[0] == 0:
            delta, keys = f32(st[1]), st[2]
            if keys & W: pos = (pos + (f32(50.0) * nd) * delta).astype(f32)
            if keys & S: pos = (pos - (f32(50.0) * nd) * delta).astype(f32)
            if keys & A: pos = (pos - (right * f32(50.0)) * delta).astype(f32)
            if keys & D: pos = (pos + (right * f32(50.0)) * delta).astype(f32)
        else:
            _, x, y, button = st
            if button:
                if last[0] > 0 and last[1] > 0:
                    yaw = f32(yaw - f32(x - last[0])); pitch = f32(min(max(f32(pitch - f32(y - last[1])), f32(-90.0)), f32(90.0)))
                last = [x, y]
            else:
                last = [-1.0, -1.0]
        d, r, u = O.camera_vectors(float(yaw), float(pitch), float(focal))
        want = np.concatenate([pos, [yaw, pitch], d, r, u]).astype(f32)
        got = np.array([float.fromhex(t) for t in line.split()], f32)
        assert bits_equal(got, want)[0], (st, got, want)


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
