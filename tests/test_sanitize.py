"""SURVEY.md 8(c) self-check (v) and ADVICE r1 (parser hardening): the oracle and the product's host-side parsers run clean under
AddressSanitizer + UndefinedBehaviorSanitizer. Both drivers are plain C++ built from the sources where they lie (no CUDA, no GPU) with
-fno-sanitize-recover=all, so a report of either sanitizer ends the process with a non-zero code."""
from __future__ import annotations

import glob
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SAN = os.path.join(ROOT, "tests", "sanitize")
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*.vox")))
GXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
FLAGS = ["-std=c++17", "-O1", "-g", "-ffp-contract=off", "-fsanitize=address,undefined", "-fno-sanitize-recover=all"]


def build(name, sources, extra=()):
    exe = os.path.join(SAN, name)
    srcs = [os.path.join(SAN, name + ".cpp")] + [os.path.join(ROOT, s) for s in sources]
    if not os.path.exists(exe) or any(os.path.getmtime(s) > os.path.getmtime(exe) for s in srcs):
        subprocess.check_call([GXX] + FLAGS + ["-o", exe] + srcs + list(extra))
    return exe


def run(exe):
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1", OMP_NUM_THREADS="2")
    return subprocess.run([exe] + GOLDEN, capture_output=True, text=True, timeout=900, env=env)


def test_oracle_is_clean_under_asan_and_ubsan():
    exe = build("oracle_asan", ["oracle/orc_render.cpp", "oracle/orc_denoise.cpp", "oracle/orc_scene.cpp", "oracle/orc_generate.cpp", "oracle/orc_output.cpp"], ["-fopenmp"])
    p = run(exe)
    assert p.returncode == 0 and "oracle sanitize ok" in p.stdout, p.stdout[-2000:] + p.stderr[-4000:]


def test_product_parsers_are_clean_under_asan_and_ubsan_on_damaged_files():
    exe = build("parser_asan", ["voxel-raytracing_b200/csrc/vrt_vox.cpp", "voxel-raytracing_b200/csrc/vrt_images.cpp", "voxel-raytracing_b200/csrc/vrt_hostside.cpp"], ["-lz"])
    p = run(exe)
    assert p.returncode == 0 and "parser sanitize ok" in p.stdout, p.stdout[-2000:] + p.stderr[-4000:]
    accepted, rejected = [int(x) for x in p.stdout.split("ok:")[1].replace("inputs accepted,", "").replace("rejected", "").split()]
    assert accepted > 1000 and rejected > 1000          # the damage is neither harmless nor always fatal: both paths of the parsers ran


def test_shared_traversal_core_is_clean_under_ubsan():
    """csrc/vrt_core.cuh & co. are compiled for the device by nvcc and for the host by the test harness: the same C++. Here the harness
    is built with UBSan and renders three poses through every back-end and both pipelines (tests/sanitize/harness_ubsan.py)."""
    src = os.path.join(ROOT, "tests", "host_harness", "harness.cpp")
    lib = os.path.join(SAN, "libhh_ubsan.so")
    deps = [src] + glob.glob(os.path.join(ROOT, "voxel-raytracing_b200", "csrc", "*.cuh")) + glob.glob(os.path.join(ROOT, "voxel-raytracing_b200", "csrc", "*.h"))
    if not os.path.exists(lib) or any(os.path.getmtime(d) > os.path.getmtime(lib) for d in deps):
        subprocess.check_call([GXX, "-std=c++17", "-O1", "-g", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared",
                               "-fsanitize=undefined", "-fno-sanitize-recover=undefined", "-x", "c++", src, "-o", lib])
    env = dict(os.environ, UBSAN_OPTIONS="print_stacktrace=1")
    p = subprocess.run([os.sys.executable, os.path.join(SAN, "harness_ubsan.py"), lib], capture_output=True, text=True, timeout=900, env=env)
    assert p.returncode == 0 and "harness under UBSan: ok" in p.stdout, p.stdout[-2000:] + p.stderr[-4000:]
