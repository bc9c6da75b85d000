import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # test infrastructure builds (CPU): oracle restatement, and the reference .vox shim when /root/reference exists
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "all"])
    if os.path.isdir("/root/reference") and not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_vox.so")):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "ref"])
    lib = os.path.join(ROOT, "voxel-raytracing_b200", "libvrt.so")
    if not os.path.exists(lib):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "voxel-raytracing_b200")])


def _has_gpu():
    try:
        import pyvrt
        c = pyvrt.Context(0)
        c.close()
        return True
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu_ctx():
    import vrtlib  # noqa: F401  (sets sys.path)
    import pyvrt
    ctx = pyvrt.Context(0)   # raises loudly if there is no device: GPU tests must not silently pass
    yield ctx
    ctx.close()
