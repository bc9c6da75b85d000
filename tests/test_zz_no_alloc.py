"""vrt_denoise_static (include/vrt.h): the denoiser for callers that cannot allocate; its intermediate image is a file-scope __device__
array of the library. No counterpart in the reference, so the check is against the
existing entry point vrt_denoise_on on the same inputs: bit-exact, the bar of the other denoiser tests for two CUDA paths.
Named test_zz_* so that pytest collects it last."""
from __future__ import annotations

import numpy as np
import pytest

import vrtlib  # noqa: F401  (sets sys.path)
import pyvrt

pytestmark = pytest.mark.gpu


def _planes(rng, w, h):
    color = rng.integers(0, 256, (h, w, 4), dtype=np.uint8); color[..., 3] = 0
    normal = np.zeros((h, w, 4), np.int8); normal[..., :3] = rng.choice([-127, 0, 127], (h, w, 3))
    pos = (rng.random((h, w, 4), dtype=np.float32) * 3); pos[..., 3] = 0
    return color, normal, pos


def test_denoise_static_matches_denoise_on_two_calls_in_a_row(gpu_ctx):
    import torch
    rng = np.random.default_rng(2024)
    # two calls in a row on the shared array: the second one smaller, with other settings, so that it runs over what the first left
    # behind; then a third at the first size again. Iteration counts of both parities (the last iteration must land in d_out).
    calls = [((640, 360), dict(iterations=3, mode=0, step=2.0)), ((333, 217), dict(iterations=2, mode=1, step=2.0)),
             ((640, 360), dict(iterations=4, mode=1, step=1.5)), ((64, 48), dict(iterations=1, mode=0, step=2.0))]
    for n_call, ((w, h), k) in enumerate(calls):
        color, normal, pos = _planes(rng, w, h)
        ds = pyvrt.DenoiserSettings(1, k["iterations"], 20.4, 1e-2, 1e-1, k["step"], k["mode"])
        tc, tn, tp = torch.from_numpy(color).cuda(), torch.from_numpy(normal).cuda(), torch.from_numpy(pos).cuda()
        want, scratch = torch.zeros_like(tc), torch.zeros_like(tc)
        got = torch.full_like(tc, 0xAB)
        torch.cuda.synchronize()
        pyvrt.denoise_device_on(gpu_ctx, tc.data_ptr(), tn.data_ptr(), tp.data_ptr(), (w, h), ds, want.data_ptr(), scratch.data_ptr())
        gpu_ctx.sync()
        free_before = torch.cuda.mem_get_info()[0]
        pyvrt.denoise_device_static(gpu_ctx, tc.data_ptr(), tn.data_ptr(), tp.data_ptr(), (w, h), ds, got.data_ptr())
        gpu_ctx.sync()
        if n_call > 0:   # (the first call resolves the array's address; from then on the device's free memory must not move)
            assert torch.cuda.mem_get_info()[0] == free_before
        a, b = got.cpu().numpy(), want.cpu().numpy()
        assert np.array_equal(a, b), ((w, h), k, int(np.abs(a.astype(int) - b.astype(int)).max()))
        assert a[..., :3].any()   # not a frame of zeros


def test_denoise_static_rejects_images_beyond_the_built_in_scratch(gpu_ctx):
    import torch
    ds = pyvrt.DenoiserSettings(1, 2, 20.4, 1e-2, 1e-1, 2.0, 0)
    t = torch.zeros(64, dtype=torch.uint8, device="cuda")   # never touched: the size check comes first
    w, h = 3840, 2161
    assert w * h > pyvrt.DENOISE_STATIC_MAX_PIXELS
    with pytest.raises(pyvrt.VrtError) as e:
        pyvrt.denoise_device_static(gpu_ctx, t.data_ptr(), t.data_ptr() + 16, t.data_ptr() + 32, (w, h), ds, t.data_ptr() + 48)
    assert e.value.code == pyvrt.E_INVALID and "VRT_DENOISE_STATIC_MAX_PIXELS" in str(e.value)


def test_denoise_static_at_the_maximum_size(gpu_ctx):
    """The whole array in use: 3840 x 2160, two iterations (the size of cfg3's frame)."""
    import torch
    w, h = 3840, 2160
    assert w * h == pyvrt.DENOISE_STATIC_MAX_PIXELS
    g = torch.Generator(device="cuda").manual_seed(5)
    tc = torch.randint(0, 256, (h, w, 4), dtype=torch.uint8, device="cuda", generator=g); tc[..., 3] = 0
    tn = (torch.randint(-1, 2, (h, w, 4), dtype=torch.int8, device="cuda", generator=g) * 127).to(torch.int8); tn[..., 3] = 0
    tp = torch.rand((h, w, 4), dtype=torch.float32, device="cuda", generator=g) * 3; tp[..., 3] = 0
    ds = pyvrt.DenoiserSettings(1, 2, 20.4, 1e-2, 1e-1, 2.0, 0)
    want, scratch, got = torch.zeros_like(tc), torch.zeros_like(tc), torch.zeros_like(tc)
    torch.cuda.synchronize()
    pyvrt.denoise_device_on(gpu_ctx, tc.data_ptr(), tn.data_ptr(), tp.data_ptr(), (w, h), ds, want.data_ptr(), scratch.data_ptr())
    pyvrt.denoise_device_static(gpu_ctx, tc.data_ptr(), tn.data_ptr(), tp.data_ptr(), (w, h), ds, got.data_ptr())
    gpu_ctx.sync()
    assert torch.equal(got, want) and bool(got[..., :3].any())
