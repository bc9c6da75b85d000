"""Small ragged cases for the kernels added at the end of round 1 (blit, instance flatten, edge-blind denoiser iteration) on small,
ragged inputs, each checked against the oracle. usage: python tools/small_cases.py (compute-sanitizer is closed on this pool: small ragged cases against the CPU reference instead)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "voxel-raytracing_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np
import torch
import pyvrt
import oracle as O

ctx = pyvrt.Context(0)
rng = np.random.default_rng(0)
for (sw, sh, tw, th) in ((33, 17, 70, 45), (64, 40, 31, 19), (1, 1, 5, 3), (40, 30, 40, 30)):
    src = rng.integers(0, 256, (sh, sw, 4), dtype=np.uint8)
    t = torch.from_numpy(src).cuda(); out = torch.zeros((th, tw, 4), dtype=torch.uint8, device="cuda"); torch.cuda.synchronize()
    for flags in (0, 1, 3):
        pyvrt.blit_device(ctx, t.data_ptr(), (sw, sh), out.data_ptr(), (tw, th), flags); ctx.sync()
        want = O.blit(src, (tw, th), srgb=bool(flags & 1))
        if flags & 2: want = want[..., [2, 1, 0, 3]]
        assert np.array_equal(out.cpu().numpy(), want), (sw, sh, tw, th, flags)
print("blit ok")
for name in ("lshape", "twoinst", "nopalette", "overlap9"):
    data = open(os.path.join(ROOT, "tests", "golden", name + ".vox"), "rb").read()
    ref = np.load(os.path.join(ROOT, "tests", "golden", name + "_ref.npz"))
    gs = pyvrt.VoxelScene.from_vox_bytes(ctx, data)
    assert np.array_equal(gs.download_volume(), ref["volume"]), name
    gs.destroy()
print("flatten ok")
h, w = 45, 71
color = rng.integers(0, 256, (h, w, 4), dtype=np.uint8); normal = rng.integers(-127, 128, (h, w, 4)).astype(np.int8)
pos = (rng.random((h, w, 4), dtype=np.float32) * 10).astype(np.float32)
tc, tn, tp = torch.from_numpy(color).cuda(), torch.from_numpy(normal).cuda(), torch.from_numpy(pos).cuda()
for mode in (0, 1):
    for iters in (1, 2):
        out = torch.zeros_like(tc); torch.cuda.synchronize()
        pyvrt.denoise_device(ctx, tc.data_ptr(), tn.data_ptr(), tp.data_ptr(), (w, h), pyvrt.DenoiserSettings(1, iters, 20.4, 1e-2, 1e-1, 2.0, mode), out.data_ptr()); ctx.sync()
        d = np.abs(out.cpu().numpy().astype(int) - O.denoise(color, normal, pos, iterations=iters, mode=mode).astype(int))
        assert d.max() <= (0 if iters == 1 else 1), (mode, iters, d.max())
print("denoise ok")
ctx.close()
