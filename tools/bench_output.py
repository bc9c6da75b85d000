"""Output stage on the GPU box: blit_kernel timed with CUDA events on the context stream (inputs resident, L2 flushed between launches by
cycling through sources larger than L2 in total), against the measured HBM copy peak; the oracle timed beside it on the host cores.
Prints one JSON object per case. usage: python tools/bench_output.py > gpurun_out/r1_output_stage.json"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "voxel-raytracing_b200"), os.path.join(ROOT, "oracle")]
import torch
import pyvrt
import oracle as O

peak = 6546.9
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", peak))
except Exception:
    pass
ctx = pyvrt.Context(0)
ext = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda:0"))
rng = np.random.default_rng(0)
for (sw, sh, tw, th) in ((3840, 2160, 3840, 2160), (2258, 1270, 3840, 2160), (3840, 2160, 1920, 1080), (7680, 4320, 7680, 4320)):
    nsrc = max(2, int(160e6 // (sw * sh * 4)) + 1)          # > 126 MB of distinct sources between two uses of the same one
    host = rng.integers(0, 256, (sh, sw, 4), dtype=np.uint8)
    srcs = [torch.from_numpy(np.roll(host, k, axis=0)).cuda() for k in range(nsrc)]
    out = torch.zeros((th, tw, 4), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    for flags, tag in ((pyvrt.BLIT_SRGB, "srgb"), (0, "unorm")):
        for k in range(3):
            pyvrt.blit_device(ctx, srcs[k % nsrc].data_ptr(), (sw, sh), out.data_ptr(), (tw, th), flags)
        ctx.sync()
        K = 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext)
        for k in range(K):
            pyvrt.blit_device(ctx, srcs[k % nsrc].data_ptr(), (sw, sh), out.data_ptr(), (tw, th), flags)
        e1.record(ext); e1.synchronize()
        ms = e0.elapsed_time(e1) / K
        scale = min(sw / tw, sh / th)
        src_px = min(sw, tw * scale + 1) * min(sh, th * scale + 1)      # source pixels the footprints cover (centre crop)
        alg = 4.0 * src_px + 4.0 * tw * th
        rec = dict(case=f"{sw}x{sh}->{tw}x{th}", store=tag, ms=ms, algorithmic_bytes=alg, achieved_gbs=alg / ms / 1e6, peak_gbs=peak, frac=alg / ms / 1e6 / peak,
                   Mpixels_s=tw * th / ms / 1e3)
        if tag == "srgb" and tw * th <= 3840 * 2160:
            t0 = time.perf_counter(); want = O.blit(host, (tw, th), srgb=True); rec["oracle_ms_all_cores"] = (time.perf_counter() - t0) * 1e3
            pyvrt.blit_device(ctx, srcs[0].data_ptr(), (sw, sh), out.data_ptr(), (tw, th), flags); ctx.sync()
            rec["identical_to_oracle"] = bool(np.array_equal(out.cpu().numpy(), want))
        print(json.dumps(rec), flush=True)
ctx.close()
