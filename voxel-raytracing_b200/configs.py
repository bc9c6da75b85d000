"""The five BASELINE.json configs as data: which scene, resolution, passes and camera path. Host-side only; used by
bench.py and the tests. Scenes named after the reference's bundled assets are STAND-INS (the real files are git-LFS
pointer stubs in the reference checkout: SURVEY.md F1) and every result on them says so."""
from __future__ import annotations

import math
import os
import numpy as np

import assets
import pyvrt

CACHE = os.environ.get("VRT_ASSET_CACHE", "/tmp/vrt_assets")

P, S, A, R, SH = pyvrt.PASS_PRIMARY, pyvrt.PASS_SHADOW, pyvrt.PASS_AO, pyvrt.PASS_REFLECT, pyvrt.PASS_SHADE

CONFIGS = {
    # name: scene spec, resolution, pass mask, aoSamples, denoise, label
    "cfg1": dict(scene=("vox", "treehouse", {}), size=(1280, 720), passes=P, ao=0, denoise=False,
                 label="stand-in for resource/treehouse.vox (LFS stub), primary rays only, 1280x720"),
    "cfg2": dict(scene=("vox", "mandelbulb", {"n": 256}), size=(1920, 1080), passes=P | S | SH, ao=0, denoise=False,
                 label="stand-in for resource/mandlebulb.vox (LFS stub), primary + hard shadow, 1920x1080"),
    "cfg3": dict(scene=("vox", "floatingcolored", {}), size=(3840, 2160), passes=P | S | A | R | SH, ao=1, denoise=True,
                 label="stand-in for resource/floatingcolored.vox + stand-in sky/noise (LFS stubs), 1 AO bounce + shadow + denoiser, 3840x2160"),
    "cfg4": dict(scene=("gen", 1, dict(seed=42, whd=(1024, 1024, 1024))), size=(3840, 2160), passes=P | S | A | R | SH, ao=4, denoise=False,
                 label="synthetic 1024^3 noise terrain (seed 42), primary+shadow+AO(4)+reflections, 3840x2160"),
    "cfg5": dict(scene=("gen", 2, dict(seed=1234, whd=(2048, 2048, 2048))), size=(7680, 4320), passes=P | S | A | R | SH, ao=4, denoise=False,
                 label="synthetic 2048^3 sparse volume (seed 1234), 256-frame orbit, 7680x4320"),
}


def standin_vox_path(name: str, **kw) -> str:
    """Writes (once) the stand-in .vox into the cache and returns its path; it is then loaded through the same
    vrt_scene_load_vox as a real file would be."""
    os.makedirs(CACHE, exist_ok=True)
    key = name + "".join(f"_{k}{v}" for k, v in sorted(kw.items()))
    path = os.path.join(CACHE, key + ".vox")
    if not os.path.exists(path):
        data = assets.write_vox(**assets.STANDINS[name](**kw))
        with open(path + f".tmp{os.getpid()}", "wb") as f:
            f.write(data)
        os.replace(path + f".tmp{os.getpid()}", path)
    return path


def build_scene(ctx: pyvrt.Context, cfg: dict, whd_override=None) -> pyvrt.VoxelScene:
    kind, what, kw = cfg["scene"]
    if kind == "vox":
        sc = pyvrt.VoxelScene.from_vox_file(ctx, standin_vox_path(what, **kw))
    else:
        w, h, d = whd_override or kw["whd"]
        sc = pyvrt.VoxelScene.generate(ctx, what, kw["seed"], w, h, d)
    sc.set_noise(assets.noise_standin())
    sc.set_skybox(assets.sky_standin(2048, 1024))
    return sc


def settings_for(cfg: dict, gbuffer=True, instrument=False) -> pyvrt.RenderSettings:
    s = pyvrt.default_settings()
    s.pass_mask = cfg["passes"]
    s.out_mask = (pyvrt.OUT_GBUFFER if gbuffer else 0) | (pyvrt.OUT_INSTRUMENT if instrument else 0)
    s.vol.aoSamples = cfg["ao"]
    s.denoise.enable = 1 if cfg["denoise"] else 0
    return s


def camera_pose(name: str, bounds, k: int, nframes: int = 256):
    """(pos, yaw_deg, pitch_deg, frame) of step k. All poses avoid exactly axis-aligned rays."""
    w, h, d = bounds
    if name == "cfg1":
        return (8.0, 8.0 + 0.125 * (k % 8), -50.0), 90.0, 0.0, k            # the reference's default camera (voxel_renderer.cpp:20)
    if name == "cfg4":
        # walk over the terrain, looking slightly down; MAX_RAY_STEPS = 512 bounds the view distance
        a = 2.0 * math.pi * (k % nframes) / nframes
        pos = (0.5 * w + 0.30 * w * math.cos(a) + 0.37, 0.62 * h + 0.11, 0.5 * d + 0.30 * d * math.sin(a) + 0.29)
        return pos, math.degrees(a) + 200.3, -35.0, k
    if name == "cfg5":
        # orbit about the volume centre: radius 0.6*w (inside the 512-step horizon of most cells), pitch -20
        a = 2.0 * math.pi * (k % nframes) / nframes
        pos = (0.5 * w + 0.6 * w * math.cos(a) + 0.41, 0.62 * h + 0.23, 0.5 * d + 0.6 * d * math.sin(a) + 0.13)
        return pos, math.degrees(a) + 180.0 + 0.37, -20.0, k
    # cfg2 / cfg3: orbit outside the model looking at its centre
    a = 2.0 * math.pi * ((k % nframes) / nframes) + 0.61
    r = 0.95 * max(w, d)
    pos = (0.5 * w + r * math.cos(a), 0.5 * h + 0.55 * h, 0.5 * d + r * math.sin(a))
    v = np.array([0.5 * w, 0.5 * h, 0.5 * d]) - np.array(pos)
    return pos, math.degrees(math.atan2(v[2], v[0])), math.degrees(math.atan2(v[1], math.hypot(v[0], v[2]))), k


def make_camera(name: str, bounds, size, k: int) -> pyvrt.Camera:
    pos, yaw, pitch, frame = camera_pose(name, bounds, k)
    return pyvrt.make_camera(pos, yaw, pitch, bounds, size, frame=frame)
