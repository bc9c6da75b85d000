"""Generates the committed golden fixtures. Run in the build container (needs /root/reference for oracle/_ref):
    python tests/golden/make_golden.py
* lshape.vox / twoinst.vox are written by the REFERENCE's own writer (ogt_vox_write_scene, via oracle/_ref) and
  their dense volumes + palettes are produced by the reference's own parser (ogt_vox_read_scene) followed by the
  flatten of voxel_scene.cpp:53-117 as restated in oracle/ref_vox_shim.cpp.
* nopalette.vox has no RGBA chunk: its palette pins the default-palette table of ogt_vox.h:474.
* render_small.npz: oracle outputs for a tiny frame, to detect accidental changes of the pinned semantics
  (this one is self-generated: the reference has no golden vectors for the shader path — PARITY UNPINNED)."""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from vrtlib import O, assets, default_camera  # noqa: E402

I3 = np.eye(3, dtype=np.float32)
pal = assets.palette_standin(11)


def save(name, data):
    with open(os.path.join(HERE, name), "wb") as f:
        f.write(data)
    rc, vol, p = O.ref_load_vox(data)
    assert rc == 0
    np.savez_compressed(os.path.join(HERE, name.replace(".vox", "_ref.npz")), volume=vol, palette=p)


# 3-voxel L shape in a 4x3x2 model (axis-swap test: SURVEY 8c (iv))
m = np.zeros((4, 3, 2), np.uint8); m[0, 0, 0] = 1; m[1, 0, 0] = 2; m[0, 0, 1] = 3
save("lshape.vox", O.ref_write_vox([m], [(0, I3, (0, 0, 0))], pal, {2: 0.25}))
# two models, three instances, one rotated 90 degrees about z and translated; overlapping writes
a = np.zeros((5, 4, 3), np.uint8); a[:, :, 0] = 4; a[2, 1, 1:] = 5
b = np.zeros((2, 2, 6), np.uint8); b[:] = 6; b[0, 0, 5] = 7
rz = np.array([[0, 1, 0], [-1, 0, 0], [0, 0, 1]], np.float32)   # ogt layout m00 m01 m02 / m10.. (row-vector convention)
save("twoinst.vox", O.ref_write_vox([a, b], [(0, I3, (0, 0, 0)), (1, I3, (1, 1, 3)), (0, rz, (7, -2, 4))], pal, {5: 0.5, 7: 1.0}))
# no RGBA chunk -> default palette
raw = O.ref_write_vox([m], [(0, I3, (0, 0, 0))], pal, None)
i = raw.find(b"RGBA"); assert i > 0
stripped = raw[:i] + raw[i + 12 + 1024:]
# patch MAIN child size
import struct
main = stripped.find(b"MAIN")
stripped = stripped[:main + 8] + struct.pack("<I", len(stripped) - (main + 12)) + stripped[main + 12:]
save("nopalette.vox", stripped)

# nine instances of three models under all the rotation kinds, overlapping on purpose: the flatten's instance order decides the result
# (the device flatten of csrc/vrt_flatten.cu must reproduce it)
rng = np.random.default_rng(21)
rots = [I3, rz, [[-1, 0, 0], [0, -1, 0], [0, 0, 1]], [[0, 0, 1], [0, 1, 0], [-1, 0, 0]], [[1, 0, 0], [0, 0, -1], [0, 1, 0]], [[0, -1, 0], [0, 0, 1], [-1, 0, 0]]]
ms = []
for k, shape in enumerate(((13, 9, 7), (6, 11, 10), (17, 5, 12))):
    mm = ((rng.random(shape) < 0.6) * rng.integers(1, 200, shape)).astype(np.uint8); mm.flat[0] = 10 + k
    ms.append(mm)
inst = [(int(i % 3), np.array(rots[int(rng.integers(0, len(rots)))], np.float32), tuple(int(t) for t in rng.integers(-7, 8, 3))) for i in range(9)]
save("overlap9.vox", O.ref_write_vox(ms, inst, pal, {12: 0.5}))

# small oracle render (self-generated)
rc, vol, p = O.load_vox(open(os.path.join(HERE, "twoinst.vox"), "rb").read())
sc, keep = O.make_scene(vol, p, assets.noise_standin(), assets.sky_standin(64, 32))
cam = default_camera((vol.shape[2], vol.shape[1], vol.shape[0]), (48, 27), pos=(-6.5, 9.25, -11.0), yaw=58.0, pitch=-27.0, frame=5)
out = O.render(sc, cam, O.make_settings())
np.savez_compressed(os.path.join(HERE, "render_small.npz"), **out)
print("golden fixtures written:", sorted(os.listdir(HERE)))
