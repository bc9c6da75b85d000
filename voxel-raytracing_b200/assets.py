"""Deterministic STAND-IN assets and a MagicaVoxel .vox writer.

Every asset the reference bundles (resource/treehouse.vox, mandlebulb.vox, floatingcolored.vox,
blue_noise_rgba.png, rustig_koppie.hdr) is a git-LFS pointer stub in /root/reference, so the
named configs run on the seeded stand-ins built here. They are INPUT DATA only (the same bytes go to the CUDA
path and to the oracle); nothing in here is an algorithm under test. Every number produced on them must be
labelled "stand-in for <name>".
"""
from __future__ import annotations

import struct
import numpy as np

# ----------------------------------------------------------------------------------------------- hashing
_M1 = np.uint32(0x7FEB352D)
_M2 = np.uint32(0x846CA68B)


def lowbias32(x):
    """Chris Wellons' lowbias32 integer hash (public domain), vectorised over uint32 arrays."""
    x = np.asarray(x, dtype=np.uint32).copy()
    with np.errstate(over="ignore"):
        x ^= x >> np.uint32(16)
        x *= _M1
        x ^= x >> np.uint32(15)
        x *= _M2
        x ^= x >> np.uint32(16)
    return x


def mix(h, v):
    h = np.asarray(h, dtype=np.uint32)
    v = np.asarray(v, dtype=np.uint32)
    with np.errstate(over="ignore"):
        return lowbias32(h ^ (v + np.uint32(0x9E3779B9) + (h << np.uint32(6)) + (h >> np.uint32(2))))


def hash3(seed, a, b, c):
    return mix(mix(mix(lowbias32(np.uint32(seed)), a), b), c)


# ----------------------------------------------------------------------------------------------- textures
def noise_standin(seed: int = 7) -> np.ndarray:
    """512x512 RGBA8 stand-in for blue_noise_rgba.png: counter-hash WHITE noise (not blue noise)."""
    y, x = np.mgrid[0:512, 0:512].astype(np.uint32)
    h = hash3(seed, x, y, 0)
    return np.stack([(h >> np.uint32(s)) & np.uint32(255) for s in (0, 8, 16, 24)], axis=-1).astype(np.uint8)


def sky_standin(w: int = 2048, h: int = 1024) -> np.ndarray:
    """w x h RGBA32F equirect stand-in for rustig_koppie.hdr: closed-form gradient plus a sun disc."""
    v, u = np.mgrid[0:h, 0:w].astype(np.float32)
    u = (u + np.float32(0.5)) / np.float32(w)
    v = (v + np.float32(0.5)) / np.float32(h)
    t = v  # v = asin(-d.y)/pi + 0.5: 0 = straight up (+y), 1 = straight down
    zen = np.array([0.18, 0.35, 0.80], np.float32)
    hor = np.array([0.75, 0.82, 0.90], np.float32)
    gnd = np.array([0.20, 0.17, 0.14], np.float32)
    up = np.clip((np.float32(0.5) - t) * 2, 0, 1)[..., None]
    dn = np.clip((t - np.float32(0.5)) * 2, 0, 1)[..., None]
    col = np.where(t[..., None] < 0.5, hor + (zen - hor) * np.sqrt(up), hor + (gnd - hor) * np.sqrt(dn))
    du = np.minimum(np.abs(u - np.float32(0.62)), 1 - np.abs(u - np.float32(0.62)))
    sun = np.exp(-((du * 2) ** 2 + (v - np.float32(0.30)) ** 2) / np.float32(0.0004))[..., None]
    col = col + sun * np.array([6.0, 5.0, 3.5], np.float32)
    out = np.concatenate([col, np.ones((h, w, 1), np.float32)], axis=-1)
    return np.ascontiguousarray(out.astype(np.float32))


# ----------------------------------------------------------------------------------------------- .vox writer
def _chunk(cid: bytes, content: bytes, children: bytes = b"") -> bytes:
    return cid + struct.pack("<II", len(content), len(children)) + content + children


def _vstr(s: str) -> bytes:
    b = s.encode("ascii")
    return struct.pack("<I", len(b)) + b


def _dict(d: dict) -> bytes:
    out = struct.pack("<I", len(d))
    for k, v in d.items():
        out += _vstr(k) + _vstr(v)
    return out


def rotation_byte(rows) -> int:
    """Pack a signed permutation matrix given as three rows (MagicaVoxel ROTATION byte, nTRN `_r`)."""
    idx, sign = [], []
    for r in rows:
        nz = [i for i, v in enumerate(r) if v != 0]
        assert len(nz) == 1
        idx.append(nz[0])
        sign.append(1 if r[nz[0]] < 0 else 0)
    return idx[0] | (idx[1] << 2) | (sign[0] << 4) | (sign[1] << 5) | (sign[2] << 6)


def write_vox(models, instances, palette_rgba=None, metal=None, version: int = 150) -> bytes:
    """models: list of uint8 arrays indexed [x, y, z] (<= 256 per axis; 0 = empty).
    instances: list of (model_index, (tx, ty, tz), rotation_rows or None).
    palette_rgba: (256, 4) uint8, entry i = colour of voxel value i (entry 0 unused). metal: {index: float}."""
    body = b""
    for m in models:
        m = np.asarray(m, dtype=np.uint8)
        sx, sy, sz = m.shape
        assert max(sx, sy, sz) <= 256
        xs, ys, zs = np.nonzero(m)
        order = np.lexsort((xs, ys, zs))  # x fastest, like ogt_vox_write_scene
        xs, ys, zs = xs[order], ys[order], zs[order]
        xyzi = np.stack([xs, ys, zs, m[xs, ys, zs]], axis=-1).astype(np.uint8)
        body += _chunk(b"SIZE", struct.pack("<III", sx, sy, sz))
        body += _chunk(b"XYZI", struct.pack("<I", len(xs)) + xyzi.tobytes())
    # scene graph: nTRN 0 -> nGRP 1 -> (nTRN 2+2i -> nSHP 3+2i)
    n = len(instances)
    body += _chunk(b"nTRN", struct.pack("<I", 0) + _dict({}) + struct.pack("<IiII", 1, -1, 0, 1) + _dict({}))
    body += _chunk(b"nGRP", struct.pack("<I", 1) + _dict({}) + struct.pack("<I", n) + b"".join(struct.pack("<I", 2 + 2 * i) for i in range(n)))
    for i, (mi, t, rot) in enumerate(instances):
        frame = {"_t": "%d %d %d" % tuple(int(v) for v in t)}
        if rot is not None:
            frame["_r"] = str(rotation_byte(rot))
        body += _chunk(b"nTRN", struct.pack("<I", 2 + 2 * i) + _dict({}) + struct.pack("<IiII", 3 + 2 * i, -1, 0, 1) + _dict(frame))
        body += _chunk(b"nSHP", struct.pack("<I", 3 + 2 * i) + _dict({}) + struct.pack("<I", 1) + struct.pack("<I", mi) + _dict({}))
    if palette_rgba is not None:
        p = np.asarray(palette_rgba, dtype=np.uint8).reshape(256, 4)
        body += _chunk(b"RGBA", np.concatenate([p[1:], p[:1]]).tobytes())  # file slot i holds the colour of value i+1
    for idx, mv in sorted((metal or {}).items()):
        body += _chunk(b"MATL", struct.pack("<i", int(idx)) + _dict({"_type": "_metal", "_metal": repr(float(mv))}))
    return b"VOX " + struct.pack("<I", version) + _chunk(b"MAIN", b"", body)


def palette_standin(seed: int) -> np.ndarray:
    i = np.arange(256, dtype=np.uint32)
    h = hash3(seed, i, 17, 3)
    p = np.stack([(h >> np.uint32(s)) & np.uint32(255) for s in (0, 8, 16)] + [np.full(256, 255, np.uint32)], axis=-1).astype(np.uint8)
    p[:, :3] = 64 + (p[:, :3].astype(np.uint16) * 3 // 4).astype(np.uint8)
    p[0] = 0
    return p


# ----------------------------------------------------------------------------------------------- stand-in scenes
def _box(m, lo, hi, v):
    m[lo[0]:hi[0], lo[1]:hi[1], lo[2]:hi[2]] = v


def treehouse_standin(seed: int = 1):
    """Stand-in for resource/treehouse.vox (real file: LFS stub). One 126x126x126 model (vox axes: z up):
    ground slab, trunk, platform with railing, cabin, hashed-leaf canopy. Returns write_vox() arguments."""
    n = 126
    m = np.zeros((n, n, n), np.uint8)
    _box(m, (0, 0, 0), (n, n, 3), 1)                      # ground
    _box(m, (56, 56, 3), (70, 70, 80), 2)                 # trunk
    _box(m, (30, 30, 44), (96, 96, 47), 3)                # platform
    for a, b in (((30, 30), (96, 32)), ((30, 94), (96, 96)), ((30, 30), (32, 96)), ((94, 30), (96, 96))):
        _box(m, (a[0], a[1], 47), (b[0], b[1], 53), 4)    # railing
    _box(m, (40, 40, 47), (76, 70, 70), 5)                # cabin
    _box(m, (43, 43, 50), (73, 67, 67), 0)                # hollow inside
    _box(m, (52, 40, 47), (60, 43, 62), 0)                # door
    _box(m, (40, 50, 55), (43, 60, 63), 0)                # window
    x, y, z = np.mgrid[0:n, 0:n, 0:n].astype(np.float32)
    r2 = (x - 63) ** 2 + (y - 63) ** 2 + ((z - 95) * np.float32(1.3)) ** 2
    h = hash3(seed, x.astype(np.uint32), y.astype(np.uint32), z.astype(np.uint32))
    leaf = (r2 < 38 ** 2) & ((h & np.uint32(3)) != 0) & (m == 0)
    m[leaf] = (6 + (h[leaf] >> np.uint32(8)) % np.uint32(3)).astype(np.uint8)
    _box(m, (20, 100, 3), (24, 104, 30), 9)               # a lamp post (metallic)
    pal = palette_standin(seed)
    pal[1] = (70, 120, 50, 255); pal[2] = (110, 80, 50, 255); pal[3] = (150, 110, 70, 255); pal[4] = (130, 95, 60, 255)
    pal[5] = (170, 60, 50, 255); pal[6] = (40, 140, 40, 255); pal[7] = (60, 160, 50, 255); pal[8] = (30, 110, 45, 255)
    pal[9] = (200, 200, 210, 255)
    return dict(models=[m], instances=[(0, (0, 0, 0), None)], palette_rgba=pal, metal={9: 0.5})


def mandelbulb_standin(n: int = 256):
    """Stand-in for resource/mandlebulb.vox: power-8 Mandelbulb on an n^3 grid (escape radius 2, 8 iterations),
    palette index = 1 + (iterations survived * 13) % 255 for points that never escape... points that escape are empty."""
    m = np.zeros((n, n, n), np.uint8)
    lin = (np.arange(n, dtype=np.float64) + 0.5) / n * 2.4 - 1.2
    for k in range(n):  # slabs along z to bound memory
        cx, cy = np.meshgrid(lin, lin, indexing="ij")
        cz = np.full_like(cx, lin[k])
        x, y, z = cx.copy(), cy.copy(), cz.copy()
        alive = np.ones(cx.shape, bool)
        minr = np.full(cx.shape, 10.0)
        for _ in range(8):
            r = np.sqrt(x * x + y * y + z * z)
            minr = np.minimum(minr, r)
            alive &= r <= 2.0
            r = np.where(alive, r, 0.0)
            th = np.arctan2(np.sqrt(x * x + y * y), z) * 8
            ph = np.arctan2(y, x) * 8
            r8 = r ** 8
            x = r8 * np.sin(th) * np.cos(ph) + cx
            y = r8 * np.sin(th) * np.sin(ph) + cy
            z = r8 * np.cos(th) + cz
        alive &= np.sqrt(x * x + y * y + z * z) <= 2.0
        m[:, :, k] = np.where(alive, 1 + (np.floor(minr * 40).astype(np.int64) % 8), 0).astype(np.uint8)
    pal = palette_standin(2)
    return dict(models=[m], instances=[(0, (0, 0, 0), None)], palette_rgba=pal, metal={})


def floating_standin(seed: int = 3, n: int = 128):
    """Stand-in for resource/floatingcolored.vox: 2x2x2 instances (distinct translations, two of them rotated)
    of one n^3 model of hashed floating coloured blobs; palette entries 3 and 7 are metallic."""
    x, y, z = np.mgrid[0:n, 0:n, 0:n].astype(np.int64)
    m = np.zeros((n, n, n), np.uint8)
    cell = 32
    cxs = np.arange(0, n, cell)
    for ix in cxs:
        for iy in cxs:
            for iz in cxs:
                h = int(hash3(seed, ix // cell, iy // cell, iz // cell))
                if h & 1:
                    continue
                r = 5 + (h >> 4) % 9
                c = (ix + r + 1 + (h >> 8) % (cell - 2 * r - 2), iy + r + 1 + (h >> 13) % (cell - 2 * r - 2), iz + r + 1 + (h >> 18) % (cell - 2 * r - 2))
                col = 1 + (h >> 24) % 12
                d2 = (x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2
                m[d2 <= r * r] = col
    rz = ((0, -1, 0), (1, 0, 0), (0, 0, 1))   # 90 degrees about z
    inst = []
    for i, (ax, ay, az) in enumerate([(a, b, c) for a in (0, 1) for b in (0, 1) for c in (0, 1)]):
        inst.append((0, (ax * (n + 8), ay * (n + 8), az * (n + 8)), rz if i in (3, 5) else None))
    pal = palette_standin(seed)
    return dict(models=[m], instances=inst, palette_rgba=pal, metal={3: 0.5, 7: 0.8})


STANDINS = {"treehouse": treehouse_standin, "mandelbulb": mandelbulb_standin, "floatingcolored": floating_standin}


def standin_vox_bytes(name: str) -> bytes:
    return write_vox(**STANDINS[name]())
