"""Seeded synthetic scenes of the shapes BASELINE.json names.

The same bytes go to the CUDA path, the CPU oracle and the bench.  No file IO, no network.

  cloud  : points on the inner surfaces of a 40 m x 6 m x 3 m corridor + 12 pillars, float32 xyz + intensity
           U[0,1), shuffled (no spatial order), numpy default_rng(42)
  poses  : serpentine path, consecutive poses 0.12 m apart (every pose passes the reference's >=0.1 m keyframe
           rule, src/PointCloudProcessor.cpp:1056), looking along the direction of travel with +-10 deg yaw/pitch
           jitter, unit quaternions, pose = (x y z qw qx qy qz) as in the odometry file (cpp:973-978), seed 43
  camera : fx = fy = 1.1763 W, cx = 0.4962 W, cy = 0.5117 H (the reference's 4818/4096 FoV ratio) with the
           reference's plumb_bob coefficients (cpp:57-60)
  images : u8 BGR, 8 seeded integer triangle waves per channel + U[0,8) noise, seed 44+i, passed once through the
           reference's 8-bit BGR->HSV->BGR round trip (cpp:723-741) with cv2 when available
  masks  : u8, ~2 % of pixels == 255 in blobs
"""
from __future__ import annotations

import math

import numpy as np

REF_INTRINSICS = (4818.200388954926, 4819.10345841615, 2032.4178620390019, 1535.1895959282901)  # fx fy cx cy (cpp:57-62)
REF_DISTORTION = (0.003043514741045163, 0.06634739187544138, -0.000217681797407554, -0.0006654964142658197, 0.0)  # cpp:60

CONFIGS = {
    # name: (N points, F frames, W, H)
    "C1": (1_000_000, 20, 1280, 720),
    "C2": (10_000_000, 200, 1920, 1080),
    "C3": (1_000_000, 20, 1280, 720),  # + masks + NID
    "C4": (100_000_000, 2000, 1920, 1080),
    "C5": (50_000_000, 10_000, 640, 480),
    "tiny": (20_000, 3, 320, 240),
}

LX, LY, LZ = 40.0, 6.0, 3.0
_PILLAR_X = (5.0, 11.0, 17.0, 23.0, 29.0, 35.0)
_PILLAR_Y = (0.6, 5.4)
_PILLAR_HALF = 0.25


def make_camera(W: int, H: int):
    """-> (intrinsics[4], distortion[5])"""
    return (1.1763 * W, 1.1763 * W, 0.4962 * W, 0.5117 * H), REF_DISTORTION


def make_cloud(n: int, seed: int = 42) -> np.ndarray:
    """n x 4 float32 (x, y, z, intensity)."""
    rng = np.random.default_rng(seed)
    # faces: (origin, edge u, edge v)
    faces = []
    faces.append(((0, 0, 0), (LX, 0, 0), (0, LY, 0)))      # floor
    faces.append(((0, 0, LZ), (LX, 0, 0), (0, LY, 0)))     # ceiling
    faces.append(((0, 0, 0), (LX, 0, 0), (0, 0, LZ)))      # wall y=0
    faces.append(((0, LY, 0), (LX, 0, 0), (0, 0, LZ)))     # wall y=LY
    faces.append(((0, 0, 0), (0, LY, 0), (0, 0, LZ)))      # end x=0
    faces.append(((LX, 0, 0), (0, LY, 0), (0, 0, LZ)))     # end x=LX
    h = _PILLAR_HALF
    for px in _PILLAR_X:
        for py in _PILLAR_Y:
            faces.append(((px - h, py - h, 0), (2 * h, 0, 0), (0, 0, LZ)))
            faces.append(((px - h, py + h, 0), (2 * h, 0, 0), (0, 0, LZ)))
            faces.append(((px - h, py - h, 0), (0, 2 * h, 0), (0, 0, LZ)))
            faces.append(((px + h, py - h, 0), (0, 2 * h, 0), (0, 0, LZ)))
    org = np.array([f[0] for f in faces], np.float64)
    eu = np.array([f[1] for f in faces], np.float64)
    ev = np.array([f[2] for f in faces], np.float64)
    area = np.linalg.norm(np.cross(eu, ev), axis=1)
    out = np.empty((n, 4), np.float32)
    chunk = 4_000_000
    for lo in range(0, n, chunk):
        m = min(chunk, n - lo)
        fi = rng.choice(len(faces), size=m, p=area / area.sum())
        a = rng.random(m)
        b = rng.random(m)
        p = org[fi] + a[:, None] * eu[fi] + b[:, None] * ev[fi]
        out[lo:lo + m, :3] = p.astype(np.float32)
        out[lo:lo + m, 3] = rng.random(m, dtype=np.float32)
    # the draw order is already iid, i.e. there is no spatial order to exploit
    return out


def _path_vertices(n_needed: int, step: float):
    """Serpentine polyline whose segment lengths are all multiples of `step`."""
    pass_len = 317 * step  # 38.04 m
    x0 = 1.0
    verts = []
    k = 0
    total = 0
    while total < n_needed + 2:
        y = 1.2 + 3 * step * (k % 10)
        z = 1.2 + step * ((k // 10) % 6)
        xa, xb = (x0, x0 + pass_len) if k % 2 == 0 else (x0 + pass_len, x0)
        verts.append((xa, y, z))
        verts.append((xb, y, z))
        total += 300  # a pass yields 317 poses plus the short lateral move: under-count so that enough vertices exist
        k += 1
    return verts


def make_poses(f: int, seed: int = 43, step: float = 0.12):
    """-> (poses f x 7 float64 [x y z qw qx qy qz], timestamps f float64)"""
    rng = np.random.default_rng(seed)
    verts = _path_vertices(f, step)
    # expand the polyline (axis-aligned moves between consecutive passes: y first, then z)
    pts = []
    dirs = []
    cur = np.array(verts[0], np.float64)
    pts.append(cur.copy())
    dirs.append(np.array([1.0, 0.0, 0.0]))
    for v in verts[1:]:
        tgt = np.array(v, np.float64)
        for axis in (0, 1, 2):
            d = tgt[axis] - cur[axis]
            nst = int(round(abs(d) / step))
            if nst == 0:
                continue
            sgn = 1.0 if d > 0 else -1.0
            base = cur.copy()
            for s in range(1, nst + 1):
                p = base.copy()
                p[axis] = base[axis] + sgn * step * s
                dv = np.zeros(3)
                dv[axis] = sgn
                pts.append(p)
                dirs.append(dv)
            cur = pts[-1].copy()
        if len(pts) >= f:
            break
    pts = np.array(pts[:f])
    dirs = np.array(dirs[:f])
    # a vertical move keeps the previous horizontal heading
    for i in range(f):
        if abs(dirs[i, 2]) > 0.5:
            dirs[i] = dirs[i - 1] if i > 0 else np.array([1.0, 0.0, 0.0])
    yaw_j = np.deg2rad(rng.uniform(-10, 10, f))
    pit_j = np.deg2rad(rng.uniform(-10, 10, f))
    poses = np.empty((f, 7), np.float64)
    for i in range(f):
        yaw = math.atan2(dirs[i, 1], dirs[i, 0]) + yaw_j[i]
        pit = pit_j[i]
        fwd = np.array([math.cos(pit) * math.cos(yaw), math.cos(pit) * math.sin(yaw), math.sin(pit)])
        up = np.array([0.0, 0.0, 1.0])
        right = np.cross(fwd, up)
        right /= np.linalg.norm(right)
        down = np.cross(fwd, right)
        R = np.stack([right, down, fwd], axis=1)  # columns: camera x (right), y (down), z (forward) in world
        poses[i, :3] = pts[i]
        poses[i, 3:] = _rot_to_quat(R)
    ts = 1.7e9 + 0.1 * np.arange(f)
    return poses, ts


def _rot_to_quat(R):
    t = np.trace(R)
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        w = 0.25 * s
        x = (R[2, 1] - R[1, 2]) / s
        y = (R[0, 2] - R[2, 0]) / s
        z = (R[1, 0] - R[0, 1]) / s
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = math.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        w = (R[2, 1] - R[1, 2]) / s
        x = 0.25 * s
        y = (R[0, 1] + R[1, 0]) / s
        z = (R[0, 2] + R[2, 0]) / s
    elif R[1, 1] > R[2, 2]:
        s = math.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        w = (R[0, 2] - R[2, 0]) / s
        x = (R[0, 1] + R[1, 0]) / s
        y = 0.25 * s
        z = (R[1, 2] + R[2, 1]) / s
    else:
        s = math.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        w = (R[1, 0] - R[0, 1]) / s
        x = (R[0, 2] + R[2, 0]) / s
        y = (R[1, 2] + R[2, 1]) / s
        z = 0.25 * s
    q = np.array([w, x, y, z])
    return q / np.linalg.norm(q)


def hsv_round_trip(bgr: np.ndarray) -> np.ndarray:
    """The reference's lossy 8-bit BGR->HSV->(x1.0)->BGR (cpp:723-741).  Input preparation, not a kernel."""
    try:
        import cv2
    except Exception:  # pragma: no cover
        return bgr
    hsv = cv2.cvtColor(bgr, cv2.COLOR_BGR2HSV)
    return cv2.cvtColor(hsv, cv2.COLOR_HSV2BGR)


_BASE_CACHE: dict = {}


def _base_pattern(k: int, W: int, H: int) -> np.ndarray:
    """H x W x 3 int16 pattern: 120 + sum of 8 integer triangle waves per channel along seeded directions.
    Integer arithmetic only, so the bytes are identical on every machine (no libm / BLAS in the generator)."""
    key = (k, W, H)
    if key in _BASE_CACHE:
        return _BASE_CACHE[key]
    rng = np.random.default_rng(4400 + k)
    yy = np.arange(H, dtype=np.int32)[:, None]
    xx = np.arange(W, dtype=np.int32)[None, :]
    img = np.empty((H, W, 3), np.int16)
    for c in range(3):
        acc = np.full((H, W), 120, np.int32)
        for _ in range(8):
            a, b = (int(v) for v in rng.integers(-7, 8, 2))
            period = int(rng.integers(64, 513)) * 2
            ph = int(rng.integers(0, period))
            amp = int(rng.integers(4, 17))
            t = (a * xx + b * yy + ph) % period                # 0 .. period-1
            tri = np.abs(t - period // 2)                      # 0 .. period/2
            acc += (amp * (4 * tri - period)) // period        # -amp .. +amp
        img[:, :, c] = acc
    if len(_BASE_CACHE) > 16:
        _BASE_CACHE.clear()
    _BASE_CACHE[key] = img
    return img


def make_image(i: int, W: int, H: int, out: np.ndarray | None = None, hsv: bool = True) -> np.ndarray:
    """H x W x 3 u8 BGR: one of 4 cached integer base patterns, rolled by a frame-dependent offset, plus U[0,8)
    per-pixel noise (seed 44+i), then the reference's 8-bit HSV round trip."""
    rng = np.random.default_rng(44 + i)
    base = _base_pattern(i % 4, W, H)
    img = np.roll(base, ((i * 13) % H, (i * 29) % W), axis=(0, 1))
    img = img + rng.integers(0, 8, size=(H, W, 3), dtype=np.int16)
    u8 = np.clip(img, 0, 255).astype(np.uint8)
    if hsv:
        u8 = hsv_round_trip(u8)
    if out is not None:
        out[...] = u8
        return out
    return u8


def make_mask(i: int, W: int, H: int) -> np.ndarray:
    """H x W u8, ~2 % of pixels == 255 in rectangular blobs, other pixels small class ids."""
    rng = np.random.default_rng(10_000 + i)
    m = rng.integers(0, 4, size=(H, W), dtype=np.uint8)
    target = 0.02 * W * H
    filled = 0
    while filled < target:
        bw = int(rng.integers(max(4, W // 40), max(8, W // 12)))
        bh = int(rng.integers(max(4, H // 40), max(8, H // 12)))
        x0 = int(rng.integers(0, max(1, W - bw)))
        y0 = int(rng.integers(0, max(1, H - bh)))
        m[y0:y0 + bh, x0:x0 + bw] = 255
        filled += bw * bh
    return m


def grey_plane(bgr: np.ndarray) -> np.ndarray:
    """Single-channel float64 plane in [0,1] for the NID cost (pixel/255 as in visual_camera_calibration.cpp:173)."""
    g = (0.114 * bgr[:, :, 0].astype(np.float64) + 0.587 * bgr[:, :, 1] + 0.299 * bgr[:, :, 2])
    return np.floor(g + 0.5).clip(0, 255) * (1.0 / 255.0)


def quat_mul(a, b):
    """Hamilton product of (x,y,z,w) quaternions."""
    ax, ay, az, aw = a
    bx, by, bz, bw = b
    return np.array([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw,
                     aw * bw - ax * bx - ay * by - az * bz])


def make_nid_candidates(n_pert: int = 64, seed: int = 45) -> np.ndarray:
    """identity + n_pert perturbations within +-0.1 m / +-1 deg, Sophus order (qx qy qz qw tx ty tz)."""
    rng = np.random.default_rng(seed)
    out = np.zeros((n_pert + 1, 7), np.float64)
    out[0] = (0, 0, 0, 1, 0, 0, 0)
    for i in range(1, n_pert + 1):
        axis = rng.normal(size=3)
        axis /= np.linalg.norm(axis)
        ang = np.deg2rad(rng.uniform(-1, 1))
        q = np.concatenate([axis * math.sin(ang / 2), [math.cos(ang / 2)]])
        t = rng.uniform(-0.1, 0.1, 3) / math.sqrt(3.0)
        out[i, :4] = q
        out[i, 4:] = t
    return out
