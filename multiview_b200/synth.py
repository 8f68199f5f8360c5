"""Seeded synthetic stereo scenes: sensor images, remap tables, calibration.

numpy only.  Used by tests, bench.py and smoke(); there are no datasets on the GPU box.
Calibration constants q, t, baseline are those of the reference's only calibration fixture
(multiview_cpp/src/testcases/geometry/bcam_file.cpp:6-45, camera C0001024_v2).
"""
from __future__ import annotations

import math

import numpy as np

# rotation-axis / rotation-angle / t / baseline of C0001024_v2
_AXIS = (-7.481678288938867771307e-01, 5.427879688477643682631e-01, 3.816099064258974826735e-01)
_ANGLE = 4.989567261148587784536e-02
BCAM_T = (-9.985351324276062756624e-01, -5.359021536139291913559e-02, 7.461777623479498230963e-03)
BCAM_BASELINE = 1.435400000000000286793e-01


def bcam_quaternion():
    n = math.sqrt(sum(a * a for a in _AXIS))
    s = math.sin(_ANGLE / 2) / n
    return (_AXIS[0] * s, _AXIS[1] * s, _AXIS[2] * s, math.cos(_ANGLE / 2))


def intrinsics(W, H):
    f = 0.55 * W
    return np.array([[f, 0, W / 2.0], [0, f, H / 2.0], [0, 0, 1.0]], np.float64)


def _gauss1d(sigma):
    r = int(math.ceil(3 * sigma))
    x = np.arange(-r, r + 1, dtype=np.float64)
    k = np.exp(-0.5 * (x / sigma) ** 2)
    return (k / k.sum()).astype(np.float32)


def _blur(a, sigma):
    k = _gauss1d(sigma)
    r = len(k) // 2
    p = np.pad(a, ((r, r), (r, r)), mode="reflect").astype(np.float32)
    t = sum(k[i] * p[:, i:i + a.shape[1]] for i in range(len(k)))
    return sum(k[i] * t[i:i + a.shape[0], :] for i in range(len(k)))


def _texture(rng, H, W):
    t = _blur(rng.random((H, W), dtype=np.float32), 1.5)
    t += 0.5 * _blur(rng.random((H, W), dtype=np.float32), 4.0)
    t -= t.min()
    return t / max(float(t.max()), 1e-6)


def _bilinear(img, x, y, fill=255.0):
    """Sample img (H,W,C float32) at float coords; outside -> fill."""
    H, W = img.shape[:2]
    x0 = np.floor(x).astype(np.int64)
    y0 = np.floor(y).astype(np.int64)
    fx = (x - x0).astype(np.float32)[..., None]
    fy = (y - y0).astype(np.float32)[..., None]
    inside = (x0 >= 0) & (x0 < W - 1) & (y0 >= 0) & (y0 < H - 1)
    x0c = np.clip(x0, 0, W - 2)
    y0c = np.clip(y0, 0, H - 2)
    v = (img[y0c, x0c] * (1 - fx) * (1 - fy) + img[y0c, x0c + 1] * fx * (1 - fy) + img[y0c + 1, x0c] * (1 - fx) * fy +
         img[y0c + 1, x0c + 1] * fx * fy)
    return np.where(inside[..., None], v, np.float32(fill))


def disparity_field(rng, W, H, D, frame=0):
    """Piecewise-planar disparity on the right-image grid: floor ramp + 3-6 fronto-parallel boxes."""
    ys = np.arange(H, dtype=np.float32)[:, None]
    d = 4.0 + (0.5 * D - 4.0) * (ys / max(H - 1, 1)) + np.zeros((H, W), np.float32)
    for _ in range(int(rng.integers(3, 7))):
        bw, bh = int(rng.integers(W // 16, W // 6)), int(rng.integers(H // 5, H // 2))
        x0 = int(rng.integers(0, max(W - bw, 1))) + 3 * frame
        y0 = int(rng.integers(0, max(H - bh, 1)))
        dv = float(rng.uniform(0.3 * D, 0.8 * D))
        x0 = min(max(x0, 0), max(W - bw, 0))
        box = d[y0:y0 + bh, x0:x0 + bw]
        box[...] = np.maximum(box, dv)
    return d


def rectified_pair(W, H, D, seed, frame=0):
    """Ideal rectified BGR pair (uint8) whose true disparity lies in [4, 0.8*D]."""
    rng = np.random.default_rng(seed)
    Wb = W + D + 32
    grey = _texture(rng, H, Wb)
    col = np.stack([grey * 0.85 + 0.15 * _texture(rng, H, Wb) for _ in range(3)], axis=-1)
    base = (col * 255.0).astype(np.float32)
    left = base[:, :W].copy()
    d = disparity_field(rng, W, H, D, frame)
    xs = np.arange(W, dtype=np.float32)[None, :] + d
    yy = np.broadcast_to(np.arange(H, dtype=np.float32)[:, None], (H, W))
    right = _bilinear(base, xs, yy, fill=127.0)
    right += rng.uniform(-3, 3, right.shape).astype(np.float32)
    left += rng.uniform(-3, 3, left.shape).astype(np.float32)
    to8 = lambda a: np.clip(np.rint(a), 0, 255).astype(np.uint8)
    return to8(left), to8(right)


def radial_maps(W, H, k1=0.12, shift=(0.0, 0.0)):
    """Analytic rect->sensor maps (CV_32FC1 mapx, mapy); corners land outside the sensor."""
    cx, cy = W / 2.0 + shift[0], H / 2.0 + shift[1]
    xs, ys = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
    nx, ny = (xs - cx), (ys - cy)
    r2 = (nx * nx + ny * ny) / (cx * cx + cy * cy)
    s = 1.0 + k1 * r2
    return (cx + nx * s).astype(np.float32), (cy + ny * s).astype(np.float32)


def sensor_from_rectified(rect, k1=0.12, shift=(0.0, 0.0)):
    """Inverse-warp a rectified image into a 'sensor' image so that remap(sensor, maps) ~ rect."""
    H, W = rect.shape[:2]
    cx, cy = W / 2.0 + shift[0], H / 2.0 + shift[1]
    us, vs = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
    du, dv = us - cx, vs - cy
    px, py = du.copy(), dv.copy()
    for _ in range(6):
        s = 1.0 + k1 * (px * px + py * py) / (cx * cx + cy * cy)
        px, py = du / s, dv / s
    out = _bilinear(rect.astype(np.float32), px + cx, py + cy, fill=255.0)
    return np.clip(np.rint(out), 0, 255).astype(np.uint8)


def _quat_from_rot(R):
    t = R[0, 0] + R[1, 1] + R[2, 2]
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        return ((R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s)
    i = int(np.argmax([R[0, 0], R[1, 1], R[2, 2]]))
    j, k = (i + 1) % 3, (i + 2) % 3
    s = math.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0) * 2
    q = [0.0, 0.0, 0.0, 0.0]
    q[i] = 0.25 * s
    q[j] = (R[j, i] + R[i, j]) / s
    q[k] = (R[k, i] + R[i, k]) / s
    q[3] = (R[k, j] - R[j, k]) / s
    return tuple(q)


def camera_extrinsics(cam, n_cams):
    """Ring of cameras 2.5-3 m up, looking 30-45 deg down at the origin.  Returns et (8 doubles:
    tx,ty,tz, qx,qy,qz,qw, scale) with world = q.rotate(X*scale) + t, as EuclideanTransform::apply."""
    ang = 2 * math.pi * cam / max(n_cams, 1) + 0.3
    rad = 3.5
    pos = np.array([rad * math.cos(ang), rad * math.sin(ang), 2.5 + 0.5 * ((cam * 7) % 5) / 4.0])
    tilt = math.radians(30 + 15 * ((cam * 3) % 4) / 3.0)
    hfwd = -np.array([math.cos(ang), math.sin(ang), 0.0])
    fwd = hfwd * math.cos(tilt) + np.array([0, 0, -1.0]) * math.sin(tilt)
    right = np.cross(fwd, np.array([0, 0, 1.0]))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    R = np.stack([right, down, fwd], axis=1)  # columns: camera x, y, z axes in world
    q = _quat_from_rot(R)
    return (pos[0], pos[1], pos[2], q[0], q[1], q[2], q[3], 1.0)


def make_camera(W, H, D, config=1, cam=0, frame=0, n_cams=1):
    """One binocular camera-frame: 2 sensor images + maps + calibration + extrinsics."""
    seed = 1000 * config + cam * 10 + frame
    left, right = rectified_pair(W, H, D, seed, frame)
    shifts = [(0.013 * W, -0.011 * H), (-0.009 * W, 0.007 * H)]
    sensors, mapx, mapy = [], [], []
    for img, sh in zip((left, right), shifts):
        sensors.append(sensor_from_rectified(img, 0.12, sh))
        mx, my = radial_maps(W, H, 0.12, sh)
        mapx.append(mx)
        mapy.append(my)
    return {
        "W": W, "H": H, "D": D, "seed": seed,
        "sensor": sensors, "mapx": mapx, "mapy": mapy,
        "rect_ideal": (left, right),
        "q": bcam_quaternion(), "t": BCAM_T, "baseline": BCAM_BASELINE, "K": intrinsics(W, H),
        "et": camera_extrinsics(cam, n_cams),
    }


HIST_AABB = (-8.0, -8.0, 8.0, 8.0)  # left, top, right, bottom (lobby-like)
HIST_SZ = 0.10
