"""Deterministic synthetic inputs for BASELINE.json's configs.

Analytic scenes rendered on the host in float64 and rounded once to float32; the same
arrays are fed to the CUDA path, the CPU oracle and the reference build.  World frame is
y-down, z-forward like the camera frame; no surface lies within the truncation band of
the world cube [0, L)^3 (Q23).
"""
import numpy as np


def intrinsics(w, h):
    f = 500.0 * w / 640.0
    return (f, f, w / 2.0, h / 2.0)


def orbit_pose(k, step_deg=1.2, radius=2.0, height=-0.3, center=(0.0, 0.0, 2.0)):
    """Camera k of the C2 orbit: circle of `radius` around `center` in the xz-plane at
    y = height, looking at the centre.  Returns (R, t) float64, camera -> world."""
    th = np.deg2rad(step_deg * k)
    c = np.array(center, dtype=np.float64)
    pos = np.array([c[0] + radius * np.sin(th), height, c[2] - radius * np.cos(th)])
    z = c - pos
    z /= np.linalg.norm(z)
    down = np.array([0.0, 1.0, 0.0])
    x = np.cross(down, z)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    R = np.stack([x, y, z], axis=1)
    return R, pos


def _rays(w, h, K, R):
    fx, fy, cx, cy = K
    u = np.arange(w, dtype=np.float64) + 0.5
    v = np.arange(h, dtype=np.float64) + 0.5
    uu, vv = np.meshgrid(u, v)
    d_cam = np.stack([(uu - cx) / fx, (vv - cy) / fy, np.ones_like(uu)], axis=-1)
    return d_cam @ R.T  # world directions, z_cam = 1 => ray parameter t is the z depth


def render_sphere_plane(w, h, K, R, t, sphere_c=(0.0, 0.0, 2.0), sphere_r=0.5, plane_y=0.5, max_depth=5.0,
                        with_color=False):
    """z-depth image of a sphere resting on the plane y = plane_y (0 = miss / out of range)."""
    d = _rays(w, h, K, R)
    o = np.asarray(t, dtype=np.float64)
    depth = np.full((h, w), np.inf)
    # plane
    with np.errstate(divide="ignore", invalid="ignore"):
        tp = (plane_y - o[1]) / d[..., 1]
    tp = np.where((d[..., 1] > 1e-12) & (tp > 0), tp, np.inf)
    depth = np.minimum(depth, tp)
    # sphere
    oc = o - np.asarray(sphere_c, dtype=np.float64)
    a = np.sum(d * d, axis=-1)
    b = 2.0 * (d @ oc)
    c = oc @ oc - sphere_r * sphere_r
    disc = b * b - 4 * a * c
    ts = np.where(disc >= 0, (-b - np.sqrt(np.maximum(disc, 0))) / (2 * a), np.inf)
    ts = np.where(ts > 0, ts, np.inf)
    depth = np.minimum(depth, ts)
    valid = np.isfinite(depth) & (depth <= max_depth) & (depth >= 0.1)
    out = np.where(valid, depth, 0.0).astype(np.float32)
    if not with_color:
        return out
    X = o + d * np.where(valid, depth, 0.0)[..., None]
    checker = ((np.floor(X[..., 0] / 0.1) + np.floor(X[..., 2] / 0.1)) % 2)
    col = np.stack([0.2 + 0.6 * checker, np.full_like(checker, 0.5), 0.2 + 0.6 * (X[..., 1] - np.floor(X[..., 1]))],
                   axis=-1)
    col = np.clip(col, 0.05, 0.95)
    col = np.where(valid[..., None], col, 0.5)
    return out, col.astype(np.float32)


def render_room(w, h, K, R, t, half=(4.0, 1.5, 4.0), max_depth=8.0):
    """Interior of an axis-aligned room centred at the origin + 3 boxes (config C4)."""
    d = _rays(w, h, K, R)
    o = np.asarray(t, dtype=np.float64)
    depth = np.full((h, w), np.inf)
    for ax in range(3):
        for sgn in (-1.0, 1.0):
            with np.errstate(divide="ignore", invalid="ignore"):
                tt = (sgn * half[ax] - o[ax]) / d[..., ax]
            tt = np.where((sgn * d[..., ax] > 1e-12) & (tt > 0), tt, np.inf)
            depth = np.minimum(depth, tt)
    boxes = [((-2.0, 0.5, 2.0), (-1.0, 1.5, 3.0)), ((1.0, 0.7, -3.0), (2.2, 1.5, -2.0)), ((2.5, 0.2, 1.0), (3.2, 1.5, 2.5))]
    for lo, hi in boxes:
        lo = np.array(lo); hi = np.array(hi)
        with np.errstate(divide="ignore", invalid="ignore"):
            t0 = (lo - o) / d
            t1 = (hi - o) / d
        tn = np.max(np.minimum(t0, t1), axis=-1)
        tf = np.min(np.maximum(t0, t1), axis=-1)
        hit = (tn <= tf) & (tn > 0)
        depth = np.minimum(depth, np.where(hit, tn, np.inf))
    valid = np.isfinite(depth) & (depth <= max_depth) & (depth >= 0.1)
    return np.where(valid, depth, 0.0).astype(np.float32)


def rotation_y(deg):
    a = np.deg2rad(deg)
    return np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]])
