"""Synthetic, pose-consistent scans for the workloads of BASELINE.md §3 that have no bundled
images (C2 real-plant geometry, C3 scale variant, C4 cylinder, C5 1024^3): cameras on a circle
(the ``[ScanPath.kwargs]`` of tests/testdata/real_plant/scan.toml in the reference), a seeded
phantom of capsules standing in for stem / organs, and its analytic silhouettes rasterised per
view.  Pure data generation - nothing here is on the timed path.  Uses torch so that the big
configurations can be rasterised on the GPU; works on CPU tensors too.
"""
import math

import numpy as np
import torch


def circle_cameras(n, radius, center_xy, z, target_z, f, W, H, start_deg=0.0):
    """``n`` pinhole cameras on a horizontal circle looking at the vertical axis through
    ``center_xy``.  Returns float64 ``[n, 16]``: fx fy cx cy | rot (row-major world->camera) | tvec,
    i.e. the ``camera_model.params[0:4]`` / ``rotmat`` / ``tvec`` the reference reads at
    plant3dvision/cl.py:293-296 (OpenCV convention: x right, y down, z forward)."""
    cams = np.zeros((n, 16), dtype=np.float64)
    for v in range(n):
        th = math.radians(start_deg) + 2.0 * math.pi * v / n
        C = np.array([center_xy[0] + radius * math.cos(th), center_xy[1] + radius * math.sin(th), z])
        fwd = np.array([center_xy[0], center_xy[1], target_z]) - C
        fwd /= np.linalg.norm(fwd)
        right = np.cross(fwd, np.array([0.0, 0.0, 1.0]))
        right /= np.linalg.norm(right)
        down = np.cross(fwd, right)
        R = np.stack([right, down, fwd])
        cams[v, 0:4] = [f, f, W / 2.0, H / 2.0]
        cams[v, 4:13] = R.reshape(-1)
        cams[v, 13:16] = -R @ C
    return cams


def phantom(bbox, n_parts=40, seed=0, n_labels=1):
    """Seeded plant-like phantom inside ``bbox`` = ((x0,x1),(y0,y1),(z0,z1)): one vertical stem
    capsule plus ``n_parts`` thin capsules branching from it.  Returns (capsules ``[P, 7]``
    float64 rows ax ay az bx by bz radius, label ids ``[P]`` in ``[0, n_labels)``)."""
    rng = np.random.default_rng(seed)
    (x0, x1), (y0, y1), (z0, z1) = bbox
    cx, cy = 0.5 * (x0 + x1), 0.5 * (y0 + y1)
    ext = min(x1 - x0, y1 - y0)
    hz = z1 - z0
    caps = [[cx, cy, z0 + 0.02 * hz, cx + 0.02 * ext, cy - 0.01 * ext, z0 + 0.9 * hz, 0.012 * ext]]
    labels = [0]
    for p in range(n_parts):
        t = rng.uniform(0.1, 0.88)
        base = np.array([cx + 0.02 * ext * t, cy - 0.01 * ext * t, z0 + (0.02 + 0.88 * t) * hz])
        ang = rng.uniform(0, 2 * math.pi)
        length = rng.uniform(0.08, 0.35) * ext
        rise = rng.uniform(-0.1, 0.4) * length
        tip = base + np.array([length * math.cos(ang), length * math.sin(ang), rise])
        tip[0] = min(max(tip[0], x0 + 0.03 * ext), x1 - 0.03 * ext)
        tip[1] = min(max(tip[1], y0 + 0.03 * ext), y1 - 0.03 * ext)
        tip[2] = min(max(tip[2], z0 + 0.02 * hz), z1 - 0.02 * hz)
        caps.append([*base, *tip, rng.uniform(0.004, 0.02) * ext])
        labels.append(1 + (p % max(n_labels - 1, 1)) if n_labels > 1 else 0)
    return np.array(caps, dtype=np.float64), np.array(labels, dtype=np.int64)


def cylinder_phantom(radius, z0, z1, center_xy=(0.0, 0.0)):
    """C4: one vertical cylinder (as a capsule)."""
    return (np.array([[center_xy[0], center_xy[1], z0, center_xy[0], center_xy[1], z1, radius]], dtype=np.float64),
            np.zeros(1, dtype=np.int64))


@torch.no_grad()
def render_masks(cams, capsules, H, W, device="cpu", soft_px=0.0, select=None, out=None):
    """Rasterise capsule silhouettes: uint8 ``[n, H, W]`` (255 inside, 0 outside; with
    ``soft_px`` > 0 a linear ramp of that many pixels at the edge, as a stand-in for a
    probability map).  A pixel is inside a capsule when its distance to the projected segment is
    below the projected radius (interpolated along the segment).  ``select``: boolean ``[P]``
    choosing which capsules to draw.  ``out``: optional preallocated uint8 tensor."""
    cams_t = torch.as_tensor(np.asarray(cams), dtype=torch.float64, device=device)
    caps = torch.as_tensor(np.asarray(capsules), dtype=torch.float64, device=device)
    if select is not None:
        caps = caps[torch.as_tensor(np.asarray(select), device=device)]
    n = cams_t.shape[0]
    if out is None:
        out = torch.zeros((n, H, W), dtype=torch.uint8, device=device)
    else:
        out.zero_()
    for v in range(n):
        fx, fy, cx, cy = [float(t) for t in cams_t[v, 0:4]]
        R = cams_t[v, 4:13].reshape(3, 3)
        t = cams_t[v, 13:16]
        A = caps[:, 0:3] @ R.T + t
        B = caps[:, 3:6] @ R.T + t
        for p in range(caps.shape[0]):
            az, bz = float(A[p, 2]), float(B[p, 2])
            if az <= 1e-6 or bz <= 1e-6:
                continue
            a = (float(A[p, 0]) / az * fx + cx, float(A[p, 1]) / az * fy + cy)
            b = (float(B[p, 0]) / bz * fx + cx, float(B[p, 1]) / bz * fy + cy)
            ra, rb = float(caps[p, 6]) * fx / az, float(caps[p, 6]) * fx / bz
            rmax = max(ra, rb) + soft_px + 1.0
            xlo, xhi = int(max(0, math.floor(min(a[0], b[0]) - rmax))), int(min(W, math.ceil(max(a[0], b[0]) + rmax) + 1))
            ylo, yhi = int(max(0, math.floor(min(a[1], b[1]) - rmax))), int(min(H, math.ceil(max(a[1], b[1]) + rmax) + 1))
            if xlo >= xhi or ylo >= yhi:
                continue
            xs = torch.arange(xlo, xhi, device=device, dtype=torch.float32) + 0.5
            ys = torch.arange(ylo, yhi, device=device, dtype=torch.float32) + 0.5
            dx, dy = b[0] - a[0], b[1] - a[1]
            L2 = dx * dx + dy * dy
            px = xs[None, :] - a[0]
            py = ys[:, None] - a[1]
            s = ((px * dx + py * dy) / L2).clamp_(0.0, 1.0) if L2 > 1e-12 else torch.zeros_like(px + py)
            ddx = px - s * dx
            ddy = py - s * dy
            dist = torch.sqrt(ddx * ddx + ddy * ddy)
            rad = ra + s * (rb - ra)
            if soft_px > 0:
                val = ((rad - dist) / soft_px + 0.5).clamp_(0.0, 1.0).mul_(255.0).round_().to(torch.uint8)
            else:
                val = (dist < rad).to(torch.uint8) * 255
            region = out[v, ylo:yhi, xlo:xhi]
            torch.maximum(region, val, out=region)
    return out


def shape_origin(bbox, voxel_size):
    """Grid shape / origin as ``Voxels.run`` derives them (plant3dvision/tasks/cl.py:127-147)."""
    (x0, x1), (y0, y1), (z0, z1) = bbox
    shape = [int((x1 - x0) / voxel_size) + 1, int((y1 - y0) / voxel_size) + 1, int((z1 - z0) / voxel_size) + 1]
    return shape, [x0, y0, z0]


# ---- the named workloads of BASELINE.md §3 -------------------------------------------------
def workload(name, scale=1.0):
    """Geometry of a named workload -> dict(bbox, voxel_size, shape, origin, cams, capsules,
    labels, H, W).  ``scale`` < 1 shrinks grid and images together (parity-test sizes)."""
    if name == "C2":  # geom_pipe_real.toml bbox + voxel_size 0.5; scan.toml circle; 1440x1080
        bbox = ((300.0, 450.0), (300.0, 450.0), (-175.0, 100.0))
        vs = 0.5 / scale
        W, H = int(1440 * scale), int(1080 * scale)
        cams = circle_cameras(60, 300.0, (375.0, 375.0), 80.0, -37.5, 1163.685 * scale, W, H)
        caps, labels = phantom(bbox, 40, seed=0, n_labels=6)
    elif name == "C4":  # synthetic cylinder, 512^3, 120 views
        n = max(int(round(512 * scale)), 8)
        vs = 0.25 * 512 / n
        bbox = ((-64.0, -64.0 + vs * (n - 1)), (-64.0, -64.0 + vs * (n - 1)), (-28.0, -28.0 + vs * (n - 1)))
        W, H = int(1440 * scale), int(1080 * scale)
        cams = circle_cameras(120, 400.0, (0.0, 0.0), 100.0, 36.0, 1400.0 * scale, W, H)
        caps, labels = cylinder_phantom(40.0, -100.0, 100.0)
    elif name == "C5":  # 1024^3, 240 views at 6000x4000
        n = max(int(round(1024 * scale)), 8)
        vs = 0.25 * 1024 / n
        bbox = ((-128.0, -128.0 + vs * (n - 1)), (-128.0, -128.0 + vs * (n - 1)), (-128.0, -128.0 + vs * (n - 1)))
        W, H = int(6000 * scale), int(4000 * scale)
        cams = circle_cameras(240, 700.0, (0.0, 0.0), 40.0, 0.0, 8000.0 * scale, W, H)
        caps, labels = phantom(bbox, 40, seed=0, n_labels=6)
    else:
        raise ValueError(name)
    shape, origin = shape_origin(bbox, vs)
    return dict(name=name, bbox=bbox, voxel_size=vs, shape=shape, origin=origin, cams=cams, capsules=caps,
                labels=labels, H=H, W=W)
