"""TEST INFRASTRUCTURE (oracle) — CPU restatement of the reference's volume -> point-cloud handoff.

Follows /root/reference/plant3dvision/proc3d.py:490-570 (``vol2pcd``) statement by statement, minus
the Open3D container (points / normals are returned as float64 arrays) and minus joblib (the per-point
function is the same; it is run in a loop or, for larger cases, vectorised with the same operations).
The two third-party calls are the reference's own dependencies, imported here directly:
``scipy.ndimage.distance_transform_edt`` and ``scipy.ndimage.gaussian_filter`` (scipy is unpinned in
the reference's requirements; this image has scipy 1.18).  ``multiclass_select`` restates
/root/reference/plant3dvision/tasks/proc3d.py:80-115 (argmax + contrast / score gates).

PARITY PINNING: pinned to the reference's own text - tests/golden/make_golden_f.py cuts `vol2pcd`, `index2point`,
`backproject_points` and the scoring loop of `SegmentedPointCloud.run` out of the reference's source files and executes
them as written; tests/test_f_rows_golden.py checks this module against the committed outputs
(tests/golden/f_rows_golden.npz): bit-identical for the per-point loop form and the scoring, <= 1e-12 for the
vectorised form (np.linalg.norm vs x*x + y*y + z*z).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
import numpy as np
from scipy.ndimage import distance_transform_edt, gaussian_filter


def index2point(indexes, origin, voxel_size):
    """proc3d.py:28-45."""
    return voxel_size * indexes + np.asarray(origin)[np.newaxis, :]


def signed_distance(volume):
    """proc3d.py:515-522: binarise, two EDTs, signed half-voxel-shifted distance (float64)."""
    volume = 1.0 * (volume > 0.5)
    dist = distance_transform_edt(volume)
    mdist = distance_transform_edt(1 - volume)
    return np.where(dist > 0.5, dist - 0.5, -mdist + 0.5)


def vol2pcd(volume, origin, voxel_size, level_set_value=0, loop=False):
    """Returns (points [M,3], normals [M,3]) as the reference's PointCloud holds them after
    ``normalize_normals()``.  ``loop=True`` runs the reference's per-point function verbatim
    (np.linalg.norm per point); the default vectorises it with the same float64 operations
    (x*x + y*y + z*z summed left to right)."""
    dist = signed_distance(volume)
    gx, gy, gz = np.gradient(dist)
    gx = gaussian_filter(gx, 1)
    gy = gaussian_filter(gy, 1)
    gz = gaussian_filter(gz, 1)
    on_edge = (dist > -level_set_value) * (dist <= -level_set_value + np.sqrt(3))
    x, y, z = np.nonzero(on_edge)
    if loop:
        pts, normals = [], []
        for i in range(len(x)):
            grad = np.array([gx[x[i], y[i], z[i]], gy[x[i], y[i], z[i]], gz[x[i], y[i], z[i]]])
            grad_norm = np.linalg.norm(grad)
            if grad_norm > 0:
                grad_normalized = grad / grad_norm
                val = dist[x[i], y[i], z[i]] + level_set_value - np.sqrt(3) / 2
                pts.append([x[i] - grad_normalized[0] * val, y[i] - grad_normalized[1] * val,
                            z[i] - grad_normalized[2] * val])
                normals.append([-grad_normalized[0], -grad_normalized[1], -grad_normalized[2]])
        pts = np.array(pts, dtype=np.float64).reshape(-1, 3)
        normals = np.array(normals, dtype=np.float64).reshape(-1, 3)
    else:
        g = np.stack([gx[x, y, z], gy[x, y, z], gz[x, y, z]], axis=1)
        grad_norm = np.sqrt(g[:, 0] * g[:, 0] + g[:, 1] * g[:, 1] + g[:, 2] * g[:, 2])
        keep = grad_norm > 0
        g, grad_norm = g[keep], grad_norm[keep]
        idx = np.stack([x, y, z], axis=1)[keep].astype(np.float64)
        gn = g / grad_norm[:, None]
        val = dist[x, y, z][keep] + level_set_value - np.sqrt(3) / 2
        pts = idx - gn * val[:, None]
        normals = -gn
    pts = index2point(pts, np.asarray(origin, dtype=np.float64), voxel_size)
    # open3d PointCloud.normalize_normals(): n / |n| per normal
    nn = np.sqrt(normals[:, 0] * normals[:, 0] + normals[:, 1] * normals[:, 1] + normals[:, 2] * normals[:, 2])
    normals = normals / nn[:, None] if len(normals) else normals
    return pts, normals


def multiclass_select(voxels, background_prior=1.0, min_contrast=10.0, min_score=0.2):
    """tasks/proc3d.py:80-115: ``voxels`` is the {label: volume} mapping of the Voxels NPZ; returns
    {label: boolean volume} for every non-background label, what the task hands to vol2pcd."""
    l = list(voxels.keys())
    res = np.zeros((*voxels[l[0]].shape, len(l)))
    for i in range(len(l)):
        res[:, :, :, i] = voxels[l[i]]
    for i in range(len(l)):
        if l[i] == 'background':
            res[:, :, :, i] *= background_prior
    res_idx = np.argmax(res, axis=3)
    out = {}
    for i in range(len(l)):
        if l[i] != 'background':
            pred_no_c = np.max(np.delete(res, i, axis=3), axis=3)
            pred_c = (res_idx == i)
            if min_contrast > 1.0:
                pred_c *= (pred_c > (min_contrast * pred_no_c))
            pred_c *= (pred_c > min_score)
            out[l[i]] = pred_c
    return out


def backproject_points(points, K, rot, tvec):
    """proc3d.py:655-659."""
    x = rot @ points.transpose() + tvec[:, np.newaxis]
    x = K @ x
    x = x / x[2, :][np.newaxis, :]
    return x[:2, :].transpose()


def score_points(points, views, labels):
    """tasks/proc3d.py:203-230 (SegmentedPointCloud.run scoring): ``views`` = [(label, camera dict, mask)], the
    per-point Python loop vectorised with the same operations.  Returns (pts_labels, scores)."""
    pts = np.asarray(points)
    scores = np.zeros((len(labels), len(pts)))
    for label, camera, mask in views:
        if label not in labels:
            continue
        if camera is None:
            continue
        rotmat = np.array(camera["rotmat"])
        tvec = np.array(camera["tvec"])
        intrinsics = camera["camera_model"]["params"]
        K = np.array([[intrinsics[0], 0, intrinsics[2]], [0, intrinsics[1], intrinsics[3]], [0, 0, 1]])
        with np.errstate(all="ignore"):
            pixels = np.asarray(backproject_points(pts, K, rotmat, tvec) + 0.5, dtype=int)
        label_idx = labels.index(label)
        ok = (pixels[:, 0] >= 0) & (pixels[:, 0] < mask.shape[1]) & (pixels[:, 1] >= 0) & (pixels[:, 1] < mask.shape[0])
        idx = np.nonzero(ok)[0]
        scores[label_idx, idx] += mask[pixels[idx, 1], pixels[idx, 0]]
    return np.argmax(scores, axis=0).flatten(), scores
