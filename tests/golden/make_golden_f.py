"""Generates tests/golden/f_rows_golden.npz: outputs of the reference's OWN text for the SURVEY §8f rows, so that the
oracles of those rows (oracle/vol2pcd_oracle.py) are pinned to the reference and not only to themselves.

Run in the build container only (`python tests/golden/make_golden_f.py`): the reference package cannot be imported
here (open3d, romitask, plantdb, skimage, networkx are absent), so the functions are cut out of the reference's source
files by `ast` where they lie under /root/reference and executed as they are written, with numpy / scipy / joblib from
this image and three small stand-ins for what the text touches besides arithmetic:

  * `o3d`      geometry.PointCloud that records `points` / `normals` and normalises normals as Open3D documents
               (n / |n|); utility.Vector3dVector = numpy array
  * `logger`   swallows messages;  `tqdm` = identity
  * for the scoring loop (tasks/proc3d.py:203-230, a slice of `SegmentedPointCloud.run`): `fs`, `self`, `io`, `proc3d`
    objects exposing exactly the attributes the slice reads

Executed text (reference file:line):
  plant3dvision/proc3d.py:28-45     index2point
  plant3dvision/proc3d.py:490-570   vol2pcd (the joblib per-point function runs as written)
  plant3dvision/proc3d.py:655-659   backproject_points
  plant3dvision/tasks/proc3d.py:182-183  SegmentedPointCloud.is_in_pict
  plant3dvision/tasks/proc3d.py:203-232  the scoring loop and the argmax of SegmentedPointCloud.run
Nothing of the reference is copied into the repository: only the numeric inputs and outputs are stored.
"""
import ast
import os
import sys
import textwrap
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference/plant3dvision"


def function_source(path, name, cls=None):
    src = open(path).read()
    tree = ast.parse(src)
    nodes = tree.body
    if cls is not None:
        nodes = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls).body
    node = next(n for n in nodes if isinstance(n, ast.FunctionDef) and n.name == name)
    return ast.get_source_segment(src, node), node.lineno, node.end_lineno


class _Logger:
    def __getattr__(self, _):
        return lambda *a, **k: None


class _PointCloud:
    def __init__(self):
        self.points = None
        self.normals = None

    def normalize_normals(self):
        n = np.asarray(self.normals, dtype=np.float64)
        if len(n):
            n = n / np.sqrt((n * n).sum(axis=1))[:, None]
        self.normals = n


def reference_namespace():
    from scipy.ndimage import distance_transform_edt, gaussian_filter
    o3d = types.SimpleNamespace(geometry=types.SimpleNamespace(PointCloud=_PointCloud),
                                utility=types.SimpleNamespace(Vector3dVector=lambda a: np.asarray(a, dtype=np.float64)))
    ns = {"np": np, "o3d": o3d, "logger": _Logger(), "tqdm": lambda it, **k: it,
          "distance_transform_edt": distance_transform_edt, "gaussian_filter": gaussian_filter}
    cites = {}
    for name in ("index2point", "vol2pcd", "backproject_points"):
        text, a, b = function_source(f"{REF}/proc3d.py", name)
        exec(compile(text, f"{REF}/proc3d.py:{name}", "exec"), ns)
        cites[name] = f"plant3dvision/proc3d.py:{a}-{b}"
    return ns, cites


def scoring_slice():
    """The statements of SegmentedPointCloud.run from `scores = np.zeros(...)` to `pts_labels = np.argmax(...)`."""
    path = f"{REF}/tasks/proc3d.py"
    text, a, _ = function_source(path, "run", cls="SegmentedPointCloud")
    lines = text.splitlines()
    start = next(i for i, l in enumerate(lines) if l.strip().startswith("scores = np.zeros"))
    end = next(i for i, l in enumerate(lines) if l.strip().startswith("pts_labels = np.argmax"))
    body = textwrap.dedent("\n".join(lines[start:end + 1]))
    in_pict, _, _ = function_source(path, "is_in_pict", cls="SegmentedPointCloud")
    return body, textwrap.dedent(in_pict), f"plant3dvision/tasks/proc3d.py:{a + start}-{a + end}"


class _File:
    def __init__(self, channel, camera, image, colmap_camera=None):
        self._md = {"channel": channel, "camera": camera, "colmap_camera": colmap_camera}
        self.image = image

    def get_metadata(self, key):
        return self._md.get(key)


def run_scoring(ns, pts, files, labels, use_colmap_poses=False):
    body, in_pict, _ = scoring_slice()
    holder = {}
    exec(in_pict, {}, holder)
    self_obj = types.SimpleNamespace(use_colmap_poses=use_colmap_poses)
    self_obj.is_in_pict = types.MethodType(holder["is_in_pict"], self_obj)
    env = {"np": np, "pts": pts, "labels": list(labels), "self": self_obj, "logger": _Logger(),
           "fs": types.SimpleNamespace(get_files=lambda: files),
           "io": types.SimpleNamespace(read_image=lambda fi: fi.image),
           "proc3d": types.SimpleNamespace(backproject_points=ns["backproject_points"])}
    with np.errstate(all="ignore"):
        exec(compile(body, "tasks/proc3d.py:scoring", "exec"), env)
    return env["scores"], env["pts_labels"]


def volumes():
    rng = np.random.default_rng(42)
    out = {}
    g = np.indices((24, 22, 26)).astype(np.float64)
    ball = ((g[0] - 11.5) ** 2 + (g[1] - 10) ** 2 + (g[2] - 13) ** 2) < 7.3 ** 2
    out["ball"] = (ball.astype(np.float64), np.array([1.0, -2.0, 0.5]), 0.7, 0.0)
    out["ball_ls1"] = (ball.astype(np.float64), np.array([0.0, 0.0, 0.0]), 1.0, 1.0)
    blobs = np.zeros((30, 28, 33), dtype=np.float64)
    for _ in range(6):
        c = rng.uniform(5, 24, 3)
        r = rng.uniform(2.2, 5.5)
        gg = np.indices(blobs.shape).astype(np.float64)
        blobs[((gg[0] - c[0]) ** 2 + (gg[1] - c[1]) ** 2 + (gg[2] - c[2]) ** 2) < r * r] = 1.0
    out["blobs"] = (blobs, np.array([-3.0, 4.0, 10.0]), 0.35, 1.0)
    # a carve-coded int32 volume (-1 / 0 / 1) touching the border, and a probability volume around the 0.5 threshold
    carve = np.where(blobs > 0, 1, -1).astype(np.int32)
    carve[:3, :, :] = 1
    out["carve_border"] = (carve, np.array([0.0, 0.0, 0.0]), 0.5, 1.0)
    prob = np.clip(0.5 + 0.5 * np.sin(np.indices((18, 20, 19)).sum(axis=0) / 3.1) + rng.normal(0, 0.05, (18, 20, 19)), 0, 1)
    out["prob"] = (prob.astype(np.float32), np.array([2.0, 2.0, 2.0]), 1.5, 0.0)
    return out


def scoring_case():
    rng = np.random.default_rng(7)
    z = np.load(os.path.join(HERE, "virtual_plant_inputs.npz"))
    cams = z["cams"]
    labels = ["stem", "leaf", "fruit"]
    files = []
    for v in range(0, 18, 3):
        c = cams[v]
        cam = {"camera_model": {"params": c[0:4].tolist() + [0.0] * 4}, "rotmat": c[4:13].reshape(3, 3).tolist(),
               "tvec": c[13:16].tolist()}
        for lab in labels + ["background"]:
            files.append(_File(lab, cam if v != 6 else None, z[f"map_{lab}"][v]))
    bbox = z["bbox"]
    pts = np.stack([rng.uniform(bbox[a][0] - 3, bbox[a][1] + 3, 400) for a in range(3)], axis=1)
    pts[:5] = cams[0, 13:16] * 0 + 1e3   # far away: outside every picture
    views = [(f._md["channel"], f._md["camera"], f.image) for f in files]
    return pts, files, views, labels


def main():
    ns, cites = reference_namespace()
    out, meta = {}, {"cites": cites}
    for name, (vol, origin, vs, lsv) in volumes().items():
        pcd = ns["vol2pcd"](vol, origin, vs, lsv)
        out[f"v2p_{name}_volume"] = vol
        out[f"v2p_{name}_args"] = np.array([*origin, vs, lsv], dtype=np.float64)
        out[f"v2p_{name}_points"] = np.asarray(pcd.points, dtype=np.float64).reshape(-1, 3)
        out[f"v2p_{name}_normals"] = np.asarray(pcd.normals, dtype=np.float64).reshape(-1, 3)
        print(name, vol.shape, "->", len(out[f"v2p_{name}_points"]), "points")
    idx = np.array([[0, 0, 0], [1, 2, 3], [10, 0, 5]], dtype=np.float64)
    out["index2point"] = ns["index2point"](idx, np.array([1.0, -2.0, 0.5]), 0.25)
    pts, files, views, labels = scoring_case()
    scores, pts_labels = run_scoring(ns, pts, files, labels)
    out["score_points"] = pts
    out["score_scores"] = scores
    out["score_labels"] = pts_labels
    out["score_view_index"] = np.array(list(range(0, 18, 3)))
    meta["cites"]["scoring"] = scoring_slice()[2]
    meta["score_label_names"] = labels
    print("scoring:", scores.shape, "labels histogram", np.bincount(pts_labels, minlength=len(labels)))
    import json
    np.savez_compressed(os.path.join(HERE, "f_rows_golden.npz"), meta=json.dumps(meta), **out)


if __name__ == "__main__":
    main()
