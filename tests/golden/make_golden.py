"""Generates tests/golden/virtual_plant_*.npz.  Run in the build container only
(`python tests/golden/make_golden.py`): it reads the reference's own fixture
/root/reference/tests/testdata/virtual_plant and runs the reference's own kernel source
(kernels/backprojection.c compiled unmodified into oracle/_ref/libbp_ref_off.so, see
oracle/Makefile) through the host conversions of cl.py:190-305.

Outputs
  virtual_plant_inputs.npz   cameras (18 x 16 float32: fx fy cx cy | rot[9] | tvec[3], from the
                             per-file "camera" metadata), bounding box, and the 18 views x
                             {background + 6 label channels} uint8 maps, 896 x 896
  virtual_plant_golden.npz   reference outputs: carving volumes (int8-coded) and averaging volumes
                             (float32) on small grids, sha256 + histograms on the larger ones,
                             and the count of voxels that differ under FMA contraction (ref_fma)
"""
import hashlib
import json
import os
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle.backprojection_oracle import BackprojectionOracle, build  # noqa: E402

FIX = "/root/reference/tests/testdata/virtual_plant"
CHANNELS = ["background", "flower", "fruit", "leaf", "pedicel", "stem"]


def shape_origin(bbox, voxel_size):
    # plant3dvision/tasks/cl.py:127-147
    (x0, x1), (y0, y1), (z0, z1) = bbox["x"], bbox["y"], bbox["z"]
    n = [int((x1 - x0) / voxel_size) + 1, int((y1 - y0) / voxel_size) + 1, int((z1 - z0) / voxel_size) + 1]
    return n, [x0, y0, z0]


def main():
    build()
    bbox = json.load(open(f"{FIX}/metadata/images.json"))["bounding_box"]
    cams, maps = [], {c: [] for c in CHANNELS}
    for v in range(18):
        md = json.load(open(f"{FIX}/metadata/images/{v:05d}_background.json"))["camera"]
        cams.append(md["camera_model"]["params"][0:4] + sum(md["rotmat"], []) + md["tvec"])
        for c in CHANNELS:
            cm = json.load(open(f"{FIX}/metadata/images/{v:05d}_{c}.json"))
            assert cm["camera"] == md and cm["channel"] == c
            maps[c].append(np.asarray(Image.open(f"{FIX}/images/{v:05d}_{c}.png")))
    cams = np.array(cams, dtype=np.float64)  # float64 as stored in JSON; consumers cast like cl.py:293-296
    for c in CHANNELS:
        maps[c] = np.stack(maps[c])
        assert maps[c].dtype == np.uint8 and maps[c].shape == (18, 896, 896)
    if len(sys.argv) <= 1:  # (adding one golden entry leaves the committed inputs file alone)
        np.savez_compressed(f"{HERE}/virtual_plant_inputs.npz", cams=cams,
                            bbox=np.array([bbox["x"], bbox["y"], bbox["z"]], dtype=np.float64),
                            **{f"map_{c}": maps[c] for c in CHANNELS})

    def run(backend, kind, vs, label="background", log=False, invert=None):
        shape, origin = shape_origin(bbox, vs)
        o = BackprojectionOracle(shape, origin, vs, type=kind, log=log, backend=backend)
        for v in range(18):
            m = maps[label][v]
            if invert if invert is not None else (kind == "carving"):
                m = np.invert(m)
            c = cams[v].astype(np.float32)
            o.process_view(c[0:4], c[4:13], c[13:16], m)
        return o.get_values().copy()

    out = {}
    meta = {}
    only = sys.argv[1] if len(sys.argv) > 1 else None  # e.g. `make_golden.py carve_vs0.06`: add one entry, keep the rest
    if only:
        with np.load(f"{HERE}/virtual_plant_golden.npz") as z:
            meta = json.loads(str(z["meta"]))
            out = {k: z[k] for k in z.files if k != "meta"}
    # carving (geom_pipe_virtual.toml: background channel, invert, camera metadata "camera"; voxel_size 0.06 is the
    # value configs/geom_pipe_virtual.toml:6 sets - 386 x 397 x 1998 voxels - and 0.1 the CI variant's)
    for vs in (0.6, 0.3, 0.1, 0.06):
        if only and only != f"carve_vs{vs}":
            continue
        vol = run("ref_off", "carving", vs)
        fma = run("ref_fma", "carving", vs)
        key = f"carve_vs{vs}"
        meta[key] = {"shape": list(vol.shape), "sha256": hashlib.sha256(vol.tobytes()).hexdigest(),
                     "hist": {str(k): int((vol == k).sum()) for k in (-1, 0, 1)},
                     "n_diff_vs_fma_contracted": int((vol != fma).sum())}
        if vs >= 0.3:
            out[key] = vol.astype(np.int8)
        print(key, meta[key])
    # averaging, 6 labels (ml_pipe_virtual.toml: voxel_size 0.3, log False; task default log True)
    for vs, log in ((0.6, False), (0.6, True), (0.3, False), (0.3, True)):
        for label in CHANNELS:
            if only and only != f"avg_vs{vs}_log{int(log)}_{label}":
                continue
            vol = run("ref_off", "averaging", vs, label=label, log=log, invert=False)
            fma = run("ref_fma", "averaging", vs, label=label, log=log, invert=False)
            key = f"avg_vs{vs}_log{int(log)}_{label}"
            meta[key] = {"shape": list(vol.shape), "sha256": hashlib.sha256(vol.tobytes()).hexdigest(),
                         "sum": float(vol.sum(dtype=np.float64)),
                         "n_diff_vs_fma_contracted": int((vol != fma).sum())}
            if vs == 0.6 and label in ("background", "stem"):
                out[key] = vol
            print(key, meta[key])
    np.savez_compressed(f"{HERE}/virtual_plant_golden.npz", meta=json.dumps(meta), **out)


if __name__ == "__main__":
    main()
