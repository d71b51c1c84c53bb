import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
CHANNELS = ["background", "flower", "fruit", "leaf", "pedicel", "stem"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def vp_inputs():
    """The reference's virtual_plant fixture (tests/testdata/virtual_plant), repacked by
    tests/golden/make_golden.py: cams float64 [18,16], bbox, uint8 maps [18,896,896] per channel."""
    z = np.load(os.path.join(GOLDEN, "virtual_plant_inputs.npz"))
    bbox = {"x": z["bbox"][0].tolist(), "y": z["bbox"][1].tolist(), "z": z["bbox"][2].tolist()}
    return {"cams": z["cams"], "bbox": bbox, "maps": {c: z[f"map_{c}"] for c in CHANNELS}}


@pytest.fixture(scope="session")
def vp_golden():
    """Outputs of the reference's own kernel source (oracle/_ref, -ffp-contract=off) on vp_inputs."""
    z = np.load(os.path.join(GOLDEN, "virtual_plant_golden.npz"))
    meta = json.loads(str(z["meta"]))
    return {"meta": meta, "arrays": {k: z[k] for k in z.files if k != "meta"}}


def vp_shape_origin(bbox, voxel_size):
    # plant3dvision/tasks/cl.py:127-147
    (x0, x1), (y0, y1), (z0, z1) = bbox["x"], bbox["y"], bbox["z"]
    n = [int((x1 - x0) / voxel_size) + 1, int((y1 - y0) / voxel_size) + 1, int((z1 - z0) / voxel_size) + 1]
    return n, [x0, y0, z0]


def split_cam(c):
    c = np.asarray(c, dtype=np.float32)
    return c[0:4], c[4:13], c[13:16]
