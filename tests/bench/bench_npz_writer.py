#!/usr/bin/env python3
"""Time `npz_parallel.savez_compressed_parallel` against `numpy.savez_compressed` (what plantdb's io.write_npz
calls, io.py:610-622) on a Voxels-like label volume: float64, mostly small values with a clipped-to-1 plateau
(SURVEY §8f-2).  CPU only.  Prints one JSON line.  Usage: bench_npz_writer.py [nx ny nz] [n_labels]"""
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    from plant3dvision_b200.npz_parallel import savez_compressed_parallel
    shape = tuple(int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (301, 301, 276)
    n_labels = int(sys.argv[4]) if len(sys.argv) > 4 else 2
    rng = np.random.default_rng(0)
    arrays = {f"label{i}": np.clip(np.exp(rng.normal(-3, 2, shape)), 0, 1) for i in range(n_labels)}
    nbytes = sum(a.nbytes for a in arrays.values())
    with tempfile.TemporaryDirectory() as d:
        t0 = time.perf_counter()
        np.savez_compressed(os.path.join(d, "ref.npz"), **arrays)
        t_ref = time.perf_counter() - t0
        t0 = time.perf_counter()
        p = savez_compressed_parallel(os.path.join(d, "par.npz"), arrays)
        t_par = time.perf_counter() - t0
        with np.load(p) as z:
            equal = all(np.array_equal(z[k], a) for k, a in arrays.items())
        out = {"workload": f"{n_labels} float64 label volumes {list(shape)} ({nbytes / 1e6:.0f} MB)", "cores": os.cpu_count(),
               "numpy_savez_compressed_s": t_ref, "savez_compressed_parallel_s": t_par, "speedup": t_ref / t_par,
               "file_bytes_numpy": os.path.getsize(os.path.join(d, "ref.npz")), "file_bytes_parallel": os.path.getsize(p),
               "arrays_equal": bool(equal)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
