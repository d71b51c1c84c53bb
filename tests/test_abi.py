"""CPU tests of the drop-in boundary: ``libbp_b200.so`` loads, exports every symbol that
``include/bp_b200.h`` declares, rejects bad arguments with the documented codes, and FAILS LOUDLY
(no CPU fallback) when there is no CUDA device.  No compute call is made here."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT
from plant3dvision_b200 import _cabi

HEADER = os.path.join(ROOT, "include", "bp_b200.h")


V2P_HEADER = os.path.join(ROOT, "include", "vol2pcd_b200.h")
MASKS_HEADER = os.path.join(ROOT, "include", "masks_b200.h")


def declared_functions(header=HEADER):
    text = open(header).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bp_[a-z0-9_]+)\s*\(", text)))


def have_cuda():
    import torch
    return torch.cuda.is_available()


def test_header_declares_what_the_binding_lists():
    assert declared_functions() == sorted(_cabi.SYMBOLS)


def test_vol2pcd_header_matches_binding_and_cites_the_reference():
    assert declared_functions(V2P_HEADER) == sorted(_cabi.V2P_SYMBOLS)
    text = open(V2P_HEADER).read()
    for cite in ("proc3d.py:490-570", "tasks/proc3d.py:80-115"):
        assert cite in text, cite
    lib = _cabi.lib()
    assert lib.bp_vol2pcd_free(None) == 0
    assert lib.bp_vol2pcd_info(None, None, None, None, None, None) == -1
    assert lib.bp_vol2pcd(None, 0, None, 0, 0, None, 0, 0, 0.5, None, 1.0, 0.0, 0.0, 0.0, None, 4) == -1


def test_masks_header_matches_binding_and_cites_the_reference():
    assert declared_functions(MASKS_HEADER) == sorted(_cabi.MASKS_SYMBOLS)
    text = open(MASKS_HEADER).read()
    for cite in ("proc2d.py:69-115", "proc2d.py:118-169", "proc2d.py:172-220", "tasks/proc2d.py:232"):
        assert cite in text, cite
    lib = _cabi.lib()
    assert lib.bp_masks_filter(0, 0, None, 0, 0, 3, 0, None, 0.3, 0, None, None, None, None, None) == -1  # no coefs
    assert lib.bp_masks_filter(0, 1, None, 4, 4, 2, 1, None, 0.3, 0, None, None, None, None, None) == -1  # 2 channels
    assert b"channels" in lib.bp_last_error()


def test_library_exports_every_declared_symbol():
    lib = _cabi.lib()
    for name in declared_functions() + declared_functions(V2P_HEADER) + declared_functions(MASKS_HEADER):
        assert hasattr(lib, name), f"libbp_b200.so does not export {name}"
    # and nothing C++-mangled leaks an unintended bp_* entry point
    import re
    # "bp_b200 <version> sm_100a src=<hash of every source file> kern=<hash of the device code>"
    assert re.fullmatch(r"bp_b200 [0-9.]+ sm_100a src=[0-9a-f]{16} kern=[0-9a-f]{16}", lib.bp_version().decode())


def test_built_library_carries_the_hash_of_the_sources_in_the_tree():
    """build() rebuilds on a content hash, not on mtime (a stale binary must never be benchmarked): the hash embedded
    in the library is the hash of the sources next to it."""
    import __graft_entry__ as g
    assert g.built_hashes() == g.source_hashes()


def test_header_cites_the_reference_interface_it_replaces():
    text = open(HEADER).read()
    for cite in ("cl.py:118-188", "cl.py:190-227", "cl.py:229-232", "cl.py:307-311",
                 "kernels/backprojection.c:57-84", "kernels/backprojection.c:36-55"):
        assert cite in text, cite


def test_bad_arguments_are_rejected_with_codes_and_messages():
    lib = _cabi.lib()
    h = ctypes.c_void_p()
    shape = (ctypes.c_int32 * 3)(4, 4, 4)
    origin = (ctypes.c_float * 3)(0, 0, 0)
    assert lib.bp_create(ctypes.byref(h), shape, 0, 4, origin, 1.0, 7, 0.0, 0) == -1  # BP_EINVAL: mode
    assert b"mode" in lib.bp_last_error()
    assert lib.bp_create(ctypes.byref(h), shape, 3, 2, origin, 1.0, 0, 0.0, 0) == -1  # bad slab
    assert b"slab" in lib.bp_last_error()
    neg = (ctypes.c_int32 * 3)(4, -1, 4)
    assert lib.bp_create(ctypes.byref(h), neg, 0, 4, origin, 1.0, 0, 0.0, 0) == -1
    assert lib.bp_create(None, shape, 0, 4, origin, 1.0, 0, 0.0, 0) == -1
    for fn in (lib.bp_clear, lib.bp_sync):
        assert fn(None) == -1
    assert lib.bp_get_values(None, None) == -1
    assert lib.bp_destroy(None) == 0  # like free(NULL)
    assert lib.bp_host_free(None) == 0
    with pytest.raises(ValueError):
        _cabi.check(-1)
    with pytest.raises(MemoryError):
        _cabi.check(-3)
    with pytest.raises(_cabi.BackprojectionError):
        _cabi.check(-2)


@pytest.mark.skipif(have_cuda(), reason="this checks the no-GPU behaviour")
def test_no_cpu_fallback_without_a_cuda_device():
    lib = _cabi.lib()
    h = ctypes.c_void_p()
    shape = (ctypes.c_int32 * 3)(4, 4, 4)
    origin = (ctypes.c_float * 3)(0, 0, 0)
    rc = lib.bp_create(ctypes.byref(h), shape, 0, 4, origin, 1.0, 0, 0.0, 0)
    assert rc == -2 and not h.value  # BP_ECUDA, no handle
    from plant3dvision_b200.cl import Backprojection
    with pytest.raises(_cabi.BackprojectionError):
        Backprojection([10, 10, 10], [0, 0, 0], 1.0)


def test_missing_library_is_a_hard_error(monkeypatch, tmp_path):
    monkeypatch.setattr(_cabi, "_lib", None)
    monkeypatch.setattr(_cabi, "LIB_PATH", str(tmp_path / "libbp_b200.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        _cabi.lib()


def test_product_package_never_imports_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's baseline legs may execute oracle/: neither the package,
    nor its alias, nor tools/ do."""
    for top in ("plant-3d-vision_b200", "plant3dvision_b200", "tools", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".h", ".cpp", ".sh")):
                    text = open(os.path.join(dirpath, f)).read()
                    assert "import oracle" not in text and "from oracle" not in text and "libbp_oracle" not in text, f


def test_unknown_type_raises_like_the_reference():
    from plant3dvision_b200.cl import Backprojection
    with pytest.raises(ValueError, match="Unknown kernel type"):
        Backprojection([10, 10, 10], [0, 0, 0], 1.0, type="nope")  # cl.py:151-152
    assert np.dtype(np.int32) == np.int32
