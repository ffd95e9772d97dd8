"""CPU: the C-ABI library loads and exports every symbol include/comd_b200.h declares."""
import ctypes
import os
import re

import pycomd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "comd_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(comd_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert header_symbols() == sorted(pycomd.ABI_SYMBOLS)


def test_library_exports_every_symbol():
    L = ctypes.CDLL(pycomd.LIB_PATH)
    for name in header_symbols():
        assert hasattr(L, name), name


def test_no_device_fails_loudly(gpu_available):
    """No CPU fallback: without a GPU, creating a context is an error, not a slow path."""
    if gpu_available:
        return
    geom = pycomd.make_geometry([36.15] * 3, 5.7875)
    try:
        pycomd.Device(geom, 1000)
    except pycomd.ComdError as e:
        assert e.code == -1
    else:
        raise AssertionError("device context created without a GPU")


def test_geometry_matches_reference_arithmetic():
    """gridSize/boxSize of initLinkCells (linkCells.c:131-140) for the BASELINE configs."""
    lat = 3.615
    for n, cutoff, grid in ((20, 2.5 * 2.315, 12), (20, 4.95, 14), (60, 5.506786, 39), (160, 5.506786, 105)):
        g = pycomd.make_geometry([n * lat] * 3, cutoff)
        assert list(g.gridSize) == [grid] * 3
        assert g.boxSize[0] >= cutoff


def test_every_option_key_is_documented_in_the_header():
    """comd_dev_set_option keys are part of the boundary: each key the library accepts is listed in the header."""
    core = open(os.path.join(ROOT, "comd-cuda-async_b200", "csrc", "comd_core.cu")).read()
    keys = set(re.findall(r'strcmp\(key, "([a-z_0-9]+)"\)', core))
    assert len(keys) >= 8
    header = open(os.path.join(ROOT, "include", "comd_b200.h")).read()
    missing = sorted(k for k in keys if '"%s"' % k not in header)
    assert not missing, missing
