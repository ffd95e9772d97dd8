"""CPU: the C-ABI library loads and exports every symbol include/comd_b200.h declares."""
import ctypes
import os
import re

import pytest

import pycomd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "comd_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(comd_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert header_symbols() == sorted(pycomd.ABI_SYMBOLS)


def test_library_exports_every_symbol():
    """Both precision builds (comd_real = double / float) export the whole interface and say which one they are."""
    for precision, size in (("f64", 8), ("f32", 4)):
        L = ctypes.CDLL(pycomd.LIB_PATHS[precision])
        for name in header_symbols():
            assert hasattr(L, name), (precision, name)
        assert L.comd_precision() == size


def test_f32_geometry_is_the_float_references_arithmetic():
    """make_geometry(precision="f32") against the link-cell geometry the reference's real_t = float build derives
    (initLinkCells, linkCells.c:122-182; fixture from libcomd_ref_f32.so, tests/golden/make_golden_f32.py)."""
    import numpy as np
    for kind in ("lj", "funcfl", "setfl"):
        g = np.load(os.path.join(ROOT, "tests", "golden", "f32_%s_n6.npz" % kind))
        ext = np.float32(6) * np.float32(g["lat"])
        geo = pycomd.make_geometry([ext] * 3, np.float32(g["cutoff"]), precision="f32")
        assert list(geo.gridSize) == list(g["cell_grid"])
        assert [geo.boxSize[a] for a in range(3)] == list(g["cell_box_size"])
        assert [geo.invBoxSize[a] for a in range(3)] == list(g["cell_inv_box_size"])
        assert pycomd.n_total_boxes(tuple(geo.gridSize)) == int(g["cell_n_total"])


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


def test_hilbert_key_walks_the_cube_through_face_neighbours():
    """The key option nl_hilbert sorts the work blocks by (CoMD's -H, linkCells.c:28-63): a bijection of the 2^b cube onto
    0 .. 8^b - 1 whose consecutive cells share a face -- the property that makes consecutive work chunks neighbours in space."""
    L = pycomd.lib()
    for bits in (1, 2, 3, 4):
        n = 1 << bits
        cells = {}
        for x in range(n):
            for y in range(n):
                for z in range(n):
                    cells[L.comd_hilbert_key(x, y, z, bits)] = (x, y, z)
        assert sorted(cells) == list(range(n ** 3))
        path = [cells[k] for k in range(n ** 3)]
        for a, b in zip(path, path[1:]):
            assert sum(abs(p - q) for p, q in zip(a, b)) == 1, (bits, a, b)


@pytest.mark.parametrize("cells,block", [((47, 47, 47), (4, 4, 2)), ((95, 95, 95), (4, 4, 4)), ((13, 9, 6), (4, 4, 4)), ((7, 7, 7), (2, 2, 1)), ((5, 3, 3), (4, 4, 4))])
def test_work_block_sets_list_every_block_exactly_once(cells, block):
    """What the list kernels walk (nl_blocks.cuh): in every order (the set's own, largest block first, Hilbert curve) the set of
    all blocks is a permutation of 0 .. n-1, interior and boundary shell partition it, the interior holds exactly the blocks that
    touch no face of the domain, and largest-first really is non-increasing in cells per block."""
    import ctypes as C
    L = pycomd.lib()
    nb = [(cells[a] + block[a] - 1) // block[a] for a in range(3)]
    n = nb[0] * nb[1] * nb[2]
    arr = (C.c_int * 3)(*cells)
    def table(which, mode):
        out = (C.c_int * n)()
        m = L.comd_block_order_table(arr, block[0], block[1], block[2], which, mode, n, out)
        assert m >= 0
        return list(out[:m])
    def is_interior(b):
        ix, iy, iz = b % nb[0], (b // nb[0]) % nb[1], b // (nb[0] * nb[1])
        return all(0 < c < m - 1 for c, m in zip((ix, iy, iz), nb))
    def size(b):
        ix, iy, iz = b % nb[0], (b // nb[0]) % nb[1], b // (nb[0] * nb[1])
        return min(block[0], cells[0] - ix * block[0]) * min(block[1], cells[1] - iy * block[1]) * min(block[2], cells[2] - iz * block[2])
    for mode in (0, 1, 2):
        everything, inner, shell = table(0, mode), table(1, mode), table(2, mode)
        assert sorted(everything) == list(range(n))
        assert sorted(inner + shell) == list(range(n))
        assert all(is_interior(b) for b in inner) and not any(is_interior(b) for b in shell)
        if mode == 1:
            for t in (everything, inner, shell):
                assert all(size(a) >= size(b) for a, b in zip(t, t[1:]))
    assert table(0, 0) == list(range(n))
