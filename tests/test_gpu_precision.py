"""GPU: both precision builds of the library (comd_real = double in libcomd_b200.so, float in
libcomd_b200_f32.so -- the reference's DOUBLE_PRECISION switch, mytype.h:11-24) against vectors the reference
itself produced in that precision (tests/golden/make_golden.py, make_golden_f32.py).

Tolerances.  FP64: the north_star gates (1e-12 step 0, 1e-6 energy at step 100).  FP32: the reference's float
build and ours sum the same ~70 float terms per atom in different orders, so per-atom results agree to a few
float ulps of the largest term: 2e-6 of max|F| / max|U| at step 0; 3e-6 for the energy sums (the reference adds its
864 per-atom energies into a float, 1.1e-6 off on Cu01; the device sums in double in both builds); 2e-5 for the total
energy after 100 steps (two float trajectories; the float reference itself drifts by up to 5e-6 over those steps).
esum_big: the 32000-atom FP64 sums of the trace fixture, added in a different order (measured 1.3e-12).
Cell assignment is bit-exact in both.
"""
import os

import numpy as np
import pytest

import pyoracle as po
import pycomd
from common import GOLDEN_DIR, LJ_EPS, LJ_SIGMA, by_gid, rel_err

pytestmark = pytest.mark.gpu

KINDS = {"lj": po.POT_LJ, "funcfl": po.POT_FUNCFL, "setfl": po.POT_SETFL}
TOL = {"f64": dict(step0=1e-12, esum=1e-12, esum_big=1e-11, e100=1e-6),
       "f32": dict(step0=2e-6, esum=3e-6, esum_big=3e-6, e100=2e-5)}
PATHS = {"link_cell": None, "list_n3": 1, "list_full": 0}


def golden(kind, precision):
    if precision == "f32":
        g = dict(np.load(os.path.join(GOLDEN_DIR, "f32_%s_n6.npz" % kind)))
        if kind != "lj":
            g["tables"] = [dict(n=int(g["tab_%s_n" % w]), x0=float(g["tab_%s_x0" % w]), inv_dx=float(g["tab_%s_inv_dx" % w]),
                                padded=g["tab_%s_padded" % w].astype(np.float64)) for w in ("phi", "rho", "F")]
        return g
    g = dict(np.load(os.path.join(GOLDEN_DIR, "forces_%s_n6.npz" % kind)))
    orc = po.Oracle(pot=KINDS[kind], nx=6, ny=6, nz=6, initial_force=False)      # tables / mass of the FP64 build
    pot = orc.potential()
    g.update(mass=pot["mass"], lat=pot["lat"], trace=np.load(os.path.join(GOLDEN_DIR, "trace_%s_n20.npz" % kind))["trace"])
    if kind != "lj":
        g["tables"] = [orc.table(w) for w in range(3)]
    return g


def device_from(g, kind, precision, path="link_cell", n=6, atoms=True):
    cutoff, lat = float(g["cutoff"]), float(g["lat"])
    skin = 0.1 * cutoff if path != "link_cell" else 0.0
    T = pycomd.REAL[precision]
    gext = [T(n) * T(lat)] * 3                                   # globalExtent = nx * lat in real_t (initLattice / CoMD.c)
    geom = pycomd.make_geometry(gext, T(T(cutoff) + T(skin)), precision=precision)
    dev = pycomd.Device(geom, 4 * n ** 3 * 3 + 16384, precision=precision)
    dev.set_mass(float(g["mass"]))
    if kind == "lj":
        dev.set_lj(LJ_SIGMA, LJ_EPS, cutoff)
    else:
        dev.set_eam(cutoff, *g["tables"])
    if path != "link_cell":
        dev.set_neighbor_list(True, skin)
        dev.set_option("nl_n3", PATHS[path])
    if atoms:
        dev.upload(g["gid"], g["r"], g["p"])
    return dev, geom


def u_scale(kind):
    return 4.0 * LJ_EPS if kind == "lj" else 1.0            # ljForce.c:256-261 keeps per-atom energies in units of 4 eps


@pytest.mark.parametrize("precision", ["f64", "f32"])
@pytest.mark.parametrize("kind", list(KINDS))
@pytest.mark.parametrize("path", list(PATHS))
def test_step0_forces_and_energies_against_the_reference_in_that_precision(kind, precision, path):
    g = golden(kind, precision)
    dev, _ = device_from(g, kind, precision, path)
    assert dev.real == pycomd.REAL[precision] and dev.L.comd_precision() == np.dtype(dev.real).itemsize
    dev.compute_force()
    d = by_gid(dev.download())
    o = by_gid({k: g[k] for k in ("gid", "r", "p", "f", "U")})
    assert d["f"].dtype == dev.real
    assert np.array_equal(d["r"], o["r"].astype(dev.real)) and np.array_equal(d["p"], o["p"].astype(dev.real))
    t = TOL[precision]
    assert rel_err(d["f"].astype(np.float64), o["f"].astype(np.float64)) <= t["step0"]
    assert rel_err(d["U"].astype(np.float64), o["U"].astype(np.float64) * u_scale(kind)) <= t["step0"]
    ep, ek = dev.energy()
    assert abs(ep - float(g["ePot"])) <= t["esum"] * abs(float(g["ePot"]))
    assert abs(ek - float(g["eKin"])) <= t["esum"] * abs(float(g["eKin"]))
    dev.close()


@pytest.mark.parametrize("kind", list(KINDS))
def test_f32_cells_are_the_float_references_cells(kind):
    """getBoxFromCoord in float (linkCells.c:448-480 with real_t = float): same cell for every atom, bit-exact."""
    g = golden(kind, "f32")
    dev, geom = device_from(g, kind, "f32")
    assert [geom.gridSize[a] for a in range(3)] == list(g["cell_grid"])
    assert np.array_equal(np.array([geom.boxSize[a] for a in range(3)]), g["cell_box_size"])
    assert np.array_equal(np.array([geom.invBoxSize[a] for a in range(3)]), g["cell_inv_box_size"])
    d = dev.download()
    want = dict(zip(g["gid"].tolist(), g["cell_of_atom"].tolist()))
    assert [want[i] for i in d["gid"].tolist()] == d["box"].tolist()
    counts = dev.cell_counts()
    assert len(counts) == int(g["cell_n_total"])
    assert np.array_equal(counts[:np.prod(g["cell_grid"])], np.bincount(g["cell_of_atom"], minlength=np.prod(g["cell_grid"])))
    dev.close()


@pytest.mark.parametrize("precision", ["f64", "f32"])
@pytest.mark.parametrize("kind", list(KINDS))
@pytest.mark.parametrize("path", ["link_cell", "list_n3"])
def test_energy_after_100_steps_in_that_precision(kind, precision, path):
    g = golden(kind, precision)
    if precision == "f64":
        # the FP64 trace fixture is the BASELINE config (20^3, perfect lattice, 600 K): atoms from the oracle's initAtoms
        orc = po.Oracle(pot=KINDS[kind], nx=20, ny=20, nz=20, temperature=600.0, initial_force=False)
        a = orc.atoms()
        g.update(gid=a["gid"], r=a["r"], p=a["p"])
        n = 20
    else:
        n = 6
    dev, _ = device_from(g, kind, precision, path, n=n)
    dev.compute_force()
    e0 = sum(dev.energy())
    for _ in range(10):
        dev.timestep(10, 1.0)
    e1 = sum(dev.energy())
    tr = g["trace"].sum(1)
    t = TOL[precision]
    assert abs(e0 - tr[0]) <= t["esum_big"] * abs(tr[0])
    assert abs(e1 - tr[-1]) <= t["e100"] * abs(tr[-1])
    assert dev.num_local() == 4 * n ** 3
    dev.close()


@pytest.mark.parametrize("kind", list(KINDS))
def test_f32_build_stays_within_float_rounding_of_the_f64_build(kind):
    """Same (float-representable) inputs and the same tables through both libraries in one process (the two builds
    share no state).  LJ: float rounding of the pair term, <= 1e-5 of max|F|.  EAM: interpolate() (eam.c:573-578)
    differentiates the table by differencing neighbouring entries; rounded to float (6e-8 relative) and divided by
    the grid step (5.5e-4 A on Cu01) that is ~1e-4 of the slope -- a property of the reference's float build, which
    the device reproduces (the step-0 test above holds it to 2e-6 of the float reference)."""
    g = golden(kind, "f32")
    g64 = golden(kind, "f64")
    res = {}
    for precision in ("f32", "f64"):
        gg = dict(g)
        if kind != "lj":
            gg["tables"] = g64["tables"]
        dev, _ = device_from(gg, kind, precision, "list_n3")
        dev.compute_force()
        res[precision] = by_gid(dev.download())
        dev.close()
    tol = 1e-5 if kind == "lj" else 1e-4
    assert rel_err(res["f32"]["f"].astype(np.float64), res["f64"]["f"]) <= tol
    assert rel_err(res["f32"]["U"].astype(np.float64), res["f64"]["U"]) <= tol


@pytest.mark.parametrize("kind", ["lj", "setfl"])
def test_f32_step_host_round_trip(kind):
    g = golden(kind, "f32")
    dev, _ = device_from(g, kind, "f32")
    dev.compute_force()
    s0 = by_gid(dev.download())
    dev.timestep(1, 1.0)
    want = by_gid(dev.download())
    dev2, _ = device_from(g, kind, "f32", atoms=False)
    cols = lambda a: [np.ascontiguousarray(a[:, k]) for k in range(3)]
    m, o, e = dev2.step_host(1.0, s0["gid"].astype(np.int32), cols(s0["r"]), cols(s0["p"]), cols(s0["f"]))
    got = by_gid(dict(gid=o["gid"][:m].copy(), r=np.stack([a[:m] for a in o["r"]], 1), p=np.stack([a[:m] for a in o["p"]], 1),
                      f=np.stack([a[:m] for a in o["f"]], 1), U=o["U"][:m].copy()))
    assert m == len(want["gid"]) and o["U"].dtype == np.float32
    for k in ("r", "p", "f", "U"):
        assert np.array_equal(got[k], want[k]), k
    dev.close(); dev2.close()
