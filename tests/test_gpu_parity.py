"""GPU: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.

Gates (BASELINE.json north_star / BASELINE.md section 4):
  * cell assignment, per-cell counts, ghost sets, atom conservation ......... bit-exact, keyed by gid
  * step-0 per-atom forces and energies (displaced lattice) .................. <= 1e-12 relative (max-norm)
  * total energy after 100 steps ............................................. <= 1e-6 relative
"""
import os

import numpy as np
import pytest

import pyoracle as po
import pycomd
from common import GOLDEN_DIR, LJ_EPS, by_gid, device_for, oracle_sim, rel_err

pytestmark = pytest.mark.gpu

KINDS = {"lj": po.POT_LJ, "funcfl": po.POT_FUNCFL, "setfl": po.POT_SETFL}
FORCE_TOL = 1e-12      # north_star: step-0 per-atom forces and energies, FP64
ENERGY100_TOL = 1e-6   # north_star: total energy at step 100


def u_scale(kind):
    # the CPU path keeps per-atom LJ energies in units of 4*epsilon (ljForce.c:256-261); the device stores eV
    return 4.0 * LJ_EPS if kind == "lj" else 1.0


@pytest.mark.parametrize("kind", list(KINDS))
@pytest.mark.parametrize("n", [6, 10])
def test_step0_forces_and_energies(kind, n):
    orc = oracle_sim(KINDS[kind], n, delta=0.1, temperature=600.0)
    a = orc.atoms()
    dev, _ = device_for(orc, (n, n, n))
    dev.upload(a["gid"], a["r"], a["p"])
    dev.compute_force()
    d = by_gid(dev.download())
    o = by_gid(a)
    assert np.array_equal(d["gid"], o["gid"])
    assert np.array_equal(d["r"], o["r"])            # positions untouched by upload + redistribute
    assert rel_err(d["f"], o["f"]) <= FORCE_TOL
    assert rel_err(d["U"], o["U"] * u_scale(kind)) <= FORCE_TOL
    ep_d, ek_d = dev.energy()
    ep_o, ek_o = orc.energies()
    assert abs(ep_d - ep_o) <= FORCE_TOL * abs(ep_o)
    assert abs(ek_d - ek_o) <= FORCE_TOL * abs(ek_o)
    dev.close()


@pytest.mark.parametrize("kind", list(KINDS))
def test_step0_forces_against_reference_golden(kind):
    """Same gate against vectors the reference itself produced (tests/golden/make_golden.py)."""
    g = np.load(os.path.join(GOLDEN_DIR, "forces_%s_n6.npz" % kind))
    orc = po.Oracle(pot=KINDS[kind], nx=6, ny=6, nz=6, initial_force=False)   # only for tables / cutoff
    dev, _ = device_for(orc, (6, 6, 6))
    dev.upload(g["gid"], g["r"], g["p"])
    dev.compute_force()
    d = by_gid(dev.download())
    o = by_gid({k: g[k] for k in ("gid", "r", "p", "f", "U")})
    assert rel_err(d["f"], o["f"]) <= FORCE_TOL
    assert rel_err(d["U"], o["U"] * u_scale(kind)) <= FORCE_TOL
    ep, ek = dev.energy()
    assert abs(ep - float(g["ePot"])) <= FORCE_TOL * abs(float(g["ePot"]))
    assert abs(ek - float(g["eKin"])) <= FORCE_TOL * abs(float(g["eKin"]))
    dev.close()


@pytest.mark.parametrize("kind", list(KINDS))
def test_known_answer_cohesive_energy(kind):
    """Perfect lattice, step 0 (CoMD.c:896-900)."""
    known = {"lj": -1.243619295058, "funcfl": -3.538079224691, "setfl": -3.539999969176}[kind]
    n = 10
    orc = oracle_sim(KINDS[kind], n)
    a = orc.atoms()
    dev, _ = device_for(orc, (n, n, n))
    dev.upload(a["gid"], a["r"], a["p"])
    dev.compute_force()
    ep, ek = dev.energy()
    assert abs(ep / len(a["gid"]) - known) < 5e-11
    assert abs(ek / len(a["gid"]) / (8.6173324e-5 * 1.5) - 600.0) < 1e-9
    dev.close()


@pytest.mark.parametrize("kind", list(KINDS))
def test_cells_counts_and_ghosts_bit_exact(kind):
    """Binning + halo against the oracle on identical positions: owner cell of every local and
    ghost atom (reference box numbering), per-cell counts, ghost gid multisets per halo cell."""
    n = 8
    # skin 0 makes the oracle size its link cells on the bare cutoff, like the device
    orc = oracle_sim(KINDS[kind], n, delta=0.3, temperature=600.0, skin=0.0)
    a = orc.atoms()
    dev, geom = device_for(orc, (n, n, n))
    dev.upload(a["gid"], a["r"], a["p"])
    og = orc.geometry()
    assert list(geom.gridSize) == list(og["grid"])
    assert np.array_equal(np.array(geom.invBoxSize[:]), og["inv_box_size"])
    d = by_gid(dev.download())
    o = by_gid(a)
    assert np.array_equal(d["gid"], o["gid"])
    assert np.array_equal(d["box"], o["box"])
    # every atom incl. ghosts: (box, gid) multiset and positions
    da, oa = dev.download_all(), orc.all_atoms()
    kd = np.lexsort((da["gid"], da["box"])); ko = np.lexsort((oa["gid"], oa["box"]))
    assert len(kd) == len(ko)
    assert np.array_equal(da["box"][kd], oa["box"][ko])
    assert np.array_equal(da["gid"][kd], oa["gid"][ko])
    assert np.array_equal(da["r"][kd], oa["r"][ko])        # PBC shift applied identically
    counts = dev.cell_counts()
    assert np.array_equal(counts, np.bincount(oa["box"], minlength=og["n_total_boxes"]))
    # independent statement of the same rule
    cells = po.OracleCells([n * orc.potential()["lat"]] * 3, orc.potential()["cutoff"])
    bx, _ = cells.bin(da["r"])
    assert np.array_equal(bx, da["box"])
    dev.close()


@pytest.mark.parametrize("kind", list(KINDS))
def test_atoms_sorted_by_cell_then_gid(kind):
    n = 6
    orc = oracle_sim(KINDS[kind], n, delta=0.2)
    a = orc.atoms()
    dev, _ = device_for(orc, (n, n, n))
    rng = np.random.default_rng(3)
    perm = rng.permutation(len(a["gid"]))               # upload order must not matter
    dev.upload(a["gid"][perm], a["r"][perm], a["p"][perm])
    dev.compute_force()
    d1 = dev.download()
    dev.upload(a["gid"], a["r"], a["p"])
    dev.compute_force()
    d2 = dev.download()
    for k in ("gid", "box", "r", "p", "f", "U"):
        assert np.array_equal(d1[k], d2[k]), k          # bit-reproducible, order independent
    dev.close()


@pytest.mark.parametrize("kind", list(KINDS))
def test_migration_and_conservation(kind):
    """Hot, fast atoms cross cell and periodic boundaries: nobody is lost or duplicated and the
    owner cell always equals the reference rule evaluated on the device's own positions."""
    n = 6
    orc = oracle_sim(KINDS[kind], n, delta=0.05, temperature=1500.0)
    a = orc.atoms()
    dev, _ = device_for(orc, (n, n, n))
    # a uniform drift (forces are Galilean invariant) carries the whole crystal through cell and
    # periodic boundaries: 0.02 A/fs * 300 fs = 6 A along x, half that along y, a quarter along z
    drift = orc.potential()["mass"] * np.array([0.02, 0.01, -0.005])
    dev.upload(a["gid"], a["r"], a["p"] + drift)
    dev.compute_force()
    cells = po.OracleCells([n * orc.potential()["lat"]] * 3, orc.potential()["cutoff"])
    moved = 0
    box0 = by_gid(dev.download())["box"]
    for _ in range(6):
        dev.timestep(25, 2.0)
        da = dev.download_all()
        bx, _ = cells.bin(da["r"])
        assert np.array_equal(bx, da["box"])
        d = by_gid(dev.download())
        assert np.array_equal(d["gid"], np.sort(a["gid"]))          # all there, exactly once
        assert dev.num_local() == len(a["gid"])
        L = n * orc.potential()["lat"]
        assert d["r"].min() >= 0.0 and d["r"].max() < L              # locals live inside the box
        moved = max(moved, int(np.sum(d["box"] != box0)))
    assert moved > len(a["gid"]) // 4                                # the test did exercise migration
    dev.close()


@pytest.mark.parametrize("kind", list(KINDS))
def test_timestep_matches_oracle_one_step(kind):
    n = 6
    orc = oracle_sim(KINDS[kind], n, delta=0.05, temperature=600.0)
    a = orc.atoms()
    dev, _ = device_for(orc, (n, n, n))
    dev.upload(a["gid"], a["r"], a["p"])
    dev.compute_force()
    dev.timestep(1, 1.0)
    orc.timestep(1)
    d = by_gid(dev.download()); o = by_gid(orc.atoms())
    assert rel_err(d["r"], o["r"]) <= 1e-15
    assert rel_err(d["p"], o["p"]) <= FORCE_TOL
    assert rel_err(d["f"], o["f"]) <= 1e-11
    dev.close()


@pytest.mark.parametrize("kind,n", [("lj", 20), ("funcfl", 20), ("setfl", 12)])
def test_energy_after_100_steps(kind, n):
    """BASELINE configs C1/C2 (20^3, 32 000 atoms, 100 steps, dt 1 fs) + a setfl case."""
    orc = oracle_sim(KINDS[kind], n)
    a = orc.atoms()
    dev, _ = device_for(orc, (n, n, n))
    dev.upload(a["gid"], a["r"], a["p"])
    dev.compute_force()
    e0 = sum(dev.energy())
    for _ in range(10):
        dev.timestep(10, 1.0)
    ep, ek = dev.energy()
    if n == 20:
        g = np.load(os.path.join(GOLDEN_DIR, "trace_%s_n20.npz" % kind))["trace"]   # reference's own trace
        ref0, ref100 = g[0].sum(), g[-1].sum()
        assert abs(g[-1][0] - ep) <= ENERGY100_TOL * abs(g[-1][0])
    else:
        ref0 = sum(orc.energies())
        orc.timestep(100)
        ref100 = sum(orc.energies())
    # a sum over 32 000 atoms: the reference's own serial summation carries ~N*eps/2 of round-off,
    # so the SUM is gated at 1e-11 (per-atom energies are gated at 1e-12 in the step-0 tests above)
    assert abs(e0 - ref0) <= 1e-11 * abs(ref0)
    assert abs((ep + ek) - ref100) <= ENERGY100_TOL * abs(ref100)
    assert dev.num_local() == len(a["gid"])
    dev.close()


def test_split_calls_equal_fused_timestep():
    """advanceVelocity / advancePosition / redistributeAtoms / computeForce called one by one (the
    reference's timestep.c:48-100 sequence) give bit-identical state to comd_dev_timestep."""
    n = 6
    orc = oracle_sim(po.POT_FUNCFL, n, delta=0.05)
    a = orc.atoms()
    out = []
    for fused in (True, False):
        dev, _ = device_for(orc, (n, n, n))
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        if fused:
            dev.timestep(3, 1.0)
        else:
            for _ in range(3):
                dev.advance_velocity(0.5); dev.advance_position(1.0); dev.redistribute()
                dev.compute_force(); dev.advance_velocity(0.5)
        out.append(by_gid(dev.download()))
        dev.close()
    for k in ("r", "p", "f", "U"):
        assert np.array_equal(out[0][k], out[1][k]), k


def test_step_host_round_trip():
    n = 6
    orc = oracle_sim(po.POT_LJ, n, delta=0.05)
    a = orc.atoms()
    dev, _ = device_for(orc, (n, n, n))
    dev.upload(a["gid"], a["r"], a["p"])
    dev.compute_force()
    d0 = dev.download()
    cols = lambda x: [np.ascontiguousarray(x[:, k]) for k in range(3)]
    m, o, e = dev.step_host(1.0, np.ascontiguousarray(d0["gid"]), cols(d0["r"]), cols(d0["p"]), cols(d0["f"]))
    assert m == len(a["gid"])
    dev2, _ = device_for(orc, (n, n, n))
    dev2.upload(a["gid"], a["r"], a["p"]); dev2.compute_force(); dev2.timestep(1, 1.0)
    d1 = by_gid(dev2.download())
    got = by_gid(dict(gid=o["gid"][:m], r=np.stack([x[:m] for x in o["r"]], 1), p=np.stack([x[:m] for x in o["p"]], 1),
                      f=np.stack([x[:m] for x in o["f"]], 1), U=o["U"][:m]))
    for k in ("r", "p", "f", "U"):
        assert np.array_equal(got[k], d1[k]), k
    assert e == dev2.energy()
    dev.close(); dev2.close()


def test_errors_are_returned_not_fatal():
    orc = oracle_sim(po.POT_LJ, 6)
    dev, _ = device_for(orc, (6, 6, 6))
    a = orc.atoms()
    with pytest.raises(pycomd.ComdError):
        big = np.zeros((dev.capacity + 1, 3))
        dev.upload(np.arange(dev.capacity + 1, dtype=np.int32), big)
    geom = pycomd.make_geometry([6 * 3.615] * 3, 2.5 * 2.315)
    d2 = pycomd.Device(geom, 4096)
    with pytest.raises(pycomd.ComdError):
        d2.compute_force()                       # no potential set
    with pytest.raises(pycomd.ComdError):
        d2.set_lj(cutoff=50.0)                   # cells smaller than the cutoff
    d2.close(); dev.close()


@pytest.mark.parametrize("kind", list(KINDS))
@pytest.mark.parametrize("tx", [1, 3, 8])
def test_tiled_kernels_match_thread_atom_kernels(kind, tx):
    """The tiled TMA/filter/process engine and the plain stencil loop are two summation orders of
    the same pair set: same forces to round-off, and both within tolerance of the oracle."""
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.12, temperature=900.0)
    a = orc.atoms()
    o = by_gid(a)
    res = []
    for variant in (0, 1):
        dev, _ = device_for(orc, (n, n, n))
        dev.set_option("force_variant", variant)
        dev.set_option("tile_tx", tx)
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        d = by_gid(dev.download())
        assert rel_err(d["f"], o["f"]) <= FORCE_TOL
        assert rel_err(d["U"], o["U"] * u_scale(kind)) <= FORCE_TOL
        res.append(d)
        dev.close()
    assert rel_err(res[1]["f"], res[0]["f"]) <= 1e-13
    assert rel_err(res[1]["U"], res[0]["U"]) <= 1e-13


# ---------------------------------------------------------------- neighbour-list path
@pytest.mark.parametrize("kind", list(KINDS))
def test_neighbor_list_step0_forces(kind):
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.1, temperature=600.0)
    a = orc.atoms()
    dev, _ = device_for(orc, (n, n, n), nl_skin=0.1)
    dev.upload(a["gid"], a["r"], a["p"])
    dev.compute_force()
    d = by_gid(dev.download()); o = by_gid(a)
    assert rel_err(d["f"], o["f"]) <= FORCE_TOL
    assert rel_err(d["U"], o["U"] * u_scale(kind)) <= FORCE_TOL
    st = dev.neighbor_stats()
    assert st["builds"] == 1
    # full list: every in-range pair is stored twice; brute-force count on the oracle's positions
    L = n * orc.potential()["lat"]
    rl = orc.potential()["cutoff"] * 1.1
    r = o["r"]
    pairs = 0
    for i0 in range(0, len(r), 256):
        dr = r[i0:i0 + 256, None, :] - r[None, :, :]
        dr -= L * np.round(dr / L)
        r2 = np.sum(dr * dr, axis=2)
        pairs += int(np.sum(r2 <= rl * rl)) - len(r[i0:i0 + 256])
    assert st["pairs"] == pairs
    dev.close()


@pytest.mark.parametrize("kind", list(KINDS))
def test_neighbor_list_trajectory_matches_link_cell_path(kind):
    """60 steps with refreshes and rebuilds: the list path and the link-cell path integrate the same
    pair set, so energies agree to round-off; the list is rebuilt a few times, not every step."""
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.0, temperature=1200.0)
    a = orc.atoms()
    res = []
    for skin in (None, 0.1):
        dev, _ = device_for(orc, (n, n, n), nl_skin=skin)
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        dev.timestep(60, 1.0)
        res.append((by_gid(dev.download()), dev.energy(), dev.neighbor_stats() if skin else None, dev.num_local()))
        dev.close()
    (d0, e0, _, n0), (d1, e1, st, n1) = res
    assert n0 == n1 == len(a["gid"])
    assert 1 < st["builds"] < 30
    assert abs(sum(e1) - sum(e0)) <= 1e-9 * abs(sum(e0))
    assert rel_err(d1["r"], d0["r"]) <= 1e-9
    assert rel_err(d1["f"], d0["f"]) <= 1e-7


@pytest.mark.parametrize("kind", list(KINDS))
def test_neighbor_list_migration_across_faces_matches_link_cell_path(kind):
    """The list-path twin of test_migration_and_conservation: a uniform drift carries every atom through cell and
    periodic faces, so each rebuild ships emigrants and the steps in between run on the in-place ghost refresh.
    An emigrant lives on as a ghost in its old owner's halo cell; the refresh must move that copy with the live
    atom and never write a local one (round 1 replayed the rebuild's send lists and reset immigrants to their
    rebuild-time position on every refresh step).  Same pair set as the link-cell path => same trajectory."""
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.05, temperature=600.0)
    a = orc.atoms()
    drift = orc.potential()["mass"] * np.array([0.02, 0.01, -0.005])
    L = n * orc.potential()["lat"]
    res = []
    for skin in (None, 0.1):
        dev, _ = device_for(orc, (n, n, n), nl_skin=skin)
        dev.upload(a["gid"], a["r"], a["p"] + drift)
        dev.compute_force()
        box0 = by_gid(dev.download())["box"]
        trace = []
        for _ in range(4):
            dev.timestep(40, 1.0)
            d = by_gid(dev.download())
            assert np.array_equal(d["gid"], np.sort(a["gid"]))            # all there, exactly once
            assert dev.num_local() == len(a["gid"])
            trace.append((d, dev.energy()))
        moved = int(np.sum(trace[-1][0]["box"] != box0))
        res.append((trace, dev.neighbor_stats() if skin else None, moved))
        dev.close()
    (t0, _, moved0), (t1, st, _) = res
    assert moved0 > len(a["gid"]) // 4                                    # the test did exercise migration
    assert 4 < st["builds"] < 160                                         # several rebuilds, refresh steps in between
    for (d0, e0), (d1, e1) in zip(t0, t1):
        assert abs(sum(e1) - sum(e0)) <= 1e-9 * abs(sum(e0))
        dr = d1["r"] - d0["r"]
        dr -= L * np.round(dr / L)                                        # the two paths re-bin (wrap) at different steps
        assert np.max(np.abs(dr)) <= 1e-8
        assert rel_err(d1["f"], d0["f"]) <= 1e-6


def test_neighbor_list_energy_after_100_steps_vs_reference():
    """C2 (EAM Cu_u6 20^3, 100 steps) through the list path against the reference's own trace (-m cpu_nl)."""
    n = 20
    orc = oracle_sim(po.POT_FUNCFL, n)
    a = orc.atoms()
    dev, _ = device_for(orc, (n, n, n), nl_skin=0.1)
    dev.upload(a["gid"], a["r"], a["p"])
    dev.compute_force()
    dev.timestep(100, 1.0)
    ep, ek = dev.energy()
    g = np.load(os.path.join(GOLDEN_DIR, "trace_funcfl_n20.npz"))["trace"]
    assert abs((ep + ek) - g[-1].sum()) <= ENERGY100_TOL * abs(g[-1].sum())
    assert abs(ep - g[-1][0]) <= ENERGY100_TOL * abs(g[-1][0])
    assert dev.num_local() == len(a["gid"])
    dev.close()


@pytest.mark.parametrize("kind", ["funcfl", "setfl"])
def test_neighbor_list_energy_on_demand_does_not_change_the_trajectory(kind):
    """The list path evaluates the EAM pair energies only where the caller can read them (after
    comd_dev_compute_force and after the last step of comd_dev_timestep).  Forces come from the same
    instructions either way, so 25 steps in one call, 25 steps with energies at every step and 25 single
    steps must end in the same bits, energies included."""
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.05, temperature=900.0)
    a = orc.atoms()
    out = []
    for mode in ("last", "every", "single"):
        dev, _ = device_for(orc, (n, n, n), nl_skin=0.1)
        if mode == "every":
            dev.set_option("energy_every_step", 1)
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        if mode == "single":
            for _ in range(25):
                dev.timestep(1, 1.0)
        else:
            dev.timestep(25, 1.0)
        out.append((by_gid(dev.download()), dev.energy()))
        dev.close()
    for d, e in out[1:]:
        for k in ("r", "p", "f", "U"):
            assert np.array_equal(d[k], out[0][0][k]), k
        assert e == out[0][1]


@pytest.mark.parametrize("kind", list(KINDS))
@pytest.mark.parametrize("block", [10101, 20302, 80804, 40116])
@pytest.mark.parametrize("n3", [0, 1])
def test_neighbor_list_work_block_shape_is_invisible(kind, block, n3):
    """The list kernels walk blocks of bx*by*bz cells per CTA; the shape (also ragged against the grid, also
    a single block for the whole domain) is a scheduling choice.  Full-list kernels (nl_n3 0): not one bit of the
    result may change.  Once-per-pair kernels (default): the shape decides which pairs are added in 2^-46 fixed
    point instead of FP64, so results agree to that resolution -- and stay bit-reproducible for a given shape."""
    n = 10
    orc = oracle_sim(KINDS[kind], n, delta=0.08, temperature=600.0)
    a = orc.atoms()
    res = []
    for b in (0, block, block):
        dev, _ = device_for(orc, (n, n, n), nl_skin=0.1)
        dev.set_option("nl_n3", n3)
        dev.set_option("nl_block", b)
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        dev.timestep(3, 1.0)
        res.append(by_gid(dev.download()))
        dev.close()
    for k in ("r", "p", "f", "U"):
        assert np.array_equal(res[1][k], res[2][k]), k                  # run-to-run: same bits
        if n3 == 0:
            assert np.array_equal(res[0][k], res[1][k]), k
        else:
            assert rel_err(res[1][k], res[0][k]) <= 2e-13, k


@pytest.mark.parametrize("kind", list(KINDS))
@pytest.mark.parametrize("n3", [0, 1])
def test_neighbor_list_hilbert_traversal_of_work_blocks_is_invisible(kind, n3):
    """Option nl_hilbert (CoMD's -H, linkCells.c:28-63,151-174): the work blocks are handed to the SMs along a 3-D
    Hilbert curve, on a grid that is NOT a power of two (the reference would refuse it).  A scheduling choice only:
    every bit of the trajectory stays."""
    n = 11
    orc = oracle_sim(KINDS[kind], n, delta=0.06, temperature=900.0)
    a = orc.atoms()
    res = []
    for hil in (0, 1):
        dev, _ = device_for(orc, (n, n, n), nl_skin=0.1)
        dev.set_option("nl_n3", n3)
        dev.set_option("nl_block", 10201)              # small blocks: a long curve (several hundred blocks)
        dev.set_option("nl_hilbert", hil)
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        dev.timestep(12, 1.0)
        res.append((by_gid(dev.download()), dev.energy()))
        dev.close()
    assert res[0][1] == res[1][1]
    for k in ("r", "p", "f", "U"):
        assert np.array_equal(res[0][0][k], res[1][0][k]), k


@pytest.mark.parametrize("kind", list(KINDS))
@pytest.mark.parametrize("n3", [0, 1])
def test_neighbor_list_sync_free_step_equals_the_synchronised_step(kind, n3):
    """comd_dev_timestep enqueues a list-mode step for the no-rebuild case without waiting for the displacement check
    (option nl_speculate, default 1): the force kernels return at once when the flag turns out set, the host re-bins
    and repeats them.  Same bits, same number of builds and refreshes as the step that waits for the verdict
    (nl_speculate 0, round 1) -- over a hot run with several rebuilds."""
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.05, temperature=1500.0)
    a = orc.atoms()
    res = []
    for spec in (0, 1):
        dev, _ = device_for(orc, (n, n, n), nl_skin=0.1)
        dev.set_option("nl_n3", n3)
        dev.set_option("nl_speculate", spec)
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        for _ in range(5):
            dev.timestep(10, 1.0)
        res.append((by_gid(dev.download()), dev.energy(), dev.neighbor_counters()))
        dev.close()
    (d0, e0, c0), (d1, e1, c1) = res
    assert c0 == c1 and c0["builds"] >= 3 and c0["refreshes"] >= 30
    assert e0 == e1
    for k in ("r", "p", "f", "U"):
        assert np.array_equal(d0[k], d1[k]), k


@pytest.mark.parametrize("kind", list(KINDS))
@pytest.mark.parametrize("n3", [0, 1])
def test_neighbor_list_graph_replayed_steps_equal_eager_steps(kind, n3):
    """Option graph_step: the steps between two list builds that need no energies are replayed from a CUDA graph captured
    once per build (kick + kick + drift + check, flag, ghost refresh, force kernels -- one launch).  Same bits as the
    eagerly enqueued steps over a hot run with several rebuilds; the rebuild steps themselves are redone eagerly."""
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.05, temperature=1500.0)
    a = orc.atoms()
    res = []
    for graph in (0, 1):
        dev, _ = device_for(orc, (n, n, n), nl_skin=0.1)
        dev.set_option("nl_n3", n3)
        dev.set_option("graph_step", graph)
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        for _ in range(5):
            dev.timestep(10, 1.0)
        res.append((by_gid(dev.download()), dev.energy(), dev.neighbor_counters(), dev.step_stats(), dev.kernel_stats()["launches"]))
        dev.close()
    (d0, e0, c0, s0, l0), (d1, e1, c1, s1, l1) = res
    assert s0["graph_steps"] == 0 and s1["graph_steps"] >= 25          # 8 of the 10 steps of a call qualify, minus the rebuild steps
    assert c0 == c1 and c0["builds"] >= 3 and s0["redone_steps"] == s1["redone_steps"] == c0["builds"] - 1
    assert l0 == l1                                                     # the replays are counted kernel by kernel
    assert e0 == e1
    for k in ("r", "p", "f", "U"):
        assert np.array_equal(d0[k], d1[k]), k


@pytest.mark.parametrize("kind", list(KINDS))
def test_neighbor_list_full_list_kernels_still_match(kind):
    """nl_n3 0 keeps round 1's full-list kernels (every pair met from both ends, no shared-memory accumulators) and
    nl_lanes 8 the eight-lanes-per-row form of the once-per-pair kernels: same parity gate, and the same trajectory
    as the default (once per pair, four lanes per row)."""
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.1, temperature=600.0)
    a = orc.atoms()
    out = []
    for opts in ({"nl_n3": 0}, {}, {"nl_lanes": 8}):
        dev, _ = device_for(orc, (n, n, n), nl_skin=0.1)
        for k, v in opts.items():
            dev.set_option(k, v)
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        d = by_gid(dev.download()); o = by_gid(a)
        assert rel_err(d["f"], o["f"]) <= FORCE_TOL
        assert rel_err(d["U"], o["U"] * u_scale(kind)) <= FORCE_TOL
        dev.timestep(30, 1.0)
        out.append((by_gid(dev.download()), dev.energy(), dev.neighbor_stats()))
        dev.close()
    (d0, e0, s0) = out[0]
    for (d1, e1, s1) in out[1:]:
        assert s0["pairs"] == s1["pairs"] and s0["builds"] == s1["builds"]
        assert abs(sum(e1) - sum(e0)) <= 1e-11 * abs(sum(e0))
        assert rel_err(d1["r"], d0["r"]) <= 1e-12
        assert rel_err(d1["f"], d0["f"]) <= 1e-9


@pytest.mark.parametrize("kind", ["lj", "setfl"])
def test_step_host_chained_calls_resume_without_resort(kind):
    """A trajectory driven through host memory on the list path: each call's results are the next call's
    inputs.  After the first call (full upload) the library must recognise its own atoms, keep the sort and
    the neighbour list (skin/2 criterion) and still land on the trajectory of a device-resident run; handing it
    shuffled atoms must take the full path again and keep integrating the same system."""
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.05, temperature=600.0)
    a = orc.atoms()
    dev1, _ = device_for(orc, (n, n, n), nl_skin=0.1)
    dev1.upload(a["gid"], a["r"], a["p"])
    dev1.compute_force()
    d0 = dev1.download()
    dev1.timestep(1, 1.0)
    for _ in range(5):
        dev1.timestep(1, 1.0)
    d1 = by_gid(dev1.download()); e1 = dev1.energy()
    dev1.close()

    cols = lambda x: [np.ascontiguousarray(x[:, k]) for k in range(3)]
    dev2, _ = device_for(orc, (n, n, n), nl_skin=0.1)
    m, o, e = dev2.step_host(1.0, d0["gid"], cols(d0["r"]), cols(d0["p"]), cols(d0["f"]))
    assert dev2.step_host_resumed() == 0
    for _ in range(5):
        m, o, e = dev2.step_host(1.0, o["gid"][:m].copy(), [x[:m].copy() for x in o["r"]], [x[:m].copy() for x in o["p"]],
                                 [x[:m].copy() for x in o["f"]])
    assert dev2.step_host_resumed() == 5
    got = by_gid(dict(gid=o["gid"][:m].copy(), r=np.stack([x[:m] for x in o["r"]], 1), p=np.stack([x[:m] for x in o["p"]], 1),
                      f=np.stack([x[:m] for x in o["f"]], 1), U=o["U"][:m].copy()))
    # the two runs build their lists from different snapshots (before / after the first drift), so the
    # cutoff-first order of a row, hence the summation order, may differ: round-off, not bits
    assert rel_err(got["r"], d1["r"]) <= 1e-14
    assert rel_err(got["p"], d1["p"]) <= 1e-12
    assert rel_err(got["f"], d1["f"]) <= 1e-11
    assert rel_err(got["U"], d1["U"]) <= 1e-12
    assert abs(sum(e) - sum(e1)) <= 1e-12 * abs(sum(e1))
    # shuffled hand-back: not the library's order any more -> full upload path, same physics to round-off
    perm = np.random.default_rng(3).permutation(m)
    g = o["gid"][:m][perm].copy()
    m2, o2, e2 = dev2.step_host(1.0, g, [x[:m][perm].copy() for x in o["r"]], [x[:m][perm].copy() for x in o["p"]],
                                [x[:m][perm].copy() for x in o["f"]])
    assert dev2.step_host_resumed() == 5 and m2 == m
    assert np.array_equal(np.sort(o2["gid"][:m2]), np.sort(g))
    assert abs(sum(e2) - sum(e1)) <= 1e-5 * abs(sum(e1))            # one more step of an energy-conserving integrator
    dev2.close()


@pytest.mark.parametrize("kind", ["lj", "setfl"])
def test_step_host_in_place_hand_back_runs_ahead_of_the_upload(kind):
    """The caller updates its pinned buffers in place (results of call k are the inputs of call k+1, same
    memory).  On one rank the library starts the step from the atoms it still holds while the upload is in
    flight and verifies the upload against them afterwards: same bits as the path that waits for the upload
    (option e2e_speculate 0).  An input that was touched in between must be noticed: the step is redone from
    the uploaded atoms and matches a library that never speculated."""
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.05, temperature=600.0)
    a = orc.atoms()
    dev0, _ = device_for(orc, (n, n, n), nl_skin=0.1)
    dev0.upload(a["gid"], a["r"], a["p"])
    dev0.compute_force()
    d0 = dev0.download()
    dev0.close()
    cols = lambda x: [np.ascontiguousarray(x[:, k]) for k in range(3)]
    snap = lambda o, m: dict(gid=o["gid"][:m].copy(), r=[x[:m].copy() for x in o["r"]], p=[x[:m].copy() for x in o["p"]],
                             f=[x[:m].copy() for x in o["f"]], U=o["U"][:m].copy())
    runs = []
    for speculate in (1, 0):
        dev, _ = device_for(orc, (n, n, n), nl_skin=0.1)
        dev.set_option("e2e_speculate", speculate)
        m, o, e = dev.step_host(1.0, d0["gid"], cols(d0["r"]), cols(d0["p"]), cols(d0["f"]))
        for _ in range(6):                         # views of the pinned result buffers: in place
            m, o, e = dev.step_host(1.0, o["gid"][:m], [x[:m] for x in o["r"]], [x[:m] for x in o["p"]], [x[:m] for x in o["f"]])
        assert dev.step_host_resumed() == 6
        before = snap(o, m)
        o["p"][0][:m] *= 1.0 + 1e-3                 # a thermostat touches the momenta between two calls
        touched = snap(o, m)
        m2, o2, e2 = dev.step_host(1.0, o["gid"][:m], [x[:m] for x in o["r"]], [x[:m] for x in o["p"]], [x[:m] for x in o["f"]])
        after = snap(o2, m2)
        runs.append((before, e, touched, after, e2, dev.step_host_resumed()))
        dev.close()
    (b1, e1, t1, a1, f1, r1), (b0, e0, t0, a0, f0, r0) = runs
    for k in ("gid", "U"):
        assert np.array_equal(b1[k], b0[k]), k
    for k in ("r", "p", "f"):
        for c in range(3):
            assert np.array_equal(b1[k][c], b0[k][c]), (k, c)
    assert e1 == e0
    assert r1 == 6 and r0 == 7                      # speculation lost once and redid the step; the waiting path resumed
    # after the touch: full path (new sort, new list) against resume -> same atoms, round-off apart
    o1 = np.argsort(a1["gid"]); o0 = np.argsort(a0["gid"])
    assert np.array_equal(a1["gid"][o1], a0["gid"][o0])
    for k, tol in (("r", 1e-14), ("p", 1e-12), ("f", 1e-11)):
        for c in range(3):
            assert rel_err(a1[k][c][o1], a0[k][c][o0]) <= tol, (k, c)
    assert rel_err(a1["U"][o1], a0["U"][o0]) <= 1e-12
    assert abs(sum(f1) - sum(f0)) <= 1e-12 * abs(sum(f0))


@pytest.mark.parametrize("kind", ["lj", "setfl"])
def test_neighbor_list_non_cubic_box(kind):
    """6 x 9 x 13 unit cells: different cell counts per axis, work blocks ragged against every face."""
    nxyz = (6, 9, 13)
    orc = oracle_sim(KINDS[kind], 0, delta=0.07, temperature=600.0, nxyz=nxyz)
    a = orc.atoms()
    dev, _ = device_for(orc, nxyz, nl_skin=0.1)
    dev.upload(a["gid"], a["r"], a["p"])
    dev.compute_force()
    d = by_gid(dev.download()); o = by_gid(a)
    assert rel_err(d["f"], o["f"]) <= FORCE_TOL
    assert rel_err(d["U"], o["U"] * u_scale(kind)) <= FORCE_TOL
    dev.timestep(10, 1.0)
    orc.timestep(10)
    ep, ek = dev.energy(); eo = orc.energies()
    assert abs((ep + ek) - sum(eo)) <= 1e-9 * abs(sum(eo))
    assert dev.num_local() == len(a["gid"])
    dev.close()


# ---------------------------------------------------------------- BASELINE.json's full sizes
@pytest.mark.parametrize("path", ["link_cell", "list"])
def test_c3_full_size_setfl_60_against_the_oracle(path):
    """BASELINE.json configs[2], EAM Cu01.eam.alloy 60^3 = 864 000 atoms, both force paths: step-0 per-atom forces and
    energies against the CPU oracle at the full size (<= 1e-12), and the total energy after 10 steps (<= 1e-9; the 100-step
    1e-6 gate is tested at 12^3 and 20^3 where the oracle needs seconds, not minutes)."""
    n = 60
    orc = oracle_sim(po.POT_SETFL, n, delta=0.1, temperature=600.0)
    a = orc.atoms()
    dev, _ = device_for(orc, (n, n, n), capacity_factor=1.6, nl_skin=0.1 if path == "list" else None)
    dev.upload(a["gid"], a["r"], a["p"])
    dev.compute_force()
    d = by_gid(dev.download()); o = by_gid(a)
    assert np.array_equal(d["gid"], o["gid"])
    assert rel_err(d["f"], o["f"]) <= FORCE_TOL
    assert rel_err(d["U"], o["U"]) <= FORCE_TOL
    dev.timestep(10, 1.0)
    orc.timestep(10)
    e_dev, e_orc = sum(dev.energy()), sum(orc.energies())
    assert abs(e_dev - e_orc) <= 1e-9 * abs(e_orc)
    assert dev.num_local() == len(a["gid"])
    dev.close()


@pytest.mark.parametrize("workload", ["lj_weak", "eam_strong"])
def test_c4_c5_full_size_properties(workload):
    """BASELINE.json configs[3] (LJ 100^3, 4.0 M atoms per GPU) and configs[4] (EAM Cu01 160^3, 16.384 M atoms) on one GPU,
    list path, through properties that need no oracle run: Newton's third law (the forces of a periodic system sum to zero --
    a missed, doubled or one-sided pair breaks it), conservation of the atom count and of the total momentum over 20 steps with
    a list rebuild, energy conservation as the reference states it (CoMD.c:914-945), and the step-0 energy per atom equal to
    the 20^3 oracle's (a perfect lattice is periodic with the unit cell: the per-atom energy does not depend on the size)."""
    from pycomd import potentials as P
    pot_kind, ncell = ("lj", 100) if workload == "lj_weak" else ("setfl", 160)
    pot = P.load(pot_kind)
    lat = pot["lat"]
    skin = 0.1 * pot["cutoff"]
    geom = pycomd.make_geometry([ncell * lat] * 3, pot["cutoff"] + skin)
    gid, r = P.fcc_block([ncell] * 3, lat, [0.0] * 3, [ncell * lat] * 3)
    p = P.maxwell_momenta(len(gid), pot["mass"], 600.0, seed=11)
    p -= p.mean(axis=0)                                        # zero total momentum, as setVcm leaves it (initAtoms.c:127-150)
    dev = pycomd.Device(geom, int(len(gid) * 1.3) + 65536)
    dev.set_mass(pot["mass"])
    if pot_kind == "lj":
        dev.set_lj(pot["sigma"], pot["epsilon"], pot["cutoff"])
    else:
        dev.set_eam(pot["cutoff"], pot["phi"], pot["rho"], pot["f"])
    dev.set_neighbor_list(True, skin)
    dev.upload(gid, r, p)
    dev.compute_force()
    e0 = dev.energy()
    small = oracle_sim(KINDS[pot_kind], 20, delta=0.0, temperature=0.0)
    assert abs(e0[0] / len(gid) - small.energies()[0] / (4 * 20 ** 3)) <= 1e-11 * abs(e0[0] / len(gid))       # positions round differently at 5x the box
    dev.timestep(20, 1.0)
    d = dev.download()
    e1 = dev.energy()
    assert len(d["gid"]) == len(gid) and dev.num_local() == len(gid)
    assert np.array_equal(np.sort(d["gid"]), np.sort(gid))
    fmax = np.max(np.abs(d["f"]))
    assert fmax > 0.1
    assert np.max(np.abs(d["f"].sum(axis=0))) <= 1e-9 * fmax * np.sqrt(len(gid))          # sum of forces: round-off only
    pscale = np.sqrt(np.mean(d["p"] ** 2))
    assert np.max(np.abs(d["p"].sum(axis=0))) <= 1e-9 * pscale * np.sqrt(len(gid))          # total momentum stays zero
    assert abs(sum(e1) - sum(e0)) <= (5e-5 if pot_kind == "lj" else 5e-6) * abs(sum(e0))   # the reference's own LJ run drifts 1.1e-5 (cut-and-shifted pair force), its EAM runs 4e-6
    assert dev.neighbor_counters()["builds"] >= 1
    dev.close()


def _expect_error_then_recover(kind, gid, r, p, orc, n, code, words, nl_block=0):
    """The once-per-pair list kernels (option nl_n3 1) must REPORT a configuration outside their limits at the next call
    that returns data -- never hand back forces from a wrapped or skipped accumulator -- and the report names the way on:
    nl_n3 0 and a fresh upload.  Following it gives the oracle's forces and energies for the very same atoms
    (ljForce.c:146-265 / eam.c:266-419 compute any configuration; so does this library, through its full-list kernels)."""
    dev, _ = device_for(orc, (n, n, n), nl_skin=0.1)
    try:
        dev.set_option("nl_n3", 1)
        dev.set_option("nl_block", nl_block)
        dev.upload(gid, r, p)
        dev.compute_force()
        with pytest.raises(pycomd.ComdError) as err:
            dev.download()
        assert err.value.code == code, str(err.value)
        for w in words + ("nl_n3",):
            assert w in str(err.value), str(err.value)
        dev.set_option("nl_n3", 0)
        dev.upload(gid, r, p)                                  # clears the sticky error bits, re-bins, rebuilds the list in full form
        dev.compute_force()
        d = by_gid(dev.download())
    finally:
        dev.close()
    orc.set_atoms(gid, r, p)
    o = by_gid(orc.atoms())
    assert np.array_equal(d["gid"], o["gid"])
    assert rel_err(d["f"], o["f"]) <= FORCE_TOL
    assert rel_err(d["U"], o["U"] * u_scale(kind)) <= FORCE_TOL
    return d


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["lj", "funcfl"])
def test_once_per_pair_kernels_report_a_pair_beyond_their_fixed_point_range(kind):
    """Two atoms 1.2 A apart: their pair force is far beyond the +-32 eV/A one addend of the shared-memory fixed-point
    accumulators can hold (force_n3.cuh).  ERRB_FIXED -> COMD_E_STATE, not a silently wrapped sum."""
    n = 8
    orc = oracle_sim(KINDS[kind], n, delta=0.0, temperature=300.0)
    a = by_gid(orc.atoms())
    lat = orc.potential()["lat"]
    r = a["r"].copy()
    centre = np.argmin(np.sum((r - 0.5 * n * lat) ** 2, axis=1))
    other = np.argsort(np.sum((r - r[centre]) ** 2, axis=1))[1]         # a nearest neighbour: pulled in to 1.2 A along x
    r[other] = r[centre] + np.array([1.2, 0.0, 0.0])
    d = _expect_error_then_recover(kind, a["gid"], r, a["p"], orc, n, -4, ("fixed-point",))
    assert np.max(np.abs(d["f"])) > 32.0                                 # the configuration really is outside the range


@pytest.mark.gpu
def test_once_per_pair_kernels_report_a_work_block_denser_than_their_accumulators():
    """A crystal squeezed to 1.45x its density over the first 4 link cells along x (and thinned out over the rest): the work
    block there holds more atoms than the launch sized its shared-memory accumulators for (1.3x the mean + 48).
    ERRB_TILE -> COMD_E_CAPACITY, not forces with that block left out."""
    kind, n, squeeze = "funcfl", 12, 1.45
    orc = oracle_sim(KINDS[kind], n, delta=0.0, temperature=300.0)
    a = by_gid(orc.atoms())
    pot = orc.potential()
    L = n * pot["lat"]
    cells = int(L / (1.1 * pot["cutoff"]))                               # link cells per axis on cutoff + skin
    edge = 4.2 * (L / cells)                                             # the squeezed slab covers the first block (4 cells) and a bit
    x0 = edge * squeeze                                                  # original extent that is mapped onto [0, edge)
    assert x0 < 0.95 * L
    r = a["r"].copy()
    x = r[:, 0]
    r[:, 0] = np.where(x < x0, x / squeeze, edge + (x - x0) * (L - edge) / (L - x0))
    assert r[:, 0].min() >= 0.0 and r[:, 0].max() < L
    # (a domain this small would get one-cell work blocks, whose fixed slack of 48 slots hides any squeeze: ask for the
    # 4x4x4-cell blocks a production-size domain runs with)
    _expect_error_then_recover(kind, a["gid"], r, a["p"], orc, n, -3, ("work block",), nl_block=40404)


@pytest.mark.gpu
@pytest.mark.parametrize("pot_kind,n3", [("lj", 1), ("funcfl", 1), ("funcfl", 0)])
def test_second_gpu_of_one_process_gets_its_own_kernel_attributes(pot_kind, n3):
    """One process, two handles on two GPUs (the ABI takes a device id; CoMD.c:105-117 maps one rank to one GPU, a host
    that drives several GPUs from one process is equally legal).  The dynamic shared-memory limit and carve-out of the list
    kernels are per-device function attributes: the handle on GPU 1 must get them too, or its launches fail once the
    accumulators exceed 48 KB.  Forces and energies on GPU 1 equal GPU 0's bit for bit.  Needs two GPUs; skipped on one."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    kind = KINDS[pot_kind]
    got = []
    for device_id in (0, 1):
        orc = oracle_sim(kind, 12, delta=0.05, temperature=600.0)
        a = orc.atoms()
        pot = orc.potential()
        skin = 0.1 * pot["cutoff"]
        geom = pycomd.make_geometry([12 * pot["lat"]] * 3, pot["cutoff"] + skin)
        dev = pycomd.Device(geom, 4 * 12 ** 3 * 3 + 16384, device_id=device_id)
        dev.set_mass(pot["mass"])
        if kind == po.POT_LJ:
            dev.set_lj(2.315, LJ_EPS, pot["cutoff"])
        else:
            dev.set_eam(pot["cutoff"], orc.table(0), orc.table(1), orc.table(2))
        dev.set_neighbor_list(True, skin)
        dev.set_option("nl_n3", n3)
        dev.set_option("nl_block", 40404)                      # 4x4x4 cells: the largest accumulator block (> 48 KB for LJ)
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        dev.timestep(3, 1.0)
        got.append((by_gid(dev.download()), dev.energy()))
        dev.close()
    (d0, e0), (d1, e1) = got
    assert np.array_equal(d0["f"], d1["f"]) and np.array_equal(d0["r"], d1["r"]) and np.array_equal(d0["U"], d1["U"])
    assert e0 == e1
