"""CPU: the plain-C host layer (comd-cuda-async_b200/host/) against the oracle -- initial state
(lattice, gid-seeded momenta at exactly T, random displacements), link-cell geometry and box
numbering, the option table, and the 2-rank shared-memory collectives (fork, no MPI, no GPU)."""
import os
import struct
import subprocess

import numpy as np
import pytest

import pyoracle as po
from common import oracle_sim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "comd-cuda-async_b200")
POTS = os.path.join(ROOT, "pots")


@pytest.fixture(scope="module")
def tools():
    subprocess.check_call(["make", "-s", "-C", PKG, "host"])
    return {k: os.path.join(PKG, k) for k in ("host_dump", "comd-launch", "CoMD-b200", "host_dump_f32", "CoMD-b200-f32")}


def read_dump(path, real=np.float64):
    with open(path, "rb") as fh:
        raw = fh.read()
    n, n_global, gx, gy, gz, n_boxes = struct.unpack_from("6i", raw, 0)
    dh = np.frombuffer(raw, np.float64, 15, 24)
    off = 24 + 15 * 8
    gid = np.frombuffer(raw, np.int32, n, off); off += 4 * n
    box = np.frombuffer(raw, np.int32, n, off); off += 4 * n
    w = np.dtype(real).itemsize
    r = np.frombuffer(raw, real, 3 * n, off).reshape(3, n).T; off += 3 * w * n
    p = np.frombuffer(raw, real, 3 * n, off).reshape(3, n).T; off += 3 * w * n
    tables = []
    while off < len(raw):                                         # --doeam: phi, rho, F as the host layer holds them
        tn = struct.unpack_from("i", raw, off)[0]; off += 4
        x0, inv_dx = struct.unpack_from("2d", raw, off); off += 16
        tables.append(dict(n=tn, x0=x0, inv_dx=inv_dx, padded=np.frombuffer(raw, real, tn + 3, off))); off += w * (tn + 3)
    return dict(n=n, n_global=n_global, grid=(gx, gy, gz), n_boxes=n_boxes, cutoff=dh[0], mass=dh[1], lat=dh[2], box_size=dh[3:6],
                inv_box_size=dh[6:9], local_min=dh[9:12], local_max=dh[12:15], gid=gid, box=box, r=r, p=p, tables=tables)


FLAGS = {"lj": [], "funcfl": ["-e"], "setfl": ["-e", "-t", "setfl"]}
KINDS = {"lj": po.POT_LJ, "funcfl": po.POT_FUNCFL, "setfl": po.POT_SETFL}


@pytest.mark.parametrize("kind", list(KINDS))
def test_initial_state_is_bit_identical_to_the_oracle(tools, tmp_path, kind):
    n = 7
    out = str(tmp_path / "d.bin")
    subprocess.check_call([tools["host_dump"], out, "-x", str(n), "-y", str(n), "-z", str(n), "-r", "0.07", "-T", "450", "-d", POTS] + FLAGS[kind],
                          cwd=tmp_path, stdout=subprocess.DEVNULL)
    d = read_dump(out)
    orc = oracle_sim(KINDS[kind], n, delta=0.07, temperature=450.0, skin=0.0)
    a = orc.atoms()
    pot = orc.potential()
    assert d["n"] == d["n_global"] == 4 * n ** 3 == len(a["gid"])
    assert d["cutoff"] == pot["cutoff"] and d["mass"] == pot["mass"] and d["lat"] == pot["lat"]     # readers, unit conversion
    g = orc.geometry()
    assert tuple(g["grid"]) == d["grid"] and g["n_total_boxes"] == d["n_boxes"]
    assert np.array_equal(g["inv_box_size"], d["inv_box_size"]) and np.array_equal(g["box_size"], d["box_size"])
    o = np.argsort(a["gid"]); h = np.argsort(d["gid"])
    assert np.array_equal(d["gid"][h], a["gid"][o])
    assert np.array_equal(d["r"][h], a["r"][o])                   # lattice + gid-seeded displacement
    assert np.array_equal(d["p"][h], a["p"][o])                   # gid-seeded momenta, COM removal, rescale to T: same sums, same order
    assert np.array_equal(d["box"][h], a["box"][o])               # getBoxFromCoord / getBoxFromTuple


@pytest.mark.parametrize("kind", list(KINDS))
def test_single_precision_host_layer_matches_the_float_reference(tools, tmp_path, kind):
    """host/ built with -DCOMD_SINGLE (real_t = float, mytype.h:8-24) against what the reference's own float build
    produced (tests/golden/f32_*.npz, made by make_golden_f32.py from oracle/_ref/libcomd_ref_f32.so): potential
    constants, lattice + displacements, momenta, the link-cell id of every atom -- bit for bit, all floats."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "f32_%s_n6.npz" % kind))
    base = ["-x", "6", "-y", "6", "-z", "6", "-r", "0.1", "-T", "600", "-d", POTS] + FLAGS[kind]
    out = str(tmp_path / "d32.bin")
    subprocess.check_call([tools["host_dump_f32"], out] + base, cwd=tmp_path, stdout=subprocess.DEVNULL)
    d = read_dump(out, np.float32)
    assert d["n"] == len(g["gid"]) == 864
    assert d["cutoff"] == float(g["cutoff"]) and d["mass"] == float(g["mass"]) and d["lat"] == float(g["lat"])
    assert np.float32(d["cutoff"]) == d["cutoff"]                 # the constants live in floats
    assert tuple(int(x) for x in g["cell_grid"]) == d["grid"] and int(g["cell_n_total"]) == d["n_boxes"]
    assert np.array_equal(g["cell_inv_box_size"], d["inv_box_size"]) and np.array_equal(g["cell_box_size"], d["box_size"])
    o = np.argsort(g["gid"]); h = np.argsort(d["gid"])
    assert np.array_equal(d["gid"][h], g["gid"][o])
    assert np.array_equal(d["r"][h], g["r"][o])
    assert np.array_equal(d["box"][h], g["cell_of_atom"][o])      # getBoxFromCoord in float, cells of one cutoff
    # momenta: float sums (centre of mass, temperature) depend on the traversal order, i.e. on the link cells; the golden
    # run is the reference's -m cpu_nl (cells of cutoff + skin), so the same method here
    subprocess.check_call([tools["host_dump_f32"], out, "-m", "thread_atom_nl"] + base, cwd=tmp_path, stdout=subprocess.DEVNULL)
    d = read_dump(out, np.float32)
    h = np.argsort(d["gid"])
    assert np.array_equal(d["r"][h], g["r"][o])
    assert np.array_equal(d["p"][h], g["p"][o])


@pytest.mark.parametrize("kind", ["funcfl", "setfl"])
def test_potential_readers_give_the_reference_tables_entry_for_entry(tools, tmp_path, kind):
    """eamReadFuncfl / eamReadSetfl + initInterpolationObject (eam.c:680-872,496-519) as restated in host/potentials.c (C)
    and pycomd/potentials.py (test plumbing): n, x0, 1/dx and every padded entry of phi, rho and F equal the oracle's,
    and -- where oracle/_ref is built -- the tables the reference's own code leaves in its EamPotential."""
    import pycomd.potentials as P
    out = str(tmp_path / "t.bin")
    subprocess.check_call([tools["host_dump"], out, "-x", "4", "-y", "4", "-z", "4", "-d", POTS] + FLAGS[kind], cwd=tmp_path, stdout=subprocess.DEVNULL)
    host = read_dump(out)["tables"]
    assert len(host) == 3
    py = P.load(kind)
    orc = po.Oracle(pot=KINDS[kind], nx=4, ny=4, nz=4, initial_force=False)
    ref = po.Ref(["-e", "-x", 4, "-y", 4, "-z", 4] + (["-t", "setfl"] if kind == "setfl" else [])) if po.have_ref() else None
    for which, name in enumerate(("phi", "rho", "f")):
        want = [orc.table(which)] + ([ref.table(which)] if ref else [])
        for w in want:
            for got in (host[which], py[name]):
                assert got["n"] == w["n"] and got["x0"] == w["x0"] and got["inv_dx"] == w["inv_dx"], (name, got["n"], w["n"])
                assert np.array_equal(np.asarray(got["padded"]), w["padded"]), name
    assert py["cutoff"] == orc.potential()["cutoff"] and py["mass"] == orc.potential()["mass"] and py["lat"] == orc.potential()["lat"]


def test_reference_host_code_links_against_the_c_abi(tmp_path, gpu_available):
    """INTEGRATION.md section 2 as a build: the reference's own host C (compiled unmodified from /root/reference by
    oracle/ref_build/Makefile) with its device entry points bound to libcomd_b200.so through oracle/ref_build/ref_bind_b200.c.
    Here (no GPU needed): it links, its undefined symbols are the C ABI's, and without a device it stops at comd_dev_create
    with the library's message -- no CPU fallback behind the reference's GPU methods.  tests/test_gpu_driver.py runs it."""
    exe = os.path.join(ROOT, "oracle", "_ref", "CoMD-ref-b200")
    if os.path.isdir("/root/reference/src-mpi"):
        subprocess.check_call(["make", "-s", "-C", PKG, "libcomd_b200.so"])
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle", "ref_build"), "../_ref/CoMD-ref-b200"])
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/CoMD-ref-b200 not built (needs /root/reference at build time)")
    und = subprocess.run(["nm", "-D", "--undefined-only", exe], capture_output=True, text=True, check=True).stdout
    for sym in ("comd_dev_create", "comd_dev_upload_atoms", "comd_dev_redistribute", "comd_dev_compute_force", "comd_dev_energy",
                "comd_dev_advance_velocity", "comd_dev_advance_position", "comd_dev_set_eam", "comd_dev_set_lj"):
        assert sym in und, sym
    if not gpu_available:
        res = subprocess.run([exe, "-e", "-x", "6", "-y", "6", "-z", "6", "-N", "10", "-n", "5", "-d", POTS, "-m", "thread_atom"],
                             cwd=tmp_path, capture_output=True, text=True)
        assert "no CUDA device" in res.stderr and "comd_dev_create" in res.stderr


def test_two_ranks_partition_the_same_lattice(tools, tmp_path):
    out = str(tmp_path / "d.bin")
    subprocess.check_call([tools["comd-launch"], "-n", "2", tools["host_dump"], out, "-x", "10", "-y", "6", "-z", "6", "-i", "2", "-r", "0.03",
                           "-e", "-d", POTS], cwd=tmp_path, stdout=subprocess.DEVNULL, timeout=120)
    parts = [read_dump("%s.%d" % (out, r)) for r in range(2)]
    orc = po.Oracle(pot=po.POT_FUNCFL, nx=10, ny=6, nz=6, delta=0.03, temperature=600.0, skin=0.0)
    a = orc.atoms()
    gid = np.concatenate([p["gid"] for p in parts]); r = np.concatenate([p["r"] for p in parts]); mom = np.concatenate([p["p"] for p in parts])
    assert parts[0]["n_global"] == parts[1]["n_global"] == len(a["gid"]) == len(gid)
    o = np.argsort(a["gid"]); h = np.argsort(gid)
    assert np.array_equal(gid[h], a["gid"][o])                    # every atom on exactly one rank
    assert np.array_equal(r[h], a["r"][o])
    # the centre-of-mass and temperature sums are reduced rank by rank here, box by box there: round-off only
    assert np.max(np.abs(mom[h] - a["p"][o])) <= 1e-13 * np.max(np.abs(a["p"]))
    for rank, p in enumerate(parts):                              # ownership = position inside the rank's brick
        assert np.all(p["r"][:, 0] >= p["local_min"][0] - 0.03) and np.all(p["r"][:, 0] < p["local_max"][0] + 0.03)
        assert p["local_min"][0] == rank * parts[0]["local_max"][0]


def test_driver_refuses_to_run_without_a_gpu_and_keeps_the_option_table(tools, tmp_path, gpu_available):
    res = subprocess.run([tools["CoMD-b200"], "--help"], cwd=tmp_path, capture_output=True, text=True)
    assert res.returncode == 2
    for opt in ("--potDir", "--potName", "--potType", "--doeam", "--nx", "--xproc", "--nSteps", "--printRate", "--dt", "--lat", "--temp",
                "--delta", "--hilbert", "--skinDistance", "--method", "--gpuAsync", "--gpuProfile", "--ljInterpolation", "--spline",
                "--usePairlist"):
        assert opt in res.stdout, opt
    res = subprocess.run([tools["CoMD-b200"], "-m", "cpu_nl", "-x", "6", "-y", "6", "-z", "6"], cwd=tmp_path, capture_output=True, text=True)
    assert res.returncode != 0 and "no CPU force path" in res.stdout
    if not gpu_available:
        res = subprocess.run([tools["CoMD-b200"], "-x", "6", "-y", "6", "-z", "6"], cwd=tmp_path, capture_output=True, text=True)
        assert res.returncode != 0 and "no CUDA device" in res.stderr      # fails loudly, no CPU fallback


@pytest.mark.parametrize("args", [["-e", "-x", "6", "-y", "6", "-z", "6", "-N", "10", "-n", "5"],
                                  ["-x", "8", "-y", "6", "-z", "6", "-T", "450", "-r", "0.05", "-D", "2.0", "-l", "3.7"],
                                  ["-e", "-t", "setfl", "-p", "Cu01.eam.alloy", "-x", "6", "-y", "6", "-z", "6", "-H", "1", "-a", "1"]])
def test_command_line_block_is_the_references_text(tools, tmp_path, args, gpu_available):
    """north_star: "the same command-line options and YAML output".  The block both drivers print (and write to their YAML
    report) right after parsing -- option names, order, units, number formatting (mycommand.c:262-320) -- is compared line by
    line between CoMD-b200 and the reference executable for the same arguments; only the value of `GPU method` may differ
    (the reference needs -m cpu_nl to run here).  Works without a device: the block is out before the device is created."""
    ref = os.path.join(ROOT, "oracle", "_ref", "CoMD-ref")
    if not os.path.exists(ref):
        pytest.skip("oracle/_ref/CoMD-ref not built (needs /root/reference at build time)")
    def block(exe, extra):
        res = subprocess.run([exe] + args + ["-d", POTS] + extra, cwd=tmp_path, capture_output=True, text=True, timeout=300)
        out = res.stdout
        a = out.index("Command Line Parameters:")
        b = out.index("Space-filling (Hilbert)", a)
        return out[a:out.index("\n", b)].splitlines()
    ours = block(tools["CoMD-b200"], [])
    theirs = block(ref, ["-m", "cpu_nl"])
    assert len(ours) == len(theirs) > 15
    for x, y in zip(ours, theirs):
        if x.strip().startswith("GPU method"):
            assert x.split(":")[0] == y.split(":")[0]
            continue
        assert x == y, (x, y)
