"""CPU, world_size 2 over gloo: the host-side logic of bench.py's parity leg at N>1 -- who owns an atom, which link
cell of the owner's brick it falls into, and the gather of per-rank results to one gid-sorted global state -- checked
against the oracle's single-rank state.  No device is touched: each rank "computes" by looking its atoms up in the
oracle's arrays, so a wrong partition, a lost or duplicated atom or a scrambled gather shows up as a mismatch."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return port


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import bench
    import pyoracle as po
    import pycomd
    from pycomd import potentials as P
    R = bench.Ranks("gloo")
    assert R.grid == (2, 1, 1) and R.coord == (rank, 0, 0)
    n = 8
    orc = po.Oracle(pot=po.POT_FUNCFL, nx=n, ny=n, nz=n, delta=0.1, temperature=600.0)
    a = orc.atoms()
    pot = P.load("funcfl")
    gext = [n * pot["lat"]] * 3
    cut = pot["cutoff"] + 0.1 * pot["cutoff"]
    own = bench.owner_of(a["r"], gext, R.grid)
    mine = own == rank
    box = bench.expected_boxes(a["r"], own, gext, R.grid, cut, po)
    rows = np.concatenate([a["gid"][mine].astype(float)[:, None], a["f"][mine], a["U"][mine][:, None], box[mine].astype(float)[:, None]], axis=1)
    allrows, counts = R.gather_rows(rows)
    tot = R.reduce([float(np.sum(mine))])[0]
    each = R.reduce_each(int(np.sum(mine)))
    mx = R.reduce([float(rank)], "max")[0]
    if rank == 0:
        o = np.argsort(allrows[:, 0], kind="stable"); g = np.argsort(a["gid"], kind="stable")
        geom = pycomd.make_geometry(gext, cut, R.grid, (1, 0, 0))
        np.savez(out, ok_gid=np.array_equal(allrows[o, 0].astype(np.int64), a["gid"][g]), ok_f=np.array_equal(allrows[o, 1:4], a["f"][g]),
                 ok_u=np.array_equal(allrows[o, 4], a["U"][g]), counts=np.array(counts), tot=tot, each=np.array(each), mx=mx,
                 nbox_max=int(box.max()), nboxes=pycomd.n_total_boxes(tuple(geom.gridSize[:])), natoms=len(a["gid"]),
                 right_half=bool(np.all(a["r"][own == 1][:, 0] >= geom.localMin[0])))
    R.barrier()
    R.dist.destroy_process_group()


def test_two_ranks_gather_matches_the_oracle_state(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "r0.npz")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    d = np.load(out)
    assert bool(d["ok_gid"]) and bool(d["ok_f"]) and bool(d["ok_u"])        # every atom exactly once, values attached to the right gid
    assert int(d["counts"].sum()) == int(d["natoms"]) == int(d["tot"]) and list(d["each"]) == list(d["counts"])
    assert d["counts"][0] == d["counts"][1]                                   # a 2x1x1 split of a cubic lattice is even
    assert float(d["mx"]) == 1.0
    assert 0 <= int(d["nbox_max"]) < int(d["nboxes"]) and bool(d["right_half"])


def test_reference_arm_process_grids():
    import bench
    assert bench._ref_grid(1) == (1, 1, 1) and bench._ref_grid(8) == (2, 2, 2)
    assert bench._ref_grid(16) == (4, 2, 2) and bench._ref_grid(32) == (4, 4, 2) and bench._ref_grid(12) == (3, 2, 2)
    for p in (6, 24, 48, 64):
        g = bench._ref_grid(p)
        assert g[0] * g[1] * g[2] == p


def test_reference_arm_prints_one_contract_line(tmp_path):
    """`bench.py --impl reference` (the arm the driver runs first): ONE JSON line with the contract's keys, the reference's own
    CPU path (oracle/_ref, -m cpu_nl) as one job of cooperating ranks on this host's cores, zero bytes over PCIe."""
    import json
    import subprocess
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "CoMD-ref")):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "1", "--warmup", "1",
                          "--workload", "eam_u6_20"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in j, key
    assert j["impl"] == "reference" and j["unit"] == "atom-steps/s" and j["higher_is_better"] is True and j["value"] > 1e5
    cb = j["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and "cooperating ranks" in cb["sample"] and "single-rank copies" in cb["sample"]
    # the line's value is the FASTER of the two ways of filling the cores, and says which one it is
    every = list((cb["mpi_job"] or {}).values()) + list((cb["independent_copies"] or {}).values())
    assert cb["value_is"] in ("mpi_job", "independent_copies") and cb["value"] == max(every) == max(cb[cb["value_is"]].values())
    assert j["e2e"] == {"value": j["value"], "unit": "atom-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in j["config"] and "model" not in j["config"]
