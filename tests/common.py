"""Shared helpers for the parity tests (the oracle is the checker, never the product)."""
import numpy as np

import pyoracle as po
import pycomd

LJ_SIGMA, LJ_EPS = 2.315, 0.167
GOLDEN_DIR = __import__("os").path.join(__import__("os").path.dirname(__import__("os").path.abspath(__file__)), "golden")


def by_gid(d, keys=("r", "p", "f", "U", "box")):
    o = np.argsort(d["gid"], kind="stable")
    out = {"gid": d["gid"][o]}
    for k in keys:
        if k in d:
            out[k] = d[k][o]
    return out


def oracle_sim(pot, n, delta=0.0, temperature=600.0, skin=0.1, nxyz=None, lj_cut_mult=2.5):
    nx, ny, nz = nxyz if nxyz else (n, n, n)
    return po.Oracle(pot=pot, nx=nx, ny=ny, nz=nz, delta=delta, temperature=temperature, skin=skin,
                     lj_cut_mult=lj_cut_mult)


def device_for(orc, nxyz, capacity_factor=3.0, proc_grid=(1, 1, 1), proc_coord=(0, 0, 0), nl_skin=None):
    """A device context with the link cells sized on the bare cutoff (on cutoff + skin with the
    neighbour-list path, CoMD.c:257-268), potential taken from the oracle."""
    pot = orc.potential()
    lat = pot["lat"]
    gext = [nxyz[0] * lat, nxyz[1] * lat, nxyz[2] * lat]
    skin = pot["cutoff"] * nl_skin if nl_skin is not None else 0.0
    geom = pycomd.make_geometry(gext, pot["cutoff"] + skin, proc_grid, proc_coord)
    n_atoms = 4 * nxyz[0] * nxyz[1] * nxyz[2]
    dev = pycomd.Device(geom, int(n_atoms * capacity_factor) + 16384)
    dev.set_mass(pot["mass"])
    if orc.pot == po.POT_LJ:
        dev.set_lj(LJ_SIGMA, LJ_EPS, pot["cutoff"])
    else:
        dev.set_eam(pot["cutoff"], orc.table(0), orc.table(1), orc.table(2))
    if nl_skin is not None:
        dev.set_neighbor_list(True, skin)
    return dev, geom


def rel_err(a, b):
    """max |a-b| normalised by max |b| (forces on a perfect lattice are round-off,
    so per-component relative error is meaningless; the norm is the largest reference entry)."""
    den = np.max(np.abs(b))
    return float(np.max(np.abs(a - b)) / (den if den > 0 else 1.0))
