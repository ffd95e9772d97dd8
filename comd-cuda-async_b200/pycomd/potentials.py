"""Host-side potential set-up mirroring the reference's readers (plumbing for tests/bench; the C
host layer in host/ carries the same logic for the CoMD-compatible driver).

  read_funcfl  <-> eamReadFuncfl (eam.c:802-872)     read_setfl <-> eamReadSetfl (eam.c:680-757)
  make_table   <-> initInterpolationObject (eam.c:496-519)

Python floats are IEEE doubles and float() / glibc's strtod are both correctly rounded, so the
tables are bit-identical to the reference's (tests/test_host_potentials.py).
"""
import os

import numpy as np

AMU_TO_INTERNAL_MASS = 1.660538921e-27 * 1.0e-10 * 1.0e-10 / (1.0e-15 * 1.0e-15 * 1.602176565e-19)   # constants.h:28-30
HARTREE_TO_EV = 27.21138505
BOHR_TO_ANGS = 0.52917721092
KB_EV = 8.6173324e-5

LJ_CU = dict(sigma=2.315, epsilon=0.167, mass=63.55 * AMU_TO_INTERNAL_MASS, lat=3.615, name="Cu", atomic_no=29)   # ljForce.c:102-120


def pot_dir():
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    for d in (os.environ.get("COMD_POT_DIR", ""), os.path.join(root, "pots"), "/root/reference/pots"):
        if d and os.path.exists(os.path.join(d, "Cu_u6.eam")):
            return d
    raise FileNotFoundError("EAM tables (Cu_u6.eam, Cu01.eam.alloy) not found; set COMD_POT_DIR or run build()")


def make_table(n, x0, dx, data):
    v = np.zeros(n + 3)
    v[1:n + 1] = data[:n]
    v[0] = v[1]
    v[n + 1] = v[n]
    v[n + 2] = v[n]
    return dict(n=int(n), x0=float(x0), inv_dx=1.0 / dx, padded=v)


def _numbers(tokens, count):
    return [float(tokens.pop(0)) for _ in range(count)]


def read_funcfl(path):
    with open(path) as fh:
        lines = fh.read().split("\n")
    name = lines[0].split()[0] if lines[0].split() else ""
    t = lines[1].split()
    atomic_no, mass, lat, lattice = int(t[0]), float(t[1]), float(t[2]), t[3]
    t = lines[2].split()
    n_rho, d_rho, n_r, d_r, cutoff = int(t[0]), float(t[1]), int(t[2]), float(t[3]), float(t[4])
    tok = " ".join(lines[3:]).split()
    f = _numbers(tok, n_rho)
    z = _numbers(tok, n_r)
    rho = _numbers(tok, n_r)
    phi = list(z)
    for i in range(1, n_r):
        r = 0.0 + i * d_r
        phi[i] *= phi[i] / r
        phi[i] *= HARTREE_TO_EV * BOHR_TO_ANGS
    phi[0] = phi[1] + (phi[1] - phi[2])
    return dict(kind="eam", name=name[:2], atomic_no=atomic_no, mass=mass * AMU_TO_INTERNAL_MASS, lat=lat, lattice=lattice,
                cutoff=cutoff, phi=make_table(n_r, 0.0, d_r, phi), rho=make_table(n_r, 0.0, d_r, rho),
                f=make_table(n_rho, 0.0, d_rho, f))


def read_setfl(path):
    with open(path) as fh:
        lines = fh.read().split("\n")
    n_elems = int(lines[3].split()[0])
    if n_elems != 1:
        raise ValueError("setfl files with several species are not supported (eam.c:703-704)")
    t = lines[4].split()
    n_rho, d_rho, n_r, d_r, cutoff = int(t[0]), float(t[1]), int(t[2]), float(t[3]), float(t[4])
    t = lines[5].split()
    atomic_no, mass, lat, lattice = int(t[0]), float(t[1]), float(t[2]), t[3]
    tok = " ".join(lines[6:]).split()
    f = _numbers(tok, n_rho)
    rho = _numbers(tok, n_r)
    phi = _numbers(tok, n_r)
    for i in range(1, n_r):
        r = 0.0 + i * d_r
        phi[i] /= r
    phi[0] = phi[1] + (phi[1] - phi[2])
    return dict(kind="eam", name="Cu", atomic_no=atomic_no, mass=mass * AMU_TO_INTERNAL_MASS, lat=lat, lattice=lattice,
                cutoff=cutoff, phi=make_table(n_r, 0.0, d_r, phi), rho=make_table(n_r, 0.0, d_r, rho),
                f=make_table(n_rho, 0.0, d_rho, f))


def load(pot_type, pot_name=None, directory=None):
    """pot_type: 'lj' | 'funcfl' | 'setfl' (mycommand.c:192-196 defaults for the file names)."""
    if pot_type == "lj":
        d = dict(LJ_CU)
        d.update(kind="lj", cutoff=2.5 * d["sigma"])
        return d
    directory = directory or pot_dir()
    if pot_type == "funcfl":
        return read_funcfl(os.path.join(directory, pot_name or "Cu_u6.eam"))
    if pot_type == "setfl":
        return read_setfl(os.path.join(directory, pot_name or "Cu01.eam.alloy"))
    raise ValueError("potential type %r not supported (eam.c:887-894)" % pot_type)


def fcc_block(nxyz, lat, local_min, local_max):
    """createFccLattice (initAtoms.c:81-124) for one rank's brick, vectorised: gid and positions of the
    FCC sites with local_min <= r < local_max, in the reference's loop order (ix, iy, iz, basis)."""
    nx, ny, nz = nxyz
    basis = np.array([[0.25, 0.25, 0.25], [0.25, 0.75, 0.75], [0.75, 0.25, 0.75], [0.75, 0.75, 0.25]])
    b = [int(np.floor(local_min[a] / lat)) for a in range(3)]
    e = [int(np.ceil(local_max[a] / lat)) for a in range(3)]
    ix, iy, iz, ib = np.meshgrid(np.arange(b[0], e[0]), np.arange(b[1], e[1]), np.arange(b[2], e[2]), np.arange(4), indexing="ij")
    ix, iy, iz, ib = ix.ravel(), iy.ravel(), iz.ravel(), ib.ravel()
    r = np.empty((ix.size, 3))
    r[:, 0] = (ix + basis[ib, 0]) * lat
    r[:, 1] = (iy + basis[ib, 1]) * lat
    r[:, 2] = (iz + basis[ib, 2]) * lat
    keep = np.ones(ix.size, bool)
    for a in range(3):
        keep &= (r[:, a] >= local_min[a]) & (r[:, a] < local_max[a])
    gid = (ib + 4 * (iz + nz * (iy + ny * ix))).astype(np.int32)
    return gid[keep], r[keep]


def maxwell_momenta(n, mass, temperature, seed):
    """Synthetic Maxwell-Boltzmann momenta with zero net momentum, rescaled to exactly T (the recipe of
    setTemperature, initAtoms.c:162-198, with a numpy stream instead of the gid-seeded LCG)."""
    rng = np.random.default_rng(seed)
    p = rng.standard_normal((n, 3)) * (mass * np.sqrt(KB_EV * temperature / mass))
    p -= p.mean(0)
    ek = np.sum(p * p) * (0.5 / mass)
    t = ek / n / KB_EV / 1.5
    p *= np.sqrt(temperature / t)
    return p
