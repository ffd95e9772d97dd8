"""Golden vectors of the reference's SINGLE-PRECISION build (real_t = float, mytype.h:11-24), produced by
oracle/_ref/libcomd_ref_f32.so = the reference's own C sources compiled with -DCOMD_SINGLE
(oracle/ref_build/Makefile).  Run where /root/reference exists:   python tests/golden/make_golden_f32.py

One process per case (the reference keeps global state).  Every value is a float widened to double.
  f32_<pot>_n6.npz   displaced lattice (-r 0.1 -T 600, 864 atoms): gid, r, p, f, U, energies at step 0 in the
                     reference's traversal order; its potential (cutoff, mass, lattice constant, the three EAM
                     tables as it stores them); the link-cell geometry it derives for cells of one cutoff and
                     the cell of every atom in it (getBoxFromCoord in float); (ePot, eKin) every 10 of 100 steps
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

POTS = {"lj": [], "funcfl": ["-e"], "setfl": ["-e", "-t", "setfl"]}


def one(kind, n):
    import numpy as np
    import pyoracle as po
    ref = po.Ref32(POTS[kind] + ["-x", n, "-y", n, "-z", n, "-r", 0.1, "-T", 600])
    a = ref.atoms()
    ep, ek = ref.energies()
    pot = ref.potential()
    out = dict(gid=a["gid"], r=a["r"].astype(np.float32), p=a["p"].astype(np.float32), f=a["f"].astype(np.float32),
               U=a["U"].astype(np.float32), ePot=ep, eKin=ek, cutoff=pot["cutoff"], mass=pot["mass"], lat=pot["lat"])
    for k in ("r", "p", "f", "U"):
        assert np.array_equal(out[k].astype(np.float64), a[k]), k      # floats all along
    cells = ref.alt_cells(pot["cutoff"])
    box, tup = ref.alt_bin(a["r"])
    out.update(cell_grid=cells["grid"], cell_box_size=cells["box_size"], cell_inv_box_size=cells["inv_box_size"],
               cell_n_total=cells["n_total_boxes"], cell_of_atom=box)
    if kind != "lj":
        for w, name in enumerate(("phi", "rho", "F")):
            t = ref.table(w)
            out["tab_%s_n" % name] = t["n"]; out["tab_%s_x0" % name] = t["x0"]; out["tab_%s_inv_dx" % name] = t["inv_dx"]
            out["tab_%s_padded" % name] = t["padded"].astype(np.float32)
            assert np.array_equal(out["tab_%s_padded" % name].astype(np.float64), t["padded"])
    tr = [(ep, ek)]
    for _ in range(10):
        ref.timestep(10)
        tr.append(ref.energies())
    out["trace"] = np.array(tr)
    np.savez_compressed(os.path.join(HERE, "f32_%s_n%d.npz" % (kind, n)), **out)


if __name__ == "__main__":
    if len(sys.argv) == 3:
        one(sys.argv[1], int(sys.argv[2]))
    else:
        for kind in POTS:
            subprocess.check_call([sys.executable, __file__, kind, "6"], stdout=subprocess.DEVNULL)
            print("golden f32", kind)
