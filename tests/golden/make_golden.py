"""Generate the committed golden fixtures from the REFERENCE ITSELF (oracle/_ref, i.e. the
reference's own C sources compiled by oracle/ref_build/Makefile).  Run in the container that
has /root/reference:   python tests/golden/make_golden.py

Each (potential, size) case runs in its own process because the reference keeps global state.
Fixtures:
  forces_<pot>_n6.npz   per-atom gid, r, p, f, U and energies at step 0 of a DISPLACED lattice
                        (-r 0.1 -T 600; 864 atoms), in the reference's own traversal order
  trace_<pot>_n<k>.npz  (ePot, eKin) every 10 steps over 100 steps of the BASELINE configs
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

POTS = {"lj": [], "funcfl": ["-e"], "setfl": ["-e", "-t", "setfl"]}


def one(kind, what, n):
    import numpy as np
    import pyoracle as po
    if what == "forces":
        ref = po.Ref(POTS[kind] + ["-x", n, "-y", n, "-z", n, "-r", 0.1, "-T", 600])
        a = ref.atoms()
        ep, ek = ref.energies()
        np.savez_compressed(os.path.join(HERE, "forces_%s_n%d.npz" % (kind, n)), gid=a["gid"], box=a["box"], r=a["r"],
                            p=a["p"], f=a["f"], U=a["U"], ePot=ep, eKin=ek, cutoff=ref.potential()["cutoff"])
    else:
        ref = po.Ref(POTS[kind] + ["-x", n, "-y", n, "-z", n])
        tr = [ref.energies()]
        for _ in range(10):
            ref.timestep(10)
            tr.append(ref.energies())
        np.savez_compressed(os.path.join(HERE, "trace_%s_n%d.npz" % (kind, n)), trace=np.array(tr), nAtoms=4 * n ** 3)


if __name__ == "__main__":
    if len(sys.argv) == 4:
        one(sys.argv[1], sys.argv[2], int(sys.argv[3]))
    else:
        for kind in POTS:
            for what, n in (("forces", 6), ("trace", 20)):
                subprocess.check_call([sys.executable, __file__, kind, what, str(n)], stdout=subprocess.DEVNULL)
                print("golden", kind, what, n)
