"""Small driver for profiling: build a lattice, then call the force path K times.
usage: python tools/prof_force.py <lj|setfl|funcfl> <unit cells per axis> [K] [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "comd-cuda-async_b200"))
import numpy as np
import pycomd
from pycomd import potentials as P

pot = sys.argv[1]; n = int(sys.argv[2]); K = int(sys.argv[3]) if len(sys.argv) > 3 else 5
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 0
potential = P.load(pot)
lat = potential["lat"]
NL = int(os.environ.get("COMD_NL", "0"))
skin = 0.1 * potential["cutoff"] if NL else 0.0
PREC = os.environ.get("COMD_PREC", "f64")            # f32: the -DCOMD_SINGLE build (libcomd_b200_f32.so)
T = pycomd.REAL[PREC]
geom = pycomd.make_geometry([T(n) * T(lat)] * 3, T(T(potential["cutoff"]) + T(skin)), precision=PREC)
gid, r = P.fcc_block([n] * 3, lat, [0.0] * 3, [n * lat] * 3)
p = P.maxwell_momenta(len(gid), potential["mass"], 600.0, seed=7)
dev = pycomd.Device(geom, int(len(gid) * 1.6) + 65536, precision=PREC)
dev.set_mass(potential["mass"])
if pot == "lj":
    dev.set_lj(potential["sigma"], potential["epsilon"], potential["cutoff"])
else:
    dev.set_eam(potential["cutoff"], potential["phi"], potential["rho"], potential["f"])
if NL:
    dev.set_neighbor_list(True, skin)
for kv in filter(None, os.environ.get("COMD_OPTS", "").split(",")):
    k, v = kv.split("="); dev.set_option(k, int(v))
dev.upload(gid, r, p)
dev.compute_force()
dev.timestep(3, 1.0)          # thermalise a little: realistic hit patterns
dev.sync(); dev.kernel_stats()
for _ in range(K):
    dev.compute_force()
dev.sync()
ms = dev.kernel_stats()["last_force_ms"]
e = dev.energy()
print("%s n=%d atoms=%d force %.4f ms  (%.1f ps/atom)  E/atom %.9f" % (pot, n, len(gid), ms, ms * 1e9 / len(gid), (e[0] + e[1]) / len(gid)))
if steps:
    dev.kernel_stats()
    dev.timer_start(); dev.timestep(steps, 1.0); t = dev.timer_stop()
    print("step %.4f ms  force inside timestep %.4f ms  opts=%s" % (t / steps, dev.kernel_stats()["last_force_ms"], os.environ.get("COMD_OPTS", "")))
if NL:
    print("neighbour list:", dev.neighbor_stats())
dev.close()
