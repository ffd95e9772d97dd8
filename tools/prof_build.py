"""Time the neighbour-list build alone: force evaluation with a forced rebuild minus one without.
usage: python tools/prof_build.py <lj|setfl|funcfl> <unit cells per axis> [K] [path of an alternative libcomd_b200.so]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "comd-cuda-async_b200"))
import numpy as np
import pycomd
from pycomd import potentials as P

pot = sys.argv[1]; n = int(sys.argv[2]); K = int(sys.argv[3]) if len(sys.argv) > 3 else 5
if len(sys.argv) > 4:
    pycomd.LIB_PATH = os.path.abspath(sys.argv[4])
potential = P.load(pot)
lat = potential["lat"]
skin = 0.1 * potential["cutoff"]
geom = pycomd.make_geometry([n * lat] * 3, potential["cutoff"] + skin)
gid, r = P.fcc_block([n] * 3, lat, [0.0] * 3, [n * lat] * 3)
p = P.maxwell_momenta(len(gid), potential["mass"], 600.0, seed=7)
dev = pycomd.Device(geom, int(len(gid) * 1.6) + 65536)
dev.set_mass(potential["mass"])
if pot == "lj":
    dev.set_lj(potential["sigma"], potential["epsilon"], potential["cutoff"])
else:
    dev.set_eam(potential["cutoff"], potential["phi"], potential["rho"], potential["f"])
dev.set_neighbor_list(True, skin)
dev.upload(gid, r, p)
dev.compute_force()
dev.timestep(5, 1.0)
dev.sync()
tf, tb = [], []
for _ in range(K):
    dev.timer_start(); dev.compute_force(); tf.append(dev.timer_stop())
    dev.set_neighbor_list(True, skin)
    dev.timer_start(); dev.compute_force(); tb.append(dev.timer_stop())
tf.sort(); tb.sort()
e = dev.energy()
print("%s n=%d atoms=%d lib=%s: force %.3f ms, force+build %.3f ms, build %.3f ms (%.2f ns/atom)  E/atom %.12f  list %s" % (
    pot, n, len(gid), os.path.basename(os.path.dirname(pycomd.LIB_PATH)) or ".", tf[K // 2], tb[K // 2], tb[K // 2] - tf[K // 2],
    (tb[K // 2] - tf[K // 2]) * 1e6 / len(gid), (e[0] + e[1]) / len(gid), dev.neighbor_stats()))
dev.close()
