"""Tiny single-precision run (LJ and EAM, link-cell and list paths) for compute-sanitizer."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("comd-cuda-async_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import numpy as np
import test_gpu_precision as T

prec = sys.argv[1] if len(sys.argv) > 1 else "f32"
for kind in ("lj", "setfl"):
    for path in ("link_cell", "list_n3", "list_full"):
        g = T.golden(kind, prec)
        dev, _ = T.device_from(g, kind, prec, path)
        dev.compute_force()
        dev.timestep(3, 1.0)
        d = T.by_gid(dev.download())
        o = T.by_gid({k: g[k] for k in ("gid", "r", "p", "f", "U")})
        print(kind, path, "E", dev.energy(), flush=True)
        dev.close()
