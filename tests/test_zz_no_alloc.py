"""comd_energy_sums_static: the energy sums for callers that cannot allocate device memory.

Who that is: code called from inside a running time-step loop (cudaMalloc / cudaFree wait for the whole GPU, so an
allocation there stalls the halo/compute overlap), and hosts whose own arena or caching allocator already holds all
of HBM, under which a library's cudaMalloc may fail.  The entry point takes device pointers the caller owns and keeps
its reduction scratch in a fixed __device__ array inside the library (include/comd_b200.h).

The file is named so that pytest collects it last: with `-x`, a failure here cannot hide another test.
The GPU test stands in for such a caller with torch tensors (torch is only the owner of the device memory here)
and checks the result against comd_dev_energy() -- the existing, allocating entry point -- twice in a row, the
second call on FEWER atoms than the first, so that partials left in the shared scratch by the first call would
show up if the second one read them.
"""
import ctypes as C

import numpy as np
import pytest

import pycomd
import pyoracle as po
from common import device_for, oracle_sim

MAX_ATOMS = 67108864            # COMD_ENERGY_STATIC_MAX_ATOMS (include/comd_b200.h)
REL_TOL = 1e-12                 # the parity suite's bound for FP64 sums that differ in summation order only


def test_size_limit_is_declared_and_enforced_before_any_device_work():
    """Runs without a GPU: a count beyond the scratch array's capacity is refused on argument checking alone,
    with a message that names the limit; so are null arrays and a non-positive mass."""
    import os
    import re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "comd_b200.h")).read()
    assert int(re.search(r"#define\s+COMD_ENERGY_STATIC_MAX_ATOMS\s+(\d+)L", hdr).group(1)) == MAX_ATOMS
    for precision in ("f64", "f32"):
        L = pycomd.lib(precision)
        fake = 256                                      # never dereferenced: the call must stop at the checks
        rc = L.comd_energy_sums_static(0, None, MAX_ATOMS + 1, 63.55, fake, fake, fake, fake, fake)
        assert rc == -2
        msg = L.comd_last_error().decode()
        assert str(MAX_ATOMS) in msg and "exceed" in msg, msg
        assert L.comd_energy_sums_static(0, None, 10, 63.55, None, fake, fake, fake, fake) == -2
        assert L.comd_energy_sums_static(0, None, 10, 63.55, fake, fake, fake, fake, None) == -2
        assert L.comd_energy_sums_static(0, None, 10, 0.0, fake, fake, fake, fake, fake) == -2
        assert L.comd_energy_sums_static(0, None, -1, 63.55, fake, fake, fake, fake, fake) == -2


def _energies_both_ways(precision, n_cells, steps):
    """(existing entry point's sums, new entry point's sums, atom count) for an EAM crystal of n_cells^3 unit cells
    after `steps` time steps.  The atoms' U and p are read back through the C ABI and handed to the new entry point
    as device memory owned by torch, on torch's current stream."""
    import torch
    orc = oracle_sim(po.POT_FUNCFL, n_cells, delta=0.05, temperature=600.0)
    a = orc.atoms()
    pot = orc.potential()
    geom = pycomd.make_geometry([n_cells * pot["lat"]] * 3, pot["cutoff"], precision=precision)
    dev = pycomd.Device(geom, 4 * n_cells ** 3 * 3 + 16384, precision=precision)
    try:
        dev.set_mass(pot["mass"])
        dev.set_eam(pot["cutoff"], orc.table(0), orc.table(1), orc.table(2))
        dev.upload(a["gid"], a["r"], a["p"])
        dev.compute_force()
        if steps:
            dev.timestep(steps, 1.0)
        want = dev.energy()                              # comd_dev_energy: per-handle cudaMalloc'ed partials
        d = dev.download()                               # local atoms only, like the sum above
    finally:
        dev.close()
    real = np.float64 if precision == "f64" else np.float32
    assert d["U"].dtype == real and d["p"].dtype == real
    cu = torch.device("cuda:0")
    U = torch.from_numpy(np.ascontiguousarray(d["U"])).to(cu)
    p = [torch.from_numpy(np.ascontiguousarray(d["p"][:, k])).to(cu) for k in range(3)]
    out = torch.full((2,), float("nan"), dtype=torch.float64, device=cu)
    L = pycomd.lib(precision)
    stream = torch.cuda.current_stream(cu).cuda_stream
    rc = L.comd_energy_sums_static(0, C.c_void_p(stream), len(d["gid"]), pot["mass"],
                                   U.data_ptr(), p[0].data_ptr(), p[1].data_ptr(), p[2].data_ptr(), out.data_ptr())
    assert rc == 0, L.comd_last_error().decode()
    torch.cuda.synchronize(cu)
    got = out.cpu().numpy()
    return want, (float(got[0]), float(got[1])), len(d["gid"])


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_static_scratch_energy_matches_the_allocating_entry_point_twice_in_a_row(precision):
    """Call 1: 6912 atoms (two 4096-atom tiles, the second one ragged) at step 0.  Call 2, straight after and through
    the same scratch array: 864 atoms (one tile) after three steps.  Each must agree with comd_dev_energy() on the
    same atoms to 1e-12 relative -- both sum the same per-atom terms in double; only the order differs."""
    for n_cells, steps in ((12, 0), (6, 3)):
        want, got, n = _energies_both_ways(precision, n_cells, steps)
        assert n == 4 * n_cells ** 3
        for name, w, g in (("potential", want[0], got[0]), ("kinetic", want[1], got[1])):
            assert np.isfinite(g), (precision, n_cells, name, g)
            assert abs(g - w) <= REL_TOL * abs(w), (precision, n_cells, name, g, w)
        assert want[1] > 0.0 and want[0] < 0.0           # a real crystal at 600 K, not two zeros agreeing


@pytest.mark.gpu
def test_static_scratch_energy_of_no_atoms_is_zero():
    """nAtoms = 0 is legal (a rank may own no atoms): both sums come back as exact zeros, whatever the scratch holds."""
    import torch
    cu = torch.device("cuda:0")
    out = torch.full((2,), float("nan"), dtype=torch.float64, device=cu)
    L = pycomd.lib("f64")
    rc = L.comd_energy_sums_static(0, C.c_void_p(torch.cuda.current_stream(cu).cuda_stream), 0, 63.55,
                                   None, None, None, None, out.data_ptr())
    assert rc == 0, L.comd_last_error().decode()
    torch.cuda.synchronize(cu)
    assert out.cpu().tolist() == [0.0, 0.0]
