"""pycomd -- thin ctypes binding of the C ABI in include/comd_b200.h.

This is plumbing for the tests and bench.py; the product is the shared library
(hand-written sm_100a kernels behind a plain C interface) and the C host layer in host/.
There is NO CPU fallback: if the CUDA library is missing or no GPU is present every
device call raises ``ComdError``.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PKG_DIR = os.path.dirname(_HERE)
LIB_PATH = os.path.join(PKG_DIR, "libcomd_b200.so")
# the same sources built with -DCOMD_SINGLE (the reference's DOUBLE_PRECISION=OFF, mytype.h:11-24): comd_real = float
LIB_PATHS = {"f64": LIB_PATH, "f32": os.path.join(PKG_DIR, "libcomd_b200_f32.so")}
REAL = {"f64": np.float64, "f32": np.float32}

_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)

COMD_OK = 0
ERRORS = {-1: "COMD_E_CUDA", -2: "COMD_E_ARG", -3: "COMD_E_CAPACITY", -4: "COMD_E_STATE", -5: "COMD_E_COMM",
          -6: "COMD_E_MISMATCH"}


class ComdError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s (%d): %s" % (ERRORS.get(code, "COMD_E_?"), code, msg))
        self.code = code


class ComdGeometry(C.Structure):
    _fields_ = [("gridSize", C.c_int * 3), ("localMin", C.c_double * 3), ("localMax", C.c_double * 3),
                ("boxSize", C.c_double * 3), ("invBoxSize", C.c_double * 3), ("globalExtent", C.c_double * 3),
                ("procGrid", C.c_int * 3), ("procCoord", C.c_int * 3)]


class ComdTable(C.Structure):
    _fields_ = [("n", C.c_int), ("x0", C.c_double), ("invDx", C.c_double), ("values", _dp)]


# every symbol include/comd_b200.h declares (tests/test_abi.py checks the .so exports them all)
ABI_SYMBOLS = [
    "comd_dev_create", "comd_dev_destroy", "comd_last_error", "comd_version", "comd_dev_set_mass", "comd_dev_set_lj",
    "comd_dev_set_eam", "comd_dev_upload_atoms", "comd_dev_download_atoms", "comd_dev_num_local", "comd_dev_num_total",
    "comd_dev_cell_counts", "comd_dev_download_all", "comd_dev_redistribute", "comd_dev_compute_force",
    "comd_dev_advance_velocity", "comd_dev_advance_position", "comd_dev_energy", "comd_dev_timestep",
    "comd_dev_set_neighbor_list", "comd_dev_neighbor_stats", "comd_comm_unique_id", "comd_comm_init",
    "comd_comm_allreduce_sum", "comd_comm_allreduce_sum_int", "comd_comm_barrier", "comd_dev_sync",
    "comd_dev_kernel_stats", "comd_dev_step_host", "comd_dev_fp64_peak", "comd_dev_timer_start", "comd_dev_timer_stop",
    "comd_dev_set_option", "comd_host_alloc", "comd_host_free", "comd_dev_step_host_stats", "comd_dev_neighbor_counters", "comd_dev_halo_bytes", "comd_precision",
    "comd_hilbert_key", "comd_dev_fp32_peak", "comd_dev_step_stats", "comd_block_order_table",
    "comd_energy_sums_static",
]

_libs = {}


def lib(precision="f64"):
    """Load libcomd_b200.so / libcomd_b200_f32.so (fails loudly when it has not been built)."""
    if precision not in _libs:
        path = LIB_PATHS[precision]
        if not os.path.exists(path):
            raise FileNotFoundError(path + " not built: run `make -C comd-cuda-async_b200` "
                                    "(or python -c 'import __graft_entry__ as g; g.build()')")
        L = C.CDLL(path)
        _rp = _dp if precision == "f64" else _fp        # comd_real*
        if L.comd_precision() != np.dtype(REAL[precision]).itemsize:
            raise RuntimeError(path + ": comd_precision() = %d" % L.comd_precision())
        L.comd_last_error.restype = C.c_char_p
        L.comd_version.restype = C.c_char_p
        L.comd_dev_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(ComdGeometry), C.c_long]
        L.comd_dev_destroy.argtypes = [C.c_void_p]
        L.comd_dev_set_mass.argtypes = [C.c_void_p, C.c_double]
        L.comd_dev_set_lj.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double]
        L.comd_dev_set_eam.argtypes = [C.c_void_p, C.c_double, C.POINTER(ComdTable), C.POINTER(ComdTable), C.POINTER(ComdTable)]
        L.comd_dev_upload_atoms.argtypes = [C.c_void_p, C.c_int, _ip, _ip] + [_rp] * 6
        L.comd_dev_download_atoms.argtypes = [C.c_void_p, _ip, C.c_int, _ip, _ip, _ip] + [_rp] * 10
        L.comd_dev_num_local.argtypes = [C.c_void_p, _ip]
        L.comd_dev_num_total.argtypes = [C.c_void_p, _ip]
        L.comd_dev_cell_counts.argtypes = [C.c_void_p, _ip, C.c_int]
        L.comd_dev_download_all.argtypes = [C.c_void_p, _ip, C.c_int, _ip, _ip, _rp, _rp, _rp]
        L.comd_dev_redistribute.argtypes = [C.c_void_p]
        L.comd_dev_compute_force.argtypes = [C.c_void_p]
        L.comd_dev_advance_velocity.argtypes = [C.c_void_p, C.c_double]
        L.comd_dev_advance_position.argtypes = [C.c_void_p, C.c_double]
        L.comd_dev_energy.argtypes = [C.c_void_p, _dp]
        L.comd_dev_timestep.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.comd_dev_set_neighbor_list.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.comd_dev_neighbor_stats.argtypes = [C.c_void_p, C.POINTER(C.c_long), _ip, _ip]
        L.comd_dev_neighbor_counters.argtypes = [C.c_void_p, C.POINTER(C.c_long), C.POINTER(C.c_long)]
        L.comd_dev_halo_bytes.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_long), C.POINTER(C.c_long)]
        L.comd_comm_unique_id.argtypes = [C.c_char_p]
        L.comd_comm_init.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int]
        L.comd_comm_allreduce_sum.argtypes = [C.c_void_p, _dp, C.c_int]
        L.comd_comm_allreduce_sum_int.argtypes = [C.c_void_p, C.POINTER(C.c_long), C.c_int]
        L.comd_comm_barrier.argtypes = [C.c_void_p]
        L.comd_dev_sync.argtypes = [C.c_void_p]
        L.comd_dev_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_long]
        L.comd_dev_kernel_stats.argtypes = [C.c_void_p, _dp, C.POINTER(C.c_long)]
        L.comd_dev_fp64_peak.argtypes = [C.c_void_p, _dp]
        L.comd_dev_fp32_peak.argtypes = [C.c_void_p, _dp]
        L.comd_hilbert_key.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_int]; L.comd_hilbert_key.restype = C.c_ulonglong
        L.comd_dev_timer_start.argtypes = [C.c_void_p]
        L.comd_dev_timer_stop.argtypes = [C.c_void_p, _dp]
        L.comd_dev_step_host.argtypes = [C.c_void_p, C.c_int, C.c_double, _ip, C.POINTER(_rp), C.POINTER(_rp), C.POINTER(_rp),
                                         _ip, _ip, C.POINTER(_rp), C.POINTER(_rp), C.POINTER(_rp), _rp, _dp]
        L.comd_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        L.comd_host_free.argtypes = [C.c_void_p]
        L.comd_dev_step_host_stats.argtypes = [C.c_void_p, C.POINTER(C.c_long)]
        # device pointers travel as integers (tensor.data_ptr()), the stream as a cudaStream_t value
        L.comd_energy_sums_static.argtypes = [C.c_int, C.c_void_p, C.c_long, C.c_double] + [C.c_void_p] * 5
        _libs[precision] = L
    return _libs[precision]


class PinnedArray:
    """A numpy view of page-locked host memory (comd_host_alloc): step_host copies such buffers
    without staging.  Keep the object alive as long as the array is in use."""

    def __init__(self, n, dtype=np.float64):
        self.nbytes = int(n) * np.dtype(dtype).itemsize
        self.ptr = C.c_void_p()
        _chk(lib().comd_host_alloc(C.byref(self.ptr), max(self.nbytes, 8)))
        buf = (C.c_char * max(self.nbytes, 8)).from_address(self.ptr.value)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(n))

    def free(self):
        if self.ptr:
            self.array = None
            lib().comd_host_free(self.ptr)
            self.ptr = None


def pinned_like(a):
    """Copy a numpy array into pinned host memory; returns (PinnedArray, view)."""
    pa = PinnedArray(a.size, a.dtype)
    pa.array[:] = a.ravel()
    return pa, pa.array


def _chk(rc, L=None):
    if rc != COMD_OK:
        raise ComdError(rc, (L or lib()).comd_last_error().decode())


def _d(a):
    if a is None:
        return None
    return a.ctypes.data_as(_fp if a.dtype == np.float32 else _dp)


def _i(a):
    return a.ctypes.data_as(_ip) if a is not None else None


def make_geometry(global_extent, cell_cutoff, proc_grid=(1, 1, 1), proc_coord=(0, 0, 0), precision="f64"):
    """Host arithmetic of initDecomposition (decomposition.c:18-50) + initLinkCells
    (linkCells.c:122-182), in IEEE doubles exactly as the C does it -- or, for the single-precision
    library, in floats as the reference's real_t = float build does it (the struct then carries those
    float values, which the library's own rounding to comd_real leaves untouched)."""
    T = REAL[precision]
    g = ComdGeometry()
    cut = T(cell_cutoff)
    for a in range(3):
        gext = T(global_extent[a])
        lext = T(gext / T(proc_grid[a]))
        g.globalExtent[a] = float(gext)
        g.procGrid[a] = proc_grid[a]
        g.procCoord[a] = proc_coord[a]
        g.localMin[a] = float(T(T(0.0) + T(T(proc_coord[a]) * lext)))
        g.localMax[a] = float(T(T(0.0) + T(T(proc_coord[a] + 1) * lext)))
        g.gridSize[a] = int(T(lext / cut))
        box = T(lext / T(g.gridSize[a]))
        g.boxSize[a] = float(box)
        g.invBoxSize[a] = float(T(1.0 / float(box)))
    return g


def n_total_boxes(grid):
    nx, ny, nz = grid
    return nx * ny * nz + 2 * ((nx + 2) * (ny + nz + 2) + ny * nz)


class Device:
    """One rank's device context (ComdDev*)."""

    def __init__(self, geom, capacity, device_id=0, precision="f64"):
        self.h = C.c_void_p()
        self.geom = geom
        self.capacity = int(capacity)
        self.precision = precision
        self.real = REAL[precision]                     # numpy type of comd_real
        self.L = lib(precision)
        self._chk(self.L.comd_dev_create(C.byref(self.h), device_id, C.byref(geom), int(capacity)))
        self._keep = []

    def _chk(self, rc):
        _chk(rc, self.L)

    def close(self):
        if self.h:
            self.L.comd_dev_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- potential
    def set_mass(self, m):
        self._chk(self.L.comd_dev_set_mass(self.h, m))

    def set_lj(self, sigma=2.315, epsilon=0.167, cutoff=2.5 * 2.315):
        self._chk(self.L.comd_dev_set_lj(self.h, sigma, epsilon, cutoff))

    def set_eam(self, cutoff, phi, rho, f):
        """phi/rho/f: dict(n, x0, inv_dx, padded) with padded = values[-1..n+1]."""
        tabs = []
        for t in (phi, rho, f):
            pad = np.ascontiguousarray(t["padded"], np.float64)
            self._keep.append(pad)
            ct = ComdTable(t["n"], t["x0"], t["inv_dx"], C.cast(pad.ctypes.data + 8, _dp))
            tabs.append(ct)
        self._chk(self.L.comd_dev_set_eam(self.h, cutoff, C.byref(tabs[0]), C.byref(tabs[1]), C.byref(tabs[2])))

    # --- atoms
    def upload(self, gid, r, p=None, species=None):
        gid = np.ascontiguousarray(gid, np.int32)
        rs = [np.ascontiguousarray(r[:, k], self.real) for k in range(3)]
        ps = [np.ascontiguousarray(p[:, k], self.real) for k in range(3)] if p is not None else [None] * 3
        sp = np.ascontiguousarray(species, np.int32) if species is not None else None
        self._chk(self.L.comd_dev_upload_atoms(self.h, len(gid), _i(gid), _i(sp), _d(rs[0]), _d(rs[1]), _d(rs[2]),
                                         _d(ps[0]), _d(ps[1]), _d(ps[2])))

    def num_local(self):
        n = C.c_int()
        self._chk(self.L.comd_dev_num_local(self.h, C.byref(n)))
        return n.value

    def num_total(self):
        n = C.c_int()
        self._chk(self.L.comd_dev_num_total(self.h, C.byref(n)))
        return n.value

    def download(self):
        cap = self.capacity
        n = C.c_int()
        gid = np.zeros(cap, np.int32); sp = np.zeros(cap, np.int32); box = np.zeros(cap, np.int32)
        arr = [np.zeros(cap, self.real) for _ in range(10)]
        self._chk(self.L.comd_dev_download_atoms(self.h, C.byref(n), cap, _i(gid), _i(sp), _i(box), *[_d(a) for a in arr]))
        m = n.value
        return dict(gid=gid[:m], species=sp[:m], box=box[:m],
                    r=np.stack([a[:m] for a in arr[0:3]], 1), p=np.stack([a[:m] for a in arr[3:6]], 1),
                    f=np.stack([a[:m] for a in arr[6:9]], 1), U=arr[9][:m].copy())

    def download_all(self):
        cap = self.capacity
        n = C.c_int()
        gid = np.zeros(cap, np.int32); box = np.zeros(cap, np.int32)
        arr = [np.zeros(cap, self.real) for _ in range(3)]
        self._chk(self.L.comd_dev_download_all(self.h, C.byref(n), cap, _i(gid), _i(box), *[_d(a) for a in arr]))
        m = n.value
        return dict(gid=gid[:m], box=box[:m], r=np.stack([a[:m] for a in arr], 1))

    def cell_counts(self):
        g = self.geom.gridSize
        nt = n_total_boxes((g[0], g[1], g[2]))
        out = np.zeros(nt, np.int32)
        self._chk(self.L.comd_dev_cell_counts(self.h, _i(out), nt))
        return out

    # --- the path
    def redistribute(self):
        self._chk(self.L.comd_dev_redistribute(self.h))

    def compute_force(self):
        self._chk(self.L.comd_dev_compute_force(self.h))

    def advance_velocity(self, dt):
        self._chk(self.L.comd_dev_advance_velocity(self.h, dt))

    def advance_position(self, dt):
        self._chk(self.L.comd_dev_advance_position(self.h, dt))

    def energy(self):
        e = np.zeros(2)
        self._chk(self.L.comd_dev_energy(self.h, _d(e)))
        return float(e[0]), float(e[1])

    def timestep(self, n, dt):
        self._chk(self.L.comd_dev_timestep(self.h, n, dt))

    def sync(self):
        self._chk(self.L.comd_dev_sync(self.h))

    def set_option(self, key, value):
        self._chk(self.L.comd_dev_set_option(self.h, key.encode(), int(value)))

    def set_neighbor_list(self, use, skin):
        self._chk(self.L.comd_dev_set_neighbor_list(self.h, 1 if use else 0, skin))

    def neighbor_stats(self):
        a = C.c_long(); b = C.c_int(); c = C.c_int()
        self._chk(self.L.comd_dev_neighbor_stats(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(pairs=a.value, max_per_atom=b.value, builds=c.value)

    def neighbor_counters(self):
        a = C.c_long(); b = C.c_long()
        self._chk(self.L.comd_dev_neighbor_counters(self.h, C.byref(a), C.byref(b)))
        return dict(builds=a.value, refreshes=b.value)

    def halo_bytes(self, kind, face):
        a = C.c_long(); b = C.c_long()
        self._chk(self.L.comd_dev_halo_bytes(self.h, kind, face, C.byref(a), C.byref(b)))
        return a.value, b.value

    def kernel_stats(self):
        ms = C.c_double(); n = C.c_long()
        self._chk(self.L.comd_dev_kernel_stats(self.h, C.byref(ms), C.byref(n)))
        return dict(last_force_ms=ms.value, launches=n.value)

    def timer_start(self):
        self._chk(self.L.comd_dev_timer_start(self.h))

    def timer_stop(self):
        t = C.c_double()
        self._chk(self.L.comd_dev_timer_stop(self.h, C.byref(t)))
        return t.value

    def fp64_peak(self):
        t = C.c_double()
        self._chk(self.L.comd_dev_fp64_peak(self.h, C.byref(t)))
        return t.value

    def step_stats(self):
        a = C.c_long(); b = C.c_long()
        self._chk(self.L.comd_dev_step_stats(self.h, C.byref(a), C.byref(b)))
        return dict(graph_steps=a.value, redone_steps=b.value)

    def fp32_peak(self):
        t = C.c_double()
        self._chk(self.L.comd_dev_fp32_peak(self.h, C.byref(t)))
        return t.value

    def step_host(self, dt, gid, r, p, f):
        """One timestep through host buffers.  r,p,f: lists of three contiguous arrays of self.real."""
        n = len(gid)
        cap = self.capacity
        P3 = (_dp if self.precision == "f64" else _fp) * 3
        rin = P3(*[_d(a) for a in r]); pin = P3(*[_d(a) for a in p]); fin = P3(*[_d(a) for a in f])
        if not hasattr(self, "_out"):
            self._pin = [PinnedArray(cap, np.int32)] + [PinnedArray(cap, self.real) for _ in range(10)]
            v = [q.array for q in self._pin]
            self._out = dict(gid=v[0], r=v[1:4], p=v[4:7], f=v[7:10], U=v[10])
        o = self._out
        rout = P3(*[_d(a) for a in o["r"]]); pout = P3(*[_d(a) for a in o["p"]]); fout = P3(*[_d(a) for a in o["f"]])
        e = np.zeros(2)
        m = C.c_int()
        self._chk(self.L.comd_dev_step_host(self.h, n, dt, _i(gid), rin, pin, fin, C.byref(m), _i(o["gid"]), rout, pout, fout,
                                      _d(o["U"]), _d(e)))
        return m.value, o, (float(e[0]), float(e[1]))

    def step_host_resumed(self):
        """step_host calls served from the previous call's sort and neighbour list."""
        n = C.c_long()
        self._chk(self.L.comd_dev_step_host_stats(self.h, C.byref(n)))
        return n.value

    # --- multi-GPU
    @staticmethod
    def comm_unique_id():
        buf = C.create_string_buffer(128)
        _chk(lib().comd_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, uid, rank, nranks):
        self._chk(self.L.comd_comm_init(self.h, uid, rank, nranks))

    def allreduce_sum(self, vals):
        a = np.ascontiguousarray(vals, np.float64).copy()
        self._chk(self.L.comd_comm_allreduce_sum(self.h, _d(a), len(a)))
        return a
