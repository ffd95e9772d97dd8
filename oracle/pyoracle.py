"""ctypes bindings for the CPU oracle (TEST INFRASTRUCTURE ONLY).

Two checkers live here:

* ``Oracle``  -- oracle/libcomd_oracle.so, the plain-C restatement of the reference's
  CPU path (oracle/comd_oracle.c).
* ``Ref``     -- oracle/_ref/libcomd_ref.so, the reference's own C sources compiled
  where they lie (oracle/ref_build/Makefile).  It is what pins ``Oracle``.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libcomd_oracle.so")
REF_DIR = os.path.join(HERE, "_ref")
REF_SO = os.path.join(REF_DIR, "libcomd_ref.so")
REF_SO_F32 = os.path.join(REF_DIR, "libcomd_ref_f32.so")   # the reference with real_t = float (DOUBLE_PRECISION=OFF)
REF_EXE = os.path.join(REF_DIR, "CoMD-ref")
REF_SRC = "/root/reference"

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _i(a):
    return a.ctypes.data_as(_ip) if a is not None else None


def pot_dir():
    """Directory holding Cu_u6.eam / Cu01.eam.alloy (run-time input data)."""
    for d in (os.path.join(HERE, "..", "pots"), os.path.join(REF_DIR, "pots"), os.path.join(REF_SRC, "pots")):
        if os.path.exists(os.path.join(d, "Cu_u6.eam")):
            return os.path.abspath(d)
    raise FileNotFoundError("EAM potential tables not found: run `make -C oracle ref` where /root/reference exists")


def build(ref=True):
    """Compile the restatement; and, where /root/reference exists, the reference itself."""
    subprocess.check_call(["make", "-s", "-C", HERE])
    if ref and os.path.isdir(os.path.join(REF_SRC, "src-mpi")):
        subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "ref_build")])


def have_ref(precision="f64"):
    return os.path.exists(REF_SO if precision == "f64" else REF_SO_F32)


POT_LJ, POT_FUNCFL, POT_SETFL = 0, 1, 2


def pot_args(kind):
    if kind == POT_LJ:
        return None
    name = "Cu_u6.eam" if kind == POT_FUNCFL else "Cu01.eam.alloy"
    return os.path.join(pot_dir(), name).encode()


class Oracle:
    """One single-rank periodic CoMD simulation on the CPU restatement."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            if not os.path.exists(ORACLE_SO):
                build(ref=False)
            L = C.CDLL(ORACLE_SO)
            L.orc_create.restype = C.c_void_p
            L.orc_create.argtypes = [C.c_int, C.c_char_p, C.c_double, C.c_int, C.c_int, C.c_int, C.c_double,
                                     C.c_double, C.c_double, C.c_double, C.c_double, C.c_int]
            L.orc_destroy.argtypes = [C.c_void_p]
            L.orc_timestep.argtypes = [C.c_void_p, C.c_int]
            L.orc_set_atoms.argtypes = [C.c_void_p, C.c_int, _ip, _dp, _dp]
            L.orc_n_local.argtypes = [C.c_void_p]
            L.orc_n_builds.argtypes = [C.c_void_p]
            L.orc_energies.argtypes = [C.c_void_p, _dp, _dp]
            L.orc_potential.argtypes = [C.c_void_p, _dp, _dp, _dp]
            L.orc_geometry.argtypes = [C.c_void_p, _ip, _dp]
            L.orc_get_atoms.argtypes = [C.c_void_p, _ip, _ip, _dp, _dp, _dp, _dp]
            L.orc_get_all_atoms.argtypes = [C.c_void_p, _ip, _ip, _dp]
            L.orc_interpolate.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp, _dp]
            L.orc_table_n.argtypes = [C.c_void_p, C.c_int]
            L.orc_table.argtypes = [C.c_void_p, C.c_int, _dp, _dp, _dp]
            L.orc_cells_only.restype = C.c_void_p
            L.orc_cells_only.argtypes = [_dp, C.c_double]
            L.orc_bin.argtypes = [C.c_void_p, C.c_int, _dp, _ip, _ip]
            L.orc_box_from_tuple.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
            L.orc_tuple_from_box.argtypes = [C.c_void_p, C.c_int, _ip]
            L.orc_mk_seed.restype = C.c_uint64
            L.orc_mk_seed.argtypes = [C.c_uint32, C.c_uint32]
            L.orc_lcg61.restype = C.c_double
            L.orc_lcg61.argtypes = [C.POINTER(C.c_uint64)]
            L.orc_gasdev.restype = C.c_double
            L.orc_gasdev.argtypes = [C.POINTER(C.c_uint64)]
            cls._lib = L
        return cls._lib

    def __init__(self, pot=POT_LJ, nx=20, ny=20, nz=20, lat=-1.0, temperature=600.0, delta=0.0,
                 skin=0.1, dt=1.0, lj_cut_mult=2.5, initial_force=True):
        L = self.lib()
        self.h = L.orc_create(pot, pot_args(pot), lj_cut_mult, nx, ny, nz, lat, temperature, delta, skin, dt,
                              1 if initial_force else 0)
        if not self.h:
            raise RuntimeError("orc_create failed (potential file?)")
        self.pot = pot

    def close(self):
        if self.h:
            self.lib().orc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n_local(self):
        return self.lib().orc_n_local(self.h)

    @property
    def n_builds(self):
        return self.lib().orc_n_builds(self.h)

    def timestep(self, n):
        self.lib().orc_timestep(self.h, n)

    def energies(self):
        a, b = C.c_double(), C.c_double()
        self.lib().orc_energies(self.h, C.byref(a), C.byref(b))
        return a.value, b.value

    def potential(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        self.lib().orc_potential(self.h, C.byref(a), C.byref(b), C.byref(c))
        return dict(cutoff=a.value, mass=b.value, lat=c.value)

    def geometry(self):
        g = np.zeros(5, np.int32)
        d = np.zeros(12, np.float64)
        self.lib().orc_geometry(self.h, _i(g), _d(d))
        return dict(grid=g[:3].copy(), n_local_boxes=int(g[3]), n_total_boxes=int(g[4]), local_min=d[0:3].copy(),
                    local_max=d[3:6].copy(), box_size=d[6:9].copy(), inv_box_size=d[9:12].copy())

    def atoms(self):
        n = self.n_local
        gid = np.zeros(n, np.int32); box = np.zeros(n, np.int32)
        r = np.zeros((n, 3)); p = np.zeros((n, 3)); f = np.zeros((n, 3)); U = np.zeros(n)
        m = self.lib().orc_get_atoms(self.h, _i(gid), _i(box), _d(r), _d(p), _d(f), _d(U))
        assert m == n
        return dict(gid=gid, box=box, r=r, p=p, f=f, U=U)

    def all_atoms(self):
        n = self.lib().orc_get_all_atoms(self.h, None, None, None)
        gid = np.zeros(n, np.int32); box = np.zeros(n, np.int32); r = np.zeros((n, 3))
        self.lib().orc_get_all_atoms(self.h, _i(gid), _i(box), _d(r))
        return dict(gid=gid, box=box, r=r)

    def set_atoms(self, gid, r, p=None):
        gid = np.ascontiguousarray(gid, np.int32)
        r = np.ascontiguousarray(r, np.float64)
        p = np.ascontiguousarray(p, np.float64) if p is not None else None
        self.lib().orc_set_atoms(self.h, len(gid), _i(gid), _d(r), _d(p))

    def interpolate(self, which, x):
        x = np.ascontiguousarray(x, np.float64)
        f = np.zeros_like(x); df = np.zeros_like(x)
        self.lib().orc_interpolate(self.h, which, len(x), _d(x), _d(f), _d(df))
        return f, df

    def table(self, which):
        n = self.lib().orc_table_n(self.h, which)
        v = np.zeros(n + 3)
        a, b = C.c_double(), C.c_double()
        self.lib().orc_table(self.h, which, C.byref(a), C.byref(b), _d(v))
        return dict(n=n, x0=a.value, inv_dx=b.value, padded=v)


class OracleCells:
    """The reference's ownership rule on an explicit cell-size cutoff (single rank)."""

    def __init__(self, global_extent, cell_cutoff):
        L = Oracle.lib()
        ge = np.ascontiguousarray(global_extent, np.float64)
        self.h = L.orc_cells_only(_d(ge), cell_cutoff)
        g = np.zeros(5, np.int32); d = np.zeros(12)
        L.orc_geometry(self.h, _i(g), _d(d))
        self.grid = g[:3].copy(); self.n_local_boxes = int(g[3]); self.n_total_boxes = int(g[4])
        self.box_size = d[6:9].copy(); self.inv_box_size = d[9:12].copy()

    def bin(self, r):
        r = np.ascontiguousarray(r, np.float64)
        n = len(r)
        box = np.zeros(n, np.int32); tup = np.zeros((n, 3), np.int32)
        Oracle.lib().orc_bin(self.h, n, _d(r), _i(box), _i(tup))
        return box, tup

    def box_from_tuple(self, ix, iy, iz):
        return Oracle.lib().orc_box_from_tuple(self.h, ix, iy, iz)

    def tuple_from_box(self, b):
        t = np.zeros(3, np.int32)
        Oracle.lib().orc_tuple_from_box(self.h, b, _i(t))
        return t


class Ref:
    """The reference's own CPU path (one simulation per process: it keeps global state)."""

    _lib = None
    SO = REF_SO

    @classmethod
    def lib(cls):
        if cls._lib is None:
            if not os.path.exists(cls.SO):
                raise FileNotFoundError(cls.SO + " missing: run `make -C oracle ref` where /root/reference exists")
            L = C.CDLL(cls.SO)
            L.ref_init.argtypes = [C.c_int, C.POINTER(C.c_char_p)]
            L.ref_set_lj_cutoff_mult.argtypes = [C.c_double]
            L.ref_geometry.argtypes = [_ip, _dp]
            L.ref_potential.argtypes = [_dp, _dp, _dp]
            L.ref_pot_handle.restype = C.c_void_p
            L.ref_get_atoms.argtypes = [_ip, _ip, _dp, _dp, _dp, _dp]
            L.ref_get_all_atoms.argtypes = [_ip, _ip, _dp]
            L.ref_timestep.argtypes = [C.c_int]
            L.ref_energies.argtypes = [_dp, _dp]
            L.ref_alt_linkcells.argtypes = [C.c_double]
            L.ref_alt_geometry.argtypes = [_ip, _dp]
            L.ref_alt_box_from_coord.argtypes = [C.c_int, _dp, _ip, _ip]
            L.ref_alt_box_from_tuple.argtypes = [C.c_int, C.c_int, C.c_int]
            L.ref_eam_interpolate.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp, _dp]
            L.ref_eam_table_n.argtypes = [C.c_void_p, C.c_int]
            L.ref_eam_table.argtypes = [C.c_void_p, C.c_int, _dp, _dp, _dp]
            L.ref_mk_seed.restype = C.c_uint64
            L.ref_mk_seed.argtypes = [C.c_uint32, C.c_uint32]
            L.ref_lcg61.restype = C.c_double
            L.ref_lcg61.argtypes = [C.POINTER(C.c_uint64)]
            L.ref_gasdev.restype = C.c_double
            L.ref_gasdev.argtypes = [C.POINTER(C.c_uint64)]
            cls._lib = L
        return cls._lib

    def __init__(self, args, lj_cut_mult=2.5):
        """args: CoMD command line without argv[0], e.g. ['-e', '-x', '10', ...]; -m cpu_nl is added."""
        L = self.lib()
        L.ref_set_lj_cutoff_mult(lj_cut_mult)
        argv = [b"CoMD-ref", b"-m", b"cpu_nl", b"-d", pot_dir().encode()] + [str(a).encode() for a in args]
        arr = (C.c_char_p * (len(argv) + 1))(*argv, None)
        cwd = os.getcwd()
        os.chdir(REF_DIR)           # the reference writes its yaml report into the cwd
        try:
            self.n = L.ref_init(len(argv), arr)
        finally:
            os.chdir(cwd)

    def geometry(self):
        g = np.zeros(3, np.int32); d = np.zeros(12)
        self.lib().ref_geometry(_i(g), _d(d))
        return dict(grid=g, local_min=d[0:3].copy(), local_max=d[3:6].copy(), box_size=d[6:9].copy(),
                    inv_box_size=d[9:12].copy())

    def potential(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        self.lib().ref_potential(C.byref(a), C.byref(b), C.byref(c))
        return dict(cutoff=a.value, mass=b.value, lat=c.value)

    def atoms(self):
        L = self.lib()
        n = L.ref_n_local()
        gid = np.zeros(n, np.int32); box = np.zeros(n, np.int32)
        r = np.zeros((n, 3)); p = np.zeros((n, 3)); f = np.zeros((n, 3)); U = np.zeros(n)
        m = L.ref_get_atoms(_i(gid), _i(box), _d(r), _d(p), _d(f), _d(U))
        assert m == n
        return dict(gid=gid, box=box, r=r, p=p, f=f, U=U)

    def all_atoms(self):
        L = self.lib()
        n = L.ref_get_all_atoms(None, None, None)
        gid = np.zeros(n, np.int32); box = np.zeros(n, np.int32); r = np.zeros((n, 3))
        L.ref_get_all_atoms(_i(gid), _i(box), _d(r))
        return dict(gid=gid, box=box, r=r)

    def timestep(self, n):
        self.lib().ref_timestep(n)

    def energies(self):
        a, b = C.c_double(), C.c_double()
        self.lib().ref_energies(C.byref(a), C.byref(b))
        return a.value, b.value

    def alt_cells(self, cell_cutoff):
        L = self.lib()
        L.ref_alt_linkcells(cell_cutoff)
        g = np.zeros(5, np.int32); d = np.zeros(12)
        L.ref_alt_geometry(_i(g), _d(d))
        return dict(grid=g[:3].copy(), n_local_boxes=int(g[3]), n_total_boxes=int(g[4]),
                    box_size=d[6:9].copy(), inv_box_size=d[9:12].copy())

    def alt_bin(self, r):
        r = np.ascontiguousarray(r, np.float64)
        n = len(r)
        box = np.zeros(n, np.int32); tup = np.zeros((n, 3), np.int32)
        self.lib().ref_alt_box_from_coord(n, _d(r), _i(box), _i(tup))
        return box, tup

    def interpolate(self, which, x):
        x = np.ascontiguousarray(x, np.float64)
        f = np.zeros_like(x); df = np.zeros_like(x)
        L = self.lib()
        L.ref_eam_interpolate(L.ref_pot_handle(), which, len(x), _d(x), _d(f), _d(df))
        return f, df

    def table(self, which):
        L = self.lib()
        h = L.ref_pot_handle()
        n = L.ref_eam_table_n(h, which)
        v = np.zeros(n + 3)
        a, b = C.c_double(), C.c_double()
        L.ref_eam_table(h, which, C.byref(a), C.byref(b), _d(v))
        return dict(n=n, x0=a.value, inv_dx=b.value, padded=v)


class Ref32(Ref):
    """The reference's single-precision build (real_t = float); the wrappers still speak double, so every value
    that comes back is a float widened exactly."""

    _lib = None
    SO = REF_SO_F32
