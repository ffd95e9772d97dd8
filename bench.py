#!/usr/bin/env python
"""bench.py -- atom-steps/s of CoMD's short-range force path on N B200s (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
  torchrun --nproc-per-node N bench.py --gpus N ...        (one rank per GPU, 3-D domain decomposition)

A "step" is one full timestep() (kick, drift, redistribute incl. atom halo, force incl. F' halo, kick)
over the whole synthetic FCC-copper system.  Prints ONE JSON line on rank 0; exits non-zero when the parity
leg fails.

Workloads (BASELINE.json configs):
  eam_strong   C5: EAM Cu01.eam.alloy (setfl), 160^3 unit cells = 16 384 000 atoms, strong scaling   [default]
  lj_weak      C4: LJ 2.5 sigma, 100^3 unit cells (4.0 M atoms) per GPU, weak scaling
  eam_u6_20    C2: EAM Cu_u6 (funcfl) 20^3 = 32 000 atoms;  lj_20  C1: LJ 20^3;  eam_60  C3: EAM Cu01 60^3

Legs of the default run, in order:
  1. parity  (`check.parity_vs_oracle`): C2 on the SAME process grid against the CPU oracle -- step-0 forces and
     energies, owner rank / link cell / per-rank counts, total energy after 100 steps, a drifting crystal that
     ships atoms across rank faces between list rebuilds, halo/compute overlap on vs off.  The oracle is the checker.
  2. the timed region: W warm-up steps, then EXACTLY K steps (CUDA events on the library's stream, max over ranks).
  3. `secondary`: C4 (LJ weak scaling) and C3 (link-cell vs neighbour-list kernels) at the same N, short timed regions.
  4. `e2e`: the same step through host buffers.
  5. `cpu_baseline`: the reference's CPU path on the host cores (bounded sample).

The oracle (oracle/) is executed only by legs 1 and 5 and by --impl reference -- as the checker / the baseline, never
as the thing measured.
"""
import argparse
import json
import os
import re
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "comd-cuda-async_b200"))

import numpy as np

WORKLOADS = {
    # name: (potential, unit cells per axis (per GPU if weak), scaling, config label)
    "eam_strong": ("setfl", 160, "strong", "C5 EAM Cu01.eam.alloy setfl 160x160x160 (16.384M atoms)"),
    "lj_weak": ("lj", 100, "weak", "C4 LJ 2.5sigma 100x100x100 unit cells per GPU (4.0M atoms/GPU)"),
    "eam_60": ("setfl", 60, "strong", "C3 EAM Cu01.eam.alloy setfl 60x60x60 (864k atoms)"),
    "eam_u6_20": ("funcfl", 20, "strong", "C2 EAM Cu_u6 funcfl 20x20x20 (32k atoms)"),
    "lj_20": ("lj", 20, "strong", "C1 LJ 2.5sigma 20x20x20 (32k atoms)"),
}
PROC_GRIDS = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}
# algorithmic work per atom-step, SURVEY.md 8(d) / BASELINE.md 3 (FMA = 2, div = sqrt = 1, full-neighbour form)
FLOP_PER_ATOM_STEP = {"lj": 78 * 27 + 22, "setfl": 54 * (56 + 37) + 20 + 22, "funcfl": 42 * 93 + 42}
HBM_B_PER_ATOM_STEP = {"lj": 96, "setfl": 192, "funcfl": 192}
# the force kernels' share of that work (what roofline.achieved counts per force launch set)
FORCE_FLOP_PER_ATOM = {"lj": 78 * 27, "setfl": 54 * (56 + 37) + 20, "funcfl": 42 * 93 + 20}
# DRAM bytes per atom per force evaluation measured by ncu --set full at 864k atoms (sum over the force kernels)
NCU_DRAM = json.load(open(os.path.join(ROOT, "profiles", "ncu_dram_per_atom.json"))) if os.path.exists(os.path.join(ROOT, "profiles", "ncu_dram_per_atom.json")) else {}

FORCE_TOL = 1e-12       # north_star: step-0 per-atom forces and energies, FP64, max-norm relative
ENERGY100_TOL = 1e-6    # north_star: total energy at step 100
DRIFT_TOL = (1e-11, 1e-12, 1e-11)   # drift case of the parity leg: total energy (relative), positions (A), forces (max-norm relative) after 60 steps


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=d.get("hbm_gbs", 6650.0), source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """SM clock + throttle reasons during the timed region (B200_PROFILING.md recipe).  In-process NVML on a thread,
    two light queries every 200 ms: an `nvidia-smi -lms 100` child polling nine fields stalled the 8-rank step
    (4.6 against 2.4 ms/step on 8 B200 with and without it, gpurun_out/r2f_bench_n8.json); nvidia-smi is only the
    fallback where pynvml does not load."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}     # nvml.h nvmlClocksEventReason*

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.nvml = None
        self.lines = []
        self.samples = []
        self.stop_flag = threading.Event()

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else self.index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "250"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _poll(self):
        n = self.nvml
        while not self.stop_flag.is_set():
            try:
                self.samples.append((float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)),
                                     int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))))
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.nvml is not None:
            self.stop_flag.set()
            self.th.join(timeout=2)
            sm = [s[0] for s in self.samples]
            reasons = sorted(k for k, b in self.BITS.items() if any(s[1] & b for s in self.samples))
            return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=self.smax, reasons=reasons, samples=len(sm), source="nvml")
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(smax) if smax else None,
                    reasons=sorted(reasons), samples=len(sm), source="nvidia-smi")


def host_cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


# ------------------------------------------------------------------------------------------------ reference CPU arm
def _ref_grid(procs):
    """R cooperating ranks -> process grid i x j x k, as cubic as the factorisation allows."""
    best = (procs, 1, 1)
    for i in range(1, procs + 1):
        if procs % i:
            continue
        for j in range(1, procs // i + 1):
            if (procs // i) % j:
                continue
            k = procs // i // j
            if max(i, j, k) - min(i, j, k) < max(best) - min(best):
                best = (i, j, k)
    return tuple(sorted(best, reverse=True))


def cpu_reference_rate(pot, procs=None):
    """The reference's own CPU path (oracle/_ref, -m cpu_nl) on a bounded sample of the same workload, timed by
    its own timestep timer.  Two ways of filling the host cores, both reported:
      * `mpi`: ONE job of R cooperating ranks (R = host cores), the reference compiled with -DDO_MPI against the
        fork + shared-memory MPI shim of oracle/ref_build/shim (its 9 MPI calls, parallel.c:69-191): a real
        decomposed run with halo exchange, a 32^3 (EAM) / 40^3 (LJ) brick of unit cells per rank;
      * `copies`: R independent single-rank copies of that brick, rates added (no halo traffic at all).
    Falls back to the restated oracle (kind 'port') where oracle/_ref does not exist."""
    ref = os.path.join(ROOT, "oracle", "_ref")
    exe = os.path.join(ref, "CoMD-ref")
    potdir = os.path.join(ref, "pots")
    procs = procs or min(host_cores(), 64)
    n, steps = (32, 10) if pot != "lj" else (40, 10)       # EAM 131k atoms / LJ 256k atoms x 10 steps ~ 5-10 s per core
    flags = {"lj": [], "funcfl": ["-e"], "setfl": ["-e", "-t", "setfl"]}[pot]
    name = {"lj": "LJ 2.5sigma", "funcfl": "EAM Cu_u6", "setfl": "EAM Cu01"}[pot]
    def job_rate(text, atoms):
        # whole-job atom-steps/s from the reference's own timestep timer (rank 0's table, performanceTimers.c:137-160:
        # "timestep  calls  avg  total"); its "Average atom rate" line carries two decimals only
        m = re.search(r"^timestep\s+(\d+)\s+([0-9.]+)\s+([0-9.]+)", text, re.M)
        if m and float(m.group(3)) > 0:
            return atoms * steps / float(m.group(3))
        m = re.search(r"Average atom rate:\s+([0-9.eE+-]+) atoms/us", text)
        return float(m.group(1)) * 1e6 if m else None
    if os.path.exists(exe):
        import shutil
        import tempfile
        t0 = time.time()
        scratch = tempfile.mkdtemp(prefix="comd_ref_")      # the reference writes a YAML report per run into its cwd
        out = dict(unit="atom-steps/s", cores=procs, kind="reference")
        copies = {}
        for bname, binary in (("-O2", exe), ("-O3 -march=x86-64-v3", exe + "-O3")):      # oracle/ref_build/Makefile; the reference's own Makefile sets no -O at all
            if not os.path.exists(binary):
                continue
            cmd = [binary, "-m", "cpu_nl", "-d", potdir, "-x", str(n), "-y", str(n), "-z", str(n), "-N", str(steps), "-n", str(steps)] + flags
            t1 = time.time()
            ps = [subprocess.Popen(cmd, cwd=scratch, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL) for _ in range(procs)]
            rates = []
            for q in ps:
                v = job_rate(q.communicate()[0].decode("utf-8", "replace"), 4 * n ** 3)
                if v:
                    rates.append(v)
            if len(rates) == procs:
                copies[bname] = dict(value=float(sum(rates)), wall_s=round(time.time() - t1, 2))
        mpi = {}
        grid = _ref_grid(procs)
        for bname, binary in (("-O2", exe + "-mpi"), ("-O3 -march=x86-64-v3", exe + "-mpi-O3")):
            if not os.path.exists(binary):
                continue
            cmd = [binary, "-m", "cpu_nl", "-d", potdir, "-x", str(n * grid[0]), "-y", str(n * grid[1]), "-z", str(n * grid[2]),
                   "-i", str(grid[0]), "-j", str(grid[1]), "-k", str(grid[2]), "-N", str(steps), "-n", str(steps)] + flags
            env = dict(os.environ, COMD_SHIM_NRANKS=str(procs))
            t1 = time.time()
            try:
                q = subprocess.run(cmd, cwd=scratch, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, env=env, timeout=600)
                v = job_rate(q.stdout.decode("utf-8", "replace"), 4 * n ** 3 * procs)
                if v and q.returncode == 0:
                    mpi[bname] = dict(value=v, wall_s=round(time.time() - t1, 2))
            except subprocess.TimeoutExpired:
                pass
        shutil.rmtree(scratch, ignore_errors=True)
        if mpi or copies:
            # the baseline is the FASTER of the two ways of filling the cores (the copies carry no halo traffic and have come out
            # ~8 % ahead in every recorded run): a speed-up quoted against this line is then against the best CPU figure measured
            best_of = {"mpi_job": max(mpi.values(), key=lambda r: r["value"])["value"] if mpi else 0.0,
                       "independent_copies": max(copies.values(), key=lambda r: r["value"])["value"] if copies else 0.0}
            out["value_is"] = max(best_of, key=best_of.get)
            src = mpi if out["value_is"] == "mpi_job" else copies
            best = max(src, key=lambda k: src[k]["value"])
            out["value"] = src[best]["value"]; out["build"] = best; out["per_core"] = out["value"] / procs
            out["sample_wall_s"] = src[best]["wall_s"]
            mpi_text = ("ONE job of %d cooperating ranks (%dx%dx%d, fork + shm MPI shim, -DDO_MPI): %s, %d^3 unit cells per rank "
                        "(%d atoms in all) x %d steps, -m cpu_nl" % (procs, grid[0], grid[1], grid[2], name, n, 4 * n ** 3 * procs, steps))
            copies_text = "%d concurrent single-rank copies of: %s %dx%dx%d unit cells (%d atoms) x %d steps, -m cpu_nl" % (
                procs, name, n, n, n, 4 * n ** 3, steps)
            if out["value_is"] == "mpi_job":
                out["sample"] = mpi_text + ("; also timed, slower: " + copies_text if copies else "")
            else:
                out["sample"] = copies_text + ("; also timed, slower: " + mpi_text if mpi else "")
            if mpi:
                out["grid"] = "%dx%dx%d" % grid
            out["mpi_job"] = {k: v["value"] for k, v in mpi.items()} or None
            out["independent_copies"] = {k: v["value"] for k, v in copies.items()} or None
            out["wall_s"] = round(time.time() - t0, 2)
            return out
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as po
    kind = {"lj": po.POT_LJ, "funcfl": po.POT_FUNCFL, "setfl": po.POT_SETFL}[pot]
    n = 16
    orc = po.Oracle(pot=kind, nx=n, ny=n, nz=n)
    t0 = time.time()
    orc.timestep(steps)
    dt = time.time() - t0
    return dict(value=4 * n ** 3 * steps / dt, unit="atom-steps/s", cores=1, kind="port",
                sample="oracle/comd_oracle.c, %d atoms x %d steps, 1 thread" % (4 * n ** 3, steps), wall_s=round(dt, 2))


def run_reference(args, pot, label, scaling):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rates = []
    base = None
    for _ in range(max(1, min(args.steps, 2))):
        base = cpu_reference_rate(pot)
        rates.append(base["value"])
    v = float(np.mean(rates))
    base["value"] = v
    ncell = WORKLOADS[args.workload][1]
    n_atoms = 4 * ncell ** 3 * (args.gpus if scaling == "weak" else 1)
    print(json.dumps({
        "impl": "reference", "metric": "atom-steps/s", "value": v, "unit": "atom-steps/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * n_atoms / v, "ms_per_step_is": "extrapolated: atoms of the workload / sampled rate (the sample's own wall time is cpu_baseline.sample_wall_s)",
        "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": {"workload": label, "host_cores": base.get("cores"),
                                                        "note": "reference CPU path (-m cpu_nl) on %s host cores, bounded sample, rate is size-independent (O(N))" % base.get("cores")},
        "cpu_baseline": base, "e2e": {"value": v, "unit": "atom-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------ rank plumbing
class Ranks:
    """torch.distributed is plumbing only: rendezvous, scalar reductions, gathers of check data."""

    def __init__(self, backend="nccl"):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        self.tdev = "cpu"
        if self.world > 1:
            import torch
            import torch.distributed as dist
            if backend == "nccl":
                torch.cuda.set_device(self.local_rank)
                dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
                self.tdev = "cuda"
            elif not dist.is_initialized():
                dist.init_process_group(backend)
            self.dist = dist
        self.grid = PROC_GRIDS[self.world]
        r = self.rank
        self.coord = (r % self.grid[0], (r // self.grid[0]) % self.grid[1], r // (self.grid[0] * self.grid[1]))

    def reduce(self, vals, op="sum"):
        if self.dist is None:
            return [float(v) for v in vals]
        import torch
        t = torch.tensor(list(vals), dtype=torch.float64, device=self.tdev)
        self.dist.all_reduce(t, op={"sum": self.dist.ReduceOp.SUM, "max": self.dist.ReduceOp.MAX, "min": self.dist.ReduceOp.MIN}[op])
        return [float(x) for x in t.cpu()]

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def gather_rows(self, arr):
        """Rows of a 2-D float64 array from every rank, concatenated in rank order (on every rank); also the
        per-rank row counts."""
        arr = np.ascontiguousarray(arr, np.float64)
        if self.dist is None:
            return arr, [len(arr)]
        import torch
        counts = [int(c) for c in self.reduce_each(len(arr))]
        pad = max(counts)
        mine = torch.zeros((pad, arr.shape[1]), dtype=torch.float64, device=self.tdev)
        mine[:len(arr)] = torch.from_numpy(arr).to(self.tdev)
        parts = [torch.zeros_like(mine) for _ in range(self.world)]
        self.dist.all_gather(parts, mine)
        return np.concatenate([p[:c].cpu().numpy() for p, c in zip(parts, counts)]), counts

    def reduce_each(self, v):
        if self.dist is None:
            return [v]
        import torch
        t = torch.zeros(self.world, dtype=torch.float64, device=self.tdev)
        t[self.rank] = float(v)
        self.dist.all_reduce(t)
        return [float(x) for x in t.cpu()]

    def new_comm(self, dev):
        """NCCL communicator of one ComdDev (the library binds NCCL itself; the id travels through torch)."""
        if self.dist is None:
            return
        import torch
        import pycomd
        uid = torch.zeros(128, dtype=torch.uint8, device=self.tdev)
        if self.rank == 0:
            uid = torch.frombuffer(bytearray(pycomd.Device.comm_unique_id()), dtype=torch.uint8).to(self.tdev)
        self.dist.broadcast(uid, 0)
        dev.comm_init(bytes(uid.cpu().numpy().tobytes()), self.rank, self.world)


def owner_of(r, gext, grid):
    """Rank that owns a position: the brick [localMin, localMax) of decomposition.c:18-50 it falls into, with
    make_geometry's own bounds (so the comparison is against the very doubles the device bins with)."""
    import pycomd
    own = np.zeros(len(r), np.int64)
    mult = [1, grid[0], grid[0] * grid[1]]
    for a in range(3):
        idx = np.zeros(len(r), np.int64)
        for c in range(grid[a]):
            co = [0, 0, 0]; co[a] = c
            g = pycomd.make_geometry(gext, 1.0, grid, tuple(co))
            idx[(r[:, a] >= g.localMin[a]) & (r[:, a] < g.localMax[a])] = c
        own += mult[a] * idx
    return own


def make_device(R, pot, potential, n_cells, skin_rel, use_nl, opts=(), gid_r_p=None, seed=1234, precision="f64"):
    """One ComdDev for this rank's brick of an n_cells FCC lattice; atoms generated in place (or taken from the
    caller: global arrays, this rank keeps what it owns)."""
    import pycomd
    from pycomd import potentials as P
    lat = potential["lat"]
    gext = [n_cells[a] * lat for a in range(3)]
    skin = skin_rel * potential["cutoff"] if use_nl else 0.0
    geom = pycomd.make_geometry(gext, potential["cutoff"] + skin, R.grid, R.coord, precision=precision)
    lmin = [geom.localMin[a] for a in range(3)]; lmax = [geom.localMax[a] for a in range(3)]
    if gid_r_p is None:
        gid, r = P.fcc_block(n_cells, lat, lmin, lmax)
        p = P.maxwell_momenta(len(gid), potential["mass"], 600.0, seed=seed + R.rank)
    else:
        g_all, r_all, p_all = gid_r_p
        mine = owner_of(r_all, gext, R.grid) == R.rank
        gid, r, p = g_all[mine], r_all[mine], p_all[mine]
    n_local = len(gid)
    ghost_frac = 1.0
    for a in range(3):
        ghost_frac *= (geom.gridSize[a] + 2.0) / geom.gridSize[a]
    cap = int(n_local * (2.0 * ghost_frac - 1.0) * 1.10) + 65536     # old ghosts + new ghosts coexist in the pool while redistributing
    dev = pycomd.Device(geom, cap, device_id=R.local_rank, precision=precision)
    dev.set_mass(potential["mass"])
    if pot == "lj":
        dev.set_lj(potential["sigma"], potential["epsilon"], potential["cutoff"])
    else:
        dev.set_eam(potential["cutoff"], potential["phi"], potential["rho"], potential["f"])
    if use_nl:
        dev.set_neighbor_list(True, skin)
    for kv in list(filter(None, os.environ.get("COMD_OPTS", "").split(","))) + list(opts):      # tuning knobs (comd_dev_set_option), A/B runs only
        k, v = kv.split("="); dev.set_option(k, int(v))
    R.new_comm(dev)
    dev.upload(gid, r, p)
    return dev, geom, n_local


def rel_err(a, b):
    den = np.max(np.abs(b))
    return float(np.max(np.abs(a - b)) / (den if den > 0 else 1.0))


# ------------------------------------------------------------------------------------------------ leg 1: parity
def gathered_state(R, dev, keys=("r", "p", "f", "U", "box")):
    """(gid-sorted) state of all ranks on every rank, plus the owner rank of every atom."""
    d = dev.download()
    cols = [d["gid"].astype(np.float64)[:, None], d["r"], d["p"], d["f"], d["U"][:, None], d["box"].astype(np.float64)[:, None],
            np.full((len(d["gid"]), 1), float(R.rank))]
    allrows, counts = R.gather_rows(np.concatenate(cols, axis=1))
    o = np.argsort(allrows[:, 0], kind="stable")
    a = allrows[o]
    return dict(gid=a[:, 0].astype(np.int64), r=a[:, 1:4], p=a[:, 4:7], f=a[:, 7:10], U=a[:, 10], box=a[:, 11].astype(np.int64),
                owner=a[:, 12].astype(np.int64), counts=counts)


def expected_boxes(r, owner, gext, grid, cell_cutoff, po):
    """Reference link-cell id (getBoxFromCoord, linkCells.c:448-480) of every atom inside its owner's brick."""
    import pycomd
    box = np.zeros(len(r), np.int64)
    for rk in range(grid[0] * grid[1] * grid[2]):
        co = (rk % grid[0], (rk // grid[0]) % grid[1], rk // (grid[0] * grid[1]))
        g = pycomd.make_geometry(gext, cell_cutoff, grid, co)
        m = owner == rk
        if not np.any(m):
            continue
        cells = po.OracleCells([g.localMax[a] - g.localMin[a] for a in range(3)], cell_cutoff)
        lmin = np.array([g.localMin[a] for a in range(3)])
        box[m] = cells.bin(r[m] - lmin)[0]
    return box


def parity_leg(R):
    """C2 (EAM Cu_u6 funcfl 20^3, 32 000 atoms; BASELINE.json configs[1]) on this job's process grid against the CPU
    oracle.  Every rank builds the same deterministic initial state (the reference's gid-seeded RNG); rank 0 advances
    the oracle.  Returns the `parity_vs_oracle` record; `pass` is the verdict."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as po                      # the checker
    from pycomd import potentials as P
    t_begin = time.time()
    n = 20
    rec = {"config": "C2 EAM Cu_u6 funcfl 20x20x20 (32000 atoms), displaced 0.1 A, 600 K, list path, process grid %dx%dx%d" % R.grid,
           "tolerances": {"force_energy_step0": FORCE_TOL, "energy_step100": ENERGY100_TOL}}
    potential = P.load("funcfl")
    orc = po.Oracle(pot=po.POT_FUNCFL, nx=n, ny=n, nz=n, delta=0.1, temperature=600.0)
    a = orc.atoms()
    o = np.argsort(a["gid"], kind="stable")
    og = {k: a[k][o] for k in ("gid", "r", "p", "f", "U")}
    lat = potential["lat"]; gext = [n * lat] * 3
    skin = 0.1 * potential["cutoff"]
    cell_cut = potential["cutoff"] + skin
    ok = True

    def rebin(dev):
        # Between list rebuilds an atom that crossed a face stays with its rank (redistributeAtomsGpuNL re-bins only on
        # rebuild steps, timestep.c:278-354); ownership by position is defined right after a re-bin, so force one.
        dev.set_neighbor_list(True, skin)
        dev.redistribute()

    # (a) step 0: forces, energies, owner rank, link cell, per-rank counts
    dev, geom, n_local = make_device(R, "funcfl", potential, [n] * 3, 0.1, True, gid_r_p=(og["gid"], og["r"], og["p"]))
    dev.compute_force()
    s = gathered_state(R, dev)
    exp_owner = owner_of(og["r"], gext, R.grid)
    rec["atoms"] = int(len(s["gid"]))
    rec["force_rel_err"] = rel_err(s["f"], og["f"]) if len(s["gid"]) == len(og["gid"]) else None
    rec["energy_rel_err"] = rel_err(s["U"], og["U"]) if len(s["gid"]) == len(og["gid"]) else None
    same = len(s["gid"]) == len(og["gid"]) and np.array_equal(s["gid"], og["gid"])
    rec["gids_exact"] = bool(same)
    rec["owner_mismatches"] = int(np.sum(s["owner"] != exp_owner)) if same else None
    rec["box_mismatches"] = int(np.sum(s["box"] != expected_boxes(og["r"], exp_owner, gext, R.grid, cell_cut, po))) if same else None
    rec["per_rank_counts"] = [int(c) for c in s["counts"]]
    rec["per_rank_counts_expected"] = [int(np.sum(exp_owner == k)) for k in range(R.world)]
    ok &= same and rec["force_rel_err"] <= FORCE_TOL and rec["energy_rel_err"] <= FORCE_TOL
    ok &= rec["owner_mismatches"] == 0 and rec["box_mismatches"] == 0 and rec["per_rank_counts"] == rec["per_rank_counts_expected"]

    # (b) 100 steps: total energy, atom conservation, per-rank counts from the oracle's final positions
    dev.timestep(100, 1.0)
    e = dev.energy()
    e_tot = R.reduce([e[0] + e[1]])[0]
    rebin(dev)
    n_after = [int(c) for c in R.reduce_each(dev.num_local())]
    if R.rank == 0:
        orc.timestep(100)
        eo = sum(orc.energies())
        fin = orc.atoms()
        cnt_exp = [int(np.sum(owner_of(fin["r"], gext, R.grid) == k)) for k in range(R.world)]
    else:
        eo, cnt_exp = 0.0, [0] * R.world
    eo = R.reduce([eo])[0]
    cnt_exp = [int(c) for c in R.reduce([float(c) for c in cnt_exp])]
    rec["energy100_rel_err"] = abs(e_tot - eo) / abs(eo)
    rec["per_rank_counts_step100"] = n_after
    rec["per_rank_counts_step100_expected"] = cnt_exp
    rec["list_builds_100_steps"] = dev.neighbor_counters()["builds"]
    ok &= rec["energy100_rel_err"] <= ENERGY100_TOL and sum(n_after) == len(og["gid"]) and n_after == cnt_exp
    dev.close()

    # (c) drift: the whole crystal moves 0.02 A/fs along x (half / a quarter of that along y / z), so atoms cross rank
    # (and periodic) faces all the time, also between list rebuilds; compared with the oracle run from the same state
    drift = potential["mass"] * np.array([0.02, 0.01, -0.005])
    pd = og["p"] + drift
    steps_d = 60
    runs = []
    # run 0: defaults (sync-free steps; N > 1: exchange overlapped, one-shot interior CTAs + boundary shell on a side stream);
    # run 1: round 1's step (host waits for the rebuild verdict every step; N > 1: overlapped, persistent interior CTAs on
    # 140 SMs); run 2 (N > 1): no overlap at all
    # run 3 (N > 1): steady-state steps replayed from a CUDA graph (on one rank that is run 0's default at this size)
    variants = [[], ["nl_speculate=0", "overlap_mode=0"]] + ([["overlap=0"], ["graph_step=1"]] if R.world > 1 else [])
    for vopts in variants:
        dev, geom, _ = make_device(R, "funcfl", potential, [n] * 3, 0.1, True, opts=vopts, gid_r_p=(og["gid"], og["r"], pd))
        dev.compute_force()
        own0 = gathered_state(R, dev)["owner"]
        dev.timestep(steps_d, 1.0)
        e = dev.energy()
        nc = dev.neighbor_counters()
        rebin(dev)
        dev.compute_force()            # the re-bin re-sorts r and p only (as the reference's does); forces on the new owners
        runs.append((gathered_state(R, dev), R.reduce([e[0] + e[1]])[0], own0, nc, dev.step_stats()))
        dev.close()
    sd, ed, own0, nc, st0 = runs[0]
    if R.rank == 0:
        orc2 = po.Oracle(pot=po.POT_FUNCFL, nx=n, ny=n, nz=n, delta=0.1, temperature=600.0)
        orc2.set_atoms(og["gid"], og["r"], pd)
        orc2.timestep(steps_d)
        f2 = orc2.atoms(); o2 = np.argsort(f2["gid"], kind="stable")
        eo2 = sum(orc2.energies())
        L = np.array(gext)
        dr = sd["r"] - f2["r"][o2]; dr -= L * np.round(dr / L)
        pos_err = float(np.max(np.abs(dr)))
        f_err = rel_err(sd["f"], f2["f"][o2])
    else:
        eo2, pos_err, f_err = 0.0, 0.0, 0.0
    eo2, pos_err, f_err = R.reduce([eo2, pos_err, f_err])
    rec["drift"] = {"steps": steps_d, "atoms_that_changed_rank": int(np.sum(sd["owner"] != own0)), "list_builds": nc["builds"], "ghost_refreshes": nc["refreshes"],
                    "energy_rel_err": abs(ed - eo2) / abs(eo2), "max_position_err_A": pos_err, "force_rel_err": f_err,
                    "owner_mismatches": int(np.sum(sd["owner"] != owner_of(sd["r"], gext, R.grid))), "atoms": int(len(sd["gid"]))}
    # 60 steps of a drifting crystal against the oracle: measured 1.0e-13 / 2.8e-14 A / 1.4e-13 at N = 1, 2 and 8 (profiles/r2_v3_*, r3_*)
    rec["drift"]["tolerances"] = {"energy_rel_err": DRIFT_TOL[0], "max_position_err_A": DRIFT_TOL[1], "force_rel_err": DRIFT_TOL[2]}
    ok &= rec["drift"]["energy_rel_err"] <= DRIFT_TOL[0] and pos_err <= DRIFT_TOL[1] and f_err <= DRIFT_TOL[2] and rec["drift"]["owner_mismatches"] == 0
    ok &= len(sd["gid"]) == len(og["gid"]) and (R.world == 1 or rec["drift"]["atoms_that_changed_rank"] > 0)
    same_bits = lambda q: bool(all(np.array_equal(sd[k], q[k]) for k in ("gid", "r", "p", "f", "owner")))
    rec["sync_free_step_bit_identical_to_synchronised"] = same_bits(runs[1][0]) and runs[1][3] == nc
    ok &= rec["sync_free_step_bit_identical_to_synchronised"]
    rec["drift"]["steps_replayed_from_cuda_graph"] = st0["graph_steps"]
    if R.world > 1:
        rec["overlap_on_off_bit_identical"] = same_bits(runs[2][0])
        rec["graph_step_bit_identical"] = same_bits(runs[3][0]) and runs[3][3] == nc
        rec["graph_step_replays"] = runs[3][4]["graph_steps"]
        ok &= rec["overlap_on_off_bit_identical"] and rec["graph_step_bit_identical"]
    rec["pass"] = bool(ok)
    rec["wall_s"] = round(time.time() - t_begin, 1)
    return rec


# ------------------------------------------------------------------------------------------------ leg 2/3: timed regions
def timed_steps(R, dev, steps, print_rate, dt=1.0):
    """EXACTLY `steps` steps in intervals of printRate (CoMD.c:122-139: energies are due at the end of every
    interval), CUDA events on the library's stream, max over ranks.  Returns (seconds, mean force ms, launches, list counters)."""
    dev.sync(); R.barrier()
    c0 = dev.neighbor_counters()
    l0 = dev.kernel_stats()["launches"]          # also resets the force-time ring
    dev.timer_start()
    left = steps
    walls = []                                   # host wall clock per timestep() call (diagnostic: a stalled interval shows here)
    while left > 0:
        w0 = time.perf_counter()
        dev.timestep(min(left, print_rate), dt)
        walls.append(round(1e3 * (time.perf_counter() - w0), 3))
        left -= print_rate
    t_local = dev.timer_stop() * 1e-3
    ks = dev.kernel_stats()
    c1 = dev.neighbor_counters()
    t_max = R.reduce([t_local], "max")[0]
    return t_max, ks["last_force_ms"], ks["launches"] - l0, dict(builds=c1["builds"] - c0["builds"], refreshes=c1["refreshes"] - c0["refreshes"],
                                                                 host_wall_ms_per_call=walls)


def run_workload(R, name, steps, warmup, print_rate, skin, use_nl, every_step_too=False, opts=(), sampler=None, precision="f64"):
    from pycomd import potentials as P
    pot, ncell, scaling, label = WORKLOADS[name]
    n_cells = [ncell * (R.grid[a] if scaling == "weak" else 1) for a in range(3)]
    potential = P.load(pot)
    dev, geom, n_local = make_device(R, pot, potential, n_cells, skin, use_nl, opts=opts, precision=precision)
    n_global = 4 * n_cells[0] * n_cells[1] * n_cells[2]
    dev.compute_force()
    e0 = dev.energy()
    if use_nl:
        # first-use costs belong to the warm-up: with the list marked stale, the first warm-up step takes the rebuild branch
        # INSIDE timestep() (re-bin, atom exchange, list build, uncached F' halo on the exchange stream) -- at N = 8 its first
        # occurrence cost 13 ms more than any later one (NCCL's lazy set-up of those message sizes, gpurun_out/r2k_ab_n8.txt)
        dev.set_neighbor_list(True, skin * potential["cutoff"])
    dev.timestep(warmup, 1.0)
    res = dict(name=name, pot=pot, label=label, scaling=scaling, n_global=n_global, n_local=n_local, geom=geom, dev=dev, e0=e0)
    if sampler is not None:                      # clocks + throttle reasons DURING the timed region only
        dev.sync(); sampler.start(); time.sleep(0.3)
    res["t"], res["force_ms"], res["launches"], res["list"] = timed_steps(R, dev, steps, print_rate)
    res["clocks"] = sampler.stop() if sampler is not None else None
    res["e1"] = dev.energy()
    if every_step_too:
        # the same K steps on eam.c's own schedule (pair energies evaluated at every step, not only where they are read)
        dev.set_option("energy_every_step", 1)
        res["t_every"], res["force_ms_every"], _, res["list_every"] = timed_steps(R, dev, steps, print_rate)
        dev.set_option("energy_every_step", 0)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="eam_strong", choices=sorted(WORKLOADS))
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"], help="precision build of the library: f64 (libcomd_b200.so, the headline) or f32 "
                    "(libcomd_b200_f32.so, the reference's -DSINGLE / DOUBLE_PRECISION=OFF build, mytype.h:8-24)")
    ap.add_argument("--e2e-steps", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the parity leg (A/B tuning runs only)")
    ap.add_argument("--no-secondary", action="store_true", help="skip the C4 / C3 secondary timed regions")
    ap.add_argument("--skin", type=float, default=0.1, help="neighbour-list skin relative to the cutoff (CoMD -S, default 0.1)")
    ap.add_argument("--print-rate", type=int, default=10, help="steps per timestep() call = CoMD's printRate (default 10, mycommand.c)")
    ap.add_argument("--neighbor-list", type=int, default=int(os.environ.get("COMD_NL", "1")),
                    help="1: neighbour-list force path (skin 0.1*cutoff, mycommand.c:210), 0: link-cell path")
    args = ap.parse_args()
    pot, ncell, scaling, label = WORKLOADS[args.workload]

    if args.impl == "reference":
        run_reference(args, pot, label, scaling)
        return 0

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("launch with torchrun --nproc-per-node %d for --gpus %d" % (args.gpus, args.gpus))
    import pycomd
    R = Ranks("nccl")
    rank = R.rank

    # ---- leg 1: parity on this process grid (before anything is timed)
    parity = None if args.no_verify else parity_leg(R)

    # ---- leg 2: warm-up, then EXACTLY K timed steps
    M = run_workload(R, args.workload, args.steps, args.warmup, args.print_rate, args.skin, args.neighbor_list, every_step_too=True, precision=args.dtype,
                     sampler=ClockSampler(R.local_rank) if rank == 0 and not os.environ.get("COMD_BENCH_NO_SAMPLER") else None)
    clocks = M["clocks"]
    dev, geom, n_local, n_global = M["dev"], M["geom"], M["n_local"], M["n_global"]
    fp64_peak = dev.fp64_peak() if args.dtype == "f64" else dev.fp32_peak()      # the FMA pipe the build computes on
    fp32_peak = dev.fp32_peak()
    t_max, t_every = M["t"], M["t_every"]
    n_after = dev.num_local()

    # ---- global check figures: energies, atom count, largest force component -- summed / maxed over ALL ranks
    d0 = dev.download()
    e0g = R.reduce([M["e0"][0] + M["e0"][1]])[0]; e1g = R.reduce([M["e1"][0] + M["e1"][1]])[0]
    n_after_g = int(R.reduce([n_after])[0])
    fmax_g = R.reduce([float(np.max(np.abs(d0["f"])))], "max")[0]
    per_rank_after = [int(c) for c in R.reduce_each(n_after)]

    # ---- leg 4: e2e, the same step through HOST buffers (H2D of gid,r,p,f + D2H of gid,r,p,f,U inside the timed region)
    cols = lambda x: [np.ascontiguousarray(x[:, k]) for k in range(3)]
    _pins = []
    def pin(a):
        pa, v = pycomd.pinned_like(np.ascontiguousarray(a)); _pins.append(pa); return v
    hg, hr, hp, hf = pin(d0["gid"]), [pin(c) for c in cols(d0["r"])], [pin(c) for c in cols(d0["p"])], [pin(c) for c in cols(d0["f"])]
    wb = np.dtype(dev.real).itemsize                # bytes per comd_real
    def chain(k, o, m):
        t0 = time.perf_counter()
        for _ in range(k):
            m, o, _ = dev.step_host(1.0, o["gid"][:m], [a[:m] for a in o["r"]], [a[:m] for a in o["p"]], [a[:m] for a in o["f"]])
        return time.perf_counter() - t0, o, m
    # a trajectory driven through host memory: the caller feeds each call's results (pinned) to the next call
    m, o, _ = dev.step_host(1.0, hg, hr, hp, hf)          # warm-up (allocates pinned staging; full upload path)
    dev.sync(); R.barrier()
    r0 = dev.step_host_resumed()
    t_e2e, o, m = chain(args.e2e_steps, o, m)             # default mode: every call waits for its upload
    e2e_resumed = dev.step_host_resumed() - r0
    t_e2e = R.reduce([t_e2e], "max")[0]
    t_spec = None
    if world == 1:
        # opt-in extra: the step starts from the atoms the device still holds while the inputs cross PCIe, and is kept
        # only if the upload turns out bit-identical -- pays only for an in-place caller; NOT the e2e headline
        dev.set_option("e2e_speculate", 1)
        _, o, m = chain(1, o, m)
        t_spec, o, m = chain(args.e2e_steps, o, m)
        dev.set_option("e2e_speculate", 0)
    h2d = n_local * (4 + 9 * wb)
    d2h = n_local * (4 + 10 * wb) + 16              # local atoms only (compacted on the device): gid + r,p,f,U, and the two energies
    dev.close()
    for pa in _pins:
        pa.free()

    # ---- leg 3: secondary timed regions (north_star's weak-scaling and kernel-variant numbers in the same record)
    secondary = None
    if not args.no_secondary:
        secondary = {}
        ssteps = 20
        def line(S, nl):
            v = S["n_global"] * ssteps / S["t"]
            fl = FORCE_FLOP_PER_ATOM[S["pot"]] * S["n_local"] / (S["force_ms"] * 1e-3) * 1e-12 if S["force_ms"] > 0 else None
            return {"workload": S["label"], "value": v, "unit": "atom-steps/s", "steps": ssteps, "ms_per_step": 1e3 * S["t"] / ssteps, "scaling": S["scaling"],
                    "force_ms_per_step": S["force_ms"], "force_frac_fp64": fl / fp64_peak if fl else None, "atoms": S["n_global"], "atoms_per_gpu": S["n_local"],
                    "force_path": "neighbour list" if nl else "link cells (27-cell stencil, TMA-staged tiles)", "list_builds_in_window": S["list"]["builds"],
                    "e_total_per_atom_after": R.reduce([S["e1"][0] + S["e1"][1]])[0] / S["n_global"]}
        for key, wl, nl in (("lj_weak_C4", "lj_weak", 1), ("eam_60_C3_list", "eam_60", 1), ("eam_60_C3_linkcell", "eam_60", 0)):
            if wl == args.workload and nl == args.neighbor_list:
                continue
            S = run_workload(R, wl, ssteps, 3, args.print_rate, args.skin, nl, precision=args.dtype)
            secondary[key] = line(S, nl)
            S["dev"].close()
        if args.dtype == "f64":
            # the headline workload through the single-precision build (-DCOMD_SINGLE): beside the FP64 line, never instead of it
            S = run_workload(R, args.workload, ssteps, 3, args.print_rate, args.skin, args.neighbor_list, precision="f32")
            ln = line(S, args.neighbor_list)
            fl = FORCE_FLOP_PER_ATOM[S["pot"]] * S["n_local"] / (S["force_ms"] * 1e-3) * 1e-12 if S["force_ms"] > 0 else None
            del ln["force_frac_fp64"]
            ln.update(dtype="f32", library="libcomd_b200_f32.so", force_frac_fp32=fl / fp32_peak if fl else None, fp32_peak_tflops=fp32_peak,
                      peak_source="FFMA microbenchmark run in this process (comd_dev_fp32_peak)",
                      tolerance="tests/test_gpu_precision.py: 2e-6 of max|F|, max|U| at step 0 against the reference's own float build")
            secondary[args.workload + "_f32"] = ln
            S["dev"].close()

    if rank != 0:
        return 0 if (parity is None or parity["pass"]) else 1
    value = n_global * args.steps / t_max
    value_every = n_global * args.steps / t_every
    peaks = measured_peaks()
    fms, fms_every = M["force_ms"], M["force_ms_every"]
    flops_force = FORCE_FLOP_PER_ATOM[pot] * n_local
    achieved_tf = flops_force / (fms * 1e-3) * 1e-12 if fms > 0 else None
    achieved_every = flops_force / (fms_every * 1e-3) * 1e-12 if fms_every > 0 else None
    hbm_gbs = value / world * HBM_B_PER_ATOM_STEP[pot] * 1e-9
    # option nl_n3 "auto" (the default): once-per-pair kernels for EAM, full-list kernels for LJ
    n3 = pot != "lj" and "nl_n3=0" not in os.environ.get("COMD_OPTS", "") or "nl_n3=1" in os.environ.get("COMD_OPTS", "")
    kern = ("k_force_n3" if n3 else "k_force_nl") if args.neighbor_list else "k_force_tile"
    dram = NCU_DRAM.get("%s_%s" % (pot, kern))
    e_sched = "pair energies evaluated only where they are read: 1 step in %d (the last of each timestep() call)" % args.print_rate if pot != "lj" or n3 else "every step"
    out = {
        "metric": "atom-steps/s", "value": value, "unit": "atom-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_max / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic FCC Cu lattice (a=3.615 A), numpy-seeded Maxwell-Boltzmann momenta at 600 K, dt=1 fs",
        "config": {"workload": label, "atoms": n_global, "atoms_per_gpu": n_local, "proc_grid": "%dx%dx%d" % R.grid,
                   "cells_per_gpu": "%dx%dx%d" % tuple(geom.gridSize[:]), "parallelism": "3-D spatial domain decomposition, 1 rank/GPU",
                   "force_path": ("neighbour list (skin %g*cutoff, rebuilt when an atom moved > skin/2)%s" % (args.skin, "; pairs inside a work block evaluated once for both atoms" if n3 else "")) if args.neighbor_list else "link cells (re-binned every step)",
                   "energies": "potential + kinetic energy every %d steps (CoMD printRate); %s" % (args.print_rate, e_sched),
                   "list_in_timed_window": M["list"], "list_in_every_step_window": M["list_every"],
                   "l2": "working set %.0f MB per GPU > 126 MB L2 (no flush needed)" % (n_local * 200e-6) if n_local * 200 > 126e6
                         else "working set fits L2: small-config line, not the headline"},
        "roofline": {"bound": "fp64" if args.dtype == "f64" else "fp32", "kernel": "force (%s)" % (("%s<LJ>" % kern if pot == "lj" else "%s<RHO> + F' halo + %s<EAMF|EAMFE>" % (kern, kern))),
                     "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s",
                     "frac": (achieved_tf / fp64_peak) if achieved_tf else None,
                     "energy_schedule": e_sched,
                     "frac_energies_every_step": (achieved_every / fp64_peak) if achieved_every else None,
                     "force_ms_per_step_energies_every_step": fms_every,
                     "peak_source": ("DFMA microbenchmark run in this process (comd_dev_fp64_peak)" if args.dtype == "f64" else "FFMA microbenchmark run in this process (comd_dev_fp32_peak)") + "; MEASURED_PEAKS.json has no FP64/FP32 FMA entry",
                     "algorithmic_flop_per_atom": FORCE_FLOP_PER_ATOM[pot], "force_ms_per_step": fms,
                     "step_frac_fp64": value / world * FLOP_PER_ATOM_STEP[pot] * 1e-12 / fp64_peak,
                     "step_frac_fp64_energies_every_step": value_every / world * FLOP_PER_ATOM_STEP[pot] * 1e-12 / fp64_peak,
                     "hbm": {"achieved": hbm_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": hbm_gbs / peaks["hbm_gbs"],
                             "peak_source": peaks["source"], "algorithmic_bytes_per_atom_step": HBM_B_PER_ATOM_STEP[pot]},
                     "traffic": dram["bytes_per_atom"] * n_local if dram else None,
                     "traffic_source": dram["source"] if dram else None},
        "e2e": {"value": n_global * args.e2e_steps / t_e2e, "unit": "atom-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": args.e2e_steps, "chained": True, "resumed_steps": e2e_resumed,
                "value_speculative": (n_global * args.e2e_steps / t_spec) if t_spec else None,
                "note": "each call's host results (pinned, updated in place) are the next call's host inputs and every call waits for its upload "
                        "before the step starts; a resumed step reuses the device-side sort and neighbour list (valid by the skin/2 criterion), "
                        "the others re-sort and rebuild.  value_speculative (one rank, opt-in option e2e_speculate): the step starts from the "
                        "atoms the device still holds while the inputs cross PCIe and is kept only if the upload is bit-identical -- "
                        "benchmark-shaped, reported for reference only"},
        "value_energies_every_step": value_every,
        "gpu_launches": int(M["launches"]),
        "clocks": clocks,
        "check": {"e_total_per_atom_before": e0g / n_global, "e_total_per_atom_after": e1g / n_global,
                  "atoms_after": n_after_g, "atoms_conserved": n_after_g == n_global, "atoms_per_rank_after": per_rank_after,
                  "max_abs_force_component": fmax_g, "global": True, "parity_vs_oracle": parity},
        "secondary": secondary,
    }
    if not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_reference_rate(pot)
    # LAST key of the line on purpose: the driver's records keep only the tail of stdout, and at N > 1 that tail is the only place
    # the C4 LJ weak-scaling number (north_star's second scaling metric, BASELINE.json configs[3]) can survive in
    w = (secondary or {}).get("lj_weak_C4")           # --no-secondary leaves `secondary` None
    out["scaling_summary"] = {
        "n_gpus": world,
        "eam_strong_C5": {"value": value, "ms_per_step": 1e3 * t_max / args.steps, "scaling": scaling} if args.workload == "eam_strong" else None,
        "lj_weak_C4": ({"value": w["value"], "ms_per_step": w["ms_per_step"], "force_ms_per_step": w["force_ms_per_step"], "atoms": w["atoms"],
                        "atoms_per_gpu": w["atoms_per_gpu"], "steps": w["steps"], "scaling": "weak", "unit": "atom-steps/s"} if w else None),
        "note": "weak efficiency at N = lj_weak_C4.value(N) / (N * lj_weak_C4.value(1)); strong = eam_strong_C5.value(N) / (N * value(1))"}
    print(json.dumps(out))
    return 0 if (parity is None or parity["pass"]) else 1


if __name__ == "__main__":
    sys.exit(main())
