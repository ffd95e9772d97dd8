#!/usr/bin/env python
"""bench.py -- atom-steps/s of CoMD's short-range force path on N B200s (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
  torchrun --nproc-per-node N bench.py --gpus N ...        (one rank per GPU, 3-D domain decomposition)

A "step" is one full timestep() (kick, drift, redistribute incl. atom halo, force incl. F' halo, kick)
over the whole synthetic FCC-copper system.  Prints ONE JSON line on rank 0.

Workloads (BASELINE.json configs):
  eam_strong   C5: EAM Cu01.eam.alloy (setfl), 160^3 unit cells = 16 384 000 atoms, strong scaling   [default]
  lj_weak      C4: LJ 2.5 sigma, 100^3 unit cells (4.0 M atoms) per GPU, weak scaling
  eam_u6_20    C2: EAM Cu_u6 (funcfl) 20^3 = 32 000 atoms;  lj_20  C1: LJ 20^3;  eam_60  C3: EAM Cu01 60^3

The oracle (oracle/) is executed only by the cpu_baseline leg and by --impl reference.
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
# DRAM bytes per atom per force evaluation measured by ncu (sum over the force kernels): (potential, list path?) -> B/atom
NCU_DRAM_B_PER_ATOM_FORCE = {("setfl", 1): (367.97 + 14.16 + 256.72 + 4.95) * 1e6 / 864000.0,    # profiles/r1_v4_nl_ncu_full_eam60.txt
                             ("lj", 1): (424.85 + 5.43) * 1e6 / 864000.0}                         # profiles/r1_v6_nl_ncu_full_lj60.txt
FORCE_FLOP_PER_ATOM = {"lj": 78 * 27, "setfl": 54 * (56 + 37) + 20, "funcfl": 42 * 93 + 20}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=d.get("hbm_gbs", 6650.0), source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
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
                    reasons=sorted(reasons), samples=len(sm))


def host_cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_reference_rate(pot, procs=None):
    """The reference's own CPU path (oracle/_ref/CoMD-ref, -m cpu_nl) on a bounded sample of the same
    workload, timed by its own timestep timer; falls back to the restated oracle (kind 'port').
    The reference compiled here is single-rank (its MPI layer is shimmed out, oracle/ref_build), so the host
    cores are filled the way its rank-per-core decomposition would fill them minus the halo traffic: one
    independent copy of the sample per core, all running at once, rates added."""
    exe = os.path.join(ROOT, "oracle", "_ref", "CoMD-ref")
    potdir = os.path.join(ROOT, "oracle", "_ref", "pots")
    procs = procs or min(host_cores(), 256)
    n, steps = (32, 10) if pot != "lj" else (40, 10)       # EAM 131k atoms / LJ 256k atoms x 10 steps ~ 5-10 s per core
    flags = {"lj": [], "funcfl": ["-e"], "setfl": ["-e", "-t", "setfl"]}[pot]
    sample = "%d concurrent single-rank copies of: %s %dx%dx%d unit cells (%d atoms) x %d steps, -m cpu_nl" % (
        procs, {"lj": "LJ 2.5sigma", "funcfl": "EAM Cu_u6", "setfl": "EAM Cu01"}[pot], n, n, n, 4 * n ** 3, steps)
    if os.path.exists(exe):
        t0 = time.time()
        builds = {}
        import shutil
        import tempfile
        scratch = tempfile.mkdtemp(prefix="comd_ref_")      # the reference writes a YAML report per run into its cwd
        for name, binary in (("-O2", exe), ("-O3 -march=x86-64-v3", exe + "-O3")):      # oracle/ref_build/Makefile; the reference's own Makefile sets no -O at all
            if not os.path.exists(binary):
                continue
            cmd = [binary, "-m", "cpu_nl", "-d", potdir, "-x", str(n), "-y", str(n), "-z", str(n), "-N", str(steps), "-n", str(steps)] + flags
            ps = [subprocess.Popen(cmd, cwd=scratch, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL) for _ in range(procs)]
            rates = []
            for q in ps:
                m = re.search(r"Average atom rate:\s+([0-9.eE+-]+) atoms/us", q.communicate()[0].decode("utf-8", "replace"))
                if m:
                    rates.append(float(m.group(1)) * 1e6)
            if len(rates) == procs:
                builds[name] = float(sum(rates))
        shutil.rmtree(scratch, ignore_errors=True)
        if builds:
            best = max(builds, key=builds.get)
            return dict(value=builds[best], unit="atom-steps/s", cores=procs, kind="reference", sample=sample, build=best,
                        by_build=builds, per_core=builds[best] / procs, wall_s=round(time.time() - t0, 2))
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
    for _ in range(max(1, min(args.steps, 3))):
        base = cpu_reference_rate(pot)
        rates.append(base["value"])
    v = float(np.mean(rates))
    base["value"] = v
    ncell = WORKLOADS[args.workload][1]
    n_atoms = 4 * ncell ** 3 * (args.gpus if scaling == "weak" else 1)
    print(json.dumps({
        "impl": "reference", "metric": "atom-steps/s", "value": v, "unit": "atom-steps/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * n_atoms / v, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": {"workload": label, "note": "reference CPU path (-m cpu_nl), bounded sample, rate is size-independent (O(N)); "
                                                                                     "ms_per_step = atoms of the workload / sampled rate"},
        "cpu_baseline": base, "e2e": {"value": v, "unit": "atom-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="eam_strong", choices=sorted(WORKLOADS))
    ap.add_argument("--e2e-steps", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--skin", type=float, default=0.1, help="neighbour-list skin relative to the cutoff (CoMD -S, default 0.1)")
    ap.add_argument("--print-rate", type=int, default=10, help="steps per timestep() call = CoMD's printRate (default 10, mycommand.c)")
    ap.add_argument("--neighbor-list", type=int, default=int(os.environ.get("COMD_NL", "1")),
                    help="1: neighbour-list force path (skin 0.1*cutoff, mycommand.c:210), 0: link-cell path")
    args = ap.parse_args()
    pot, ncell, scaling, label = WORKLOADS[args.workload]

    if args.impl == "reference":
        run_reference(args, pot, label, scaling)
        return

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun --nproc-per-node %d for --gpus %d" % (args.gpus, args.gpus))
    import pycomd
    from pycomd import potentials as P

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    grid = PROC_GRIDS[world]
    n_cells = [ncell * (grid[a] if scaling == "weak" else 1) for a in range(3)]
    potential = P.load(pot)
    lat = potential["lat"]
    gext = [n_cells[a] * lat for a in range(3)]
    coord = (rank % grid[0], (rank // grid[0]) % grid[1], rank // (grid[0] * grid[1]))
    skin = args.skin * potential["cutoff"] if args.neighbor_list else 0.0
    geom = pycomd.make_geometry(gext, potential["cutoff"] + skin, grid, coord)
    lmin = [geom.localMin[a] for a in range(3)]; lmax = [geom.localMax[a] for a in range(3)]
    gid, r = P.fcc_block(n_cells, lat, lmin, lmax)
    n_local = len(gid)
    n_global = 4 * n_cells[0] * n_cells[1] * n_cells[2]
    p = P.maxwell_momenta(n_local, potential["mass"], 600.0, seed=1234 + rank)
    ghost_frac = 1.0
    for a in range(3):
        ghost_frac *= (geom.gridSize[a] + 2.0) / geom.gridSize[a]
    cap = int(n_local * (2.0 * ghost_frac - 1.0) * 1.10) + 65536     # old ghosts + new ghosts coexist in the pool while redistributing
    dev = pycomd.Device(geom, cap, device_id=local_rank)
    dev.set_mass(potential["mass"])
    if pot == "lj":
        dev.set_lj(potential["sigma"], potential["epsilon"], potential["cutoff"])
    else:
        dev.set_eam(potential["cutoff"], potential["phi"], potential["rho"], potential["f"])
    if args.neighbor_list:
        dev.set_neighbor_list(True, skin)
    for kv in filter(None, os.environ.get("COMD_OPTS", "").split(",")):      # tuning knobs (comd_dev_set_option), A/B runs only
        k, v = kv.split("="); dev.set_option(k, int(v))
    if world > 1:
        import torch
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid = torch.frombuffer(bytearray(pycomd.Device.comm_unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(uid, 0)
        dev.comm_init(bytes(uid.cpu().numpy().tobytes()), rank, world)
    dev.upload(gid, r, p)
    dev.compute_force()
    e0 = dev.energy()
    dt = 1.0

    def barrier():
        dev.sync()
        if dist is not None:
            dist.barrier()

    # ---- warm-up, then EXACTLY K timed steps, CUDA events via the library's stream (comd_dev_sync on both sides)
    dev.timestep(args.warmup, dt)
    barrier()
    fp64_peak = dev.fp64_peak()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    barrier()
    l0 = dev.kernel_stats()["launches"]
    force_ms = []
    dev.kernel_stats()                       # reset the force-time ring
    dev.timer_start()                        # CUDA events on the library's stream (the one the kernels run on)
    left = args.steps                        # K steps in intervals of printRate, as CoMD.c:122-139 drives timestep():
    while left > 0:                          # potential energies are due at the end of every interval
        dev.timestep(min(left, args.print_rate), dt)
        left -= args.print_rate
    t_local = dev.timer_stop() * 1e-3
    force_ms.append(dev.kernel_stats()["last_force_ms"])
    l1 = dev.kernel_stats()["launches"]
    if dist is not None:
        import torch
        t = torch.tensor([t_local], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_max = float(t.item())
    else:
        t_max = t_local
    clocks = sampler.stop() if rank == 0 else None
    e1 = dev.energy()
    # the same K steps on eam.c's own schedule (pair energies evaluated at every step, not only where they are
    # read): reported beside `value` so that nobody has to take the on-demand energies on trust
    dev.set_option("energy_every_step", 1)
    barrier()
    dev.timer_start()
    left = args.steps
    while left > 0:
        dev.timestep(min(left, args.print_rate), dt)
        left -= args.print_rate
    t_every = dev.timer_stop() * 1e-3
    dev.set_option("energy_every_step", 0)
    if dist is not None:
        import torch
        t = torch.tensor([t_every], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_every = float(t.item())
    n_after = dev.num_local()

    # ---- e2e: the same step through HOST buffers (H2D of gid,r,p,f + D2H of gid,r,p,f,U inside the timed region)
    d0 = dev.download()
    cols = lambda x: [np.ascontiguousarray(x[:, k]) for k in range(3)]
    # host buffers of the caller live in pinned memory (comd_host_alloc), as the e2e contract asks
    _pins = []
    def pin(a):
        pa, v = pycomd.pinned_like(np.ascontiguousarray(a)); _pins.append(pa); return v
    hg, hr, hp, hf = pin(d0["gid"]), [pin(c) for c in cols(d0["r"])], [pin(c) for c in cols(d0["p"])], [pin(c) for c in cols(d0["f"])]
    # a trajectory driven through host memory: the caller feeds each call's results (pinned) to the next call
    m, o, _ = dev.step_host(dt, hg, hr, hp, hf)          # warm-up (allocates pinned staging; full upload path)
    barrier()
    r0 = dev.step_host_resumed()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        m, o, _ = dev.step_host(dt, o["gid"][:m], [a[:m] for a in o["r"]], [a[:m] for a in o["p"]], [a[:m] for a in o["f"]])
    t_e2e = time.perf_counter() - t0
    e2e_resumed = dev.step_host_resumed() - r0
    # the same chain with the library waiting for each upload before it starts the step (what every call did before
    # the speculative start existed, and what calls on more than one rank still do)
    dev.set_option("e2e_speculate", 0)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        m, o, _ = dev.step_host(dt, o["gid"][:m], [a[:m] for a in o["r"]], [a[:m] for a in o["p"]], [a[:m] for a in o["f"]])
    t_e2e_wait = time.perf_counter() - t0
    if dist is not None:
        import torch
        t = torch.tensor([t_e2e], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_e2e = float(t.item())
        t = torch.tensor([t_e2e_wait], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_e2e_wait = float(t.item())
    h2d = n_local * (4 + 9 * 8)
    d2h = n_local * (4 + 10 * 8) + 16              # local atoms only (compacted on the device): gid + r,p,f,U, and the two energies

    if rank != 0:
        return
    value = n_global * args.steps / t_max
    peaks = measured_peaks()
    fms = float(np.mean(force_ms))
    flops_force = FORCE_FLOP_PER_ATOM[pot] * n_local
    achieved_tf = flops_force / (fms * 1e-3) * 1e-12 if fms > 0 else None
    hbm_gbs = value / world * HBM_B_PER_ATOM_STEP[pot] * 1e-9
    out = {
        "metric": "atom-steps/s", "value": value, "unit": "atom-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_max / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "f64", "data": "synthetic FCC Cu lattice (a=3.615 A), numpy-seeded Maxwell-Boltzmann momenta at 600 K, dt=1 fs",
        "config": {"workload": label, "atoms": n_global, "atoms_per_gpu": n_local, "proc_grid": "%dx%dx%d" % grid,
                   "cells_per_gpu": "%dx%dx%d" % tuple(geom.gridSize[:]), "parallelism": "3-D spatial domain decomposition, 1 rank/GPU",
                   "force_path": ("neighbour list (skin %g*cutoff, rebuilt when an atom moved > skin/2)" % args.skin) if args.neighbor_list else "link cells (re-binned every step)",
                   "energies": "potential + kinetic energy every %d steps (CoMD printRate)" % args.print_rate,
                   "l2": "working set %.0f MB per GPU > 126 MB L2 (no flush needed)" % (n_local * 200e-6) if n_local * 200 > 126e6
                         else "working set fits L2: small-config line, not the headline"},
        "roofline": {"bound": "fp64", "kernel": "force (%s)" % (("k_force_nl<LJ>" if pot == "lj" else "k_force_nl<RHO> + F' halo + k_force_nl<EAMF|EAMFE>") if args.neighbor_list
                                                                 else ("k_force_tile<LJ>" if pot == "lj" else "k_force_tile<EAM1> + F' halo + k_force_tile<EAM3>")),
                     "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s",
                     "frac": (achieved_tf / fp64_peak) if achieved_tf else None,
                     "peak_source": "DFMA microbenchmark run in this process (comd_dev_fp64_peak); MEASURED_PEAKS.json has no FP64 entry",
                     "algorithmic_flop_per_atom": FORCE_FLOP_PER_ATOM[pot], "force_ms_per_step": fms,
                     "step_frac_fp64": value / world * FLOP_PER_ATOM_STEP[pot] * 1e-12 / fp64_peak,
                     "hbm": {"achieved": hbm_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": hbm_gbs / peaks["hbm_gbs"],
                             "peak_source": peaks["source"], "algorithmic_bytes_per_atom_step": HBM_B_PER_ATOM_STEP[pot]},
                     "traffic": (NCU_DRAM_B_PER_ATOM_FORCE.get((pot, int(bool(args.neighbor_list)))) or 0) * n_local or None,
                     "traffic_source": "dram__bytes_read+write per force evaluation from ncu --set full at 864k atoms (profiles/r1_v4_nl_ncu_full_eam60.txt, r1_v6_nl_ncu_full_lj60.txt), scaled per atom"
                                       if NCU_DRAM_B_PER_ATOM_FORCE.get((pot, int(bool(args.neighbor_list)))) else None},
        "e2e": {"value": n_global * args.e2e_steps / t_e2e, "unit": "atom-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": args.e2e_steps, "chained": True, "resumed_steps": e2e_resumed,
                "value_upload_first": n_global * args.e2e_steps / t_e2e_wait,
                "note": "each call's host results (pinned, updated in place) are the next call's host inputs; a resumed step reuses the "
                        "device-side sort and neighbour list (valid by the skin/2 criterion), the others re-sort and rebuild.  On one rank the "
                        "step starts from the atoms the device still holds while the inputs cross PCIe and is kept only if the uploaded "
                        "inputs equal them bit for bit (otherwise redone from the upload); value_upload_first = the same chain with that "
                        "switched off (every step waits for its upload)"},
        "value_energies_every_step": n_global * args.steps / t_every,
        "gpu_launches": int(l1 - l0),
        "clocks": clocks,
        "check": {"e_total_per_atom_before": (e0[0] + e0[1]) / n_local, "e_total_per_atom_after": (e1[0] + e1[1]) / n_local,
                  "atoms_local_after": n_after, "rank0_only": world > 1},
    }
    if not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_reference_rate(pot)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
