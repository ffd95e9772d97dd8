"""GPU: the CoMD-compatible C driver (host/CoMD.c -> C ABI -> sm_100a kernels) next to the
reference's own executable (oracle/_ref/CoMD-ref, built from /root/reference's sources, -m cpu_nl)
on the BASELINE configs C1 / C2: same command line, same printed energy trace."""
import os
import re
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "comd-cuda-async_b200")
REF = os.path.join(ROOT, "oracle", "_ref", "CoMD-ref")
POTS = os.path.join(ROOT, "pots")
LINE = re.compile(r"^\s+(\d+)\s+([0-9.]+)\s+(-?[0-9.]+)\s+(-?[0-9.]+)\s+([0-9.]+)\s+([0-9.]+)\s+([0-9.]+)\s+(\d+)\s*$")


def trace(out):
    rows = []
    for ln in out.splitlines():
        m = LINE.match(ln)
        if m:
            rows.append((int(m.group(1)), float(m.group(3)), float(m.group(4)), float(m.group(5)), int(m.group(8))))
    return rows


def run(exe, args, cwd):
    res = subprocess.run([exe] + args, cwd=cwd, capture_output=True, timeout=600)
    # the reference prints an uninitialised species name for setfl files (eam.c:680-757 never sets it)
    out = res.stdout.decode("utf-8", "replace")
    assert res.returncode == 0, out[-2000:] + res.stderr.decode("utf-8", "replace")[-2000:]
    return out


@pytest.mark.parametrize("flags,method", [([], "thread_atom"), (["-e"], "thread_atom"), (["-e"], "thread_atom_nl"), (["-e", "-t", "setfl"], "warp_atom_nl")])
def test_driver_trace_matches_reference_executable(tmp_path, flags, method):
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref/CoMD-ref not built (needs /root/reference at build time)")
    subprocess.check_call(["make", "-s", "-C", PKG, "host"])
    n = "12" if "setfl" in flags else "20"
    common = ["-x", n, "-y", n, "-z", n, "-N", "100", "-n", "10", "-d", POTS] + flags
    ours = trace(run(os.path.join(PKG, "CoMD-b200"), common + ["-m", method], tmp_path))
    ref = trace(run(REF, common + ["-m", "cpu_nl"], tmp_path))
    assert len(ours) == len(ref) == 11
    for (s0, et0, eu0, ek0, n0), (s1, et1, eu1, ek1, n1) in zip(ours, ref):
        assert s0 == s1 and n0 == n1 == 4 * int(n) ** 3
        tol = 2e-12 if s0 == 0 else 1e-6                       # printed with 12 decimals; 1e-6 over 100 steps (north_star)
        assert abs(et0 - et1) <= tol * abs(et1) + 1.5e-12, (s0, et0, et1)
        assert abs(eu0 - eu1) <= tol * abs(eu1) + 1.5e-12, (s0, eu0, eu1)
        assert abs(ek0 - ek1) <= max(tol, 1e-6 if s0 else 0) * abs(ek1) + 1.5e-12, (s0, ek0, ek1)


@pytest.mark.parametrize("flags,method", [([], "thread_atom_nl"), (["-e"], "thread_atom"), (["-e", "-t", "setfl"], "thread_atom_nl")])
def test_single_precision_driver_trace_matches_the_float_reference(tmp_path, flags, method):
    """CoMD-b200-f32 (host/ and csrc/ built with -DCOMD_SINGLE, real_t = float end to end, mytype.h:8-24) beside the
    reference's own float build (oracle/_ref/CoMD-ref-f32, -m cpu_nl) and beside its double build.
    The float reference adds every pair's energy into ONE float accumulator (ljForce.c:222, eam.c:343,369): on 6912 atoms
    its printed energies sit 5e-5 (LJ) to 1.1e-3 (Cu_u6) away from its own double build and wander by 3e-4 along the
    run (measured here, CoMD-ref vs CoMD-ref-f32) -- so 2e-3 is what a comparison WITH it can resolve.  This build sums per
    atom in float and over atoms in double, which lands on the double reference: 2e-5 on the total and potential energy
    over 100 steps (tests/test_gpu_precision.py states the per-atom float tolerances)."""
    ref32 = REF + "-f32"
    if not os.path.exists(ref32):
        pytest.skip("oracle/_ref/CoMD-ref-f32 not built (needs /root/reference at build time)")
    subprocess.check_call(["make", "-s", "-C", PKG, "host"])
    n = "12"
    common = ["-x", n, "-y", n, "-z", n, "-N", "100", "-n", "10", "-d", POTS] + flags
    out32 = run(os.path.join(PKG, "CoMD-b200-f32"), common + ["-m", method], tmp_path)
    # the float build quotes its roofline against the FFMA pipe it computes on, not the DFMA peak (host/CoMD.c, host/output.c)
    assert " FP32 roofline:" in out32 and "measured FFMA peak" in out32 and "FP64 roofline" not in out32
    ours = trace(out32)
    for exe, tol in ((ref32, 2e-3), (REF, 2e-5)):
        ref = trace(run(exe, common + ["-m", "cpu_nl"], tmp_path))
        assert len(ours) == len(ref) == 11
        for (s0, et0, eu0, ek0, n0), (s1, et1, eu1, ek1, n1) in zip(ours, ref):
            assert s0 == s1 and n0 == n1 == 4 * int(n) ** 3
            assert abs(et0 - et1) <= tol * abs(et1), (exe, s0, et0, et1)
            assert abs(eu0 - eu1) <= tol * abs(eu1), (exe, s0, eu0, eu1)
            assert abs(ek0 - ek1) <= tol * abs(eu1), (exe, s0, ek0, ek1)


@pytest.mark.parametrize("flags,method", [([], "thread_atom"), (["-e"], "thread_atom"), (["-e"], "thread_atom_nl"), (["-e", "-t", "setfl"], "warp_atom_nl")])
def test_reference_host_code_drives_the_library_through_its_own_control_flow(tmp_path, flags, method):
    """The drop-in proof of INTEGRATION.md section 2: oracle/_ref/CoMD-ref-b200 = the reference's UNMODIFIED host C (main,
    initSimulation, timestep, redistributeAtomsGpu[NL], eamForceGpu / ljForce, kineticEnergyGpu, command line, print-out)
    with its device entry points bound to libcomd_b200.so by oracle/ref_build/ref_bind_b200.c.  Run with the reference's GPU
    methods, its trace must be the trace of the reference's CPU path (-m cpu_nl): same gates as CoMD-b200 above."""
    bound = REF + "-b200"
    if not os.path.exists(bound) or not os.path.exists(REF):
        pytest.skip("oracle/_ref/CoMD-ref-b200 not built (needs /root/reference at build time)")
    n = "12"
    common = ["-x", n, "-y", n, "-z", n, "-N", "100", "-n", "10", "-d", POTS] + flags
    ours = trace(run(bound, common + ["-m", method], tmp_path))
    ref = trace(run(REF, common + ["-m", "cpu_nl"], tmp_path))
    assert len(ours) == len(ref) == 11
    for (s0, et0, eu0, ek0, n0), (s1, et1, eu1, ek1, n1) in zip(ours, ref):
        assert s0 == s1 and n0 == n1 == 4 * int(n) ** 3
        tol = 2e-12 if s0 == 0 else 1e-6
        assert abs(et0 - et1) <= tol * abs(et1) + 1.5e-12, (s0, et0, et1)
        assert abs(eu0 - eu1) <= tol * abs(eu1) + 1.5e-12, (s0, eu0, eu1)
        assert abs(ek0 - ek1) <= max(tol, 1e-6 if s0 else 0) * abs(ek1) + 1.5e-12, (s0, ek0, ek1)


def test_driver_async_flag_gives_the_same_trace(tmp_path):
    subprocess.check_call(["make", "-s", "-C", PKG, "host"])
    args = ["-x", "10", "-y", "10", "-z", "10", "-N", "30", "-n", "10", "-e", "-d", POTS]
    a = trace(run(os.path.join(PKG, "CoMD-b200"), args, tmp_path))
    b = trace(run(os.path.join(PKG, "CoMD-b200"), args + ["-a", "1"], tmp_path))
    assert a == b and len(a) == 4


@pytest.mark.parametrize("method", ["thread_atom", "thread_atom_nl"])
def test_driver_reproduces_the_trace_the_reference_published(tmp_path, method):
    """The one full-size energy trace the reference ships (its own GPU kernels on 16 K20 ranks, EAM Cu_u6 80^3,
    100 steps; tests/golden/ref_log_eam_cu_u6_80x80x80.txt) against the same command line on one B200.
    Step 0 agrees in all 12 printed decimals.  Later steps agree to 4e-6 (total), 5.4e-6 (potential) and
    2.5e-4 (kinetic) relative: the published run is the reference's GPU path, whose trajectory drifts away from
    the reference's CPU path -- the oracle of this repository -- by that much (the CPU path's total energy reads
    -3.460522629, -3.460524223, -3.460527823 ... at steps 10, 20, 30, which is what this build prints to 1e-8;
    the GPU log reads -3.460520822, -3.460517911, -3.460516645).  So the bounds below are what two implementations
    of the same model owe each other over 100 steps, not the 1e-6 of the CPU-path parity tests above."""
    subprocess.check_call(["make", "-s", "-C", PKG, "host"])
    gold = trace(open(os.path.join(ROOT, "tests", "golden", "ref_log_eam_cu_u6_80x80x80.txt")).read())
    ours = trace(run(os.path.join(PKG, "CoMD-b200"), ["-e", "-x", "80", "-y", "80", "-z", "80", "-d", POTS, "-m", method], tmp_path))
    assert len(gold) == len(ours) == 11
    for (s0, et0, eu0, ek0, n0), (s1, et1, eu1, ek1, n1) in zip(ours, gold):
        assert s0 == s1 and n0 == n1 == 2048000
        if s0 == 0:
            assert abs(et0 - et1) <= 1.5e-12 and abs(eu0 - eu1) <= 1.5e-12 and abs(ek0 - ek1) <= 1.5e-12, (et0, et1, eu0, eu1, ek0, ek1)
            continue
        assert abs(et0 - et1) <= 1e-5 * abs(et1), (s0, et0, et1)
        assert abs(eu0 - eu1) <= 1e-5 * abs(eu1), (s0, eu0, eu1)
        assert abs(ek0 - ek1) <= 5e-4 * abs(ek1), (s0, ek0, ek1)
    # energy conservation as the reference states it (CoMD.c:914-945): initial vs final total energy
    assert abs(ours[-1][1] - ours[0][1]) <= 5e-6 * abs(ours[0][1])
