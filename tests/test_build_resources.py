"""CPU: what the compiler made of the kernels -- read from the in-tree build (ptxas logs, cuobjdump), no GPU needed.
Guards the properties the measured numbers depend on: sm_100a code only, the force kernels inside their register
budgets without spills to speak of, and the Blackwell instructions the design relies on actually present."""
import glob
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "comd-cuda-async_b200")
LIB = os.path.join(PKG, "libcomd_b200.so")

ENTRY = re.compile(r"Compiling entry function '(\S+)' for 'sm_100a'\nptxas info\s+: Function properties for \S+\n"
                   r"\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\nptxas info\s+: Used (\d+) registers")


def kernels():
    logs = glob.glob(os.path.join(PKG, "build", "*.ptxas.log"))
    if not logs:
        pytest.skip("no ptxas logs: run __graft_entry__.build() first")
    out = {}
    for f in logs:
        for m in ENTRY.finditer(open(f).read()):
            out[m.group(1)] = dict(stack=int(m.group(2)), spill=int(m.group(3)) + int(m.group(4)), regs=int(m.group(5)))
    return out


def test_force_kernels_fit_their_register_budgets():
    k = kernels()
    nl = {n: v for n, v in k.items() if "k_force_nl" in n}
    tile = {n: v for n, v in k.items() if "k_force_tile" in n}
    assert len(nl) == 4 and len(tile) == 4, sorted(k)          # tile: LJ, EAM1, EAM3 and the density pass that writes the hit list
    for n, v in nl.items():
        # one CTA per SM: 1024 threads <= 64 registers (rhobar pass), 896 <= 72 (LJ, pass 3), 768 <= 80 (pass 3 + energies)
        # (k_force_nl<EAMF>, since round 2 the nl_n3 0 option path only, sits at its 72-register ceiling: 40 bytes of spill
        #  traffic per thread outside the pair loop since the block-order table joined the kernel parameters)
        assert v["regs"] <= 80 and v["spill"] <= 48, (n, v)
    kind = lambda n: int(re.search(r"ILi(\d)E", n).group(1))
    n3 = {n: v for n, v in k.items() if "k_force_n3" in n and "Lb1E" in n}           # the default kernels of the EAM list path (fast tables)
    assert len(n3) >= 6
    for n, v in n3.items():
        every_step = kind(n) in (4, 5, 7)                          # rhobar pass, force pass, LJ forces: run every step
        assert v["regs"] <= 72 and v["spill"] <= (0 if every_step else 32), (n, v)
    budget = {4: 64, 0: 72, 5: 72, 6: 80}
    for n, v in nl.items():
        assert v["regs"] <= budget[kind(n)], (n, v)
    for n, v in tile.items():
        assert v["spill"] == 0, (n, v)
    assert k[[n for n in k if "k_nl_build" in n][0]]["spill"] == 0


@pytest.mark.skipif(shutil.which("cuobjdump") is None, reason="cuobjdump not on PATH")
def test_library_is_sm100a_and_uses_the_instructions_the_design_names():
    if not os.path.exists(LIB):
        pytest.skip("library not built")
    elfs = subprocess.run(["cuobjdump", "-lelf", LIB], capture_output=True, text=True).stdout
    archs = set(re.findall(r"\.(sm_\w+)\.cubin", elfs))
    assert archs == {"sm_100a"}, archs
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    for mnemonic, why in (("LDG.E.ENL2.256", "32-byte position / table-record gathers of the list kernels"),
                          ("UBLKCP", "TMA bulk copies staging the stencil rows of the tiled link-cell kernels"),
                          ("SYNCS.ARRIVE.TRANS64", "mbarrier transaction counts behind those copies"),
                          ("CREDUX", "warp-wide reductions in one instruction (__reduce_max_sync)"),
                          ("DFMA", "FP64 pipe")):
        assert mnemonic in sass, (mnemonic, why)
    assert "HMMA" not in sass and "wgmma" not in sass.lower()       # FP64 pair arithmetic: no tensor-core detour
