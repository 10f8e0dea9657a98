"""Code-generation guard (no GPU needed): the pair loops of the shipped direct-N^2 kernel must be the instruction
streams that were measured -- all-packed FP32, 12 packed instructions per two pairs in the general
loop and 11 in the one-mass loop, two MUFU.RSQ, one LDS.128 per source, no local-memory traffic inside a loop, and the
register budget that lets six CTAs share an SM.  A toolchain that compiles the same source differently shows up here
rather than as an unexplained bench regression."""
import collections
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "nbody_b200", "libnbody_b200.so")
KERNEL = "_ZN2nb19direct_force_kernelILi256ELi2ELi1024ELb0EEEvPK6float4S3_PfPjiiiiiify"

pytestmark = pytest.mark.skipif(shutil.which("cuobjdump") is None or not os.path.exists(LIB),
                                reason="needs cuobjdump and the built library")


def _loops(kernel):
    s = subprocess.run(["cuobjdump", "-sass", "-fun", kernel, LIB], capture_output=True, text=True).stdout
    ins = [(int(a, 16), t) for a, t in re.findall(r"/\*([0-9a-f]{4})\*/\s+(.*?);", s)]
    assert ins, "kernel not found in the library"
    loops = []
    for a, txt in ins:
        m = re.search(r"BRA(?:\.\w+)*\s+(?:!?UP\d+,\s*)?0x([0-9a-f]+)", txt)
        if not m or int(m.group(1), 16) >= a:
            continue
        t = int(m.group(1), 16)
        c = collections.Counter()
        for b, x in ins:
            if t <= b <= a:
                parts = x.split()
                op = parts[1] if parts[0].startswith("@") else parts[0]
                c[op.split(".")[0]] += 1
        loops.append(c)
    return loops


def test_direct_pair_loops_are_the_measured_instruction_streams():
    loops = _loops(KERNEL)
    # innermost pair loops: the two with exactly one LDS per two MUFU (64 sources per trip)
    pair = [c for c in loops if c["MUFU"] == 128 and c["LDS"] == 64]
    assert len(pair) == 2, [dict(c) for c in loops]
    general = max(pair, key=lambda c: c["FMUL2"])
    one_mass = min(pair, key=lambda c: c["FMUL2"])
    for c in pair:
        assert c["FFMA2"] == 384 and c["FADD2"] == 192
        assert c["LDL"] == 0 and c["STL"] == 0, "local-memory traffic inside a pair loop"
        assert c["FFMA"] == 0 and c["FMUL"] == 0 and c["FADD"] == 0, "scalar FP32 mixed into the packed stream"
    assert general["FMUL2"] == 192   # (G m) / r^3: three packed multiplies per two pairs
    assert one_mass["FMUL2"] == 128  # 1 / r^3: two


def test_direct_kernel_register_budget_allows_six_ctas_per_sm():
    out = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    m = re.search(re.escape(KERNEL) + r":\s*\n\s*REG:(\d+)\s+STACK:(\d+)\s+SHARED:(\d+)", out)
    assert m, "resource usage of the direct kernel not found"
    reg, stack, shared = map(int, m.groups())
    assert reg <= 40, reg                      # 65536 / (6 x 256 threads) = 42.7 -> 40
    assert 6 * shared <= 227 * 1024, shared    # six CTAs' tiles fit the SM's shared memory
    assert stack <= 32, stack                  # the Kahan terms only, outside the pair loops
