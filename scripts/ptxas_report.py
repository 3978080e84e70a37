#!/usr/bin/env python
"""Registers / spills / shared memory per kernel from `nvcc -Xptxas -v` (forces a rebuild of libfluid_b200.so).
usage: python scripts/ptxas_report.py [substring-of-kernel-name]"""
import os
import re
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_fluid_simulation_b200 import _build  # noqa: E402

want = sys.argv[1] if len(sys.argv) > 1 else ""
import io, contextlib
buf = io.StringIO()
with contextlib.redirect_stdout(buf):
    _build.build(force=True, verbose=True)
lines = buf.getvalue().splitlines()
for i, ln in enumerate(lines):
    m = re.search(r"Function properties for (\S+)", ln)
    if not m or want not in m.group(1):
        continue
    name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
    spill = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", lines[i + 1])
    regs = None
    for ln2 in lines[i + 1:i + 5]:
        regs = regs or re.search(r"Used (\d+) registers", ln2)
    print(f"{name:70s} regs {(regs.group(1) if regs else '-'):>3s} stack {spill.group(1):>3s} spill st/ld {spill.group(2)}/{spill.group(3)}")
