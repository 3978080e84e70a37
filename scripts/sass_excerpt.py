#!/usr/bin/env python
"""SASS excerpts of the neighbour passes' loops from the built library (profiles/README.md recipe).
usage: python scripts/sass_excerpt.py <mangled-name-substring> <must-contain-mnemonic> [min-len] [max-len]
Prints the innermost backward-branch loop of that kernel whose body contains the mnemonic, plus every UBLKCP / SYNCS line."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "gpu_fluid_simulation_b200", "libfluid_b200.so")
want, must = sys.argv[1], sys.argv[2]
lo, hi = (int(sys.argv[3]) if len(sys.argv) > 3 else 8), (int(sys.argv[4]) if len(sys.argv) > 4 else 400)
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout.splitlines()
fn, body = None, []
for ln in out:
    m = re.search(r"Function : (\S+)", ln)
    if m:
        fn = m.group(1)
        continue
    if fn and want in fn and re.search(r"/\*[0-9a-f]{4}\*/\s+\S", ln):
        body.append(ln)
ins = []
for ln in body:
    m = re.search(r"/\*([0-9a-f]{4})\*/\s+(.*?);", ln)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
addr = {a: k for k, (a, _) in enumerate(ins)}
loops = []
for k, (a, t) in enumerate(ins):
    m = re.search(r"BRA(?:\.\S+)?\s+(?:P\d+, )?0x([0-9a-f]+)", t)
    if m and int(m.group(1), 16) in addr and addr[int(m.group(1), 16)] < k:
        j = addr[int(m.group(1), 16)]
        if lo <= k - j + 1 <= hi and any(must in x for _, x in ins[j:k + 1]):
            loops.append((k - j + 1, j, k))
print(f"// {want}: {len(ins)} SASS instructions; staging / barrier instructions:")
for a, t in ins:
    if "UBLKCP" in t or "SYNCS" in t:
        print(f"//   /*{a:04x}*/ {t}")
if not loops:
    sys.exit("no loop with " + must)
n, j, k = min(loops)
print(f"// innermost loop containing {must}: {n} instructions, /*{ins[j][0]:04x}*/ .. /*{ins[k][0]:04x}*/")
for a, t in ins[j:k + 1]:
    print(f"/*{a:04x}*/  {t}")
