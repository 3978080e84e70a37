#!/usr/bin/env python
"""Blocks of SASS with equal execution counts from `ncu -i rep --page source --csv --kernel-name regex:K`: warp / thread
instruction shares, active threads per instruction and stall-sample shares per block (which loop costs what)."""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
launch = int(sys.argv[3]) if len(sys.argv) > 3 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
sections, cur, hdr = [], None, None
for r in rows:
    if r and r[0] == "Address":
        hdr = r
        cur = []
        sections.append(cur)
    elif cur is not None and len(r) == len(hdr):
        cur.append(r)
data = sections[launch]
ia, it, iss, isrc = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
tot = sum(int(r[ia]) for r in data); tott = sum(int(r[it]) for r in data); tots = sum(int(r[iss]) for r in data)
print(f"{kern}: {len(data)} SASS instructions, {tot} warp instructions, {tott / tot:.2f} active threads per instruction, {tots} stall samples")
prev, start, blocks = None, 0, []
for k, r in enumerate(data):
    c = int(r[ia])
    if prev is not None and abs(c - prev) > 0.15 * max(c, prev, 1):
        blocks.append((start, k - 1))
        start = k
    prev = c
blocks.append((start, len(data) - 1))
print("| SASS range | instructions | executions of each | share of warp instructions | active threads / instruction | share of stall samples | first instruction |")
print("|---|---|---|---|---|---|---|")
for a, b in blocks:
    w = sum(int(r[ia]) for r in data[a:b + 1]); t = sum(int(r[it]) for r in data[a:b + 1]); s = sum(int(r[iss]) for r in data[a:b + 1])
    if w / tot > 0.004:
        print(f"| {a}-{b} | {b - a + 1} | {int(data[a][ia])} | {100 * w / tot:.1f} % | {t / max(w, 1):.1f} | {100 * s / max(tots, 1):.1f} % | `{data[a][isrc].strip()[:60]}` |")
