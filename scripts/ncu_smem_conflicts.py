#!/usr/bin/env python
"""Shared-memory wavefronts per SASS instruction from `ncu --page source`: which loads/stores take more wavefronts than the
ideal, and how many. usage: python scripts/ncu_smem_conflicts.py report.ncu-rep kernel-regex [launch]"""
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
        hdr, cur = r, []
        sections.append(cur)
    elif cur is not None and len(r) == len(hdr):
        cur.append(r)
data = sections[launch]
c = {k: hdr.index(k) for k in ("Source", "Instructions Executed", "Avg. Threads Executed", "L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal",
                               "L1 Wavefronts Shared Excessive", "L1 Conflicts Shared N-Way", "Access Size")}
tot_w = sum(int(r[c["L1 Wavefronts Shared"]] or 0) for r in data)
tot_x = sum(int(r[c["L1 Wavefronts Shared Excessive"]] or 0) for r in data)
print(f"{kern}: shared wavefronts {tot_w}, ideal {tot_w - tot_x}, excessive {tot_x} ({100.0 * tot_x / max(tot_w, 1):.1f} %)")
print("| SASS # | instruction | executed | avg threads | wavefronts | ideal | excessive | n-way |")
print("|---|---|---|---|---|---|---|---|")
for k, r in enumerate(data):
    w = int(r[c["L1 Wavefronts Shared"]] or 0)
    if w and w >= 0.01 * tot_w:
        print(f"| {k} | `{r[c['Source']].strip()[:60]}` | {r[c['Instructions Executed']]} | {r[c['Avg. Threads Executed']]} | {w} | {r[c['L1 Wavefronts Shared Ideal']]} | "
              f"{r[c['L1 Wavefronts Shared Excessive']]} | {r[c['L1 Conflicts Shared N-Way']]} |")
