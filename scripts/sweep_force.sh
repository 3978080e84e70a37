#!/bin/bash
# tuning sweep of the force kernel's compile-time knobs on a GPU box: rebuilds the library per variant and runs a short bench.
# usage (from the repo root, under gpurun): bash scripts/sweep_force.sh "<flags variant 1>" "<flags variant 2>" ...
out=gpurun_out/sweep_force.txt
: > $out
for v in "$@"; do
  FLUID_NVCC_EXTRA="$v" python gpu_fluid_simulation_b200/_build.py --force > /dev/null || { echo "$v BUILD FAILED" >> $out; continue; }
  python bench.py --workload ${SWEEP_WORKLOAD:-16m-u} --steps 10 --warmup 3 --no-cpu-baseline --no-e2e ${SWEEP_ARGS} 2>&1 | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v', d['ms_per_step'], d['roofline']['phases_ms'])" >> $out
done
python gpu_fluid_simulation_b200/_build.py --force > /dev/null
cat $out
