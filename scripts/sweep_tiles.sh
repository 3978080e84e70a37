#!/bin/bash
# CTA size of the neighbour passes (FLUID_SPH_THREADS) on the uniform box and on the late, crowded dam break; run under gpurun
run() {  # $1 label, rest: bench args
  l=$1; shift
  timeout 150 python bench.py "$@" --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$l', d['ms_per_step'], d['config']['sort_path'], d['roofline']['phases_ms'])"
}
(timeout 300 python -m pytest tests/test_gpu_sort_paths.py tests/test_gpu_parity.py -x -q 2>&1 | tail -1)
for v in "-DFLUID_SPH_THREADS=256" "-DFLUID_SPH_THREADS=128 -DFLUID_FORCE_MIN_CTAS=8 -DFLUID_SPH_CAP=1024"; do
  FLUID_NVCC_EXTRA="$v" python gpu_fluid_simulation_b200/_build.py --force > /dev/null
  echo "== $v"
  run 1m-d-late --workload 1m-d --steps 300 --warmup 300
  run 1m-d --workload 1m-d --steps 200 --warmup 20
  run 1m-u --workload 1m-u --steps 200 --warmup 20
  run 64m-u --steps 10 --warmup 3
done
python gpu_fluid_simulation_b200/_build.py --force > /dev/null
