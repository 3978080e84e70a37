#!/bin/bash
# Round-2 ncu captures (run under gpurun on ONE GPU; every target is first run plainly and must exit 0).
# Outputs under gpurun_out/r2_ncu/.
set -u
out=gpurun_out/r2_ncu
mkdir -p $out
NCU="ncu --clock-control none"
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-also > $out/bench_plain_64m.json 2> $out/bench_plain_64m.err || exit 1
$NCU --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum -c 70 --csv --log-file $out/launches_64m-u.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-also > $out/ncu_launches_64m.log 2>&1
python scripts/run_scene.py --workload 16m-u --ticks 5 > $out/run_16m.log 2>&1 || exit 2
$NCU --set full --import-source on -k regex:"k_density|k_force_integrate|k_rank_gather_cells|k_cell_slots" -s 12 -c 4 -f -o $out/full_16m-u \
    python scripts/run_scene.py --workload 16m-u --ticks 5 > $out/ncu_full_16m.log 2>&1
python scripts/run_scene.py --workload 1m-d --ticks 326 > $out/run_1md.log 2>&1 || exit 3
$NCU --set full --import-source on -k regex:"k_density|k_force_integrate" -s 644 -c 4 -f -o $out/full_1m-d_tick323 \
    python scripts/run_scene.py --workload 1m-d --ticks 326 > $out/ncu_full_1md.log 2>&1
$NCU --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum -k regex:"k_" -s 1600 -c 400 --csv --log-file $out/launches_1m-d_late.csv \
    python scripts/run_scene.py --workload 1m-d --ticks 326 > $out/ncu_launches_1md.log 2>&1
ls -la $out
