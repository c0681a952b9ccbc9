#!/bin/bash
# Round-2 evidence run on one B200 (under gpurun): GPU test-suite, the bench line, the ncu launch list of one full-size step and
# one `--set full` capture of the two hot kernels at 2^20 x 665.  Outputs go to gpurun_out/ (copied to profiles/ by hand).
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 3 --warmup 3 > gpurun_out/r2_bench_d.json 2> gpurun_out/r2_bench_d.err
tail -2 gpurun_out/r2_bench_d.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_d.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e_full_host'], d['small_configs'], d['cpu_baseline']['value'], d['root_ok'])"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_H.csv \
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-small > gpurun_out/r2_ncu_launch.log 2>&1
tail -2 gpurun_out/r2_ncu_launch.log | cut -c1-300
wc -l gpurun_out/r2_launches_H.csv
ncu --set full --clock-control none --import-source on -k regex:"absorb_kernel|ntt_tile_kernel" -s 10 -c 11 -o gpurun_out/r2_prof \
    python bench.py --log-rows 20 --steps 1 --warmup 1 --no-e2e --no-cpu --no-small > gpurun_out/r2_ncu_full.log 2>&1
tail -2 gpurun_out/r2_ncu_full.log | cut -c1-300
ls -la gpurun_out/r2_prof.ncu-rep
