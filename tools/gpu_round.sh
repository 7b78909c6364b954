#!/bin/bash
# One GPU-box pass: parity tests, bench, ncu launch list, ncu DRAM traffic + full set of the PCG kernel.
# Usage (from the repo root on the GPU box): bash tools/gpu_round.sh <tag>
TAG=${1:-run}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,memory.total --format=csv > $O/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench exit $?" >> $O/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
# launch list (per-launch durations, cold-cache serialised) of the same command
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > $O/ncu_launches.log 2>&1
# DRAM traffic of the PCG kernel at the bench size (few passes)
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none \
    -k regex:pcg -c 1 --csv --log-file $O/pcg_traffic.csv python tools/pcg_probe.py 2832 > $O/ncu_traffic.log 2>&1
# full set at 1/4 size (replays ~40x)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pcg -c 1 -o $O/pcg_full_nx1416 -f \
    python tools/pcg_probe.py 1416 > $O/ncu_full.log 2>&1
ls -la $O
