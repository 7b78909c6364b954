#!/bin/bash
# round 2, 1-GPU pass of the restored tree: GPU suite, bench line, ncu launch list of the same command
O=gpurun_out/${1:-r2f}; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,memory.total --format=csv > $O/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench exit $?"
cut -c1-2500 $O/bench.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-other > $O/ncu_launches.log 2>&1; echo "ncu exit $?"
