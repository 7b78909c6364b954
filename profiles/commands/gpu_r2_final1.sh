#!/bin/bash
# round 2, final 1-GPU pass: GPU suite, bench line, ncu launch list, ncu DRAM traffic of the solve kernel
O=gpurun_out/${1:-r2fin}; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,memory.total --format=csv > $O/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench exit $?"
cut -c1-900 $O/bench.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-other > $O/ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none \
    -k regex:pcg_kernel -c 1 --csv --log-file $O/pcg_traffic.csv python tools/pcg_probe.py 2832 > $O/ncu_traffic.log 2>&1; echo "ncu traffic exit $?"
tail -3 $O/pcg_traffic.csv | cut -c1-400
