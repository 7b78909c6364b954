#!/bin/bash
# round 2, third pass: whole GPU suite (new look-back scan), FP64/MUFU/gather rates, ncu of the matrix-free solve
O=gpurun_out/${1:-r2c}; mkdir -p $O
./build/fp64_rate > $O/fp64_rate.txt 2>&1; cat $O/fp64_rate.txt
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -8 $O/pytest_gpu.log
MPSB_CG_ITERATION_CAP=8 timeout 300 python tools/step_probe.py 2832 3 > $O/step_probe.log 2>&1; cat $O/step_probe.log
MPSB_MATRIX_FREE=1 MPSB_CG_ITERATION_CAP=12 timeout 600 ncu --set full --clock-control none --import-source on -k regex:pcg_kernel -c 1 \
   -o $O/pcg_mf_full -f python tools/pcg_probe.py 2832 > $O/ncu_mf.log 2>&1; tail -2 $O/ncu_mf.log
