#!/bin/bash
# round 2, first pass: the large-N reference pins + ncu of the non-solve kernels as they stood at the end of round 1
O=gpurun_out/${1:-r2a}; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -s -k "large_generated" > $O/pytest_large.log 2>&1; echo "pytest exit $?" >> $O/pytest_large.log
grep -E "iterations|passed|failed|Error|assert" $O/pytest_large.log | tail -40
MPSB_CG_ITERATION_CAP=8 timeout 300 python tools/step_probe.py 2832 3 > $O/step_probe.log 2>&1; cat $O/step_probe.log
MPSB_CG_ITERATION_CAP=8 timeout 900 ncu --set full --clock-control none --import-source on \
   -k regex:'neighbor_kernel|gradient_kernel|key_kernel|radix_hist|radix_scatter|exclusive_scan|matrix_kernel|explicit_kernel|flags_kernel|cell_range' \
   --launch-skip 12 -c 44 -o $O/nonsolve_full -f python tools/step_probe.py 2832 2 > $O/ncu_nonsolve.log 2>&1
tail -3 $O/ncu_nonsolve.log; ls -la $O
