#!/bin/bash
# round 2: L2-resident share of the matrix stream (MPSB_L2_RESIDENT_MB) at the bench size and at the per-rank size of
# an 8-GPU run (nx=1000: ~0.5 M rows, 145 MB of matrix), after the parity suite
O=gpurun_out/${1:-r2g}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
for nx in 2832 1000; do
  for mb in ${MBS:-0 48 80 104 120}; do
    MPSB_L2_RESIDENT_MB=$mb timeout 300 python tools/pcg_probe.py $nx > $O/probe_nx${nx}_mb$mb.log 2>&1
    echo "== nx=$nx resident=$mb MB: $(grep 'us/it' $O/probe_nx${nx}_mb$mb.log | tail -1 | cut -c1-200)"
  done
done
