#!/bin/bash
# single-reduction recurrence: its tests first (bounded), then the whole suite, then the solve in both modes
O=gpurun_out/${1:-r2w}; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "single_reduction and not nx1416 and not nx2832" > $O/pytest_sr.log 2>&1; rc=$?; echo "pytest sr exit $rc" >> $O/pytest_sr.log
tail -6 $O/pytest_sr.log
[ $rc -ne 0 ] && exit 1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
for NX in 1000 1416 2832; do for sr in 0 1; do
  MPSB_CG_ITERATION_CAP=6000 MPSB_CG_SINGLE_REDUCTION=$sr timeout 120 python tools/pcg_probe.py $NX > $O/probe_sr${sr}_nx$NX.log 2>&1
  echo "== sr=$sr nx=$NX: $(grep 'us/it' $O/probe_sr${sr}_nx$NX.log | tail -1 | cut -c1-200)"
done; done
