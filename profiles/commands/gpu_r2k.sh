#!/bin/bash
# all-to-all flag-line barriers: parity suite, then the solve at several sizes
O=gpurun_out/${1:-r2k}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
for NX in 708 1000 1416 2832; do
  timeout 300 python tools/pcg_probe.py $NX > $O/probe_nx$NX.log 2>&1
  echo "== nx=$NX: $(grep 'us/it' $O/probe_nx$NX.log | tail -1 | cut -c1-220)"
  grep "per-CTA cycles/it sync" $O/probe_nx$NX.log
done
