#!/bin/bash
# round 2, second pass: large-N pins on all three sweep-A variants, matrix-free operator tests, MF vs stored timing
O=gpurun_out/${1:-r2b}; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "large_generated or matrix_free" > $O/pytest_large.log 2>&1; echo "pytest exit $?" >> $O/pytest_large.log
grep -E "iterations|passed|failed|Error|assert|matrix-free entries" $O/pytest_large.log | tail -60
for mf in 0 1; do
  MPSB_MATRIX_FREE=$mf timeout 300 python tools/pcg_probe.py 2832 > $O/pcg_probe_mf$mf.log 2>&1
  echo "== MPSB_MATRIX_FREE=$mf"; grep "us/it" $O/pcg_probe_mf$mf.log | tail -2; grep "per-CTA" $O/pcg_probe_mf$mf.log
done
