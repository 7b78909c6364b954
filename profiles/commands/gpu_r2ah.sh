#!/bin/bash
O=gpurun_out/${1:-r2ah}; mkdir -p $O
for NX in 708 1000 1416 2832; do
  timeout 300 python tools/pcg_probe.py $NX > $O/probe_nx$NX.log 2>&1
  echo "== nx=$NX: $(grep 'us/it' $O/probe_nx$NX.log | tail -1 | cut -c1-220)"
done
