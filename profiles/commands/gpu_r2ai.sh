#!/bin/bash
O=gpurun_out/${1:-r2ai}; mkdir -p $O
for NX in 2832 1416; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29620 \
      tools/pcg_probe_multi.py $NX > $O/probe_nx$NX.log 2>&1
  echo "== nx=$NX"; grep "us/it" $O/probe_nx$NX.log
done
