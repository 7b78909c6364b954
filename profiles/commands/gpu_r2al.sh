#!/bin/bash
# dual ring geometry on 2 GPUs: correctness (nx=800 against the reference fixture, dambreak with collisions) and per-iteration time
O=gpurun_out/${1:-r2al}; mkdir -p $O
for case in "nx=800 1" "dambreak@200 3"; do
  set -- $case
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
      tools/multi_gpu_check.py $1 $2 > $O/check_$1.log 2>&1
  echo "== $1 exit $?"; grep "^{" $O/check_$1.log | tail -1 | cut -c1-700
done
for NX in 1416 2000; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29620 \
      tools/pcg_probe_multi.py $NX > $O/probe_nx$NX.log 2>&1
  echo "== nx=$NX"; grep "us/it" $O/probe_nx$NX.log
  MPSB_CG_SMALL_ROWS=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 \
      tools/pcg_probe_multi.py $NX > $O/probe_v0_nx$NX.log 2>&1
  echo "== nx=$NX (20x12 ring forced)"; grep "us/it" $O/probe_v0_nx$NX.log
done
