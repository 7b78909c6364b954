#!/bin/bash
# round 2, multi-GPU pass: whole GPU suite on an N-GPU box (the multi-GPU tests run), slab-decomposition checks, bench at N
O=gpurun_out/${1:-r2d}; mkdir -p $O
N=${2:-2}
nvidia-smi -L > $O/gpus.txt
for case in "dambreak 6" "nx=400 2"; do
  set -- $case
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
      tools/multi_gpu_check.py $1 $2 > $O/check_$1.log 2>&1
  echo "== $1 exit $?"; grep "^{" $O/check_$1.log | tail -1 | cut -c1-1500; grep -E "Error|error|Traceback|mps_b200 error" $O/check_$1.log | head -5
done
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -25 $O/pytest_gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 \
    bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err
echo "bench exit $?"; cat $O/bench_n$N.json | cut -c1-3000; tail -5 $O/bench_n$N.err
