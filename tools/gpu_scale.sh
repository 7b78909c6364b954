#!/bin/bash
# Multi-GPU pass on an N-GPU box: split-solve parity tests, then bench at 2, 4, 8 ranks (as the driver launches it).
TAG=${1:-scale}; MAXN=${2:-8}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > $O/pytest_multi.log 2>&1; echo "pytest exit $?" >> $O/pytest_multi.log
tail -5 $O/pytest_multi.log
for N in 2 4 8; do
  [ $N -le $MAXN ] || continue
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+N)) \
      bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err
  echo "N=$N exit $?"; tail -c 1500 $O/bench_n$N.json | head -c 1200; echo
done
