#!/bin/bash
# Parity tests + bench on the GPU box. Usage: bash tools/gpu_quick.sh <tag> [extra bench args]
TAG=${1:-quick}; shift
O=gpurun_out/$TAG; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -15 $O/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 "$@" > $O/bench.json 2> $O/bench.err; echo "bench exit $?"
cat $O/bench.json
