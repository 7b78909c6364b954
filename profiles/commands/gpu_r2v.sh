#!/bin/bash
# N-GPU pass: slab-decomposition checks (incl. sharded transfers), multi-GPU tests, bench at N (parity block, sharded e2e)
O=gpurun_out/${1:-r2v}; mkdir -p $O
N=${2:-4}
nvidia-smi -L > $O/gpus.txt
for case in "dambreak@200 4" "dambreak_fine 2" "nx=800 1"; do
  set -- $case
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
      tools/multi_gpu_check.py $1 $2 > $O/check_$1.log 2>&1
  echo "== $1 exit $?"; grep "^{" $O/check_$1.log | tail -1 | cut -c1-1400; grep -E "Error|error|Traceback" $O/check_$1.log | head -5
done
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q -x > $O/pytest_multi.log 2>&1; echo "pytest exit $?" >> $O/pytest_multi.log
tail -4 $O/pytest_multi.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 \
    bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err
echo "bench exit $?"; python - <<PY
import json
d=json.load(open("$O/bench_n$N.json"))
for k in ("value","ms_per_step","us_per_cg_iteration","phase_ms_per_step","cg_cycle_split","sync_us_per_iteration","parity_vs_1gpu","e2e","clocks"): print(k, d[k])
print(d["roofline"]["frac"], d["roofline"]["streamed_gbs"])
PY
tail -3 $O/bench_n$N.err
