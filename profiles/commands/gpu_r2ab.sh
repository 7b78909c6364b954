#!/bin/bash
# final 2-GPU pass: multi-GPU tests (slab checks, C++ host sharded + replicated), bench at N = 2
O=gpurun_out/${1:-r2ab}; mkdir -p $O
timeout 1200 python -m pytest tests/test_multi_gpu.py -m gpu -q -x > $O/pytest_multi.log 2>&1; echo "pytest exit $?" >> $O/pytest_multi.log
tail -4 $O/pytest_multi.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 \
    bench.py --gpus 2 --steps 5 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err
echo "bench exit $?"; python - <<PY
import json
d=json.load(open("$O/bench_n2.json"))
for k in ("value","ms_per_step","us_per_cg_iteration","phase_ms_per_step","cg_cycle_split","sync_us_per_iteration","parity_vs_1gpu","e2e","clocks"): print(k, d[k])
PY
bash tools/cli_two_gpu.sh ${1:-r2ab}_cli
