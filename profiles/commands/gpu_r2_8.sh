#!/bin/bash
# 8-GPU pass: slab checks, bench (default), bench with the single-reduction recurrence (short), per-rank probe
O=gpurun_out/${1:-r2_8}; mkdir -p $O
N=${2:-8}
nvidia-smi -L > $O/gpus.txt
for case in "dambreak@200 3" "nx=800 1"; do
  set -- $case
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
      tools/multi_gpu_check.py $1 $2 > $O/check_$1.log 2>&1
  echo "== $1 exit $?"; grep "^{" $O/check_$1.log | tail -1 | cut -c1-900; grep -E "Error|error|Traceback" $O/check_$1.log | head -5
done
show() { python - <<PY
import json
try:
    d=json.load(open("$1"))
    print("$1", {k:d[k] for k in ("value","ms_per_step","us_per_cg_iteration","cg_iterations_per_step","sync_us_per_iteration")}, d["phase_ms_per_step"], d["cg_cycle_split"], (d.get("parity_vs_1gpu") or {}).get("ok"), (d.get("e2e") or {}).get("value"), d["clocks"])
except Exception as e: print("no json $1", e)
PY
}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 \
    bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err
echo "bench exit $?"; show $O/bench_n$N.json
MPSB_CG_SINGLE_REDUCTION=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29545 \
    bench.py --gpus $N --steps 3 --warmup 2 --no-e2e --no-parity > $O/bench_sr_n$N.json 2> $O/bench_sr_n$N.err
echo "bench sr exit $?"; show $O/bench_sr_n$N.json
MPSB_L2_RESIDENT_KB=0 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29546 \
    bench.py --gpus $N --steps 3 --warmup 2 --no-e2e --no-parity > $O/bench_nol2_n$N.json 2> $O/bench_nol2_n$N.err
echo "bench no-l2-policy exit $?"; show $O/bench_nol2_n$N.json
