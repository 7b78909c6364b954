#!/bin/bash
# A/B on one 2-GPU box: round-1 library (PPE rows split only) vs this tree (slab decomposition), same bench command
O=gpurun_out/${1:-r2e}; mkdir -p $O
N=${2:-2}
nvidia-smi topo -m > $O/topo.txt 2>&1; head -8 $O/topo.txt
run() { # tag, extra env
  env $2 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $3 \
      bench.py --gpus $N --steps 3 --warmup 2 --no-e2e $4 > $O/bench_$1.json 2> $O/bench_$1.err
  echo "== $1 exit $?"; python - <<PY
import json
try:
    d=json.load(open("$O/bench_$1.json"))
    print({k:d[k] for k in ("ms_per_step","us_per_cg_iteration","phase_ms_per_step","cg_cycle_split","sync_us_per_iteration")}, d["clocks"])
except Exception as e: print("no json", e)
PY
}
run r1lib "MPSB_LIB_PATH=$PWD/build/exp/lib_r1.so" 29611 --no-parity
run new_noparity "X=1" 29612 --no-parity
run new "X=1" 29613 ""
for lib in build/exp/lib_r1.so mps_b200/libmps_b200.so; do
  MPSB_LIB_PATH=$PWD/$lib timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29620 \
      tools/pcg_probe_multi.py 2832 > $O/probe_$(basename $lib .so).log 2>&1
  echo "== probe $lib"; grep "us/it" $O/probe_$(basename $lib .so).log
done
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -x > $O/pytest_multi.log 2>&1; tail -3 $O/pytest_multi.log
