#!/bin/bash
# final N-GPU bench with the dual ring geometry
O=gpurun_out/${1:-r2ao}; mkdir -p $O
N=${2:-4}
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 \
    bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err
echo "bench exit $?"; python - <<PY
import json
d=json.load(open("$O/bench_n$N.json"))
for k in ("value","ms_per_step","us_per_cg_iteration","phase_ms_per_step","cg_cycle_split","sync_us_per_iteration","e2e","clocks"): print(k, d[k])
print("parity", d["parity_vs_1gpu"]["ok"], d["parity_vs_1gpu"]["max_rel"], d["parity_vs_1gpu"]["cg_iterations_n_gpus_vs_1"])
PY
