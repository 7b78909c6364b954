#!/bin/bash
# tiled neighbour build: parity suite, then timing of both builds at 4 M particles, then 2-GPU check if 2 GPUs
O=gpurun_out/${1:-r2t}; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
for tiles in 0 1; do
  MPSB_NEIGHBOR_TILES=$tiles timeout 300 python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu --no-other > $O/bench_tiles$tiles.json 2> $O/bench_tiles$tiles.err
  python - <<PY
import json
d=json.load(open("$O/bench_tiles$tiles.json")); print("tiles=$tiles", d["phase_ms_per_step"], d["us_per_cg_iteration"])
PY
done
