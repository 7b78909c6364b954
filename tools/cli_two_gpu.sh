#!/bin/bash
# dambreak_fine through the C++ CLI on 1 and on 2 GPUs (BASELINE configs[2]): ms/step from time_step_report.csv.
O=gpurun_out/${1:-cli2}; mkdir -p $O
python - <<PY
import sys; sys.path.insert(0, ".")
from mps_b200 import cases as C
S = C.load_settings(C.case_paths("dambreak_fine")[0])
for tag in ("one", "two"):
    import os; os.makedirs("$O/" + tag, exist_ok=True)
    C.write_settings_yaml(S, "$O/" + tag + "/settings.yml", finish_time=300.5 * S.dt, output_interval=1.0)
PY
./mps_b200/host/mps -s $O/one/settings.yml --quiet > $O/one.log 2>&1
for r in 0 1; do RANK=$r WORLD_SIZE=2 LOCAL_RANK=$r MASTER_ADDR=127.0.0.1 MASTER_PORT=29870 ./mps_b200/host/mps -s $O/two/settings.yml --quiet > $O/two_$r.log 2>&1 & done; wait
for t in one two; do python - <<PY
import csv
rows = list(csv.DictReader(open("$O/$t/result/time_step_report.csv")))
ms = [float(r["GPU Step (ms)"]) for r in rows[20:]]; sol = [float(r["PPE Solve (ms)"]) for r in rows[20:]]; it = [int(r["CG Iterations"]) for r in rows[20:]]
print("$t", "steps", len(rows), "gpu ms/step %.3f solve ms %.3f cg it %.1f us/it %.2f" % (sum(ms)/len(ms), sum(sol)/len(sol), sum(it)/len(it), 1e3*sum(sol)/sum(it)))
PY
done
