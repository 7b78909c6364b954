"""Shipped configurations (BASELINE.json configs[0..2] and [4]) on one B200 beside the compiled reference on the
box's host cores: particle-steps/s, ms/step, PPE-solve ms/step, CG iterations, us/iteration. One JSON line
per case. These are latency-bound sizes (458 - 6 146 particles): the GPU number is launch/barrier latency,
not bandwidth.
    python tools/bench_configs.py [--gpu-steps 400] [--cpu-seconds 8]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from mps_b200 import capi, cases as C
from oracle import refbind as R

ap = argparse.ArgumentParser()
ap.add_argument("--gpu-steps", type=int, default=400)
ap.add_argument("--cpu-seconds", type=float, default=8.0)
ap.add_argument("--cases", default="dambreak,hydrostatic,dambreak_fine,dambreak_fine_original")
ap.add_argument("--cpu-only", action="store_true", help="time only the reference (e.g. under OMP_NUM_THREADS=1)")
args = ap.parse_args()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for case in args.cases.split(","):
    S, t, p, v = C.load_case(case)
    n = int(np.count_nonzero(t != C.GHOST))
    row = {"case": case, "particles": n}
    wall = None
    if not args.cpu_only:
        dev = capi.DeviceMPS(S, t, p, v)
        dev.load_initial_state()
        for _ in range(20):
            dev.step_forward(False)
        its, solve, tot = [], [], []
        t0 = time.perf_counter()
        for _ in range(args.gpu_steps):
            dev.step_forward(False)
            st = dev.last_stats
            its.append(st.cg_iterations); solve.append(st.ms_solve); tot.append(st.ms_total)
        wall = time.perf_counter() - t0
        dev.close()
        row.update({"gpu_steps": args.gpu_steps, "gpu_ms_per_step_wall": 1e3 * wall / args.gpu_steps,
               "gpu_ms_per_step_events": float(np.mean(tot)), "gpu_solve_ms_per_step": float(np.mean(solve)),
               "cg_iterations": float(np.mean(its)), "gpu_us_per_cg_iteration": 1e3 * float(np.sum(solve)) / max(1, int(np.sum(its))),
               "gpu_particle_steps_per_s": n * args.gpu_steps / wall})
    if R.have_ref():
        yml, _ = C.case_paths(case)
        ref = R.RefMPS(yml, os.path.join(ROOT, S.input_csv_path))
        for _ in range(3):
            ref.step(False)
        k, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < args.cpu_seconds:
            ref.step(False); k += 1
        sec = time.perf_counter() - t0
        ref.close()
        threads = int(os.environ.get("OMP_NUM_THREADS", "0")) or len(os.sched_getaffinity(0))
        row.update({"cpu_kind": "reference (oracle/_ref, OpenMP, %d threads)" % threads, "cpu_steps": k,
                    "cpu_ms_per_step": 1e3 * sec / k, "cpu_particle_steps_per_s": n * k / sec})
        if wall is not None:
            row["speedup"] = (n * args.gpu_steps / wall) / (n * k / sec)
    print(json.dumps(row), flush=True)
