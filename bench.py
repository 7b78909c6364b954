#!/usr/bin/env python
"""bench.py — the MPS time step (reference MPS::stepForward, src/mps.cpp:111-130) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--nx NX]

Workload (BASELINE.json configs[3], the configuration the metric "particle-steps/sec and PPE solve
ms/step at 1/2/4/8 B200; % of HBM roofline" is quoted on): the synthetic scaled dam break —
dambreak_fine geometry and settings, l0 = 0.4/NX, dt/l0 kept — NX = 2832 -> 4 042 868 particles
(4 010 112 fluid), generated in memory. A "step" is one MPS::stepForward of the running simulation.

One JSON line on stdout (rank 0). `value` = non-ghost particles x steps / device time with the state
resident in HBM; `e2e` = the same through the host-buffer C ABI (upload from pinned host memory, step,
download) every step; `roofline` describes the dominant kernel (the persistent PCG solve);
`cpu_baseline` times the compiled reference (oracle/_ref) on this box's host cores on a bounded sample.

--impl reference: the reference's own CPU implementation (oracle/_ref/libmps_ref.so, the unmodified
sources compiled by oracle/Makefile; the oracle port if that file is absent), same metric and unit.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "particle_steps_per_sec"
UNIT = "particle-steps/s"
DEFAULT_NX = 2832
CPU_SAMPLE_NX = 400  # 84 518 particles: ~10 s per reference step on one Xeon core


def workload_config(nx, n_particles, parallelism, nx_timed=None, n_timed=None):
    """`nx_fluid` / `particles` name BASELINE configs[3]; `nx_fluid_timed` / `particles_timed` are what THIS line
    actually ran (they differ only on the CPU arm, which times a reduced-size sample of the same generator)."""
    nx_timed = nx if nx_timed is None else nx_timed
    n_timed = n_particles if n_timed is None else n_timed
    return {
        "workload": f"synthetic scaled dam break (BASELINE configs[3]): dambreak_fine geometry/settings, "
                    f"l0=0.4/{nx}, dt/l0=0.125 s/m",
        "nx_fluid": nx,
        "particles": int(n_particles),
        "nx_fluid_timed": int(nx_timed),
        "particles_timed": int(n_timed),
        "reduced_size_sample": bool(nx_timed != nx),
        "extrapolated": False,
        "parallelism": parallelism,
        "l2_policy": "working set (neighbour lists + PPE, >4 GB at nx=2832) exceeds the 126 MB L2; no flush needed",
    }


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i",
                 str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smmax, power, reasons = [], None, [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smmax = float(f[2])
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": smmax,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
def cpu_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def make_cpu_mps(nx):
    """The reference CPU implementation of the path on the scaled dam break: (object, kind, n_active)."""
    import numpy as np

    from mps_b200 import cases as C
    from oracle import refbind as R

    S, t, p, v = C.scaled_dambreak(nx)
    if R.have_ref():
        d = tempfile.mkdtemp(prefix="mpsb_bench_")
        yml = C.write_settings_yaml(S, os.path.join(d, "settings.yml"))
        return R.RefMPS(yml, arrays=(t, p, v)), "reference", int(np.count_nonzero(t != C.GHOST))
    if not R.have_oracle():
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, capture_output=True)
    return R.OracleMPS(S.to_c(), t, p, v), "port", int(np.count_nonzero(t != C.GHOST))


def set_omp_threads(k):
    """OpenMP thread count of the already-loaded libgomp (the one libmps_ref.so is linked against)."""
    import ctypes

    try:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(int(k))
        return True
    except OSError:
        return False


def time_cpu_steps(nx, steps, warmup, budget_s=None, threads=None):
    mps, kind, n_active = make_cpu_mps(nx)
    if threads is not None:
        set_omp_threads(threads)
    for _ in range(warmup):
        mps.step(False)
    times = []
    t_begin = time.perf_counter()
    for k in range(steps):
        t0 = time.perf_counter()
        mps.step(False)
        times.append(time.perf_counter() - t0)
        if budget_s is not None and time.perf_counter() - t_begin > budget_s:
            break
    mps.close()
    return kind, n_active, times


def cpu_model_name():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def cpu_sample_text(nx, n, kind, nsteps):
    what = ("the unmodified reference (oracle/_ref/libmps_ref.so; it solves the PPE twice per step, "
            "src/mps.cpp:457,460)" if kind == "reference" else "the oracle port (oracle/mps_oracle.cpp)")
    return (f"{nsteps} step(s) of {what} on the same generator at nx_fluid={nx} ({n} particles) - a REDUCED-SIZE "
            f"sample: CG iterations grow with 1/l0, so CPU particle-steps/s at the full size is lower still "
            f"(full_size_measured_once)")


def full_size_cpu_record(nx):
    """The reference timed ONCE at this size on the CPU (tests/golden/make_large_golden.py records the wall time of
    every phase next to the parity fixture). The fixture times one solve; the reference performs two identical
    ones per step (src/mps.cpp:457,460), so its step = 2 x solve + the other phases. Static, dated, not re-run."""
    import numpy as np

    path = os.path.join(ROOT, "tests", "golden", f"large_nx{nx}.npz")
    if not os.path.exists(path):
        return None
    z = np.load(path)
    sec = json.loads(str(z["seconds"]))
    other = sum(v for k, v in sec.items() if k not in ("solve_once", "solve_reference_twice"))
    step_s = 2.0 * sec["solve_once"] + other
    n = int(z["n"])
    return {"nx_fluid": int(z["nx"]), "particles": n, "seconds_per_reference_step": step_s,
            "seconds_one_solve": sec["solve_once"], "cg_iterations": int(z["cg_iterations"]),
            "value": n / step_s, "unit": UNIT, "cpu_model": str(z["cpu_model"]), "omp_threads": int(z["cpu_threads"]),
            "when": str(z["when"]), "source": f"tests/golden/large_nx{nx}.npz (make_large_golden.py)",
            "note": "measured once where the fixture was generated (not on this box, not in this run); the Eigen "
                    "solve is serial, so the thread count only touches the ~4 % outside it"}


def cpu_points(sizes, budget_s):
    """The reference at several sizes with OMP_NUM_THREADS = 1 and all cores."""
    allt = cpu_threads()
    pts, kind = [], None
    per = budget_s / max(1, 2 * len(sizes))
    for nx in sizes:
        for th in (1, allt):
            kind, n, times = time_cpu_steps(nx, 3, 0, budget_s=per, threads=th)
            sec = sum(times)
            pts.append({"nx_fluid": nx, "particles": n, "threads": th, "steps": len(times),
                        "ms_per_step": 1e3 * sec / len(times), "value": n * len(times) / sec})
    set_omp_threads(allt)
    return kind, pts


CPU_MODEL_TEXT = ("measured, not modelled: the reference's step is 2 serial Eigen solves (96-99 % of it); CG iterations "
                  "999 / 1 912 / 2 880 / 4 221 and 12.9 k / 6.0 k / 2.6 k / 1.7 k particle-steps/s at nx = 400 / 800 / 1 416 / "
                  "2 832 (profiles/r2_cpu_reference_scaling.json) - CPU throughput FALLS with size, so a ratio against a "
                  "reduced-size CPU sample understates the same-size ratio; round 1's value*nx_cpu/nx guess is retired")


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    nx_cpu = args.cpu_nx
    # keep the whole run within a few minutes: cost/step ~ 8 s * (nx/400)^3 on one core (the solve is serial)
    total = max(1, args.steps + args.warmup)
    while nx_cpu > 80 and 8.0 * (nx_cpu / 400.0) ** 3 * total > 240.0:
        nx_cpu -= 40
    allt = cpu_threads()
    kind, n, times = time_cpu_steps(nx_cpu, args.steps, args.warmup, threads=allt)
    from mps_b200 import cases as C

    n_full = len(C.scaled_dambreak(args.nx)[1])
    sec = sum(times)
    value = n * len(times) / sec
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(times), "warmup": args.warmup, "ms_per_step": 1e3 * sec / len(times), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        # the size this line TIMED is nx_fluid_timed / particles_timed; nx_fluid / particles name the workload the
        # b200 arm runs in full. A ratio of the two arms is therefore cross-size and a LOWER bound (see
        # cpu_baseline.full_size_measured_once for the reference at the full size).
        "config": workload_config(args.nx, n_full, "cpu", nx_timed=nx_cpu, n_timed=n),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": allt, "kind": kind, "cpu_model": cpu_model_name(),
                         "sample": cpu_sample_text(nx_cpu, n, kind, len(times)),
                         "nx_fluid_timed": nx_cpu, "particles_timed": n,
                         "full_size_measured_once": full_size_cpu_record(args.nx), "size_model": CPU_MODEL_TEXT},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(out)
    return 0


def other_configs(gpu_steps, with_cpu=True, cpu_seconds=1.5):
    """BASELINE configs[0], [1], [2], [4] (the shipped, latency-bound sizes) through the same C ABI on this GPU,
    beside the compiled reference on this box's cores with 1 thread and with all of them. Outside the timed
    region of `value`. Returns a list of rows."""
    import numpy as np

    from mps_b200 import capi, cases as C
    from oracle import refbind as R

    rows = []
    names = {"hydrostatic": "configs[0]", "dambreak": "configs[1]", "dambreak_fine": "configs[2]",
             "dambreak_fine_original": "configs[4]"}
    for case, cfg in names.items():
        S, t, p, v = C.load_case(case)
        n = int(np.count_nonzero(t != C.GHOST))
        dev = capi.DeviceMPS(S, t, p, v)
        dev.load_initial_state()
        for _ in range(20):
            dev.step_forward(False)
        its, solve = [], []
        t0 = time.perf_counter()
        for _ in range(gpu_steps):
            dev.step_forward(False)
            st = dev.last_stats
            its.append(st.cg_iterations)
            solve.append(st.ms_solve)
        wall = time.perf_counter() - t0
        dev.close()
        row = {"config": cfg, "case": case, "particles": n, "gpu_steps": gpu_steps,
               "gpu_ms_per_step": 1e3 * wall / gpu_steps, "gpu_particle_steps_per_s": n * gpu_steps / wall,
               "gpu_solve_ms_per_step": float(np.mean(solve)), "cg_iterations_per_step": float(np.mean(its)),
               "gpu_us_per_cg_iteration": 1e3 * float(np.sum(solve)) / max(1, int(np.sum(its))),
               "timing": "host wall clock around the C-ABI calls (one graph launch + stats read-back per step)"}
        if with_cpu and R.have_ref():
            yml, _ = C.case_paths(case)
            for th in (1, cpu_threads()):
                ref = R.RefMPS(yml, os.path.join(ROOT, S.input_csv_path))
                set_omp_threads(th)
                for _ in range(2):
                    ref.step(False)
                k, t0 = 0, time.perf_counter()
                while time.perf_counter() - t0 < cpu_seconds:
                    ref.step(False)
                    k += 1
                sec = time.perf_counter() - t0
                ref.close()
                row["cpu_ms_per_step_%s" % ("1_thread" if th == 1 else "all_threads")] = 1e3 * sec / k
            row["cpu_threads_all"] = cpu_threads()
            row["speedup_vs_best_cpu"] = min(row["cpu_ms_per_step_1_thread"], row["cpu_ms_per_step_all_threads"]) / row["gpu_ms_per_step"]
        rows.append(row)
    # 3-D cases at sizes where bandwidth starts to matter (the reference ships none): the generated open
    # tank of tests/golden/tank3d_*, 4x finer, at radius ratio 2.1 (~37 neighbours) and at the ratio every settings.yml
    # of the reference ships, 3.1 (~120 neighbours, K = 384 slots, ~1.2 kB of PPE row per particle)
    for ratio, n_fluid, n_tank in ((2.1, (32, 40, 24), (64, 56, 24)), (3.1, (64, 64, 40), (128, 80, 40))):
        try:
            rows.append(tank3d_row(ratio, n_fluid, n_tank, gpu_steps, with_cpu and ratio == 2.1, cpu_seconds))
        except Exception as e:  # a row of context must not take the headline line down with it
            rows.append({"config": "3-D (SURVEY 8 f-4, generated)", "radius_ratio": ratio, "error": str(e)[:300]})
    return rows


def tank3d_row(ratio, n_fluid, n_tank, gpu_steps, with_cpu, cpu_seconds):
    import tempfile

    import numpy as np

    from mps_b200 import capi, cases as C
    from oracle import refbind as R

    S, t, p, v = C.tank3d(n_fluid=n_fluid, n_tank=n_tank, ratio=ratio, jitter=0.05)
    n = int(np.count_nonzero(t != C.GHOST))
    dev = capi.DeviceMPS(S, t, p, v)
    dev.load_initial_state()
    for _ in range(3):
        dev.step_forward(False)
    steps3 = max(5, gpu_steps // 10)
    its, solve, tot = [], [], []
    t0 = time.perf_counter()
    for _ in range(steps3):
        dev.step_forward(False)
        st = dev.last_stats
        its.append(st.cg_iterations)
        solve.append(st.ms_solve)
        tot.append(st.ms_total)
    wall = time.perf_counter() - t0
    st = dev.last_stats
    it_all, solve_ms = max(1, int(np.sum(its))), float(np.sum(solve))
    bytes_it = 12.0 * st.nnz + 108.0 * st.n_inner  # algorithmic bytes per CG iteration
    row = {"config": "3-D (SURVEY 8 f-4, generated)", "radius_ratio": ratio,
           "case": "tank3d %dx%dx%d cells, fluid %dx%dx%d" % (n_tank + n_fluid), "particles": n,
           "gpu_steps": steps3, "gpu_ms_per_step": 1e3 * wall / steps3, "gpu_particle_steps_per_s": n * steps3 / wall,
           "gpu_solve_ms_per_step": float(np.mean(solve)), "cg_iterations_per_step": float(np.mean(its)),
           "gpu_us_per_cg_iteration": 1e3 * solve_ms / it_all,
           "max_neighbors": int(st.max_neighbors), "nnz": int(st.nnz), "n_inner": int(st.n_inner),
           "algorithmic_bytes_per_cg_iteration": bytes_it,
           "solve_algorithmic_gbs": bytes_it * it_all / (solve_ms * 1e-3) / 1e9 if solve_ms > 0 else None,
           "solve_in_cluster": int(st.solve_in_cluster), "ppe_columns_16bit": int(st.ppe_columns_16bit)}
    dev.close()
    if with_cpu and R.have_ref():
        yml = C.write_settings_yaml(S, os.path.join(tempfile.mkdtemp(prefix="mpsb_bench3d_"), "settings.yml"))
        ref = R.RefMPS(yml, arrays=(t, p, v))
        set_omp_threads(cpu_threads())
        ref.step(False)
        k, t0 = 0, time.perf_counter()
        while k < 1 or time.perf_counter() - t0 < cpu_seconds:
            ref.step(False)
            k += 1
        row["cpu_ms_per_step_all_threads"] = 1e3 * (time.perf_counter() - t0) / k
        row["cpu_threads_all"] = cpu_threads()
        row["speedup_vs_best_cpu"] = row["cpu_ms_per_step_all_threads"] / row["gpu_ms_per_step"]
        ref.close()
    return row



def parity_vs_one_gpu(dev, S, t, p, v, local, world):
    """N > 1: one step from the initial state on the N-rank handle and on a single-GPU handle of the same
    library on this rank's device. Returns the block for the JSON line; `ok` False fails the run."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from mps_b200 import capi

    single = capi.DeviceMPS(S, t, p, v, device=local)
    single.load_initial_state()
    single.step_forward(True)
    dev.step_forward(True)
    a, b = dev.download(), single.download()
    it = (dev.last_stats.cg_iterations, single.last_stats.cg_iterations)
    single.close()
    dev.upload({"type": t, "position": p, "velocity": v})  # the timed steps start from the initial state at every N
    dev.load_initial_state()
    flags = bool(np.array_equal(a["type"], b["type"]) and np.array_equal(a["boundary_condition"], b["boundary_condition"]))
    worst = {}
    for name in ("pressure", "velocity", "position", "number_density_ratio"):
        scale = max(float(np.abs(b[name]).max()), 1e-300)
        worst[name] = float(np.abs(a[name] - b[name]).max()) / scale
    mine = torch.from_numpy(np.concatenate([a["pressure"], a["velocity"].ravel(), a["position"].ravel()])).cuda()
    ref = mine.clone()
    dist.broadcast(ref, src=0)
    same = torch.tensor([int(torch.equal(mine.view(torch.int64), ref.view(torch.int64)))], device="cuda")
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    ok = bool(same.item()) and flags and all(x <= 1e-10 for x in worst.values())
    okt = torch.tensor([int(ok)], device="cuda")
    dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    return {"ok": bool(okt.item()), "ranks_bit_identical": bool(same.item()), "flags_equal": flags, "max_rel": worst,
            "cg_iterations_n_gpus_vs_1": list(it), "tolerance": 1e-10}


def parity_in_motion(rank, world, local):
    """N > 1, before the timed region: the slab path with things happening in it, on this run's own ranks
    (tools/multi_gpu_check.py, the checks of tests/test_multi_gpu.py): dambreak after 200 steps (collisions replayed
    across ranks), after 320 steps (particles leave the domain; event ids), and nx=800 against the compiled
    reference's fixture. Per case: teacher-forced steps of the N-rank handle vs a single-GPU handle (flags, counters
    and event ids identical, fields <= 1e-10), three free-running steps, sharded transfers, ranks bit-identical.
    Returns a compact block; `ok` False fails the run."""
    had = os.environ.get("MPSB_MIN_SLAB_PARTICLES")
    os.environ["MPSB_MIN_SLAB_PARTICLES"] = "0"  # shard the small cases too: this is about the sharded path
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import multi_gpu_check as M

    rows, ok = [], True
    try:
        for what, steps in (("dambreak@200", 50), ("dambreak@320", 8), ("nx=800", 1)):
            r = M.run_case(what, steps, rank, world, local)
            good = bool(r["ok"])
            if what == "dambreak@200":
                good &= r["collisions_replayed"] > 0
            if what == "dambreak@320":
                good &= r["particles_ghosted"] > 0
            if what.startswith("nx="):
                good &= bool(r["vs_compiled_reference"] and r["vs_compiled_reference"]["ok"])
            ok &= good
            rows.append({"case": what, "steps": steps, "ok": good, "collisions": r["collisions_replayed"],
                         "ghosted": r["particles_ghosted"], "ranks_bitwise": r["ranks_bit_identical"],
                         "flags_counters_events": bool(r["flags_equal"] and r["counters_equal"] and r["event_ids_equal"]),
                         "max_rel": max(r["max_rel_diff_vs_single_gpu"].values()),
                         "free_run_max_rel": max(r["free_run_3_steps_max_rel_diff"].values()),
                         "vs_reference_fixture": None if r["vs_compiled_reference"] is None else bool(r["vs_compiled_reference"]["ok"])})
    finally:
        if had is None:
            os.environ.pop("MPSB_MIN_SLAB_PARTICLES", None)
        else:
            os.environ["MPSB_MIN_SLAB_PARTICLES"] = had
    return {"ok": bool(ok), "tolerance": 1e-10, "cases": rows}


def same_size_pair(nx, steps, cpu_point):
    """The GPU on the very workload the CPU sample ran (same generator at nx, the same first `steps` steps from the
    initial state, all host threads on the CPU side): device-resident and end to end through host buffers. One
    apples-to-apples pair measured in this run on this box; the headline stays the full-size GPU line."""
    import numpy as np
    import torch

    from mps_b200 import capi, cases as C

    S, t, p, v = C.scaled_dambreak(nx)
    stream = torch.cuda.current_stream()

    def fresh():
        d = capi.DeviceMPS(S, t, p, v)
        d.set_stream(stream.cuda_stream)
        d.load_initial_state()
        return d

    d = fresh()  # first use of this size: module / graph set-up outside the timed steps
    for _ in range(3):
        d.step_forward(False)
    d.close()
    d = fresh()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(steps):
        d.step_forward(False)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    n = d.last_stats.n_active
    d.close()
    d = fresh()
    st = d.download(["type", "position", "velocity", "pressure"])
    pin = {k: torch.from_numpy(a).pin_memory() for k, a in st.items()}
    host = {k: pin[k].numpy() for k in pin}
    inputs = {k: host[k] for k in ("type", "position", "velocity")}
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        d.upload(inputs)
        d.step_forward(False)
        d.download_into(host)
    torch.cuda.synchronize()
    sec = time.perf_counter() - t0
    d.close()
    gpu_value, gpu_e2e = n * steps / (ms * 1e-3), n * steps / sec
    return {"nx_fluid": nx, "particles": int(n), "steps": steps, "gpu_value": gpu_value, "gpu_e2e": gpu_e2e,
            "gpu_ms_per_step": ms / steps, "cpu_value": cpu_point["value"], "cpu_threads": cpu_point["threads"],
            "cpu_ms_per_step": cpu_point["ms_per_step"], "cpu_steps": cpu_point["steps"],
            "ratio": gpu_value / cpu_point["value"], "e2e_ratio": gpu_e2e / cpu_point["value"], "unit": UNIT,
            "how": "both sides: the first steps from the initial state of the same generated case; CPU = the compiled "
                   "reference with all host threads (cpu_baseline.points), GPU = CUDA events / host clock around "
                   "upload+step+download from pinned buffers"}


# ------------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import numpy as np
    import torch

    from mps_b200 import capi, cases as C

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the MPS step has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist

        # keep stdout to the ONE JSON line: NCCL_DEBUG=VERSION (set on some boxes) prints a banner there
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    moving = None
    if world > 1 and not args.no_parity:
        moving = parity_in_motion(rank, world, local)  # its handles are closed again before the bench's own is made
    S, t, p, v = C.scaled_dambreak(args.nx)
    n_active0 = int(np.count_nonzero(t != C.GHOST))
    dev = capi.DeviceMPS(S, t, p, v, device=local)
    if world > 1:
        dev.attach_comm_torch(rank, world)
    stream = torch.cuda.current_stream()
    dev.set_stream(stream.cuda_stream)
    dev.load_initial_state()
    parity = None
    if world > 1 and not args.no_parity:
        parity = parity_vs_one_gpu(dev, S, t, p, v, local, world)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        dev.step_forward(False)
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    launches0 = dev.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stats = []
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        dev.step_forward(False)
        stats.append(dev.last_stats.as_dict())
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = dev.launch_count - launches0
    cyc = dev.cg_cycles()  # CTA 0 of this rank, last solve: SM cycles in sweep A / B / C / barriers+reductions(+exchanges)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        tms = torch.tensor([ms], device="cuda", dtype=torch.float64)
        torch.distributed.all_reduce(tms, op=torch.distributed.ReduceOp.MAX)
        ms = float(tms.item())
    n_active = stats[-1]["n_active"]
    value = n_active * args.steps / (ms * 1e-3)

    # roofline of the dominant kernel (pcg_kernel, one launch per step): algorithmic bytes per launch =
    # I * (12 nnz + 108 n) + 32 n, n = Inner rows, nnz incl. the diagonal
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    solve_bytes = sum(s["cg_iterations"] * (12.0 * s["nnz"] + 108.0 * s["n_inner"]) + 32.0 * s["n_inner"] for s in stats)
    solve_ms = sum(s["ms_solve"] for s in stats)
    # N > 1: each rank streams its own row block, so the per-GPU figure (1/world of the bytes) is the one
    # that compares with one GPU's HBM peak
    achieved = solve_bytes / world / (solve_ms * 1e-3) / 1e9 if solve_ms > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "pcg_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        if tj.get("nx_fluid") == args.nx and world == 1:
            traffic = tj.get("dram_bytes_per_launch")
    # what the kernel actually streams: off-diagonals at 10 B (16-bit column offsets) or 12 B, the diagonal
    # at 8 B, vectors as in the algorithmic figure; slice padding (a few %) not counted
    col16 = all(s.get("ppe_columns_16bit", 0) for s in stats)
    per_entry = 10.0 if col16 else 12.0
    streamed = sum(s["cg_iterations"] * (per_entry * (s["nnz"] - s["n_inner"]) + 8.0 * s["n_inner"] + 108.0 * s["n_inner"])
                   for s in stats) / world
    roofline = {"kernel": "pcg_kernel (persistent Jacobi-PCG, 1 launch per step)", "bound": "hbm",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": peak_src, "per_gpu": True, "bytes_per_launch": solve_bytes / world / len(stats),
                "ms_per_launch": solve_ms / len(stats),
                "column_stream": "16-bit offsets (10 B/entry)" if col16 else "int32 (12 B/entry)",
                "streamed_bytes_per_launch": streamed / len(stats),
                "streamed_gbs": streamed / (solve_ms * 1e-3) / 1e9 if solve_ms > 0 else 0.0,
                "note": "achieved = ALGORITHMIC bytes (12 nnz + 108 n per iteration, SURVEY 8d) / time; the kernel "
                        "streams fewer bytes than that when columns travel as 16-bit offsets, so frac can exceed 1; "
                        "streamed_gbs / peak is the fraction of HBM bandwidth actually drawn"}
    mean = lambda k: sum(s[k] for s in stats) / len(stats)  # noqa: E731

    # ---- e2e through the host-buffer C ABI: upload (pinned) -> step -> download, every step ----
    e2e = None
    if not args.no_e2e:
        st = dev.download(["type", "position", "velocity", "pressure"])
        pin = {k: torch.from_numpy(a).pin_memory() for k, a in st.items()}  # pinned host buffers
        host = {k: pin[k].numpy() for k in pin}
        inputs = {k: host[k] for k in ("type", "position", "velocity")}
        n_e2e = max(1, min(args.steps, args.e2e_steps if args.e2e_steps else args.steps))
        # several ranks: each rank moves only ITS block of particle ids over PCIe (mpsb_upload_shard /
        # mpsb_download_shard); the GPUs all-gather the uploaded blocks over NVLink. Bytes are per rank.
        shard = world > 1
        lo, hi = dev.shard_range() if shard else (0, len(t))
        frac = (hi - lo) / max(1, len(t))
        h2d = sum(a.nbytes for a in inputs.values()) * frac
        d2h = sum(a.nbytes for a in host.values()) * frac
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            dev.upload(inputs, shard=shard)           # H2D of the step's inputs
            dev.step_forward(False)
            dev.download_into(host, shard=shard)      # D2H of the step's result (state + pressure) into the same buffers
        barrier()
        sec = time.perf_counter() - t0
        if world > 1:
            tt = torch.tensor([sec], device="cuda", dtype=torch.float64)
            torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
            sec = float(tt.item())
        e2e = {"value": dev.last_stats.n_active * n_e2e / sec, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "steps": n_e2e,
               "io": ("sharded by particle id: bytes are PER RANK (1/%d of the state); uploaded blocks are all-gathered "
                      "between the GPUs over NVLink" % world) if shard else "whole state through one GPU"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        # bounded sample (~20 s): two sizes x {1 thread, all cores}; `value` is the larger size with all cores
        sizes = sorted({max(80, args.cpu_nx // 2), args.cpu_nx})
        kind, pts = cpu_points(sizes, budget_s=20.0)
        head = [q for q in pts if q["nx_fluid"] == args.cpu_nx and q["threads"] == cpu_threads()][0]
        cpu = {"value": head["value"], "unit": UNIT, "cores": cpu_threads(), "kind": kind, "cpu_model": cpu_model_name(),
               "sample": cpu_sample_text(args.cpu_nx, head["particles"], kind, head["steps"]),
               "nx_fluid_timed": args.cpu_nx, "particles_timed": head["particles"], "points": pts,
               "full_size_measured_once": full_size_cpu_record(args.nx), "size_model": CPU_MODEL_TEXT}

    same = None
    if cpu is not None:
        same = same_size_pair(args.cpu_nx, head["steps"], head)

    other = None
    if rank == 0 and world == 1 and not args.no_other:
        other = other_configs(args.other_steps, with_cpu=not args.no_cpu)

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args.nx, len(t), "1 GPU" if world == 1 else f"{world} GPUs: every phase slab-decomposed by row blocks of the cell-major order "
                                      "(halo particles, migration, Krylov exchanges over NVLink peer memory)"),
            # bulky blocks first, compact ones last: a reader of the tail of this line sees the numbers, not the prose
            "other_configs": other, "cpu_baseline": cpu, "roofline": roofline,
            "ppe_solve_ms_per_step": mean("ms_solve"), "cg_iterations_per_step": mean("cg_iterations"),
            "phase_ms_per_step": {k[3:]: mean(k) for k in ("ms_neighbors", "ms_explicit", "ms_collision", "ms_assemble",
                                                           "ms_solve", "ms_correct", "ms_total")},
            "n_inner": stats[-1]["n_inner"], "nnz": stats[-1]["nnz"],
            "solve_options": {"recurrence": "single reduction (Chronopoulos-Gear, opt-in)" if stats[-1].get("cg_single_reduction") else
                              "textbook PCG = Eigen's loop (default)",
                              "matrix_free": bool(stats[-1].get("ppe_matrix_free")),
                              "l2_policy_of_matrix_stream": "evict_first + 16 MB evict_last for row blocks <= 2.5 M rows (multi-GPU splits); off at "
                                                            "4 M rows on one GPU" if os.environ.get("MPSB_L2_RESIDENT_KB") is None else
                                                            "MPSB_L2_RESIDENT_KB=" + os.environ["MPSB_L2_RESIDENT_KB"]},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "same_size": same,
            "parity_vs_1gpu": parity, "parity_in_motion": moving,
            "cg_cycle_split": {"A_spmv": cyc[0] / max(1, sum(cyc)), "B_update": cyc[1] / max(1, sum(cyc)),
                               "C_direction": cyc[2] / max(1, sum(cyc)), "sync": cyc[3] / max(1, sum(cyc)),
                               "of": "rank 0, CTA 0, last solve (SM cycles)"},
            "sync_us_per_iteration": (cyc[3] / max(1, sum(cyc))) * 1e3 * solve_ms / max(1, sum(s["cg_iterations"] for s in stats)),
            "us_per_cg_iteration": 1e3 * solve_ms / max(1, sum(s["cg_iterations"] for s in stats)),
        }
        _emit(out)
    dev.close()
    if world > 1:
        torch.distributed.destroy_process_group()
    if parity is not None and not parity["ok"]:
        sys.stderr.write("bench.py: N-GPU result differs from one GPU beyond 1e-10: %s\n" % json.dumps(parity))
        return 3
    if moving is not None and not moving["ok"]:
        sys.stderr.write("bench.py: the slab path in motion missed its bars: %s\n" % json.dumps(moving))
        return 3
    return 0


_JSON_FD = None


def _claim_stdout():
    """The contract is ONE JSON line on stdout. Libraries write there too (NCCL prints its version banner at
    NCCL_DEBUG=VERSION and WARN alike), so file descriptor 1 is pointed at stderr for the whole run and the
    JSON line goes to a private duplicate of the original stdout."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def _emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, data)


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nx", type=int, default=DEFAULT_NX, help="fluid columns of the scaled dam break (multiple of 8)")
    ap.add_argument("--cpu-nx", type=int, default=CPU_SAMPLE_NX, help="size of the bounded CPU sample")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the end-to-end leg (0 = as many as --steps)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-other", action="store_true", help="skip the other_configs block (BASELINE configs 0, 1, 2, 4)")
    ap.add_argument("--other-steps", type=int, default=200)
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the one-step parity check against one GPU")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
