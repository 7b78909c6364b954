"""Pins the step — above all the PPE solve — against the UNMODIFIED compiled reference at the sizes
the benchmark runs (BASELINE.json configs[3]).

    python tests/golden/make_large_golden.py 800 800j 1416 2832      # CPU only, minutes to ~hours

For `scaled_dambreak(nx)` (suffix `j`: the same lattice with a seeded jitter and an analytic velocity
field, `mps_b200.cases.jittered`, so list distances are not lattice values) it runs ONE step of
oracle/_ref/libmps_ref.so phase by phase and writes tests/golden/large_nx<tag>.npz — small: digests,
norms, counts and a strided sample, not the state:

  nbr_total / nbr_digest     neighbour CSR (by id, reference order) after the third list build: sha256
  bc_digest, n_inner, ...    flags after setBoundaryCondition
  ppe_digest, nnz            CSR row_start / col / val / rhs after ensureDirichletBoundaryConnection
  cg_iterations, cg_error    Eigen CG as the reference constructs it (src/mps.cpp:452-457)
  p_norm2, p_norminf         of the raw solve
  sample_ids + sample_*      every 64th particle + every particle of the lowest fluid rows: raw solve,
                             and pressure / minimum pressure / velocity / position / flags after the step
  courant, seconds_*         Courant number; wall time per phase of THIS cpu run (cpu model recorded)

The reference solves twice with identical input (src/mps.cpp:457,460) and its harness solve is a third;
Eigen's CG is serial and deterministic, so above nx = 1000 the pressures of the recorded solve are
assigned (src/mps.cpp:466-469 does exactly `pi.pressure = pressures[pi.id]`) instead of solving twice
more. At nx <= 1000 solvePoissonEquation itself runs and is asserted equal to the recorded solve.
"""
import hashlib
import json
import os
import platform
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from mps_b200 import cases as C  # noqa: E402
from oracle import refbind as R  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def sha(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return platform.processor()


def main():
    for tag in sys.argv[1:]:
        jitter = tag.endswith("j")
        nx = int(tag.rstrip("j"))
        S, t, p, v = C.scaled_dambreak(nx)
        if jitter:
            p, v = C.jittered(S, t, p, seed=nx)
        n = len(t)
        yml = C.write_settings_yaml(S, os.path.join(tempfile.mkdtemp(prefix="mpsb_large_"), "settings.yml"))
        ref = R.RefMPS(yml, arrays=(t, p, v))
        sec = {}

        def run(name, phases):
            t0 = time.time()
            for ph in phases:
                ref.phase(ph)
            sec[name] = sec.get(name, 0.0) + time.time() - t0

        run("neighbors", [R.PH_NEIGHBORS])
        run("explicit", [R.PH_GRAVITY, R.PH_VISCOSITY, R.PH_MOVE])
        run("neighbors", [R.PH_NEIGHBORS])
        run("collision", [R.PH_COLLISION])
        run("neighbors", [R.PH_NEIGHBORS])
        nstart, nids, ndist = ref.neighbors()
        nbr_total, nbr_digest = len(nids), sha(nstart, nids, ndist)
        nbr_max = int(np.diff(nstart).max())
        del nstart, nids, ndist
        run("flags_source", [R.PH_BC, R.PH_SOURCE])
        bc = ref.get_state(["boundary_condition"])["boundary_condition"]
        run("matrix", [R.PH_MATRIX])
        run("dirichlet", [R.PH_DIRICHLET])
        pstart, pcol, pval, prhs = ref.ppe()
        nnz, ppe_digest = len(pcol), sha(pstart, pcol, pval, prhs)
        rhs_norm2 = float(np.sqrt(np.sum(prhs * prhs)))
        del pstart, pcol, pval
        print(tag, "n", n, "nnz", nnz, "assembled in", {k: round(s, 1) for k, s in sec.items()}, flush=True)
        t0 = time.time()
        x, iters, err, info = ref.cg_solve()
        sec["solve_once"] = time.time() - t0
        print(tag, "cg", iters, "err", err, "info", info, "%.1f s" % sec["solve_once"], flush=True)
        if nx <= 1000:
            run("solve_reference_twice", [R.PH_SOLVE])
            assert np.array_equal(ref.get_state(["pressure"])["pressure"], x)
        else:
            ref.set_state({"pressure": x})
        run("pressure_post", [R.PH_CLAMP, R.PH_MINP])
        run("gradient_move", [R.PH_GRADIENT, R.PH_MOVE2])
        run("number_density", [R.PH_NUMBER_DENSITY])
        t0 = time.time()
        courant = ref.phase(R.PH_COURANT)
        sec["courant"] = time.time() - t0
        after = ref.get_state()

        fluid = np.flatnonzero(t == C.FLUID)
        ymin = p[fluid, 1].min()
        bottom = fluid[p[fluid, 1] < ymin + 2.5 * S.particle_distance]
        ids = np.unique(np.concatenate([np.arange(0, n, 64), bottom])).astype(np.int64)
        out = dict(
            nx=nx, jitter=int(jitter), n=n, step=0, export=1,
            nbr_total=nbr_total, nbr_digest=nbr_digest, nbr_max=nbr_max,
            bc_digest=sha(bc.astype(np.int32)), n_inner=int((bc == C.BC_INNER).sum()),
            n_surface=int((bc == C.BC_SURFACE).sum()),
            nnz=nnz, ppe_digest=ppe_digest, rhs_norm2=rhs_norm2,
            cg_iterations=iters, cg_error=err, cg_info=info,
            p_norm2=float(np.sqrt(np.sum(x * x))), p_norminf=float(np.abs(x).max()),
            after_type_digest=sha(after["type"]), after_bc_digest=sha(after["boundary_condition"]),
            debug_events=ref.debug_log_count, courant=courant,
            sample_ids=ids, sample_solve=x[ids], sample_pressure=after["pressure"][ids],
            sample_minimum_pressure=after["minimum_pressure"][ids], sample_velocity=after["velocity"][ids],
            sample_position=after["position"][ids], sample_acceleration=after["acceleration"][ids],
            sample_number_density_ratio=after["number_density_ratio"][ids],
            sample_boundary_condition=after["boundary_condition"][ids],
            seconds=json.dumps(sec), cpu_model=cpu_model(), cpu_threads=int(os.environ.get("OMP_NUM_THREADS", "0")),
            cpu_count=os.cpu_count(), when=time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()),
        )
        path = os.path.join(HERE, f"large_nx{tag}.npz")
        np.savez_compressed(path, **out)
        print(tag, "written", path, "%.0f kB" % (os.path.getsize(path) / 1024), "cg", iters,
              "seconds", {k: round(s, 1) for k, s in sec.items()}, flush=True)
        ref.close()


if __name__ == "__main__":
    main()
