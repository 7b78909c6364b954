"""Multi-GPU parity check, launched one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tools/multi_gpu_check.py [case|nx=<NX>] [steps]

Every rank runs `steps` steps of the case slab-decomposed over the ranks (every phase on the rank's own rows; halo
particles, migration, halo flags / pressures, collision records and the Krylov exchanges over NVLink) and
checks its final state against (0) the compiled reference's fixture where one exists (nx=400 / 800 / 1416 / 2832), (a) the other ranks (must be bit-identical: every rank takes the same
branches) and (b) a single-GPU run of the same library on the same device (pressure / velocity /
position within 1e-10 relative — only the reduction order differs). Prints one JSON line on rank 0 and
exits non-zero on a mismatch. Used by tests/test_multi_gpu.py.
"""
import json
import os
import sys

import numpy as np

os.environ.setdefault("MPSB_MIN_SLAB_PARTICLES", "0")  # this check is about the sharded path: shard the small cases too
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run_case(what, steps, rank, world, local):
    """One case on an already initialised NCCL process group (this rank's device current). Returns the report
    (the same dict on every rank; `ok` is the minimum over the ranks). bench.py calls this at N > 1 as well."""
    import torch
    import torch.distributed as dist

    from mps_b200 import capi, cases as C

    prerun = 0
    if "@" in what:  # "dambreak@200": the 1-GPU handle first runs 200 steps, so collisions / ghost events are live
        what, prerun = what.split("@")[0], int(what.split("@")[1])
    S, t, p, v = C.scaled_dambreak(int(what[3:])) if what.startswith("nx=") else C.load_case(what)

    multi = capi.DeviceMPS(S, t, p, v, device=local)
    multi.attach_comm_torch(rank, world)
    single = capi.DeviceMPS(S, t, p, v, device=local)
    multi.load_initial_state()
    single.load_initial_state()
    # (0) against the COMPILED REFERENCE where a large-N fixture exists (tests/golden/large_nx<NX>.npz): one whole
    # step of the N-rank handle from the initial state, flags by digest, sampled fields within 1e-10
    vs_reference = None
    fixture = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                           f"large_nx{what[3:]}.npz") if what.startswith("nx=") else ""
    if fixture and os.path.exists(fixture):
        import hashlib

        z = np.load(fixture)
        multi.step_forward(True)
        got, ids = multi.download(), z["sample_ids"]
        sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()  # noqa: E731
        vs_reference = {"flags_digest_equal": sha(got["boundary_condition"]) == str(z["after_bc_digest"]) and
                        sha(got["type"]) == str(z["after_type_digest"]),
                        "cg_iterations": [multi.last_stats.cg_iterations, int(z["cg_iterations"])],
                        "n_inner_nnz_equal": (multi.last_stats.n_inner, multi.last_stats.nnz) == (int(z["n_inner"]), int(z["nnz"]))}
        for name in ("pressure", "velocity", "position", "minimum_pressure", "number_density_ratio"):
            want = z["sample_" + name]
            scale = max(float(np.abs(want).max()), 1e-300)
            vs_reference[name] = float(np.abs(got[name][ids] - want).max()) / scale
        vs_reference["ok"] = bool(vs_reference["flags_digest_equal"] and vs_reference["n_inner_nnz_equal"] and
                                  all(vs_reference[k] <= 1e-10 for k in ("pressure", "velocity", "position",
                                                                         "minimum_pressure", "number_density_ratio")))
    for _ in range(prerun):
        single.step_forward(False)
    worst = {"pressure": 0.0, "velocity": 0.0, "position": 0.0}
    flags_equal, iters = True, []
    collisions = ghosted = matrix_free_multi = 0
    events_equal = True
    for k in range(steps):
        # teacher-forced per step from the single-GPU state (the dynamics are chaotic)
        st = single.download(["type", "position", "velocity", "pressure"])
        multi.upload(st)
        cm, cs = multi.step_forward(False), single.step_forward(False)
        a, b = multi.download(), single.download()
        flags_equal &= bool(np.array_equal(a["type"], b["type"]) and
                            np.array_equal(a["boundary_condition"], b["boundary_condition"]))
        for name in worst:
            scale = max(float(np.abs(b[name]).max()), 1e-300)
            worst[name] = max(worst[name], float(np.abs(a[name] - b[name]).max()) / scale)
        iters.append((multi.last_stats.cg_iterations, single.last_stats.cg_iterations))
        collisions += multi.last_stats.n_collisions
        ghosted += multi.last_stats.n_ghosted
        # the debug-log events: each rank lists the events of ITS rows; their union is the single-GPU list
        for kind in (0, 1):
            mine_ev = [None] * world
            dist.all_gather_object(mine_ev, [int(i) for i in multi.events(kind)[0]])
            union = sorted(i for part in mine_ev for i in part)
            events_equal &= union == sorted(int(i) for i in single.events(kind)[0])
            events_equal &= len(union) == (multi.last_stats.n_ghosted if kind == 0 else multi.last_stats.n_unconnected)
        matrix_free_multi |= multi.last_stats.ppe_matrix_free
        flags_equal &= (multi.last_stats.n_collisions, multi.last_stats.n_ghosted) == \
                       (single.last_stats.n_collisions, single.last_stats.n_ghosted)
    # (2) free-running: the N-rank handle carries its own state from step to step (halo particles and migration
    # with no upload in between); three steps stay within the chaos floor of the 1-GPU run
    free = {"pressure": 0.0, "velocity": 0.0, "position": 0.0}
    counters_equal = True
    for k in range(3):
        multi.step_forward(k == 2)
        single.step_forward(k == 2)
        sm, ss = multi.last_stats, single.last_stats
        counters_equal &= (sm.n_inner, sm.n_surface, sm.nnz, sm.n_collisions, sm.n_ghosted, sm.n_unconnected, sm.max_neighbors) == \
                          (ss.n_inner, ss.n_surface, ss.nnz, ss.n_collisions, ss.n_ghosted, ss.n_unconnected, ss.max_neighbors)
    a, b = multi.download(), single.download()
    flags_equal &= bool(np.array_equal(a["type"], b["type"]) and np.array_equal(a["boundary_condition"], b["boundary_condition"]))
    for name in free:
        scale = max(float(np.abs(b[name]).max()), 1e-300)
        free[name] = float(np.abs(a[name] - b[name]).max()) / scale
    # (3) sharded transfers: a rank's mpsb_download_shard writes exactly its block of ids (and the right values);
    # after mpsb_upload_shard of the single-GPU state the device state is whole on every rank
    lo, hi = multi.shard_range()
    assert (lo, hi) == capi.shard_block(len(t), rank, world), ((lo, hi), capi.shard_block(len(t), rank, world))
    part = {k: np.full_like(a[k], 7) for k in ("type", "position", "velocity", "pressure")}
    multi.download_into(part, shard=True)
    shard_ok = all(np.array_equal(part[k][lo:hi], a[k][lo:hi]) and (part[k][:lo] == 7).all() and (part[k][hi:] == 7).all()
                   for k in part)
    multi.upload({k: b[k] for k in ("type", "position", "velocity", "pressure")}, shard=True)
    a2 = multi.download(["type", "position", "velocity", "pressure"])
    shard_ok &= all(np.array_equal(a2[k], b[k]) for k in a2)
    multi.upload({k: a[k] for k in ("type", "position", "velocity", "pressure")})
    # every rank must hold the same bits
    mine = torch.from_numpy(np.concatenate([a["pressure"], a["velocity"].ravel(), a["position"].ravel()])).cuda()
    ref = mine.clone()
    dist.broadcast(ref, src=0)
    same = torch.tensor([int(torch.equal(mine.view(torch.int64), ref.view(torch.int64)))], device="cuda")
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    ok = bool(same.item()) and flags_equal and counters_equal and events_equal and not matrix_free_multi and shard_ok and all(x <= 1e-10 for x in worst.values()) and \
        all(x <= 1e-8 for x in free.values()) and (vs_reference is None or vs_reference["ok"])
    okt = torch.tensor([int(ok)], device="cuda")
    dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    multi.close()
    single.close()
    return {"case": what + (f"@{prerun}" if prerun else ""), "world": world, "steps": steps, "ranks_bit_identical": bool(same.item()),
            "flags_equal": flags_equal, "counters_equal": counters_equal, "event_ids_equal": events_equal,
            "matrix_free_on_multi": int(matrix_free_multi), "sharded_transfers_ok": bool(shard_ok),
            "max_rel_diff_vs_single_gpu": worst, "free_run_3_steps_max_rel_diff": free, "vs_compiled_reference": vs_reference,
            "collisions_replayed": collisions, "particles_ghosted": ghosted,
            "cg_iterations_multi_vs_single": iters[-3:], "ok": bool(okt.item())}


def main():
    import torch
    import torch.distributed as dist

    what = sys.argv[1] if len(sys.argv) > 1 else "dambreak"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rep = run_case(what, steps, rank, world, local)
    if rank == 0:
        print(json.dumps(rep), flush=True)
    dist.destroy_process_group()
    return 0 if rep["ok"] else 1


if __name__ == "__main__":
    sys.exit(main())
