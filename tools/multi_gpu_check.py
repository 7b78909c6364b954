"""Multi-GPU parity check, launched one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tools/multi_gpu_check.py [case|nx=<NX>] [steps]

Every rank runs `steps` steps of the case with the PPE solve split over the ranks (k_pcg.cu, MULTI) and
checks its final state against (a) the other ranks (must be bit-identical: every rank takes the same
branches) and (b) a single-GPU run of the same library on the same device (pressure / velocity /
position within 1e-10 relative — only the reduction order differs). Prints one JSON line on rank 0 and
exits non-zero on a mismatch. Used by tests/test_multi_gpu.py.
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist

    from mps_b200 import capi, cases as C

    what = sys.argv[1] if len(sys.argv) > 1 else "dambreak"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    S, t, p, v = C.scaled_dambreak(int(what[3:])) if what.startswith("nx=") else C.load_case(what)

    multi = capi.DeviceMPS(S, t, p, v, device=local)
    multi.attach_comm_torch(rank, world)
    single = capi.DeviceMPS(S, t, p, v, device=local)
    multi.load_initial_state()
    single.load_initial_state()
    worst = {"pressure": 0.0, "velocity": 0.0, "position": 0.0}
    flags_equal, iters = True, []
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
    # every rank must hold the same bits
    mine = torch.from_numpy(np.concatenate([a["pressure"], a["velocity"].ravel(), a["position"].ravel()])).cuda()
    ref = mine.clone()
    dist.broadcast(ref, src=0)
    same = torch.tensor([int(torch.equal(mine.view(torch.int64), ref.view(torch.int64)))], device="cuda")
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    ok = bool(same.item()) and flags_equal and all(x <= 1e-10 for x in worst.values())
    okt = torch.tensor([int(ok)], device="cuda")
    dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"case": what, "world": world, "steps": steps, "ranks_bit_identical": bool(same.item()),
                          "flags_equal": flags_equal, "max_rel_diff_vs_single_gpu": worst,
                          "cg_iterations_multi_vs_single": iters[-3:], "ok": bool(okt.item())}), flush=True)
    multi.close()
    single.close()
    dist.destroy_process_group()
    return 0 if okt.item() else 1


if __name__ == "__main__":
    sys.exit(main())
