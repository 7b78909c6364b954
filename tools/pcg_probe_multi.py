"""Diagnostic (one process per GPU under torchrun): time the split PCG solve alone on the scaled dam break and
print every rank's per-sweep cycle split.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29555 \
        tools/pcg_probe_multi.py [nx]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from mps_b200 import capi, cases as C

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 2832
rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
S, t, p, v = C.scaled_dambreak(nx)
dev = capi.DeviceMPS(S, t, p, v, device=local)
dev.attach_comm_torch(rank, world)
dev.load_initial_state()
for ph in (capi.PH_NEIGHBORS, capi.PH_EXPLICIT, capi.PH_NEIGHBORS, capi.PH_COLLISION, capi.PH_NEIGHBORS,
           capi.PH_FLAGS_SOURCE, capi.PH_MATRIX, capi.PH_DIRICHLET):
    st = dev.run_phase(ph)
for rep in range(3):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    st = dev.run_phase(capi.PH_SOLVE)
    dt = time.perf_counter() - t0
    cyc = dev.cg_cycles()
    it = max(st.cg_iterations, 1)
    line = (f"rank {rank}/{world} nx={nx} it={st.cg_iterations} {1e6*dt/it:.1f} us/it  cycles/it A {cyc[0]/it:.0f} "
            f"B {cyc[1]/it:.0f} C {cyc[2]/it:.0f} sync {cyc[3]/it:.0f}")
    for r in range(world):
        dist.barrier()
        if r == rank and rep == 2:
            print(line, flush=True)
dev.close()
dist.destroy_process_group()
