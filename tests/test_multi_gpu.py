"""Multi-GPU path: host-side logic on CPU (gloo, world size 2) and, with >= 2 GPUs, the split PPE solve
against the single-GPU solve (tools/multi_gpu_check.py, one process per GPU over NCCL/NVLink)."""
import json
import os
import subprocess
import sys

import pytest

from mps_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_row_blocks_partition_the_padded_rows():
    for n in (1, 458, 2958, 6146, 4041990):
        np_rows = capi.padded_rows(n)
        assert np_rows % 256 == 0 and 0 <= np_rows - n < 256
        for world in (1, 2, 4, 8):
            edges = [capi.row_block(np_rows, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == np_rows
            for (a, b), (c, d) in zip(edges, edges[1:]):
                assert b == c and (b - a) == (d - c) and a % 32 == 0
    with pytest.raises(ValueError):
        capi.row_block(256, 0, 3)
    with pytest.raises(ValueError):
        capi.row_block(100, 0, 2)


def test_shard_blocks_partition_the_particle_ids():
    """mpsb_upload_shard / mpsb_download_shard: every id belongs to exactly one rank's block, blocks are ordered and
    only the last non-empty one can be short (it carries the padding)."""
    for n in (0, 1, 255, 458, 2958, 6146, 84518, 4041990):
        for world in (1, 2, 4, 8):
            blocks = [capi.shard_block(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            covered = 0
            for (a, b), (c, d) in zip(blocks, blocks[1:]):
                assert a <= b == c <= d
            for a, b in blocks:
                covered += b - a
            assert covered == n
            per = capi.padded_rows(n) // world
            assert all(b - a == per for a, b in blocks if b < n)


def _gloo_shard_worker(rank, world, port, out):
    """Host-side model of the sharded upload: each rank contributes its id block, an all-gather makes the array whole."""
    import numpy as np
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 6146
    full = np.arange(3 * n, dtype=np.float64).reshape(n, 3) * 0.5
    per = capi.padded_rows(n) // world
    lo, hi = capi.shard_block(n, rank, world)
    mine = torch.zeros(per, 3, dtype=torch.float64)
    mine[: hi - lo] = torch.from_numpy(full[lo:hi])          # what crosses "PCIe" on this rank
    parts = [torch.zeros(per, 3, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(parts, mine)                               # what crosses "NVLink"
    whole = torch.cat(parts)[:n].numpy()
    with open(os.path.join(out, f"s{rank}"), "w") as f:
        f.write("ok" if np.array_equal(whole, full) else "bad")
    dist.destroy_process_group()


def test_sharded_upload_model_gloo_world2(tmp_path):
    import torch.multiprocessing as mp

    port = 29100 + os.getpid() % 300
    mp.spawn(_gloo_shard_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert [open(tmp_path / f"s{r}").read() for r in range(2)] == ["ok", "ok"]


def _gloo_worker(rank, world, port, out):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    calls = []

    def make():
        calls.append(rank)
        return bytes(range(128))

    uid = capi.distribute_unique_id(rank, make)
    ok = uid == bytes(range(128)) and calls == ([0] if rank == 0 else [])
    # a malformed id must be rejected on every rank
    try:
        capi.distribute_unique_id(rank, lambda: b"short")
        ok = False
    except ValueError:
        pass
    with open(os.path.join(out, f"r{rank}"), "w") as f:
        f.write("ok" if ok else "bad")
    dist.destroy_process_group()


def test_unique_id_distribution_gloo_world2(tmp_path):
    import torch.multiprocessing as mp

    port = 29500 + os.getpid() % 400
    mp.spawn(_gloo_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert [open(tmp_path / f"r{r}").read() for r in range(2)] == ["ok", "ok"]


def _gpu_count():
    try:
        import torch

        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("case,steps", [("dambreak", 6), ("dambreak@200", 50), ("dambreak@320", 8), ("dambreak_fine", 4),
                                        ("nx=400", 2), ("nx=800", 1)])
def test_split_solve_matches_single_gpu(case, steps):
    n = _gpu_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 4 if n >= 4 and case != "dambreak" else 2
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tools", "multi_gpu_check.py"), case, str(steps)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert r.returncode == 0 and lines, r.stdout[-2000:] + r.stderr[-2000:]
    rep = json.loads(lines[-1])
    assert rep["ok"] and rep["ranks_bit_identical"] and rep["flags_equal"] and rep["counters_equal"], rep
    assert rep["event_ids_equal"], rep  # every rank lists its own rows' events; their union is the 1-GPU list
    if case == "dambreak@200":
        assert rep["collisions_replayed"] > 0, rep  # the cross-rank collision replay really ran
    if case == "dambreak@320":
        assert rep["particles_ghosted"] > 0, rep
    if case.startswith("nx="):
        assert rep["vs_compiled_reference"]["ok"], rep


@pytest.mark.gpu
def test_matrix_free_switch_is_ignored_on_several_gpus():
    """MPSB_MATRIX_FREE=1 recomputes the off-diagonals from positions, and positions of halo columns are not exchanged:
    an attached handle must keep streaming the stored entries (ppe_matrix_free 0) and stay at parity with the
    single-GPU handle of the same process, which does run matrix-free."""
    if _gpu_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    port = 29600 + (os.getpid() + 7) % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tools", "multi_gpu_check.py"), "nx=400", "2"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, MPSB_MATRIX_FREE="1", MPSB_NO_CLUSTER_SOLVE="1"))
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert r.returncode == 0 and lines, r.stdout[-2000:] + r.stderr[-2000:]
    rep = json.loads(lines[-1])
    assert rep["ok"] and rep["matrix_free_on_multi"] == 0 and rep["vs_compiled_reference"]["ok"], rep


@pytest.mark.gpu
def test_cpp_host_runs_one_process_per_gpu(tmp_path):
    """./mps launched as two processes (RANK/WORLD_SIZE/LOCAL_RANK/MASTER_* in the environment, C++ rendezvous, no
    Python): rank 0's final checkpoint against a single-process run of the same case — flags identical, fields
    within 1e-10 (only the reduction order of the split solve differs)."""
    if _gpu_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    import numpy as np

    from mps_b200 import cases as C
    from tests.test_host import MPS_EXE, read_state

    S = C.load_settings(C.case_paths("dambreak_fine")[0])

    def launch(tag, world, **extra):
        d = tmp_path / tag
        d.mkdir()
        yml = C.write_settings_yaml(S, str(d / "settings.yml"), finish_time=8.5 * S.dt, output_interval=1.0)
        procs = []
        for rank in range(world):
            env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                       MASTER_PORT=str(29900 + os.getpid() % 90), **extra)
            procs.append(subprocess.Popen([MPS_EXE, "-s", yml, "--quiet", "--state-out", str(d / "final.bin")], cwd=ROOT,
                                          stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env))
        outs = [p.communicate(timeout=300)[0] for p in procs]
        assert all(p.returncode == 0 for p in procs), outs
        return read_state(d / "final.bin")

    one, two = launch("one", 1), launch("two", 2, MPSB_MIN_SLAB_PARTICLES="0")  # 6 146 particles: force the sharded path
    rep2 = launch("replicated", 2)  # default below 32 768 particles: replicas of the single-GPU step, bit for bit
    for k in ("type", "boundary_condition", "position", "velocity", "pressure"):
        assert np.array_equal(one[k], rep2[k]), k
    assert np.array_equal(one["type"], two["type"]) and np.array_equal(one["boundary_condition"], two["boundary_condition"])
    for k in ("position", "velocity", "pressure"):
        scale = max(float(np.abs(one[k]).max()), 1e-300)
        assert float(np.abs(one[k] - two[k]).max()) / scale <= 1e-10, k
