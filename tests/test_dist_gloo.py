"""-m "not gpu": the N>1 host logic with world_size 2 on the gloo backend (CPU).

Each rank renders only its interleaved row bands (rt_set_partition semantics) with the emulated device code, the bands
are gathered to rank 0 and must be bit-identical to the 1-rank frame; per-rank ray counts all-reduce to the frame total.
On the GPU box the gather is the fused peer store into rank 0's framebuffer (bench.py); here gloo stands in for NVLink.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import importlib
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = importlib.import_module("unity-raytracer_b200.scenes")
    from emu import emu as E
    s, p = sc.random_soup(seed=5), sc.soup_params(96, 70, 2, 1)
    e = E.EmuTracer(); e.set_scene(s)
    e.set_partition(rank, world, 8)
    g = e.render(p)
    rows = np.arange(70)
    mine = np.repeat(((rows // 8) % world) == rank, 96)
    frame = torch.from_numpy(np.where(mine[:, None], g["rgba8"], 0).astype(np.uint8))
    ids = torch.from_numpy(np.where(mine[:, None], g["hit_id"], 0).astype(np.int32))
    rays = torch.tensor([g["stats"]["rays_primary"], g["stats"]["rays_shadow"], g["stats"]["rays_mirror"]], dtype=torch.int64)
    dist.reduce(frame, 0, op=dist.ReduceOp.SUM)        # disjoint bands: sum == gather
    dist.reduce(ids, 0, op=dist.ReduceOp.SUM)
    dist.all_reduce(rays)
    if rank == 0:
        e.set_partition(0, 1, 32)
        full = e.render(p)
        ok = (np.array_equal(frame.numpy(), full["rgba8"]) and np.array_equal(ids.numpy(), full["hit_id"])
              and rays.tolist() == [full["stats"]["rays_primary"], full["stats"]["rays_shadow"], full["stats"]["rays_mirror"]])
        q.put(bool(ok))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_band_partition_gloo(emu):
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    assert q.get() is True


def _shared_frame_worker(rank, world, port, q):
    """bench.py's N-GPU e2e path on the CPU: ONE host frame in POSIX shared memory (bench.SharedFrame), every rank writes the
    rows of its own bands into it (what rt_render does with a partition set), rank 0 compares with the 1-rank frame."""
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import importlib
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    bench = importlib.import_module("bench")
    sc = importlib.import_module("unity-raytracer_b200.scenes")
    from emu import emu as E
    X, Y, band = 96, 70, 8
    s, p = sc.random_soup(seed=5), sc.soup_params(X, Y, 2, 1)
    spec = {"rgba32f": ((X * Y, 4), np.float32), "rgba8": ((X * Y, 4), np.uint8), "hit_id": ((X * Y, 3), np.int32), "hit_t": ((X * Y,), np.float32)}
    shared = bench.SharedFrame(None, "gloo_test", spec, rank, world, dist)
    e = E.EmuTracer(); e.set_scene(s)
    e.set_partition(rank, world, band)
    g = e.render(p)
    mine = np.repeat(((np.arange(Y) // band) % world) == rank, X)
    for k, a in shared.arrays.items():
        a[mine] = g[k][mine]                                  # copy_rows_out: only the rows this rank owns
        a.flush()
    dist.barrier()
    ok = True
    if rank == 0:
        e.set_partition(0, 1, 32)
        full = e.render(p)
        ok = all(np.array_equal(np.asarray(shared.arrays[k]), full[k], equal_nan=True) for k in spec)
    paths = list(shared.paths)
    shared.close(world, dist)
    dist.barrier()
    if rank == 0:
        q.put(bool(ok and not any(os.path.exists(pth) for pth in paths)))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_ranks_fill_one_shared_host_frame_gloo(emu):
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_shared_frame_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    assert q.get() is True
