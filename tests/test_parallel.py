"""World-size-2 gloo tests (CPU) of the window sharding + gather + alignment driver."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import vda_oracle as O
from vda_b200 import parallel
from vda_b200.video_depth import window_frame_indices


def test_shard_windows_partition():
    for n in (1, 2, 5, 24, 47):
        for world in (1, 2, 4, 8):
            shards = [parallel.shard_windows(n, world, r) for r in range(world)]
            assert sum(shards, []) == list(range(n))
            assert max(len(s) for s in shards) == (n + world - 1) // world
    assert [len(parallel.shard_windows(47, 8, r)) for r in range(8)] == [6, 6, 6, 6, 6, 6, 6, 5]


def _fake_raw(ids, n_frames, h=9, w=7):
    """Deterministic per-window 'depth': a function of the source frame index with a per-window
    scale/shift, so that alignment has something to undo."""
    wins = window_frame_indices(n_frames)
    out = torch.empty(len(ids) * 32, h, w)
    yy, xx = torch.meshgrid(torch.arange(h, dtype=torch.float32), torch.arange(w, dtype=torch.float32), indexing="ij")
    for j, k in enumerate(ids):
        s, t = 1.0 + 0.1 * ((k * 7) % 5), 0.05 * ((k * 3) % 4)
        for i, f in enumerate(wins[k]):
            base = 2.0 + torch.sin(0.3 * xx + 0.1 * f) * torch.cos(0.2 * yy - 0.05 * f)
            out[j * 32 + i] = s * base + t
    return out


def _align_cpu(raw):
    return torch.from_numpy(np.stack(O.align_windows([raw[i].numpy().copy() for i in range(raw.shape[0])])))


def _worker(rank, world, port, n_frames, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert parallel.active()
        got = parallel.distributed_infer(n_frames, window_frame_indices, lambda ids: _fake_raw(ids, n_frames), _align_cpu)
        q.put((rank, got.numpy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_frames", [(2, 100), (2, 23), (3, 30)])
def test_distributed_driver_equals_single_process(world, n_frames):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n_win = len(window_frame_indices(n_frames))
    want = _align_cpu(_fake_raw(list(range(n_win)), n_frames))[:n_frames].numpy()
    for r in range(world):
        assert res[r].shape == want.shape
        np.testing.assert_array_equal(res[r], want)


def _worker_rank0(rank, world, port, n_frames, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        prev = parallel.set_result_rank(0)
        assert prev is None and parallel.wants_result() == (rank == 0)
        got = parallel.distributed_infer(n_frames, window_frame_indices, lambda ids: _fake_raw(ids, n_frames), _align_cpu)
        q.put((rank, None if got is None else got.numpy()))
        parallel.set_result_rank(None)
    finally:
        dist.destroy_process_group()


def test_result_on_one_rank_only():
    """set_result_rank(0): rank 0 gets the same aligned maps, the other ranks take part and return None."""
    world, n_frames = 2, 60
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_rank0, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n_win = len(window_frame_indices(n_frames))
    want = _align_cpu(_fake_raw(list(range(n_win)), n_frames))[:n_frames].numpy()
    np.testing.assert_array_equal(res[0], want)
    assert res[1] is None
