"""World-size-2/3 gloo tests (CPU) of the sharded alignment scan (vda_b200.parallel.WindowScan): round-robin
windows, 3-map all-gather, neighbour tail exchange, per-rank blending, and the three result modes -- compared with
the oracle's sequential numpy scan (video_depth.py:120-160) on the same per-window maps, bit for bit."""
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

H, W = 9, 7


def test_round_robin_partition_and_final_ranges():
    for n in (1, 2, 5, 24, 47):
        for world in (1, 2, 4, 8):
            shards = [parallel.shard_windows(n, world, r) for r in range(world)]
            assert sorted(sum(shards, [])) == list(range(n))
            assert max(len(s) for s in shards) == parallel.n_rounds(n, world)
            for r, s in enumerate(shards):
                assert all(k % world == r for k in s)
        # the final ranges of all windows tile [0, 32 + 22 (n-1)) exactly once
        cover = []
        for k in range(n):
            lo, m = parallel.final_range(k, n)
            cover += list(range(lo, lo + m))
        assert cover == list(range(32 + 22 * (n - 1)))
    assert [len(parallel.shard_windows(47, 8, r)) for r in range(8)] == [6, 6, 6, 6, 6, 6, 6, 5]


class CpuOps:
    """WindowScan arithmetic restated with the oracle's numpy float32 operations (test infrastructure)."""

    def empty(self, *shape):
        return torch.zeros(*shape)

    def scale_shift_table(self, n):
        t = torch.zeros(n, 2)
        t[:, 0] = 1.0
        return t

    def scale_shift(self, pred, target, ss, metric):
        s, t = (1.0, 0.0) if metric else O.scale_and_shift(pred.numpy().reshape(-1, pred.shape[-1]),
                                                            target.numpy().reshape(-1, target.shape[-1]))
        ss[0], ss[1] = float(s), float(t)

    def affine_clamp(self, src, dst, ss):
        s, t = np.float32(ss[0].item()), np.float32(ss[1].item())
        dst.copy_(torch.from_numpy(O._affine_clamp(src.numpy().copy(), s, t)))

    def affine_clamp_blend(self, pre, post, ss):
        s, t = np.float32(ss[0].item()), np.float32(ss[1].item())
        fixed = [O._affine_clamp(post[i].numpy().copy(), s, t) for i in range(post.shape[0])]
        out = O.interpolate_frames([pre[i].numpy().copy() for i in range(pre.shape[0])], fixed)
        pre.copy_(torch.from_numpy(np.stack(out)))


def _fake_raw(k, n_frames):
    """Deterministic 'depth' of window k: a function of the source frame index with a per-window scale/shift (so
    the alignment has something to undo) and some negative values after alignment (so the clamp matters)."""
    wins = window_frame_indices(n_frames)
    out = torch.empty(32, H, W)
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    s, t = 1.0 + 0.1 * ((k * 7) % 5), 0.05 * ((k * 3) % 4)
    for i, f in enumerate(wins[k]):
        base = 0.8 + torch.sin(0.3 * xx + 0.1 * f) * torch.cos(0.2 * yy - 0.05 * f)
        out[i] = s * base + t
    return out


def _oracle(n_frames, metric=False):
    n_win = len(window_frame_indices(n_frames))
    depth_list = [m.numpy().copy() for k in range(n_win) for m in _fake_raw(k, n_frames)]
    return np.stack(O.align_windows(depth_list, metric))[:n_frames]


def _scan_all(n_frames, metric=False, group=None):
    """Runs this rank's side of the scan; returns (frames array with NaN where not delivered, delivered mask)."""
    n_win = len(window_frame_indices(n_frames))
    scan = parallel.WindowScan(n_win, H, W, CpuOps(), metric, group=group)
    out = np.full((32 + 22 * (n_win - 1), H, W), np.nan, np.float32)
    for j in range(parallel.n_rounds(n_win, scan.world)):
        k = j * scan.world + scan.rank
        raw = _fake_raw(k, n_frames) if k < n_win else None
        for lo, maps in scan.push(j, raw):
            assert np.isnan(out[lo:lo + maps.shape[0]]).all(), "a frame was delivered twice"
            out[lo:lo + maps.shape[0]] = maps.numpy()
    return out[:n_frames]


@pytest.mark.parametrize("n_frames,metric", [(100, False), (23, False), (20, False), (55, True)])
def test_single_rank_scan_equals_oracle(n_frames, metric):
    got = _scan_all(n_frames, metric, group=parallel.SOLO)
    np.testing.assert_array_equal(got, _oracle(n_frames, metric))


def _worker(rank, world, port, n_frames, mode, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert parallel.active()
        if mode == "sharded":
            parallel.set_result_sharded()
        else:
            parallel.set_result_rank(None if mode == "all" else int(mode))
        assert parallel.wants_result() == (mode in ("all", "sharded") or int(mode) == rank)
        q.put((rank, _scan_all(n_frames)))
    finally:
        dist.destroy_process_group()


def _run(world, n_frames, mode):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, mode, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=180) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return res


@pytest.mark.parametrize("world,n_frames", [(2, 100), (2, 23), (3, 130), (2, 20), (4, 100)])
def test_sharded_scan_every_rank_gets_the_oracle_result(world, n_frames):
    res = _run(world, n_frames, "all")
    want = _oracle(n_frames)
    for r in range(world):
        np.testing.assert_array_equal(res[r], want)


def test_result_on_one_rank_only():
    """set_result_rank(1): rank 1 gets every frame, the other rank takes part in the exchange and gets none."""
    res = _run(2, 120, "1")
    np.testing.assert_array_equal(res[1], _oracle(120))
    assert np.isnan(res[0]).all()


def test_result_on_rank0_with_idle_ranks_in_the_last_round():
    """4 ranks, 5 windows: ranks 1..3 have no window in round 1 but still take part in its exchange and gather."""
    res = _run(4, 100, "0")
    np.testing.assert_array_equal(res[0], _oracle(100))
    assert all(np.isnan(res[r]).all() for r in (1, 2, 3))


def test_result_sharded_frames_partition_the_video():
    """set_result_sharded(): each rank delivers exactly the final frames of its own windows; together they are the
    oracle's result."""
    world, n_frames = 3, 150
    res = _run(world, n_frames, "sharded")
    want = _oracle(n_frames)
    have = np.stack([~np.isnan(res[r]).all(axis=(1, 2)) for r in range(world)])
    assert (have.sum(0) == 1).all()
    n_win = len(window_frame_indices(n_frames))
    for r in range(world):
        mine = np.zeros(n_frames, bool)
        for k in parallel.shard_windows(n_win, world, r):
            lo, m = parallel.final_range(k, n_win)
            mine[lo:min(lo + m, n_frames)] = True
        assert (have[r] == mine).all()
        np.testing.assert_array_equal(res[r][mine], want[mine])
