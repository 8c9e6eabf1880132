"""Window alignment scan of `infer_video_depth` (video_depth.py:120-160) and its window-level data parallelism
(one process per GPU, torch.distributed).  The reference has no distributed code; this module is the only
addition to its driver semantics, and the same code runs with one rank.

What the sequential alignment really depends on (video_depth.py:126-128, :138-158):
  * the scale/shift of window k is a least-squares fit of its slots 0 and 1 against `ref_align` = [window 0's
    slot 0, window k-1's ALIGNED slot 12]  ->  3 maps per window (slots 0, 1, 12) feed the whole scalar recurrence;
  * its slots 2..9 are cross-faded with window k-1's aligned slots 24..31  ->  8 maps from one neighbour;
  * everything else (slots 10..31) is an element-wise affine + clamp with that window's own scale/shift.
So windows are dealt ROUND-ROBIN: in round j rank r runs the forward of window k = j*R + r.  After the forward
  1. the 3 key maps are all-gathered (3.2 MB per window at 518x518) and every rank advances the identical scalar
     scan over the R windows of the round (one 2-map reduction + one 1-map affine per window);
  2. the raw slots 24..31 go to rank r+1 (ring send/recv), which aligns them with window k-1's scale/shift,
     known to every rank;
  3. each rank blends and aligns ITS OWN window and owns the 22 output frames that become final with it.
Nothing GB-sized is exchanged and nothing waits for the last forward: a window's frames are final one tiny exchange
after its forward, so their download overlaps the next round's compute.  Where the final frames go afterwards is the
caller's choice (set_result_rank / set_result_sharded): every rank (all-gather of each round's frames), one rank
(gather), or each rank keeps only the frames it owns.

The arithmetic is injected (`ops`): the product passes the CUDA kernels (vda_b200.video_depth.GpuAlignOps); the
world-size-2 gloo tests on CPU pass a torch restatement (tests/test_parallel.py).
"""
import os as _os

import torch
import torch.distributed as dist

INFER_LEN, OVERLAP, INTERP_LEN = 32, 10, 8       # video_depth.py:29-33
STEP = INFER_LEN - OVERLAP
KEY_SLOTS = (0, 1, 12)                           # pred slots 0, 1; slot 12 becomes the next window's ref_align[1]

# Which ranks receive the aligned result: "all" (SPMD callers: every process returns the full array, the default),
# an int (only that rank gets the maps; the others compute, exchange and return None) or "sharded" (every rank
# keeps just the frames of its own windows).  Set with set_result_rank() / set_result_sharded() or the environment
# variable VDA_RESULT_RANK ("all" | "sharded" | rank number).
_ev = _os.environ.get("VDA_RESULT_RANK", "all")
_RESULT = _ev if _ev in ("all", "sharded") else int(_ev)


SOLO = "solo"    # `group` sentinel of WindowScan: a single-rank scan even if a process group is initialised


def set_result_rank(rank):
    """rank: None (every rank gets the result) or the one rank that does.  Returns the previous setting."""
    global _RESULT
    prev, _RESULT = (None if _RESULT == "all" else _RESULT), ("all" if rank is None else int(rank))
    return prev


def set_result_sharded():
    """Every rank keeps only the final frames of its own windows (no result exchange at all)."""
    global _RESULT
    prev, _RESULT = _RESULT, "sharded"
    return prev


def result_mode():
    return _RESULT


_FORCE_SOLO = False


class solo:
    """Context manager: inside it this process runs whole videos by itself even if a process group is initialised
    (comparison runs in multi-GPU tests)."""

    def __enter__(self):
        global _FORCE_SOLO
        self.prev, _FORCE_SOLO = _FORCE_SOLO, True

    def __exit__(self, *exc):
        global _FORCE_SOLO
        _FORCE_SOLO = self.prev


def active(group=None):
    return (not _FORCE_SOLO) and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


def wants_result(group=None):
    """True if this rank returns aligned frames (always True outside a multi-rank group)."""
    return _RESULT in ("all", "sharded") or not active(group) or dist.get_rank(group) == _RESULT


def world_rank(group=None):
    return (dist.get_world_size(group), dist.get_rank(group)) if active(group) else (1, 0)


def n_rounds(n_windows, world):
    return (n_windows + world - 1) // world


def shard_windows(n_windows, world, rank):
    """Window ids of `rank`, one per round: rank, rank + world, ... (see the module docstring)."""
    return list(range(rank, n_windows, world))


def final_range(k, n_windows):
    """(first output frame, frame count) that become final when window k has been aligned: a window's last 8
    frames are still cross-faded with the next window (utils/util.py:65-74), except after the last one."""
    if k == 0:
        return 0, (INFER_LEN if n_windows == 1 else INFER_LEN - INTERP_LEN)
    return STEP * k + OVERLAP - INTERP_LEN, (STEP + INTERP_LEN if k == n_windows - 1 else STEP)


class WindowScan:
    """Incremental alignment of the windows of one video.  push(j, raw) takes this rank's raw maps of round j
    ([32, h, w] f32, or None if the rank has no window in the round) and returns [(first_frame, maps [m, h, w])]:
    the output frames that became final and that this rank is to deliver.  The returned tensors are views of
    buffers that are reused two rounds later.  No host synchronisation on CUDA tensors."""

    def __init__(self, n_windows, h, w, ops, metric=False, group=None):
        self.n, self.h, self.w, self.ops, self.metric = n_windows, h, w, ops, bool(metric)
        self.world, self.rank = (1, 0) if group == SOLO else world_rank(group)
        self.group = None if group == SOLO else group
        self.mode = _RESULT if self.world > 1 else "sharded"
        E = ops.empty
        self.keys_send = E(3, h, w)
        self.keys_all = E(self.world, 3, h, w) if self.world > 1 else self.keys_send.view(1, 3, h, w)
        self.tails = [E(INTERP_LEN, h, w) for _ in range(2)]     # raw slots 24..31 of the preceding window
        self.ref = E(2, h, w)                                    # ref_align (video_depth.py:126-128)
        self.ss = ops.scale_shift_table(n_windows)               # [n, 2] rows (scale, shift), row 0 = (1, 0)
        self.aligned = [E(INFER_LEN, h, w) for _ in range(2)]
        self.gathered = None
        if self.world > 1 and (self.mode == "all" or self.mode == self.rank):
            self.gathered = [E(self.world, INFER_LEN, h, w) for _ in range(2)]
        self.zero_tail = None
        self.round = 0

    # ---------------------------------------------------------------- exchange
    def _exchange(self, j, raw):
        ks = self.keys_send
        if raw is not None:
            ks[0:2].copy_(raw[0:2])
            ks[2].copy_(raw[KEY_SLOTS[2]])
        tail = self.tails[j & 1]
        if self.world == 1:
            tail.copy_(raw[INFER_LEN - INTERP_LEN:])
            return
        dist.all_gather_into_tensor(self.keys_all.view(-1, self.h, self.w), ks, group=self.group)   # concatenated form
        if raw is None:
            if self.zero_tail is None:
                self.zero_tail = self.ops.empty(INTERP_LEN, self.h, self.w).zero_()
            send = self.zero_tail
        else:
            send = raw[INFER_LEN - INTERP_LEN:]
        nxt, prv = (self.rank + 1) % self.world, (self.rank - 1) % self.world
        if self.group is not None:
            nxt, prv = dist.get_global_rank(self.group, nxt), dist.get_global_rank(self.group, prv)
        ops = [dist.P2POp(dist.isend, send, nxt, self.group), dist.P2POp(dist.irecv, tail, prv, self.group)]
        for work in dist.batch_isend_irecv(ops):
            work.wait()

    # ---------------------------------------------------------------- the scalar recurrence, identical on every rank
    def _scan_round(self, j):
        o = self.ops
        for k in range(j * self.world, min((j + 1) * self.world, self.n)):
            keys = self.keys_all[k - j * self.world]
            if k == 0:
                self.ref[0].copy_(keys[0])
                self.ref[1].copy_(keys[2])
            else:
                o.scale_shift(keys[0:2], self.ref, self.ss[k], self.metric)       # utils/util.py:40-62
                o.affine_clamp(keys[2], self.ref[1], self.ss[k])                  # video_depth.py:154-158

    # ---------------------------------------------------------------- this rank's window
    def _align_own(self, j, k, raw):
        o, out = self.ops, self.aligned[j & 1]
        if k == 0:
            out.copy_(raw)                                                        # video_depth.py:124-128
            return
        # the preceding window ran on rank-1 in this round, or (rank 0) on the last rank in the previous round
        tail = self.tails[(j - 1) & 1] if self.rank == 0 else self.tails[j & 1]
        if k == 1:
            out[0:INTERP_LEN].copy_(tail)              # window 0 enters the sequence as it is, unclamped (:124-125)
        else:
            o.affine_clamp(tail, out[0:INTERP_LEN], self.ss[k - 1])
        o.affine_clamp_blend(out[0:INTERP_LEN], raw[OVERLAP - INTERP_LEN:OVERLAP], self.ss[k])   # :144-149
        o.affine_clamp(raw[OVERLAP:], out[INTERP_LEN:INTERP_LEN + STEP], self.ss[k])             # :150-152

    def push(self, j, raw):
        assert j == self.round, "rounds must be pushed in order"
        self.round += 1
        k = j * self.world + self.rank
        if k >= self.n:
            raw = None
        assert raw is not None or self.world > 1
        self._exchange(j, raw)
        self._scan_round(j)
        if raw is not None:
            self._align_own(j, k, raw)
        mine = self.aligned[j & 1]
        if self.mode == "sharded":
            return [self._item(k, mine)] if raw is not None else []
        if self.mode == "all":
            dist.all_gather_into_tensor(self.gathered[j & 1].view(-1, self.h, self.w), mine, group=self.group)
        else:
            dst = self.mode if self.group is None else dist.get_global_rank(self.group, self.mode)
            dist.gather(mine, list(self.gathered[j & 1].unbind(0)) if self.gathered is not None else None, dst=dst,
                        group=self.group)
            if self.gathered is None:
                return []
        first = j * self.world
        return [self._item(kk, self.gathered[j & 1][kk - first]) for kk in range(first, min(first + self.world, self.n))]

    def _item(self, k, buf):
        lo, m = final_range(k, self.n)
        return lo, buf[:m]
