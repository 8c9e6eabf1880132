"""Window-level data parallelism for `infer_video_depth` (one process per GPU, torch.distributed).

Windows are independent forward passes over raw frame indices, so rank r runs the
forward of a contiguous block of windows and nothing is exchanged until the per-window depth maps
exist.  They are then all-gathered (NCCL over NVLink on GPUs; the same code runs over gloo on CPU
tensors in tests) and every rank runs the identical sequential scale/shift scan + cross-fade
(video_depth.py:120-160), which is O(windows) tiny reductions.  The reference has no distributed
code at all; this module is the only addition to its driver semantics.
"""
import torch
import torch.distributed as dist

INFER_LEN = 32

# Which ranks receive the aligned result.  None = every rank (SPMD callers: each process returns the full array, the
# default).  An int = only that rank runs the alignment scan and downloads the maps; the others take part in the
# exchange and return None.  With 8 ranks on one host, 8 copies of a GB-sized result through the same host memory
# cost more than the 6 windows each rank computes (profiles/r01c_video_8gpu.jsonl).  Set with set_result_rank()
# or the environment variable VDA_RESULT_RANK ("all" | rank number).
import os as _os

_RESULT_RANK = None if _os.environ.get("VDA_RESULT_RANK", "all") == "all" else int(_os.environ["VDA_RESULT_RANK"])


def set_result_rank(rank):
    """rank: None (every rank gets the result) or the one rank that does.  Returns the previous setting."""
    global _RESULT_RANK
    prev, _RESULT_RANK = _RESULT_RANK, (None if rank is None else int(rank))
    return prev


def wants_result(group=None):
    """True if this rank returns the aligned result (always True outside a multi-rank group)."""
    return _RESULT_RANK is None or not active(group) or dist.get_rank(group) == _RESULT_RANK


def shard_windows(n_windows, world, rank):
    """Contiguous block of window ids owned by `rank`: ceil(n/world) per rank, trailing ranks may be
    short or empty.  Contiguity keeps each rank's set of needed source frames compact."""
    per = (n_windows + world - 1) // world
    lo = min(rank * per, n_windows)
    return list(range(lo, min(lo + per, n_windows)))


def gather_windows(local_raw, n_windows, group=None):
    """local_raw: [len(shard)*32, H, W] per-window depth of this rank's windows (any device the
    group's backend supports) -> [n_windows*32, H, W] on every rank, in window order."""
    world = dist.get_world_size(group)
    if world == 1:
        return local_raw
    per = (n_windows + world - 1) // world
    h, w = local_raw.shape[1:]
    padded = torch.zeros(per * INFER_LEN, h, w, dtype=local_raw.dtype, device=local_raw.device)
    padded[: local_raw.shape[0]].copy_(local_raw)
    out = torch.empty(world * per * INFER_LEN, h, w, dtype=local_raw.dtype, device=local_raw.device)
    dist.all_gather_into_tensor(out, padded, group=group)
    return out[: n_windows * INFER_LEN]


def distributed_infer(n_frames, window_ids_fn, run_windows_fn, align_fn, group=None):
    """Generic sharded driver.
    window_ids_fn(n_frames) -> list of windows; run_windows_fn(ids) -> [len(ids)*32, H, W] raw depth;
    align_fn(raw_all) -> aligned [>= n_frames, H, W].  Returns aligned[:n_frames] on every rank that wants the
    result (see set_result_rank), None on the others."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n_windows = len(window_ids_fn(n_frames))
    mine = shard_windows(n_windows, world, rank)
    raw = run_windows_fn(mine)
    raw_all = gather_windows(raw, n_windows, group)
    if not wants_result(group):
        return None
    return align_fn(raw_all)[:n_frames]


def active(group=None):
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
