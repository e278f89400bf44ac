"""Track sharding for multi-GPU runs: tracks are independent (own LSTM state, own buffers), so every rank
demixes its own slice and there is NO collective on the data path.  The only communication is the
aggregation of the measurement (sum of audio seconds, max of elapsed time) after the timed region.
Used by bench.py; covered on CPU by tests/test_multi_gloo.py (gloo, world_size 2)."""
from __future__ import annotations


def shard_tracks(n_tracks: int, rank: int, world: int) -> range:
    """Contiguous, balanced partition: the first n_tracks % world ranks take one extra track."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, extra = divmod(n_tracks, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def aggregate(audio_seconds_local: float, elapsed_ms_local: float, dist=None, device=None):
    """(total audio seconds over all ranks, max elapsed ms over ranks).  `dist` is torch.distributed when a
    process group is initialised, else None (single process)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return audio_seconds_local, elapsed_ms_local
    import torch

    t = torch.tensor([audio_seconds_local, elapsed_ms_local], dtype=torch.float64, device=device)
    s = t.clone()
    dist.all_reduce(s, op=dist.ReduceOp.SUM)
    m = t.clone()
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    return float(s[0].item()), float(m[1].item())
