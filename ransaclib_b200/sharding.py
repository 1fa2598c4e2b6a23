"""Host logic of the two multi-GPU partitionings. Device-agnostic torch code so it
runs under gloo on CPU (tests) and NCCL on GPUs (bench.py).

* hypothesis sharding: contiguous hypothesis id ranges per rank (the lowest-id tie-break of
  RansacLib/ransac.h:195,288 then survives the cross-rank reduction), ONE all_gather of a
  (score, global id, model) record per rank per round;
* query sharding for batched localization: query q on rank q mod world, no collective.
"""
from __future__ import annotations

import torch

DBL_MAX = torch.finfo(torch.float64).max


def hypothesis_range(k_total: int, rank: int, world: int):
    """[lo, hi) of the hypothesis ids rank `rank` scores; contiguous and ascending in rank."""
    base, rem = divmod(k_total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def query_indices(num_queries: int, rank: int, world: int):
    return list(range(rank, num_queries, world))


def pack_record(score, local_id, id_offset, model):
    """(score, global id as f64, model...) — ids < 2^53 are exact in f64. local_id < 0 (no valid
    hypothesis on this rank) is sent as score = DBL_MAX so it can never win."""
    rec = torch.empty(2 + model.numel(), dtype=torch.float64, device=model.device)
    valid = local_id >= 0
    rec[0] = torch.where(valid, score, torch.full_like(score, DBL_MAX))
    rec[1] = (local_id + id_offset).to(torch.float64)
    rec[2:] = model
    return rec


def select_winner(gathered: torch.Tensor):
    """gathered: (world, 2 + model_size). Lexicographic (score, id) minimum; NaN scores never win.
    Returns (row index tensor, has_winner bool tensor)."""
    score = gathered[:, 0]
    usable = score < DBL_MAX  # False for NaN and for 'no model'
    s = torch.where(usable, score, torch.full_like(score, float("inf")))
    smin = s.min()
    ids = torch.where(s == smin, gathered[:, 1], torch.full_like(score, float("inf")))
    return torch.argmin(ids), usable.any()


def all_gather_records(rec: torch.Tensor, world: int):
    """The single collective of a scoring round."""
    if world == 1:
        return rec.view(1, -1)
    import torch.distributed as dist

    out = torch.empty((world, rec.numel()), dtype=rec.dtype, device=rec.device)
    dist.all_gather_into_tensor(out, rec)
    return out
