"""Data-parallel sharding of image pairs over ranks (one process per GPU).

Pairs are independent, so there is no collective inside a step; the only communication is one final
gather of fixed-size match records (NCCL over NVLink on the box, gloo in the CPU tests).
(reference: the eval loop is single-process, utils/export_predictions.py:30-34; this is its sharded form)
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def pairs_for_rank(num_pairs: int, rank: int, world: int):
    """Round-robin assignment: pair i belongs to rank i % world."""
    return list(range(rank, num_pairs, world))


def gather_match_records(matches0: torch.Tensor, scores0: torch.Tensor, pair_ids, num_pairs: int, world: int):
    """All-gather per-rank records (pair id, matches0 as int32, matching_scores0) into full tables.

    matches0 / scores0: (P_local, K) on the rank's device; pair_ids: the global ids of those pairs.
    Returns (matches (num_pairs, K) int32, scores (num_pairs, K) fp32) on every rank; rows of pairs no
    rank owned stay at -1 / 0.  Ranks may own different numbers of pairs: records are padded to
    ceil(num_pairs / world) rows so that one fixed-size collective suffices.
    """
    k = matches0.shape[1]
    dev = matches0.device
    rows = -(-num_pairs // world)
    rec_m = torch.full((rows, k + 1), -1, dtype=torch.int32, device=dev)
    rec_s = torch.zeros((rows, k), dtype=torch.float32, device=dev)
    n = len(pair_ids)
    if n:
        rec_m[:n, 0] = torch.tensor(pair_ids, dtype=torch.int32, device=dev)
        rec_m[:n, 1:] = matches0.to(torch.int32)
        rec_s[:n] = scores0
    if world > 1:
        all_m = [torch.empty_like(rec_m) for _ in range(world)]
        all_s = [torch.empty_like(rec_s) for _ in range(world)]
        dist.all_gather(all_m, rec_m)
        dist.all_gather(all_s, rec_s)
        rec_m, rec_s = torch.cat(all_m), torch.cat(all_s)
    out_m = torch.full((num_pairs, k), -1, dtype=torch.int32, device=dev)
    out_s = torch.zeros((num_pairs, k), dtype=torch.float32, device=dev)
    valid = rec_m[:, 0] >= 0
    ids = rec_m[valid, 0].long()
    out_m[ids] = rec_m[valid, 1:]
    out_s[ids] = rec_s[valid]
    return out_m, out_s
