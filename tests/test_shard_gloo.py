"""World-size-2 test of the pair sharding + final gather on the gloo backend (CPU)."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from glue_factory_b200 import shard


def _fake_matches(pair_id, k):
    g = torch.Generator().manual_seed(pair_id)
    return torch.randint(-1, k, (k,), generator=g), torch.rand(k, generator=g)


def _worker(rank, world, port, num_pairs, k, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ids = shard.pairs_for_rank(num_pairs, rank, world)
    ms = [_fake_matches(i, k) for i in ids]
    m = torch.stack([a for a, _ in ms]) if ms else torch.zeros(0, k, dtype=torch.int64)
    s = torch.stack([b for _, b in ms]) if ms else torch.zeros(0, k)
    gm, gs = shard.gather_match_records(m, s, ids, num_pairs, world)
    if rank == 0:
        torch.save((gm, gs), out)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_and_gather(tmp_path):
    num_pairs, k = 5, 16  # odd count: the ranks own 3 and 2 pairs
    out = str(tmp_path / "gathered.pt")
    mp.spawn(_worker, args=(2, 29531, num_pairs, k, out), nprocs=2, join=True)
    gm, gs = torch.load(out)
    for i in range(num_pairs):
        m, s = _fake_matches(i, k)
        assert torch.equal(gm[i].long(), m) and torch.equal(gs[i], s)


def test_partition_is_disjoint_and_complete():
    for world in (1, 2, 4, 8):
        owned = sorted(i for r in range(world) for i in shard.pairs_for_rank(37, r, world))
        assert owned == list(range(37))


def test_single_rank_gather_is_identity():
    m = torch.arange(12).view(3, 4) - 1
    s = torch.rand(3, 4)
    gm, gs = shard.gather_match_records(m, s, [0, 1, 2], 3, 1)
    assert torch.equal(gm.long(), m) and torch.equal(gs, s)
