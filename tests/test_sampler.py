"""K11 index order must equal torch's DistributedSampler."""
import pytest
import torch
from torch.utils.data import DistributedSampler

from pytorch_distributed_rnn_b200.data.sampler import ShardedEpochSampler, epoch_indices


class _DS(torch.utils.data.Dataset):
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return i


@pytest.mark.parametrize("n,world", [(6912, 1), (6912, 2), (6912, 8), (6912, 12), (100, 3), (7, 4)])
@pytest.mark.parametrize("epoch", [0, 3])
def test_matches_distributed_sampler(n, world, epoch):
    for rank in range(world):
        ref = DistributedSampler(_DS(n), num_replicas=world, rank=rank)
        ref.set_epoch(epoch)
        assert list(ref) == epoch_indices(n, world, rank, epoch).tolist()


def test_union_of_rank_batches_is_local_batch():
    # the property behind the reference's world-size invariance
    n, bs = 6912, 1440
    local = epoch_indices(n, 1, 0, 0)
    for world in (2, 4, 8, 12):
        per = bs // world
        shards = [epoch_indices(n, world, r, 0) for r in range(world)]
        for step in range(n // bs):
            union = torch.cat([s[step * per:(step + 1) * per] for s in shards])
            assert sorted(union.tolist()) == sorted(local[step * bs:(step + 1) * bs].tolist())


def test_sampler_object():
    s = ShardedEpochSampler(10, 2, 1, shuffle=False)
    assert list(s) == [1, 3, 5, 7, 9] and len(s) == 5
