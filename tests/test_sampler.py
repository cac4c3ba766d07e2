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


# ---- property tests (hypothesis): the sharding contract for arbitrary sizes ----
from hypothesis import given, settings, strategies as st  # noqa: E402


@settings(max_examples=60, deadline=None)
@given(n=st.integers(1, 400), world=st.integers(1, 9), epoch=st.integers(0, 5))
def test_shards_partition_the_padded_epoch(n, world, epoch):
    """Every rank gets ceil(n / world) indices; together they cover the dataset, and the padding only repeats
    indices from the head of the epoch's permutation (torch's DistributedSampler contract)."""
    shards = [epoch_indices(n, world, r, epoch).tolist() for r in range(world)]
    per = -(-n // world)
    assert all(len(s) == per for s in shards)
    everything = [i for s in shards for i in s]
    assert set(everything) == set(range(n))
    assert len(everything) - n == per * world - n
    ref = DistributedSampler(_DS(n), num_replicas=world, rank=0)
    ref.set_epoch(epoch)
    assert list(ref) == shards[0]


@settings(max_examples=40, deadline=None)
@given(n=st.integers(2, 300), epoch_a=st.integers(0, 20), epoch_b=st.integers(0, 20))
def test_epoch_seeding_is_deterministic_and_epoch_dependent(n, epoch_a, epoch_b):
    a1, a2 = epoch_indices(n, 1, 0, epoch_a), epoch_indices(n, 1, 0, epoch_a)
    assert a1.tolist() == a2.tolist()
    if epoch_a != epoch_b and n > 8:
        assert a1.tolist() != epoch_indices(n, 1, 0, epoch_b).tolist()   # 1/n! chance of a false alarm
