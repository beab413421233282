import itertools

import pytest
import torch
from torch.utils.data.distributed import DistributedSampler as TorchSampler

from distributed_compute_pytorch_b200 import DistributedSampler


class _DS:
    def __init__(self, n): self.n = n
    def __len__(self): return self.n


@pytest.mark.parametrize("n,replicas", [(10, 2), (11, 4), (7, 8), (60000, 4), (1, 2), (128, 3)])
def test_matches_library_sampler(n, replicas):
    ds = _DS(n)
    for rank, shuffle, seed, drop_last, epoch in itertools.product(range(replicas), (True, False), (0, 5), (False, True), (0, 3)):
        if rank not in (0, replicas - 1) and n > 1000:
            continue
        ours = DistributedSampler(ds, num_replicas=replicas, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
        ref = TorchSampler(ds, num_replicas=replicas, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
        ours.set_epoch(epoch); ref.set_epoch(epoch)
        assert len(ours) == len(ref)
        assert list(ours) == list(ref)


def test_shards_cover_dataset_and_pad_by_wrapping():
    ds = _DS(10)
    shards = [list(DistributedSampler(ds, num_replicas=4, rank=r, shuffle=False)) for r in range(4)]
    assert all(len(s) == 3 for s in shards)
    flat = sorted(i for s in shards for i in s)
    assert set(flat) == set(range(10)) and len(flat) == 12
    assert shards[2] == [2, 6, 0] and shards[3] == [3, 7, 1]  # padded with the head of the list


def test_same_order_without_set_epoch():
    s = DistributedSampler(_DS(100), num_replicas=2, rank=0)
    assert list(s) == list(s)  # the reference never calls set_epoch (ref:main.py:127-131)
    a = list(s); s.set_epoch(1)
    assert list(s) != a


def test_bad_rank():
    with pytest.raises(ValueError):
        DistributedSampler(_DS(4), num_replicas=2, rank=2)
