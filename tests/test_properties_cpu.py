"""Property tests (hypothesis) for the index arithmetic everything else rests on: batch shards of the
single-process engine, the distributed sampler, the gradient-bucket plan, the dataset scatter and the
collective-variant policy.  Pure host code; runs without a GPU."""
import torch
import torch.nn as nn
from hypothesis import given, settings, strategies as st

import distributed_training_dl_b200 as dtdl
from distributed_training_dl_b200.parallel.arena import ParamArena
from distributed_training_dl_b200.parallel.data_parallel import DataParallel


@settings(max_examples=200, deadline=None)
@given(world=st.integers(1, 16), extra=st.integers(0, 500))
def test_shard_bounds_tile_the_batch_and_differ_by_at_most_one(world, extra):
    dp = DataParallel.__new__(DataParallel)
    dp.world = world
    n = world + extra                                   # a batch smaller than the GPU count is refused
    b = dp.shard_bounds(n)
    assert len(b) == world and b[0][0] == 0 and b[-1][1] == n
    assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
    sizes = [hi - lo for lo, hi in b]
    assert max(sizes) - min(sizes) <= 1 and min(sizes) >= 1
    assert sizes == sorted(sizes, reverse=True)          # the larger shards come first


@settings(max_examples=60, deadline=None)
@given(n=st.integers(1, 300), world=st.integers(1, 9), shuffle=st.booleans(), epoch=st.integers(0, 3),
       drop_last=st.booleans())
def test_sampler_ranks_partition_the_padded_epoch(n, world, shuffle, epoch, drop_last):
    ds = list(range(n))
    parts = []
    for rank in range(world):
        s = dtdl.DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=5, drop_last=drop_last)
        s.set_epoch(epoch)
        idx = list(s)
        assert len(idx) == len(s)
        parts.append(idx)
    assert len({len(p) for p in parts}) == 1             # every rank runs the same number of steps
    flat = [i for p in parts for i in p]
    assert all(0 <= i < n for i in flat)
    if drop_last and n >= world:
        assert len(set(flat)) == len(flat) == (n // world) * world
    elif not drop_last:
        assert set(flat) == set(range(n))                # nothing is lost; padding repeats samples
        assert len(flat) - n < world


@settings(max_examples=40, deadline=None)
@given(sizes=st.lists(st.integers(1, 300_000), min_size=1, max_size=24),
       cap=st.sampled_from([0.25, 1.0, 4.0]), first=st.sampled_from([0.05, 0.25, 1.0]))
def test_bucket_plan_covers_every_parameter_once(sizes, cap, first):
    params = [nn.Parameter(torch.zeros(s)) for s in sizes]
    arena = ParamArena(params, "cpu", bucket_cap_mb=cap, first_bucket_mb=first)
    off = 0
    for b in arena.buckets:
        assert b.offset == off and b.numel > 0 and b.offset % 8 == 0
        off += b.numel
    assert off == arena.numel
    seen = [id(p) for b in arena.buckets for p in b.params]
    assert sorted(seen) == sorted(id(p) for p in params)
    for p in params:                                     # a parameter's storage lies inside its bucket
        b = arena.buckets[arena.bucket_of[id(p)]]
        start = (p.data_ptr() - arena.param_flat.data_ptr()) // 4
        assert b.offset <= start and start + p.numel() <= b.offset + b.numel
    # buckets fill in reverse registration order (the order backward produces gradients in)
    first_bucket_ids = {id(p) for p in arena.buckets[0].params}
    assert id(params[-1]) in first_bucket_ids


@settings(max_examples=60, deadline=None)
@given(n=st.integers(1, 200), world=st.integers(1, 8))
def test_scatter_dataset_shards_are_equal_sized_and_complete(n, world):
    """ChainerMN's scatter_dataset: every rank gets an equal-sized shard (so all ranks run the same
    number of iterations); together the shards hold every sample, repeating fewer than `world`."""
    from distributed_training_dl_b200.utils.data import scatter_dataset

    class FakeComm:
        def __init__(self, rank, size):
            self.rank, self.size = rank, size

        def broadcast_object(self, obj, src=0):
            return data                                  # what rank 0 would have sent

    data = list(range(n))
    shards = [list(scatter_dataset(data if r == 0 else None, FakeComm(r, world), shuffle=True, seed=2))
              for r in range(world)]
    assert len({len(s) for s in shards}) == 1
    flat = [v for s in shards for v in s]
    assert set(flat) == set(data) and len(flat) - n < world


@settings(max_examples=100, deadline=None)
@given(world=st.sampled_from([2, 3, 4, 6, 8]), nbytes=st.integers(4, 1 << 31), nvls=st.booleans())
def test_variant_policy_is_total(world, nbytes, nvls):
    """Every message size maps to a variant; without multicast never to the multicast variant, and then
    one-shot only below the two-shot threshold (a large message through the one-shot schedule would
    move W times the bytes)."""
    from distributed_training_dl_b200.parallel.peer import PeerComm
    pc = PeerComm.__new__(PeerComm)
    pc.world, pc.use_nvls, pc.nvls_min_bytes, pc.two_shot_min_bytes = world, nvls, 0, 256 * 1024
    v = pc.choose_two_shot(nbytes)
    assert v in (0, 1, 2)
    if not nvls:
        assert v != 2
        if nbytes > 256 * 1024:
            assert v == 1
    key = max(w for w in PeerComm.variant_table() if w <= world)
    if nvls and PeerComm.variant_table()[key] == [(None, 2)]:
        assert v == 2
