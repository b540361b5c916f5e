"""CPU stand-ins for the fused symmetric-memory aggregation kernel: the in-memory multi-rank communicator computes what the kernel's
P2P path computes, and the flag-protocol checker accepts the kernel's barrier scheme and rejects weakened variants."""
import pytest
import torch

from neuroimagedisttraining_b200 import ops
from neuroimagedisttraining_b200.parallel.sym_model import (FakeSymmetricWorld, check_flag_protocol, chunk_bounds,
                                                            fake_reduce_mask_broadcast)


@pytest.mark.parametrize("world,n,n_mask", [(1, 64, 0), (2, 1000, 600), (3, 4 * 333, 4 * 333), (8, 4 * 1001, 2048), (16, 64, 8)])
def test_fake_communicator_matches_rank_ordered_sum(world, n, n_mask):
    g = torch.Generator().manual_seed(world)
    w = FakeSymmetricWorld(world, n)
    for r in range(world):
        w.partial(r).copy_(torch.randn(n, generator=g))
    mask = (torch.rand(n_mask, generator=g) > 0.5).float() if n_mask else None
    outs = fake_reduce_mask_broadcast(w, mask)
    want = ops.reference.weighted_aggregate(torch.empty(n), [w.partial(r) for r in range(world)], [1.0] * world, mask)
    for o in outs:                                   # every rank holds the same bits, equal to the rank-ordered fp32 sum
        assert torch.equal(o, want)
    covered = sorted(chunk_bounds(n, world, r) for r in range(world))
    assert covered[0][0] == 0 and max(hi for _, hi in covered) == n
    assert all(a[1] == b[0] or b[0] == b[1] for a, b in zip(covered, covered[1:]))      # contiguous, later chunks may be empty
    assert w.epoch == 1


@pytest.mark.parametrize("world", [2, 4, 8])
def test_flag_protocol_of_the_kernel_is_race_free(world):
    rep = check_flag_protocol(world=world, launches=3, schedules=30, seed=world)
    assert rep.ok, rep.violations[:3]
    rep = check_flag_protocol(world=world, launches=3, schedules=30, seed=world + 100, stall_rank=world - 1)
    assert rep.ok, rep.violations[:3]


def test_flag_protocol_survives_epoch_wraparound():
    first = (1 << 31) - 2                            # 2e+1 / 2e+2 cross 2^32 during the second launch
    assert check_flag_protocol(world=4, launches=4, first_epoch=first, schedules=20, seed=3, stall_rank=1).ok
    broken = check_flag_protocol(world=4, launches=4, first_epoch=first, schedules=20, seed=3, stall_rank=1, wrap_safe=False)
    assert not broken.ok                             # an unsigned >= comparison deadlocks or lets a rank through early


def test_checker_rejects_weakened_protocols():
    no_exit = check_flag_protocol(world=4, launches=3, schedules=40, seed=1, exit_barrier=False, stall_rank=2)
    assert not no_exit.ok and any("result piece" in v or "partial" in v for v in no_exit.violations)
    no_entry = check_flag_protocol(world=4, launches=2, schedules=40, seed=2, entry_barrier=False, stall_rank=0)
    assert not no_entry.ok and any("partial" in v for v in no_entry.violations)
