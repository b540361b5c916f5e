"""The weighting of the data-parallel engine's cross-rank sums (ops/engine_dp.py) on synthetic partial sums, ranks = threads, CPU:
with ragged shares every rank works in its own loss scale B / B_r, so a rank's partial sums enter the total with weight
f_r = B_r / B and the total is handed back divided by the receiving rank's f."""
import threading
import types

import torch

from neuroimagedisttraining_b200.ops.engine_dp import DataParallelB200Engine, ThreadGroup, share


def test_weighted_partial_sums_over_ranks():
    world, C, B = 3, 16, 8
    g = torch.Generator().manual_seed(0)
    rows = [5, 3, 7]
    parts = [torch.randn(rows[r] * 2 * C, generator=g) * 1e3 for r in range(world)]
    f = [(share(B, world, r)[1] - share(B, world, r)[0]) / B for r in range(world)]
    assert abs(sum(f) - 1.0) < 1e-12 and f == [3 / 8, 3 / 8, 2 / 8]
    group = ThreadGroup(world)
    outs = [torch.zeros(4 * C) for _ in range(world)]
    got = [None] * world

    def work(r):
        fake = types.SimpleNamespace(dp=group.member(r))
        got[r] = DataParallelB200Engine._global_rows(fake, parts[r], rows[r], C, outs[r], f[r])

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join(60) for t in th]
    total = sum(f[r] * parts[r].double().view(rows[r], 2 * C).sum(0) for r in range(world))
    for r in range(world):
        assert got[r] == 2                                            # two fp32 rows: value = hi + lo
        val = outs[r][:2 * C].double() + outs[r][2 * C:].double()
        torch.testing.assert_close(val, total / f[r], rtol=1e-12, atol=1e-9)
    # equal shares, weight 1: a plain sum
    group = ThreadGroup(2)
    outs = [torch.zeros(4 * C) for _ in range(2)]

    def work2(r):
        DataParallelB200Engine._global_rows(types.SimpleNamespace(dp=group.member(r)), parts[r], 3, C, outs[r], 1.0)

    th = [threading.Thread(target=work2, args=(r,)) for r in range(2)]
    [t.start() for t in th]
    [t.join(60) for t in th]
    want = sum(parts[r][:3 * 2 * C].double().view(3, 2 * C).sum(0) for r in range(2))
    torch.testing.assert_close(outs[0][:2 * C].double() + outs[0][2 * C:].double(), want, rtol=1e-12, atol=1e-9)
    assert torch.equal(outs[0], outs[1])
