"""Data-parallel training of one client on the sm_100a engine (ops/engine_dp.py) against training on one engine.
The ranks are threads on ONE GPU here (``ThreadGroup``: the cross-rank sums go through host barriers), so the single-GPU test
tier runs the real multi-rank arithmetic: per-layer BatchNorm sums forward and backward, loss-scale weighting of ragged splits,
the fallback for batches smaller than the group.  What is compared and why: tests/dp_check.py.  The NCCL form of the same checks
is tests/dist_worker_dp_gpu.py."""
import os
import sys
import threading

import pytest
import torch

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
DEV = torch.device("cuda:0")


def _on_threads(world, fn):
    """fn(engine) on `world` DataParallelB200Engine members, one thread each -> list of results."""
    import dp_check
    from neuroimagedisttraining_b200.ops.engine_dp import DataParallelB200Engine, ThreadGroup
    group = ThreadGroup(world)
    engines = [DataParallelB200Engine(dp_check.make_model(), DEV, group.member(r), max_batch=8) for r in range(world)]
    outs, errors = [None] * world, []

    def work(r):
        try:
            torch.cuda.set_device(DEV)
            outs[r] = fn(engines[r])
        except BaseException as e:      # noqa: BLE001 - a failing rank must not leave the others in the barrier
            errors.append(e)
            group._barrier.abort()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(600)
    assert not errors, errors
    return outs


@pytest.mark.parametrize("world", [1, 2, 3, 4])
def test_identical_shares_reproduce_one_engine_on_one_share(world):
    import dp_check
    from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
    x, y = dp_check.make_inputs(DEV)
    ref = dp_check.run_sharp_plain(B200Engine(dp_check.make_model(), DEV, max_batch=8), x, y)
    outs = _on_threads(world, lambda eng: dp_check.run_sharp_dp(eng, x, y, world))
    for r in range(world):
        ok, errs = dp_check.check_sharp(outs[r], ref, world)
        assert ok, (world, r, errs)
        assert torch.equal(outs[r]["w"], outs[0]["w"])


@pytest.mark.parametrize("world", [2, 3])
def test_split_batches_train_like_whole_batches(world):
    import dp_check
    from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
    x, y = dp_check.make_inputs(DEV)
    ref = dp_check.programme(B200Engine(dp_check.make_model(), DEV, max_batch=8), x, y, dp_check.BATCHES, 6)
    outs = _on_threads(world, lambda eng: dp_check.programme(eng, x, y, dp_check.BATCHES, 6))
    for r in range(world):
        ok, errs = dp_check.check_coarse(outs[r], ref)
        assert ok, (world, r, errs)
        # every rank ends with the same model (same sums, same update) and the same evaluation
        assert torch.equal(outs[r]["w"], outs[0]["w"]), (world, r)
        assert torch.equal(outs[r]["eval"], outs[0]["eval"])


def test_share_covers_the_batch():
    from neuroimagedisttraining_b200.ops.engine_dp import share
    for B in range(1, 20):
        for w in range(1, 9):
            parts = [share(B, w, r) for r in range(w)]
            assert parts[0][0] == 0 and parts[-1][1] == B
            assert all(parts[i][1] == parts[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in parts]
            assert max(sizes) - min(sizes) <= 1
