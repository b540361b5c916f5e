"""Multi-process paths.  CPU (gloo, world 2): orchestration invariance.  GPU (>= 2 devices): the fused
symmetric-memory aggregation kernel against NCCL."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(script, nproc, extra_env=None, timeout=600, args=()):
    env = dict(os.environ)
    env.update(extra_env or {})
    port = 29500 + (os.getpid() % 2000)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", script), *args]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)


def test_salientgrads_two_ranks_cpu_matches_single_process():
    r = _torchrun("dist_worker_fl_cpu.py", 2, {"CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESULT ok=True" in r.stdout, r.stdout[-2000:]


def test_dispfl_over_three_ranks_matches_single_process():
    """DisPFL with its clients spread over ranks (owners train / prune and broadcast once per round): personal models, masks, accuracy
    series and the Hamming matrix equal the one-process run, with client drop-out and with neighbour aggregation."""
    r = _torchrun("dist_worker_dispfl_cpu.py", 3, {"CUDA_VISIBLE_DEVICES": ""}, timeout=400)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESULT ok=True" in r.stdout, r.stdout[-2000:]


def test_dpsgd_and_fedfomo_over_three_ranks_match_single_process():
    r = _torchrun("dist_worker_dpsgd_cpu.py", 3, {"CUDA_VISIBLE_DEVICES": ""}, timeout=400)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESULT ok=True" in r.stdout, r.stdout[-2000:]


def test_local_ditto_fedavg_turboaggregate_subavg_over_two_ranks_match_single_process():
    r = _torchrun("dist_worker_others_cpu.py", 2, {"CUDA_VISIBLE_DEVICES": ""}, timeout=400)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESULT ok=True" in r.stdout, r.stdout[-2000:]


def test_syncbn_over_ranks_matches_full_batch_batchnorm():
    """SyncBN (the reference's only collective call site, batchnorm_utils.py:150-237) as two all-reduces: a batch split unevenly
    over 3 ranks gives the full-batch BatchNorm3d outputs, gradients and running statistics."""
    r = _torchrun("dist_worker_syncbn.py", 3, {"CUDA_VISIBLE_DEVICES": ""}, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESULT ok=True" in r.stdout, r.stdout[-2000:]


@pytest.mark.parametrize("world", [2, 3])
def test_depth_slab_decomposition_matches_the_unsplit_module(world):
    """One client's volumes split along depth over ranks (halo exchange per layer, BatchNorm statistics across ranks): outputs, all
    gradients and running statistics equal the single-process module, also when the last layers leave some ranks without planes."""
    r = _torchrun("dist_worker_spatial.py", world, {"CUDA_VISIBLE_DEVICES": ""}, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESULT ok=True" in r.stdout, r.stdout[-2000:]


@pytest.mark.gpu
def test_fused_symmetric_aggregation_matches_nccl():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    r = _torchrun("dist_worker_sym_agg.py", min(n, 8))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESULT ok=True" in r.stdout, r.stdout[-2000:]
