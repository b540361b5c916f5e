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


@pytest.mark.gpu
def test_dpsgd_neighbour_averaging_over_the_peer_arena_matches_broadcasts():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    r = _torchrun("dist_worker_dpsgd_gpu.py", min(n, 3), timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESULT ok=True" in r.stdout, r.stdout[-2000:]


@pytest.mark.gpu
def test_native_data_parallel_engine_over_nccl_matches_whole_batches():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    r = _torchrun("dist_worker_dp_gpu.py", min(n, 3), timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESULT ok=True" in r.stdout, r.stdout[-2000:]


def test_spatial_parallel_cli_run_equals_the_single_process_run(tmp_path):
    """``--spatial_parallel`` under torchrun (2 gloo ranks): both ranks train every client together on depth slabs; the checkpointed
    global model and the accuracy series equal the plain one-process run; only the real rank 0 writes files."""
    import json
    common = ["--model", "tiny3DCNN", "--in_shape", "32,32,32", "--client_num_in_total", "3", "--synthetic_subjects", "30", "--batch_size", "4",
              "--epochs", "1", "--comm_round", "2", "--device", "cpu", "--frac", "1.0", "--lr", "0.02", "--checkpoint_every", "1"]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    outs = {}
    for tag, launcher in (("one", [sys.executable, "-m"]),
                          ("two", [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                                   "--master-port", str(29800 + os.getpid() % 100), "-m"])):
        d = tmp_path / tag
        cmd = launcher + ["neuroimagedisttraining_b200.cli.main_sailentgrads"] + common + ["--log_dir", str(d / "LOG"), "--checkpoint_path", str(d / "ck.pt")]
        if tag == "two":
            cmd.append("--spatial_parallel")
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=400, env=env, cwd=ROOT)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
        files = os.listdir(d / "LOG" / "ABCD")
        assert sum(f.endswith(".jsonl") for f in files) == 1
        summary = [json.loads(l) for l in open(d / "LOG" / "ABCD" / next(f for f in files if f.endswith(".jsonl")))][-1]
        outs[tag] = (torch.load(d / "ck.pt", weights_only=False), summary)
    (ck1, s1), (ck2, s2) = outs["one"], outs["two"]
    assert ck1["round"] == ck2["round"] == 1
    for k, v in ck1["state_dict"].items():
        torch.testing.assert_close(ck2["state_dict"][k].float(), v.float(), rtol=1e-4, atol=1e-5, msg=lambda m: f"{k}: {m}")
    assert all(torch.equal(ck1["mask"][k], ck2["mask"][k]) for k in ck1["mask"])
    assert s1["global_test_acc"] == s2["global_test_acc"] and s1["person_test_acc"] == s2["person_test_acc"]


def test_data_parallel_cli_run_equals_the_single_process_run(tmp_path):
    """``--data_parallel`` under torchrun (2 and 3 gloo ranks, PyTorch-op engine): every rank takes its share of each mini-batch
    (batch 4 on 3 ranks = 2 + 1 + 1, the last batches of a client are smaller still), BatchNorm statistics and gradients are summed
    over the ranks; global model, mask and accuracies equal the plain one-process run; only the real rank 0 writes files."""
    import json
    common = ["--model", "tiny3DCNN", "--in_shape", "32,32,32", "--client_num_in_total", "3", "--synthetic_subjects", "30", "--batch_size", "4",
              "--epochs", "1", "--comm_round", "2", "--device", "cpu", "--frac", "1.0", "--lr", "0.02", "--checkpoint_every", "1"]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    outs = {}
    for tag, n in (("one", 1), ("two", 2), ("three", 3)):
        launcher = [sys.executable, "-m"] if n == 1 else [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                                                          "--master-addr", "127.0.0.1", "--master-port", str(29900 + os.getpid() % 90 + n), "-m"]
        d = tmp_path / tag
        cmd = launcher + ["neuroimagedisttraining_b200.cli.main_sailentgrads"] + common + ["--log_dir", str(d / "LOG"), "--checkpoint_path", str(d / "ck.pt")]
        if n > 1:
            cmd.append("--data_parallel")
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=400, env=env, cwd=ROOT)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
        files = os.listdir(d / "LOG" / "ABCD")
        assert sum(f.endswith(".jsonl") for f in files) == 1
        summary = [json.loads(l) for l in open(d / "LOG" / "ABCD" / next(f for f in files if f.endswith(".jsonl")))][-1]
        outs[tag] = (torch.load(d / "ck.pt", weights_only=False), summary)
    ck1, s1 = outs["one"]
    for tag in ("two", "three"):
        ck2, s2 = outs[tag]
        assert ck1["round"] == ck2["round"] == 1
        for k, v in ck1["state_dict"].items():
            torch.testing.assert_close(ck2["state_dict"][k].float(), v.float(), rtol=2e-4, atol=2e-5, msg=lambda m: f"{tag} {k}: {m}")
        assert all(torch.equal(ck1["mask"][k], ck2["mask"][k]) for k in ck1["mask"])
        assert s1["global_test_acc"] == s2["global_test_acc"] and s1["person_test_acc"] == s2["person_test_acc"]

