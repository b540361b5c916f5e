"""The experiment entry points run end to end on CPU (BASELINE.json config 1: FedAvg, tiny 3-D CNN, 4 clients x 8 synthetic 32^3
volumes, 2 rounds), keep the reference's flag names and defaults and write the log / JSONL / checkpoint files."""
import argparse
import json
import os

import pytest
import torch

from neuroimagedisttraining_b200.cli import main_dispfl, main_fedavg, main_others, main_sailentgrads

TINY = ["--model", "tiny3DCNN", "--in_shape", "32,32,32", "--client_num_in_total", "4", "--synthetic_subjects", "40", "--batch_size", "4",
        "--epochs", "1", "--comm_round", "2", "--device", "cpu", "--frac", "1.0"]


def _dirs(tmp_path):
    return ["--log_dir", str(tmp_path / "LOG"), "--results_dir", str(tmp_path / "results")]


def test_flag_defaults_match_the_reference_entry_scripts():
    sg = main_sailentgrads.add_args(argparse.ArgumentParser()).parse_args([])
    assert (sg.model, sg.dataset, sg.partition_method, sg.partition_alpha, sg.batch_size, sg.client_optimizer) == ("3DCNN", "ABCD", "dir", 0.3, 16, "sgd")
    assert (sg.lr, sg.lr_decay, sg.wd, sg.epochs, sg.client_num_in_total, sg.frac, sg.momentum, sg.comm_round) == (0.01, 0.998, 5e-4, 2, 4, 1.0, 0, 200)
    assert (sg.frequency_of_the_test, sg.gpu, sg.ci, sg.dense_ratio, sg.anneal_factor, sg.seed, sg.cs, sg.itersnip_iteration) == (1, 0, 0, 0.5, 0.5, 1024, "v0", 1)
    assert (sg.active, sg.public_portion, sg.erk_power_scale, sg.tag, sg.snip_mask, sg.logfile) == (1.0, 0, 1, "test", True, "logfile")
    for flag in ("stratified_sampling", "dis_gradient_check", "strict_avg", "static", "uniform", "save_masks", "different_initial",
                 "record_mask_diff", "diff_spa", "global_test"):
        assert getattr(sg, flag) is False
    assert main_sailentgrads.add_args(argparse.ArgumentParser()).parse_args(["--snip_mask", "False"]).snip_mask is False
    fa = main_fedavg.add_args(argparse.ArgumentParser()).parse_args([])
    assert (fa.lr, fa.epochs, fa.frac, fa.seed, fa.client_num_in_total, fa.comm_round) == (0.001, 5, 0.9, 0, 4, 200)
    dp = main_dispfl.add_args(argparse.ArgumentParser()).parse_args([])
    assert (dp.lr, dp.epochs, dp.client_num_in_total, dp.frac, dp.comm_round, dp.cs) == (0.001, 5, 21, 0.1, 10, "random")
    dt = main_others.add_args(argparse.ArgumentParser(), "ditto").parse_args([])
    assert (dt.local_epochs, dt.lamda) == (3, 0.5)
    ps = main_others.add_args(argparse.ArgumentParser(), "dpsgd").parse_args([])
    assert (ps.cs, ps.type) == ("ring", "epoch")
    sa = main_others.add_args(argparse.ArgumentParser(), "subavg").parse_args([])
    assert (sa.each_prune_ratio, sa.dense_ratio, sa.dist_thresh, sa.acc_thresh) == (0.05, 0.5, 1e-4, 0.5)


def test_fedavg_cli_config1_with_checkpoint_and_resume(tmp_path):
    ck = str(tmp_path / "ck.pt")
    stat = main_fedavg.main(TINY + _dirs(tmp_path) + ["--checkpoint_every", "1", "--checkpoint_path", ck, "--lr", "0.01"])
    assert len(stat["global_test_acc"]) >= 2 and all(0.0 <= a <= 1.0 for a in stat["global_test_acc"])
    logs = os.listdir(tmp_path / "LOG" / "ABCD")
    assert any(f.endswith(".log") for f in logs) and any(f.endswith(".jsonl") for f in logs)
    lines = [json.loads(l) for l in open(tmp_path / "LOG" / "ABCD" / next(f for f in logs if f.endswith(".jsonl")))]
    evals, line = [l for l in lines if l["event"] == "eval"], lines[-1]
    assert [e["round"] for e in evals] == [0, 1, -1] and [e["global_test_acc"] for e in evals] == [float(a) for a in stat["global_test_acc"]]
    assert line["event"] == "summary" and line["global_test_acc"] == [float(a) for a in stat["global_test_acc"]] and "phase_time_s" in line
    ckpt = torch.load(ck, weights_only=False)
    assert ckpt["round"] == 1 and any(k.startswith("features.0.") for k in ckpt["state_dict"]) and "stat_info" in ckpt
    from neuroimagedisttraining_b200.models import Tiny3DCNN
    Tiny3DCNN(1, in_shape=(32, 32, 32)).load_state_dict(ckpt["state_dict"], strict=True)      # a plain torch state_dict of the module
    # one more round from the checkpoint: starts at round 2, does not repeat rounds 0-1
    stat2 = main_fedavg.main(TINY[:-6] + ["--epochs", "1", "--comm_round", "3", "--device", "cpu", "--frac", "1.0"] + _dirs(tmp_path)
                             + ["--resume", ck, "--lr", "0.01"])
    assert len(stat2["round_time_s"]) - len(stat["round_time_s"]) == 1 or len(stat2["global_test_acc"]) == len(stat["global_test_acc"]) + 1


def test_salientgrads_cli_masks_half_of_the_prunable_weights(tmp_path):
    stat = main_sailentgrads.main(TINY + _dirs(tmp_path) + ["--dense_ratio", "0.5", "--itersnip_iteration", "2"])
    # two rounds + the closing evaluation with round = -1 (sailentgrads_api.py:147)
    assert len(stat["global_test_acc"]) == 3 and len(stat["person_test_acc"]) == 3
    assert stat["sum_comm_params"] > 0


def test_dispfl_cli_runs_with_dropout_of_clients(tmp_path):
    stat = main_dispfl.main(TINY[:-2] + ["--frac", "0.5", "--active", "0.8", "--dense_ratio", "0.5", "--save_masks"] + _dirs(tmp_path))
    assert len(stat["old_mask_test_acc"]) == 2 and len(stat["final_masks"]) == 4 and len(stat["mask_dis_matrix"]) == 4


@pytest.mark.parametrize("algo", ["ditto", "dpsgd", "fedfomo", "local", "subavg", "turboaggregate"])
def test_other_algorithms_cli_on_synthetic_cifar(tmp_path, algo):
    argv = [algo, "--model", "cnn_cifar10", "--dataset", "cifar10", "--client_num_in_total", "3", "--frac", "1.0", "--batch_size", "8",
            "--epochs", "1", "--comm_round", "1", "--device", "cpu", "--lr", "0.05", "--partition_method", "homo", "--synthetic_subjects", "240"] + _dirs(tmp_path)
    stat = main_others.main(argv)
    assert isinstance(stat, dict) and "round_time_s" in stat


def test_reference_job_matrix_parses_with_our_entry_points(tmp_path):
    """Every launcher of the reference (87 scripts, ``scripts/jobs/job_matrix.py``) maps to a command our argument parsers accept."""
    import importlib.util
    import sys
    spec = importlib.util.spec_from_file_location("job_matrix", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                                             "scripts", "jobs", "job_matrix.py"))
    jm = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(jm)
    assert len(jm.JOBS) == 87 and len({r[0] for r in jm.JOBS}) == 87
    assert {r[1] for r in jm.JOBS} == {"salientgrads", "fedavg", "dispfl", "ditto", "dpsgd", "fedfomo", "local", "subavg"}
    for row in jm.JOBS:
        cmd = jm.command(row)
        assert cmd[:2] == [sys.executable, "-m"]
        mod, flags = cmd[2], cmd[3:]
        if mod.endswith("main_others"):
            algo, flags = flags[0], flags[1:]
            ns = main_others.add_args(argparse.ArgumentParser(), algo).parse_args(flags)
        else:
            ns = {"main_sailentgrads": main_sailentgrads, "main_fedavg": main_fedavg, "main_dispfl": main_dispfl}[mod.rsplit(".", 1)[1]] \
                .add_args(argparse.ArgumentParser()).parse_args(flags)
        assert ns.model == "resnet18" and ns.client_num_in_total == 100 and ns.seed == 2022 and ns.dataset in ("cifar10", "cifar100", "tiny")
    row = jm.find("salientgradssparsitywith20iteration90sps")
    f = jm.flags_of(row)
    assert (f["dense_ratio"], f["itersnip_iteration"], f["batch_size"]) == ("0.1", "20", "16")
    multi = jm.command(row, gpus=8)
    assert "torch.distributed.run" in multi and "--nproc-per-node=8" in multi
    assert "frac" not in jm.flags_of(jm.find("DisPFL/dispfljob"))          # that script leaves --frac to the default
    jm.main(["--emit-slurm", str(tmp_path / "slurm"), "--algo", "subavg"])
    assert len(os.listdir(tmp_path / "slurm")) == 15


@pytest.mark.parametrize("mod,extra", [(main_sailentgrads, ["--dense_ratio", "0.5"]), (main_fedavg, []), (main_dispfl, ["--dense_ratio", "0.5"])])
def test_main_entry_points_also_take_image_datasets(tmp_path, mod, extra):
    """The reference's job scripts run SalientGrads / FedAvg / DisPFL on CIFAR (its SalientGrads entry point then forces ABCD,
    ``main_sailentgrads.py:272``; here the dataset flag is honoured)."""
    stat = mod.main(["--model", "cnn_cifar10", "--dataset", "cifar10", "--client_num_in_total", "4", "--frac", "0.5", "--batch_size", "16",
                     "--epochs", "1", "--comm_round", "1", "--device", "cpu", "--lr", "0.05", "--synthetic_subjects", "160"] + extra + _dirs(tmp_path))
    assert "round_time_s" in stat and len(stat["round_time_s"]) >= 1
