#!/usr/bin/env python
"""The reference's experiment matrix as data: one row per launcher under ``fedml_experiments/standalone/*/`` (87 SLURM / bash
scripts with an active command; all CIFAR-10/100/Tiny-ImageNet, resnet18, 100 clients, seed 2022).  A row is (reference script,
algorithm, flags that differ from ``COMMON``); ``"_unset"`` lists common flags the script leaves to the entry point's default.

    python scripts/jobs/job_matrix.py --list [--algo ditto] [--dataset cifar100]
    python scripts/jobs/job_matrix.py --show sailentgrads/Jobs/salientgradssparsitywith20iteration90sps
    python scripts/jobs/job_matrix.py --run  fedavg/cifar10 -- --data_dir /data/cifar10 --comm_round 2
    python scripts/jobs/job_matrix.py --emit-slurm out_dir [--gpus 8] [--header my_header.txt]

Commands are this framework's entry points (``python -m neuroimagedisttraining_b200.cli...``; ``torchrun`` when ``--gpus`` > 1 for the
client-parallel algorithms).  Flags after ``--`` are appended to the command.
"""
from __future__ import annotations

import argparse
import os
import shlex
import subprocess
import sys

COMMON = {"model": "resnet18", "client_num_in_total": "100", "lr": "0.1", "lr_decay": "0.998", "seed": "2022", "comm_round": "500",
          "frac": "0.1", "epochs": "5", "partition_method": "dir", "partition_alpha": "0.3"}

ENTRY = {"salientgrads": "neuroimagedisttraining_b200.cli.main_sailentgrads", "fedavg": "neuroimagedisttraining_b200.cli.main_fedavg",
         "dispfl": "neuroimagedisttraining_b200.cli.main_dispfl"}
OTHERS = "neuroimagedisttraining_b200.cli.main_others"          # ditto | dpsgd | fedfomo | local | subavg
MULTI_RANK = {"salientgrads", "fedavg"}                          # algorithms whose clients are spread over ranks

JOBS = [
    ("DisPFL/CIFAR100dispflsparsitywithoutiteration70sps", "dispfl", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.3"}),
    ("DisPFL/CIFAR100dispflsparsitywithoutiteration80sps", "dispfl", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.2"}),
    ("DisPFL/CIFAR100dispflsparsitywithoutiteration90sps", "dispfl", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.1"}),
    ("DisPFL/CIFAR100dispflsparsitywithoutiteration95sps", "dispfl", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.05"}),
    ("DisPFL/cifar10", "dispfl", {"dataset": "cifar10", "batch_size": "128", "dense_ratio": "0.5", "anneal_factor": "0.5", "cs": "random", "dis_gradient_check": True, "different_initial": True}),
    ("DisPFL/cifar100", "dispfl", {"dataset": "cifar100", "batch_size": "128", "dense_ratio": "0.5", "anneal_factor": "0.5", "cs": "random", "dis_gradient_check": True, "different_initial": True}),
    ("DisPFL/dispfl100job", "dispfl", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.5", "anneal_factor": "0.5", "cs": "random", "dis_gradient_check": True, "different_initial": True}),
    ("DisPFL/dispfljob", "dispfl", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.5", "anneal_factor": "0.5", "cs": "random", "dis_gradient_check": True, "different_initial": True, "_unset": ["frac"]}),
    ("DisPFL/dispflsparsitywithoutiteration70sps", "dispfl", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.3"}),
    ("DisPFL/dispflsparsitywithoutiteration80sps", "dispfl", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.2"}),
    ("DisPFL/dispflsparsitywithoutiteration90sps", "dispfl", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.1"}),
    ("DisPFL/dispflsparsitywithoutiteration95sps", "dispfl", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.05"}),
    ("DisPFL/tiny", "dispfl", {"dataset": "tiny", "batch_size": "128", "dense_ratio": "0.5", "anneal_factor": "0.5", "cs": "random", "dis_gradient_check": True, "different_initial": True}),
    ("ditto/CIFAR100dittosparsitywithoutiteration70sps", "salientgrads", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.3"}),
    ("ditto/CIFAR100salientgradssparsitywithoutiteration80sps", "salientgrads", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.2"}),
    ("ditto/CIFAR100salientgradssparsitywithoutiteration90sps", "salientgrads", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.1"}),
    ("ditto/CIFAR100salientgradssparsitywithoutiteration95sps", "salientgrads", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.05"}),
    ("ditto/cifar10", "ditto", {"dataset": "cifar10", "batch_size": "16", "epochs": "2", "lamda": "0.5", "local_epochs": "3"}),
    ("ditto/cifar100", "ditto", {"dataset": "cifar100", "batch_size": "128", "lamda": "0.5", "epochs": "2", "local_epochs": "3"}),
    ("ditto/dittojob", "ditto", {"dataset": "cifar100", "batch_size": "128", "epochs": "2", "lamda": "0.5", "local_epochs": "3"}),
    ("ditto/dittojob100", "ditto", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "epochs": "2", "lamda": "0.5", "local_epochs": "3"}),
    ("ditto/dittopart", "ditto", {"dataset": "cifar10", "partition_method": "my_part", "batch_size": "16", "epochs": "2", "lamda": "0.5", "local_epochs": "3"}),
    ("ditto/dittopart100", "ditto", {"dataset": "cifar100", "partition_method": "my_part", "partition_alpha": "0.2", "batch_size": "16", "epochs": "2", "lamda": "0.5", "local_epochs": "3"}),
    ("ditto/dittosparsitywithoutiteration70sps", "ditto", {"dataset": "cifar10", "batch_size": "16", "lamda": "0.3"}),
    ("ditto/dittosparsitywithoutiteration80sps", "ditto", {"dataset": "cifar10", "batch_size": "16", "lamda": "0.2"}),
    ("ditto/dittosparsitywithoutiteration90sps", "ditto", {"dataset": "cifar10", "batch_size": "16", "lamda": "0.1"}),
    ("ditto/dittosparsitywithoutiteration95sps", "ditto", {"dataset": "cifar10", "batch_size": "16", "lamda": "0.05"}),
    ("ditto/tiny", "ditto", {"dataset": "tiny", "batch_size": "128", "comm_round": "300", "epochs": "3", "lamda": "0.5", "local_epochs": "2"}),
    ("dpsgd/cifar10", "dpsgd", {"dataset": "cifar10", "batch_size": "128", "cs": "random", "type": "epoch"}),
    ("dpsgd/cifar100", "dpsgd", {"dataset": "cifar100", "batch_size": "128", "cs": "random", "type": "epoch"}),
    ("dpsgd/dpsgdjob", "dpsgd", {"dataset": "cifar10", "batch_size": "16", "cs": "random", "type": "epoch"}),
    ("dpsgd/dpsgdjob100", "dpsgd", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "cs": "random", "type": "epoch"}),
    ("dpsgd/dpsgdpart", "dpsgd", {"dataset": "cifar10", "partition_method": "my_part", "batch_size": "16", "cs": "random", "type": "epoch"}),
    ("dpsgd/dpsgdpart100", "dpsgd", {"dataset": "cifar10", "partition_method": "my_part", "partition_alpha": "0.2", "batch_size": "16", "cs": "random", "type": "epoch"}),
    ("dpsgd/tiny", "dpsgd", {"dataset": "tiny", "batch_size": "128", "cs": "random", "type": "epoch"}),
    ("fedavg/cifar10", "fedavg", {"dataset": "cifar10", "batch_size": "128"}),
    ("fedavg/cifar100", "fedavg", {"dataset": "cifar100", "batch_size": "128"}),
    ("fedavg/fedavgjobs", "fedavg", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16"}),
    ("fedavg/fedavgpart", "fedavg", {"dataset": "cifar10", "partition_method": "my_part", "batch_size": "16"}),
    ("fedavg/fedavgpart100", "fedavg", {"dataset": "cifar100", "partition_method": "my_part", "partition_alpha": "0.2", "batch_size": "16"}),
    ("fedavg/tiny", "fedavg", {"dataset": "tiny", "batch_size": "128"}),
    ("fedfomo/cifar10", "fedfomo", {"dataset": "cifar10", "batch_size": "128"}),
    ("fedfomo/cifar100", "fedfomo", {"dataset": "cifar100", "batch_size": "16"}),
    ("fedfomo/fedfomojobs", "fedfomo", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16"}),
    ("fedfomo/tiny", "fedfomo", {"dataset": "tiny", "batch_size": "128"}),
    ("local/cifar10", "local", {"dataset": "cifar10", "batch_size": "128", "frac": "1.0"}),
    ("local/cifar100", "local", {"dataset": "cifar100", "batch_size": "128", "frac": "1.0"}),
    ("local/tiny", "local", {"dataset": "tiny", "batch_size": "128", "frac": "1.0", "comm_round": "300"}),
    ("sailentgrads/Jobs/CIFAR100salientgradssparsitywithoutiteration70sps", "salientgrads", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.3"}),
    ("sailentgrads/Jobs/CIFAR100salientgradssparsitywithoutiteration80sps", "salientgrads", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.2"}),
    ("sailentgrads/Jobs/CIFAR100salientgradssparsitywithoutiteration90sps", "salientgrads", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.1"}),
    ("sailentgrads/Jobs/CIFAR100salientgradssparsitywithoutiteration95sps", "salientgrads", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.05"}),
    ("sailentgrads/Jobs/cifar10", "salientgrads", {"dataset": "cifar10", "batch_size": "128", "dense_ratio": "0.5"}),
    ("sailentgrads/Jobs/cifar100", "fedavg", {"dataset": "cifar100", "batch_size": "128"}),
    ("sailentgrads/Jobs/sailentgradsjob", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.1"}),
    ("sailentgrads/Jobs/salientgradscifar100", "salientgrads", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.5"}),
    ("sailentgrads/Jobs/salientgradspart", "salientgrads", {"dataset": "cifar10", "partition_method": "my_part", "batch_size": "16", "dense_ratio": "0.5"}),
    ("sailentgrads/Jobs/salientgradspart100", "salientgrads", {"dataset": "cifar100", "partition_method": "my_part", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.5"}),
    ("sailentgrads/Jobs/salientgradssparsitywith100iteration70sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.3", "itersnip_iteration": "100"}),
    ("sailentgrads/Jobs/salientgradssparsitywith100iteration80sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.2", "itersnip_iteration": "100"}),
    ("sailentgrads/Jobs/salientgradssparsitywith100iteration90sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.1", "itersnip_iteration": "100"}),
    ("sailentgrads/Jobs/salientgradssparsitywith20iteration70sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.3", "itersnip_iteration": "20"}),
    ("sailentgrads/Jobs/salientgradssparsitywith20iteration80sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.2", "itersnip_iteration": "20"}),
    ("sailentgrads/Jobs/salientgradssparsitywith20iteration90sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.1", "itersnip_iteration": "20"}),
    ("sailentgrads/Jobs/salientgradssparsitywith50iteration70sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.3", "itersnip_iteration": "50"}),
    ("sailentgrads/Jobs/salientgradssparsitywith50iteration80sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.2", "itersnip_iteration": "50"}),
    ("sailentgrads/Jobs/salientgradssparsitywith50iteration90sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.1", "itersnip_iteration": "50"}),
    ("sailentgrads/Jobs/salientgradssparsitywithoutiteration70sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.3"}),
    ("sailentgrads/Jobs/salientgradssparsitywithoutiteration80sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.2"}),
    ("sailentgrads/Jobs/salientgradssparsitywithoutiteration90sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.1"}),
    ("sailentgrads/Jobs/salientgradssparsitywithoutiteration95sps", "salientgrads", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.05"}),
    ("sailentgrads/Jobs/tiny", "fedavg", {"dataset": "tiny", "batch_size": "128"}),
    ("subavg/CIFAR100subavgsparsitywithoutiteration70sps", "subavg", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.3"}),
    ("subavg/CIFAR100subavgsparsitywithoutiteration95sps", "subavg", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.05"}),
    ("subavg/CIFAR100subavgssparsitywithoutiteration80sps", "subavg", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.2"}),
    ("subavg/CIFAR100subavgssparsitywithoutiteration90sps", "subavg", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.1"}),
    ("subavg/cifar10", "subavg", {"dataset": "cifar10", "batch_size": "128", "dense_ratio": "0.5"}),
    ("subavg/cifar100", "subavg", {"dataset": "cifar100", "batch_size": "128", "dense_ratio": "0.5"}),
    ("subavg/subavgdjob", "subavg", {"dataset": "cifar100", "batch_size": "16", "dense_ratio": "0.5"}),
    ("subavg/subavgjob100", "subavg", {"dataset": "cifar100", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.5"}),
    ("subavg/subavgpart", "subavg", {"dataset": "cifar10", "partition_method": "my_part", "batch_size": "16", "dense_ratio": "0.5"}),
    ("subavg/subavgpart100", "subavg", {"dataset": "cifar100", "partition_method": "my_part", "partition_alpha": "0.2", "batch_size": "16", "dense_ratio": "0.5"}),
    ("subavg/subavgsparsitywithoutiteration70sps", "subavg", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.3"}),
    ("subavg/subavgsparsitywithoutiteration80sps", "subavg", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.2"}),
    ("subavg/subavgsparsitywithoutiteration90sps", "subavg", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.1"}),
    ("subavg/subavgsparsitywithoutiteration95sps", "subavg", {"dataset": "cifar10", "batch_size": "16", "dense_ratio": "0.05"}),
    ("subavg/tiny", "subavg", {"dataset": "tiny", "batch_size": "128", "dense_ratio": "0.5"}),]

DEFAULT_HEADER = """#!/bin/bash
#SBATCH -N 1
#SBATCH -n {gpus}
#SBATCH -c 8
#SBATCH --gres=gpu:{gpus}
#SBATCH -t 1-00
#SBATCH -J {job}
#SBATCH -e error%A.err
#SBATCH -o out%A.out
export OMP_NUM_THREADS=1
"""


def flags_of(row) -> dict:
    _, _, over = row
    f = {k: v for k, v in COMMON.items() if k not in over.get("_unset", ())}
    f.update({k: v for k, v in over.items() if k != "_unset"})
    return f


def command(row, gpus: int = 1, extra=()) -> list:
    name, algo, _ = row
    mod = ENTRY.get(algo, OTHERS)
    if gpus > 1 and algo in MULTI_RANK:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={gpus}", "--master-addr", "127.0.0.1", "-m", mod]
    else:
        cmd = [sys.executable, "-m", mod]
    if mod == OTHERS:
        cmd.append(algo)
    for k, v in flags_of(row).items():
        cmd += [f"--{k}"] if v is True else [f"--{k}", str(v)]
    return cmd + list(extra)


def find(name: str):
    hits = [r for r in JOBS if r[0] == name] or [r for r in JOBS if r[0].endswith("/" + name) or r[0].split("/")[-1] == name]
    if len(hits) != 1:
        raise SystemExit(f"{name!r} matches {len(hits)} jobs; use --list")
    return hits[0]


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    extra = []
    if "--" in argv:
        i = argv.index("--")
        argv, extra = argv[:i], argv[i + 1:]
    p = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    g = p.add_mutually_exclusive_group(required=True)
    g.add_argument("--list", action="store_true")
    g.add_argument("--show")
    g.add_argument("--run")
    g.add_argument("--emit-slurm", metavar="DIR")
    p.add_argument("--algo")
    p.add_argument("--dataset")
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--header", help="file with a site-specific SLURM header ({gpus}, {job} are substituted)")
    a = p.parse_args(argv)
    rows = [r for r in JOBS if (a.algo is None or r[1] == a.algo) and (a.dataset is None or flags_of(r).get("dataset") == a.dataset)]
    if a.list:
        for r in rows:
            f = flags_of(r)
            print(f"{r[0]:75s} {r[1]:12s} {f.get('dataset', ''):9s} {f['partition_method']}{f['partition_alpha']:>5s} bs{f.get('batch_size', '-'):>4s}"
                  + (f" dense {f['dense_ratio']}" if "dense_ratio" in f else "") + (f" itersnip {f['itersnip_iteration']}" if "itersnip_iteration" in f else ""))
        print(f"{len(rows)} jobs")
        return rows
    if a.show:
        print(" ".join(shlex.quote(c) for c in command(find(a.show), a.gpus, extra)))
        return None
    if a.run:
        return subprocess.call(command(find(a.run), a.gpus, extra))
    header = open(a.header).read() if a.header else DEFAULT_HEADER
    os.makedirs(a.emit_slurm, exist_ok=True)
    for r in rows:
        path = os.path.join(a.emit_slurm, r[0].replace("/", "__") + ".sh")
        with open(path, "w") as f:
            f.write(header.format(gpus=a.gpus, job=r[0].split("/")[-1]))
            f.write(" ".join(shlex.quote(c) for c in command(r, a.gpus, extra)).replace(shlex.quote(sys.executable), "python") + "\n")
        os.chmod(path, 0o755)
    print(f"wrote {len(rows)} scripts to {a.emit_slurm}")
    return None


if __name__ == "__main__":
    sys.exit(main() and 0)
