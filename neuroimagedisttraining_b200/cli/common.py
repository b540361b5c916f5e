"""Shared pieces of the experiment entry points (reference: ``fedml_experiments/standalone/<algo>/main_<algo>.py``):
argument groups with the reference's flag names and defaults, the ``identity`` string that names
logs / results, seeding, data + model + trainer construction, and the B200 extras
(``--engine``, ``--agg``, ``--resume``, ``--checkpoint_every``, ``--synthetic_subjects`` ...)."""
from __future__ import annotations

import argparse
import os
import random
from typing import Optional

import numpy as np
import torch

from ..data import (SyntheticABCD, HDF5VolumeStore, load_partition_data_abcd, load_partition_data_abcd_rescale,
                    load_partition_data_images)
from ..fl import SlabTrainer
from ..models import create_model
from ..parallel import init_distributed
from ..utils.logger import logger_config, JsonlMetrics


def str2bool(v) -> bool:
    """The reference declares ``--snip_mask type=bool`` (any non-empty string is True, ``main_sailentgrads.py:125``);
    this accepts the usual spellings instead."""
    if isinstance(v, bool):
        return v
    return str(v).lower() in ("1", "true", "t", "yes", "y")


def add_common_args(p: argparse.ArgumentParser, *, model="3DCNN", dataset="ABCD", lr=0.01, epochs=2, frac=1.0, clients=4,
                    comm_round=200, seed=1024, batch_size=16) -> argparse.ArgumentParser:
    p.add_argument("--model", type=str, default=model, metavar="N", help="network architecture: 3DCNN, 3DResNet, tiny3DCNN, resnet18, vgg11, cnn_cifar10, ...")
    p.add_argument("--dataset", type=str, default=dataset, metavar="N", help="ABCD | cifar10 | cifar100 | tiny")
    p.add_argument("--data_dir", type=str, default=None, help="HDF5 file (ABCD) or dataset directory; omitted -> synthetic data")
    p.add_argument("--partition_method", type=str, default="dir", metavar="N", help="site (ABCD) | dir | n_cls | my_part | homo | hetero")
    p.add_argument("--partition_alpha", type=float, default=0.3, metavar="PA")
    p.add_argument("--batch_size", type=int, default=batch_size, metavar="N")
    p.add_argument("--client_optimizer", type=str, default="sgd")
    p.add_argument("--lr", type=float, default=lr, metavar="LR")
    p.add_argument("--lr_decay", type=float, default=0.998, metavar="LR_decay")
    p.add_argument("--wd", type=float, default=5e-4)
    p.add_argument("--epochs", type=int, default=epochs, metavar="EP")
    p.add_argument("--client_num_in_total", type=int, default=clients, metavar="NN")
    p.add_argument("--frac", type=float, default=frac, metavar="NN")
    p.add_argument("--momentum", type=float, default=0, metavar="NN")
    p.add_argument("--comm_round", type=int, default=comm_round)
    p.add_argument("--frequency_of_the_test", type=int, default=1)
    p.add_argument("--gpu", type=int, default=0)
    p.add_argument("--ci", type=int, default=0)
    p.add_argument("--seed", type=int, default=seed)
    p.add_argument("--tag", type=str, default="test")
    p.add_argument("--logfile", type=str, default="logfile")
    # ---- additions of this framework
    g = p.add_argument_group("b200")
    g.add_argument("--engine", default="auto", choices=["auto", "b200", "torch"], help="compute engine (auto: sm_100a kernels when the model has a plan)")
    g.add_argument("--agg", default="auto", choices=["auto", "fused", "nccl"], help="cross-rank aggregation: fused symmetric-memory kernel or NCCL baseline")
    g.add_argument("--staging", default="auto", choices=["auto", "client", "stream", "resident"],
                   help="how volumes reach the GPU: client = each loader's volumes uploaded once per use, batches gathered on the device "
                        "(default on CUDA); stream = batch-by-batch copies; resident = whole store kept on the device")
    g.add_argument("--client_lanes", type=int, default=3, help="native engines: clients of a rank trained concurrently, each on its own "
                                                               "engine and CUDA stream (1 = strictly one after another, like the reference)")
    g.add_argument("--placement", default="lpt", choices=["lpt", "roundrobin"], help="client -> rank placement: by training-set size or client_idx %% world")
    g.add_argument("--balance_eval", type=int, default=0, help="1: spread the round's evaluation jobs over ranks by load (personal models are "
                                                               "copied to the evaluating rank)")
    g.add_argument("--device", default=None, help="cpu | cuda (default: cuda if available)")
    g.add_argument("--resume", default=None, help="checkpoint to resume from")
    g.add_argument("--checkpoint_path", default=None)
    g.add_argument("--checkpoint_every", type=int, default=0, help="save every k rounds (0 = never)")
    g.add_argument("--log_dir", default="LOG")
    g.add_argument("--results_dir", default="results")
    g.add_argument("--synthetic_subjects", type=int, default=0, help="without --data_dir: number of synthetic subjects / training images (default 20 per client for ABCD, 2000 images)")
    g.add_argument("--in_shape", default="121,145,121", help="volume shape D,H,W")
    g.add_argument("--augment", action="store_true", help="image datasets: random crop (padding 4) + horizontal flip on training batches, as the "
                                                          "reference's transforms; on by default when --data_dir points at real images")
    g.add_argument("--no_augment", action="store_true", help="image datasets: no training augmentation even on real data")
    g.add_argument("--cutout", type=int, default=0, help="with --augment: side of the Cutout square (the reference keeps Cutout(16) commented out)")
    g.add_argument("--labels_file", default=None, help="ABCD: HDF5 file with y / site when they are not in --data_dir (the reference keeps "
                                                       "volumes in alldatain8bitsnormalized.h5 and labels in dataset_all_labels_site.h5)")
    g.add_argument("--defense_type", default="none", choices=["none", "norm_diff_clipping", "weak_dp"],
                   help="FedAvg / SalientGrads: robust-aggregation defence applied to every client upload")
    g.add_argument("--norm_bound", type=float, default=5.0)
    g.add_argument("--stddev", type=float, default=0.025)
    g.add_argument("--spatial_parallel", action="store_true",
                   help="under torchrun: all ranks train every client together, each holding a depth slab of the activations "
                        "(for volumes / batches that do not fit one device); PyTorch engine")
    g.add_argument("--preload", action="store_true",
                   help="ABCD from --data_dir: read the whole HDF5 file into (pinned) host memory once and serve batches from there "
                        "(row copies, zero-run-skipping uploads, device caches) instead of re-reading the file per batch")
    g.add_argument("--data_parallel", action="store_true",
                   help="under torchrun: all ranks train every client together, each taking 1/world of every mini-batch (BatchNorm "
                        "statistics and gradients summed over the ranks): the sm_100a engine on GPUs, PyTorch ops with --device cpu / "
                        "--engine torch; for few, large clients")
    g.add_argument("--rescale", action="store_true", help="ABCD: cut the site-ordered data into client_num_in_total equal shards")
    g.add_argument("--correct_eval_loss", action="store_true", help="evaluate BCEWithLogits(logit) instead of the reference's BCEWithLogits(sigmoid(logit))")
    return p


def add_sparse_args(p: argparse.ArgumentParser, cs="v0") -> argparse.ArgumentParser:
    p.add_argument("--dense_ratio", type=float, default=0.5, help="fraction of weights KEPT (90%% sparsity = 0.1)")
    p.add_argument("--anneal_factor", type=float, default=0.5)
    p.add_argument("--cs", type=str, default=cs)
    p.add_argument("--active", type=float, default=1.0)
    p.add_argument("--public_portion", type=float, default=0)
    p.add_argument("--erk_power_scale", type=float, default=1)
    p.add_argument("--dis_gradient_check", action="store_true")
    p.add_argument("--strict_avg", action="store_true")
    p.add_argument("--static", action="store_true")
    p.add_argument("--uniform", action="store_true")
    p.add_argument("--save_masks", action="store_true")
    p.add_argument("--different_initial", action="store_true")
    p.add_argument("--record_mask_diff", action="store_true")
    p.add_argument("--diff_spa", action="store_true")
    p.add_argument("--global_test", action="store_true")
    return p


def seed_everything(seed: int) -> None:
    """np seed fixes the partition, torch seed the initial weights (``main_sailentgrads.py:264-268``)."""
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


def build_identity(args, prefix: str, extra: Optional[list] = None) -> str:
    part = args.partition_method if args.partition_method == "homo" else args.partition_method + str(args.partition_alpha)
    ident = f"{prefix}-{args.dataset}-{part}-mdl{args.model}"
    for e in (extra or []):
        ident += e
    ident += f"-cm{args.comm_round}-total_clnt{args.client_num_in_total}-neighbor{args.client_num_per_round}"
    if hasattr(args, "dense_ratio"):
        ident += f"-dr{args.dense_ratio}"
    if hasattr(args, "active"):
        ident += f"-active{args.active}"
    ident += f"-seed{args.seed}-lr{args.lr}-batchsize{args.batch_size}"
    return ident


def setup(args, prefix: str, extra_identity: Optional[list] = None, loss_kind: Optional[str] = None, val_split: float = 0.0):
    """-> (ctx, device, logger, dataset 8-tuple, trainer).  Everything an entry point needs before building its API."""
    # the reference computes int(total * frac) (main_sailentgrads.py:234), which is 0 for e.g. 4 clients at frac 0.1 and then
    # fails in the first aggregation; at least one client is sampled here
    args.client_num_per_round = max(1, int(args.client_num_in_total * args.frac))
    args.identity = build_identity(args, prefix, extra_identity)
    ctx = init_distributed(device=args.device)
    spatial = bool(getattr(args, "spatial_parallel", False)) and ctx.is_distributed
    data_par = bool(getattr(args, "data_parallel", False)) and ctx.is_distributed
    if spatial and data_par:
        raise SystemExit("--spatial_parallel and --data_parallel are alternatives")
    dp_group = None
    dp_native = data_par and ctx.device.type == "cuda" and getattr(args, "engine", "auto") in ("auto", "b200") and args.dataset == "ABCD"
    if dp_native:
        from ..ops.engine_dp import DistGroup
        dp_group = DistGroup(ctx.group)
    if spatial or data_par:
        from ..parallel import ReplicatedContext
        ctx = ReplicatedContext.of(ctx)
    device = ctx.device
    real_rank = getattr(ctx, "real_rank", ctx.rank)
    log_path = os.path.join(args.log_dir, args.dataset, args.identity + (f".rank{real_rank}" if real_rank else "") + ".log")
    logger = logger_config(log_path=log_path, logging_name=args.identity, console=ctx.is_main)
    logger.info(args)
    logger.info(device)
    seed_everything(args.seed)
    in_shape = tuple(int(s) for s in args.in_shape.split(","))
    store = None
    if args.dataset == "ABCD":
        if args.data_dir:
            store = HDF5VolumeStore(args.data_dir, args.labels_file, preload=bool(getattr(args, "preload", False)))
        else:
            n = args.synthetic_subjects or 20 * args.client_num_in_total
            n_sites = args.client_num_in_total if not args.rescale else min(21, args.client_num_in_total)
            store = SyntheticABCD(n, shape=in_shape, n_sites=n_sites, seed=args.seed, pin=device.type == "cuda")
        loader = load_partition_data_abcd_rescale if args.rescale else load_partition_data_abcd
        dataset = loader(store, args.partition_method, args.partition_alpha, args.client_num_in_total, args.batch_size, logger, seed=args.seed)
        class_num, loss_kind = 1, loss_kind or "bce"
    else:
        dataset = load_partition_data_images(args.dataset, args.data_dir, args.partition_method if args.partition_method != "site" else "dir",
                                             args.partition_alpha, args.client_num_in_total, args.batch_size, logger, seed=args.seed, val_split=val_split, augment=(args.augment or bool(args.data_dir)) and not args.no_augment, cutout=args.cutout,
                                             **({"synthetic_size": (args.synthetic_subjects, max(args.synthetic_subjects // 4, args.client_num_in_total))}
                                                if args.synthetic_subjects else {}))
        class_num, loss_kind = {"cifar10": 10, "cifar100": 100, "tiny": 200}[args.dataset], loss_kind or "ce"
    model = create_model(args.model, class_num, in_shape, args.dataset)
    logger.info(model)
    engine = None
    if spatial:
        from ..fl.engine import SpatialTorchEngine
        engine = SpatialTorchEngine(model, device, loss_kind=loss_kind, volumetric=store is not None, input_scale=getattr(store, "scale", 1.0))
    elif dp_native:
        from ..ops.engine_dp import DataParallelB200Engine
        engine = DataParallelB200Engine(model, device, dp_group, max_batch=args.batch_size, input_scale=store.scale, seed=args.seed)
    elif data_par:      # CPUs / --engine torch / image datasets: the same split on PyTorch ops (SyncBN + gradient all-reduce)
        from ..fl.engine import DataParallelTorchEngine
        engine = DataParallelTorchEngine(model, device, ctx.group, loss_kind=loss_kind, volumetric=store is not None,
                                         input_scale=getattr(store, "scale", 1.0))
    elif device.type == "cuda" and args.engine in ("auto", "b200") and store is not None:
        try:
            from ..ops.engines import make_native_engine
            engine = make_native_engine(model, device, max_batch=args.batch_size, input_scale=store.scale, seed=args.seed)
        except ImportError as e:
            if args.engine == "b200":
                raise
            logger.info(f"native engine unavailable for {args.model}: {e}; using the PyTorch engine for forward/backward")
    trainer = SlabTrainer(model, args, logger, device=device, store=store, engine=engine, loss_kind=loss_kind)
    trainer.double_sigmoid_eval = not args.correct_eval_loss
    if engine is not None and getattr(engine, "native", False) and not spatial and type(engine).__name__ == "B200Engine":
        # concurrent client slots: FedAvg / SalientGrads train (and evaluate) this many clients of a rank at once
        from ..ops.engines import make_native_engine
        for _ in range(max(0, int(args.client_lanes) - 1)):
            trainer.add_lane(make_native_engine(model, device, max_batch=args.batch_size, input_scale=store.scale, seed=args.seed))
    args.metrics_path = os.path.join(args.log_dir, args.dataset, args.identity + ".jsonl")
    return ctx, device, logger, dataset, trainer


def attach_metrics(api, args, ctx):
    """One JSON line per evaluation while the run is going (``event: eval``), plus the summary line :func:`finish` appends."""
    if ctx.is_main:
        api._jsonl = JsonlMetrics(args.metrics_path)
        api.metrics_sink = api._jsonl.write
    return api


def finish(api, args, logger, ctx) -> dict:
    stat = api.finalize_stats() if "phase_time_s" not in api.stat_info else api.stat_info
    if ctx.is_main:
        sink = getattr(api, "_jsonl", None) or JsonlMetrics(os.path.join(args.log_dir, args.dataset, args.identity + ".jsonl"))
        sink.write(
            event="summary", identity=args.identity, global_test_acc=[float(x) for x in stat.get("global_test_acc", [])],
            person_test_acc=[float(x) for x in stat.get("person_test_acc", [])], sum_comm_params=int(stat.get("sum_comm_params", 0)),
            sum_training_flops=float(stat.get("sum_training_flops", 0)), phase_time_s=stat.get("phase_time_s"))
        logger.info({"sum_comm_params": stat.get("sum_comm_params"), "sum_training_flops": stat.get("sum_training_flops"),
                     "phase_time_s": stat.get("phase_time_s")})
    return stat
