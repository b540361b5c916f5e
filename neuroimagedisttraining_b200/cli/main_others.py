"""Entry points of the remaining algorithms (reference: ``fedml_experiments/standalone/{ditto,dpsgd,fedfomo,local,subavg}/main_*.py``
and the entry script TurboAggregate never had).  ``python -m neuroimagedisttraining_b200.cli.main_others <algo> [flags]``."""
import argparse
import sys

from ..fl import DittoAPI, DPSGDAPI, FedFomoAPI, LocalAPI, SubAvgAPI, TurboAggregateAPI
from .common import add_common_args, attach_metrics, finish, setup

ALGOS = {"ditto": DittoAPI, "dpsgd": DPSGDAPI, "fedfomo": FedFomoAPI, "local": LocalAPI, "subavg": SubAvgAPI, "turboaggregate": TurboAggregateAPI}


def add_args(parser, algo: str):
    add_common_args(parser, model="resnet18", dataset="cifar10", lr=0.1, epochs=5, frac=0.1, clients=100, comm_round=500, seed=2022, batch_size=128)
    if algo == "ditto":
        parser.add_argument("--local_epochs", type=int, default=3)
        parser.add_argument("--lamda", type=float, default=0.5)
    if algo == "dpsgd":
        parser.add_argument("--cs", type=str, default="ring")
        parser.add_argument("--type", type=str, default="epoch")
    if algo == "fedfomo":
        parser.add_argument("--cs", type=str, default="random")
    if algo == "subavg":
        parser.add_argument("--each_prune_ratio", type=float, default=0.05)
        parser.add_argument("--dense_ratio", type=float, default=0.5)
        parser.add_argument("--dist_thresh", type=float, default=1e-4)
        parser.add_argument("--acc_thresh", type=float, default=0.5)
    if algo == "turboaggregate":
        parser.add_argument("--ta_group_size", type=int, default=2)
        parser.add_argument("--ta_scale", type=int, default=1 << 16)
    return parser


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] not in ALGOS:
        raise SystemExit(f"usage: main_others {{{','.join(ALGOS)}}} [flags]")
    algo = argv.pop(0)
    args = add_args(argparse.ArgumentParser(description=f"{algo}-standalone"), algo).parse_args(argv)
    ctx, device, logger, dataset, trainer = setup(args, algo, val_split=0.1 if algo == "fedfomo" else 0.0)
    kw = {"dist": ctx}
    if algo == "turboaggregate":
        kw["agg_mode"] = args.agg
        args.final_finetune = False
    api = ALGOS[algo](dataset, device, args, trainer, logger, **kw)
    attach_metrics(api, args, ctx)
    if args.resume:
        api.resume(args.resume)
    api.train()
    return finish(api, args, logger, ctx)


if __name__ == "__main__":
    main()
