"""DisPFL entry point (reference: ``fedml_experiments/standalone/DisPFL/main_dispfl.py:31-274``; defaults lr 0.001, epochs 5,
21 clients, frac 0.1, 10 rounds, cs random)."""
import argparse

from ..fl import DisPFLAPI
from .common import add_common_args, attach_metrics, add_sparse_args, finish, setup


def add_args(parser):
    add_common_args(parser, lr=0.001, epochs=5, frac=0.1, clients=21, comm_round=10, seed=1024)
    add_sparse_args(parser, cs="random")
    parser.add_argument("--neighbor_aggregation", action="store_true",
                        help="enable the neighbour aggregation the reference has commented out (dispfl_api.py:138-142)")
    return parser


def main(argv=None):
    args = add_args(argparse.ArgumentParser(description="DisPFL-standalone")).parse_args(argv)
    extra = [f"-cs{args.cs}", "-masks" if args.save_masks else "", "-diff_spa" if args.diff_spa else "",
             "-uniform_init" if args.uniform else "-ERK_init", "-diff_init" if args.different_initial else "-same_init",
             "-g" if args.global_test else "", "-RSM" if args.static else "-DST"]
    ctx, device, logger, dataset, trainer = setup(args, "DisPFL", extra)
    api = DisPFLAPI(dataset, device, args, trainer, logger, dist=ctx)
    attach_metrics(api, args, ctx)
    if args.resume:
        api.resume(args.resume)
    api.train()
    return finish(api, args, logger, ctx)


if __name__ == "__main__":
    main()
