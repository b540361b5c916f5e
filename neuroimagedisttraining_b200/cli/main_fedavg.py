"""FedAvg entry point (reference: ``fedml_experiments/standalone/fedavg/main_fedavg.py:39-226``; defaults lr 0.001,
epochs 5, frac 0.9, seed 0).  Unlike the reference (``:203``) it also runs without CUDA."""
import argparse

from ..fl import FedAvgAPI
from .common import add_common_args, attach_metrics, finish, setup


def add_args(parser):
    return add_common_args(parser, lr=0.001, epochs=5, frac=0.9, clients=4, seed=0)


def main(argv=None):
    args = add_args(argparse.ArgumentParser(description="FedAvg-standalone")).parse_args(argv)
    ctx, device, logger, dataset, trainer = setup(args, "fedavg")
    api = FedAvgAPI(dataset, device, args, trainer, logger, dist=ctx, agg_mode=args.agg)
    attach_metrics(api, args, ctx)
    if args.resume:
        api.resume(args.resume)
    api.train()
    return finish(api, args, logger, ctx)


if __name__ == "__main__":
    main()
