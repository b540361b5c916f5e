"""SalientGrads entry point (reference: ``fedml_experiments/standalone/sailentgrads/main_sailentgrads.py:31-280``).
``python -m neuroimagedisttraining_b200.cli.main_sailentgrads --client_num_in_total 21 --dense_ratio 0.5 ...``; multi-GPU:
``torchrun --nproc-per-node N -m neuroimagedisttraining_b200.cli.main_sailentgrads ...``."""
import argparse

from ..fl import SalientGradsAPI
from .common import add_common_args, attach_metrics, add_sparse_args, finish, setup, str2bool


def add_args(parser):
    add_common_args(parser, lr=0.01, epochs=2, frac=1.0, clients=4, seed=1024)
    add_sparse_args(parser, cs="v0")
    parser.add_argument("--itersnip_iteration", type=int, default=1)
    parser.add_argument("--stratified_sampling", default=False, action="store_true")
    parser.add_argument("--snip_mask", type=str2bool, default=True)
    return parser


def main(argv=None):
    args = add_args(argparse.ArgumentParser(description="SailentGrads-standalone")).parse_args(argv)
    args.dataset = "ABCD" if args.dataset == "ABCD" else args.dataset     # the reference forces ABCD (:272); other datasets are allowed here
    extra = ["customizedlowbatch", f"-cs{args.cs}", "-uniform_init" if args.uniform else "-ERK_init",
             "-diff_init" if args.different_initial else "-same_init", "-RSM" if args.static else "-DST"]
    ctx, device, logger, dataset, trainer = setup(args, "SailentGrads", extra)
    args.identity += f"-iteration{args.itersnip_iteration}-stratified{args.stratified_sampling}"
    api = SalientGradsAPI(dataset, device, args, trainer, logger, dist=ctx, agg_mode=args.agg)
    attach_metrics(api, args, ctx)
    if args.resume:
        api.resume(args.resume)
    api.train()
    return finish(api, args, logger, ctx)


if __name__ == "__main__":
    main()
