"""torchrun worker (gloo): rank 1 fails inside ``raise_dist_error`` while rank 0 waits in a collective -- the job must end, not hang."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from neuroimagedisttraining_b200.utils.context import raise_dist_error


def main():
    dist.init_process_group("gloo")
    rank = dist.get_rank()
    with raise_dist_error(exit_code=7):
        if rank == 1:
            time.sleep(0.5)
            raise RuntimeError("injected fault on rank 1")
        t = torch.zeros(1)
        dist.all_reduce(t)              # rank 1 never joins
        time.sleep(120)
    print("UNREACHABLE", flush=True)


if __name__ == "__main__":
    main()
