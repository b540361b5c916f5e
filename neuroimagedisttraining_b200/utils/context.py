"""Fail-fast context for multi-process runs: any exception in one rank tears the whole job down instead of leaving the
other ranks blocked in a collective (reference: ``raise_MPI_error`` -> ``MPI.COMM_WORLD.Abort()``,
``fedml_api/utils/context.py:10-19``; here the process group is destroyed and the process exits non-zero, which makes
``torchrun`` kill the remaining ranks).  Also a watchdog for kernels that wait on peers."""
from __future__ import annotations

import logging
import os
import sys
import threading
import traceback
from contextlib import contextmanager


@contextmanager
def raise_dist_error(exit_code: int = 1):
    try:
        yield
    except Exception as e:  # noqa: BLE001
        logging.error("rank %s: %s\n%s", os.environ.get("RANK", "0"), e, traceback.format_exc())
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                dist.destroy_process_group()
        finally:
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(exit_code)


raise_MPI_error = raise_dist_error   # reference name


class Watchdog:
    """Kills the process if ``pet()`` is not called for ``timeout_s`` -- a hung collective or a peer that died must not
    leave a GPU spinning."""

    def __init__(self, timeout_s: float = 600.0):
        self.timeout_s = timeout_s
        self._ev = threading.Event()
        self._stop = False
        self._thr = threading.Thread(target=self._run, daemon=True)

    def start(self):
        self._thr.start()
        return self

    def pet(self):
        self._ev.set()

    def stop(self):
        self._stop = True
        self._ev.set()

    def _run(self):
        while not self._stop:
            if not self._ev.wait(self.timeout_s):
                logging.error("watchdog: no progress for %.0f s, aborting rank %s", self.timeout_s, os.environ.get("RANK", "0"))
                os._exit(124)
            self._ev.clear()
