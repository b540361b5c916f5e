"""Simulated FL client (reference: ``fedml_api/standalone/sailentgrads/client.py:14-119``,
``fedavg/client.py:9-65``).  A client owns its data loaders, its sample count and a personal
parameter slab (its last local weights, used for the personalised evaluation); the trainer is
shared."""
from __future__ import annotations

from typing import Optional

import numpy as np

from ..runtime.slab import ParamSlab
from .trainer import SlabTrainer


class Client:
    def __init__(self, client_idx, local_training_data, local_test_data, local_sample_number, args, device,
                 model_trainer: SlabTrainer, logger=None):
        self.logger = logger
        self.client_idx = client_idx
        self.local_training_data = local_training_data
        self.local_test_data = local_test_data
        self.local_sample_number = local_sample_number
        if logger is not None:
            logger.info("self.local_sample_number = " + str(self.local_sample_number))
        self.args = args
        self.device = device
        self.model_trainer = model_trainer
        self.slab: Optional[ParamSlab] = None     # personal weights (device resident)

    def update_local_dataset(self, client_idx, local_training_data, local_test_data, local_sample_number):
        self.client_idx = client_idx
        self.local_training_data = local_training_data
        self.local_test_data = local_test_data
        self.local_sample_number = local_sample_number

    def get_sample_number(self):
        return self.local_sample_number

    def ensure_slab(self, like: ParamSlab) -> ParamSlab:
        if self.slab is None:
            self.slab = like.clone()
        return self.slab

    # ------------------------------------------------------------------------- saliency
    def generate_sailency_scores_from_each_client(self, itersnip_iteration: int = 1, stratified_sampling: bool = False,
                                                  w_global: Optional[ParamSlab] = None, out: Optional[ParamSlab] = None) -> ParamSlab:
        """Mean over ``itersnip_iteration`` random mini-batches of ``|dL/dm|`` (IterSNIP;
        ``client.py:30-53``).  Every iteration draws the first batch of a freshly shuffled pass.
        ``stratified_sampling`` draws class-balanced mini-batches instead (the reference's variant of
        this flag is broken for ABCD; this implements its intent)."""
        tr = self.model_trainer
        if w_global is not None:
            tr.attach(w_global)
        scores = out if out is not None else ParamSlab(tr.layout, device=tr.device, params_only=True)
        n_it = int(itersnip_iteration)
        loader = self.local_training_data
        for it in range(n_it):
            if stratified_sampling and hasattr(loader, "indices"):
                batch = self._stratified_batch(loader, it)
                x, y = next(iter(tr._batches(_OneBatch(batch, loader.batch_size))))
            else:
                x, y = next(iter(tr._batches(_FirstBatch(loader))))
            tr.saliency_batch(x, y, scores, scale=1.0 / n_it)
        return scores

    def _stratified_batch(self, loader, it: int):
        rng = np.random.RandomState(42 + it)
        y = loader.labels.numpy()
        bs = min(loader.batch_size, len(y))
        pick = []
        classes = np.unique(y)
        for c in classes:
            pool = np.where(y == c)[0]
            k = max(1, int(round(bs * len(pool) / len(y))))
            pick.append(rng.choice(pool, min(k, len(pool)), replace=False))
        sel = np.concatenate(pick)[:bs]
        return loader.indices[sel], loader.labels[sel], loader.sites[sel]

    # ---------------------------------------------------------------------------- train
    def train(self, w_global: ParamSlab, round_idx: int, masks=None, **kw):
        """Local training starting from ``w_global``; the result stays in ``self.slab`` (device).
        Returns ``(slab, training_flops, num_comm_params)`` like the reference (``client.py:80-104``);
        ``num_comm_params`` is a device scalar (no host sync)."""
        gen = self.train_steps(w_global, round_idx, masks, **kw)
        try:
            while True:
                next(gen)
        except StopIteration as stop:
            return stop.value

    def train_steps(self, w_global: ParamSlab, round_idx: int, masks=None, trainer: Optional[SlabTrainer] = None, **kw):
        """:meth:`train` as a generator yielding after every enqueued SGD step (see ``SlabTrainer.train_steps``); ``trainer``: the
        lane to run on (default: the client's own trainer)."""
        tr = trainer if trainer is not None else self.model_trainer
        if hasattr(tr.engine, "step"):
            # the dropout stream of the native engines is a counter: tie it to (round, client) so that the masks do not depend on
            # how many steps OTHER clients ran on this engine before (sequential and concurrent schedules then agree bit for bit)
            tr.engine.step = (((int(round_idx) + 2) * 1000003 + int(self.client_idx)) & 0xFFFFFFF) << 12
        comm_out = kw.pop("comm_out", None)        # int64 device accumulator: nnz(download) + nnz(upload) are added by the count kernels
        slab = self.ensure_slab(w_global)
        slab.copy_(w_global)
        tr.attach(slab)
        tr.set_id(self.client_idx)
        if comm_out is not None and slab.flat.is_cuda:
            from .. import ops
            ops.count_nonzero(w_global.flat, out=comm_out)
            num_comm_params = comm_out
        else:
            num_comm_params = tr.count_communication_params(w_global)
        yield from tr.train_steps(self.local_training_data, self.device, self.args, round_idx, masks, **kw)
        tr.engine.detach_sync()
        training_flops = self.args.epochs * self.local_sample_number
        if num_comm_params is comm_out and comm_out is not None:
            ops.count_nonzero(slab.flat, out=comm_out)
        else:
            num_comm_params = num_comm_params + tr.count_communication_params(slab)
        return slab, training_flops, num_comm_params

    def local_test(self, w_per: ParamSlab, b_use_test_dataset: bool = True):
        data = self.local_test_data if b_use_test_dataset else self.local_training_data
        self.model_trainer.attach(w_per)
        return self.model_trainer.test(data, self.device, self.args)


class _FirstBatch:
    """Iterating yields only the first batch of a (re-shuffled) pass: ``next(iter(loader))``."""

    def __init__(self, loader):
        self.loader = loader
        self.batch_size = loader.batch_size
        if hasattr(loader, "indices"):
            self.indices = loader.indices

    def __iter__(self):
        for b in self.loader:
            yield b
            return


class _OneBatch:
    def __init__(self, batch, batch_size):
        self.batch = batch
        self.batch_size = batch_size
        self.indices = batch[0]

    def __iter__(self):
        yield self.batch
