"""ModelTrainer ABC and the slab-backed trainer used by every algorithm.

API surface follows ``fedml_core/trainer/model_trainer.py:8-58`` and
``fedml_api/standalone/sailentgrads/my_model_trainer.py:22-280`` (``get_model_params``,
``set_model_params``, ``train``, ``test``, ``get_trainable_params``, ``get_model_sps``,
``calculate_sparsities``, ``init_masks``, ``screen_gradients``, ``count_communication_params``),
but the state lives in device-resident :class:`ParamSlab` s instead of CPU state_dict copies.
"""
from __future__ import annotations

import contextlib
from abc import ABC, abstractmethod
from collections import OrderedDict
from typing import Dict, Optional, Union

import numpy as np
import torch
import torch.nn as nn

from .. import ops
from ..runtime.slab import ParamSlab, make_mask_slab
from ..runtime import xfer
from ..runtime.staging import BatchStager, make_stager
from .engine import EvalCounters, LazyMean, StepHParams, TorchEngine


class ModelTrainer(ABC):
    """Abstract trainer (same contract as the reference's ABC)."""

    def __init__(self, model, args=None):
        self.model = model
        self.id = 0
        self.args = args

    def set_id(self, trainer_id):
        self.id = trainer_id

    @abstractmethod
    def get_model_params(self):
        ...

    @abstractmethod
    def set_model_params(self, model_parameters):
        ...

    @abstractmethod
    def train(self, train_data, device, args=None):
        ...

    @abstractmethod
    def test(self, test_data, device, args=None):
        ...

    def test_on_the_server(self, train_data_local_dict, test_data_local_dict, device, args=None) -> bool:
        return False

    def count_training_flops_per_sample(self):
        from ..utils.flops import count_training_flops
        return count_training_flops(self.model, getattr(self.args, "dataset", "ABCD"))

    def count_full_flops_per_sample(self):
        from ..utils.flops import count_training_flops
        return count_training_flops(self.model, getattr(self.args, "dataset", "ABCD"), full=True)

    def count_inference_flops(self, w):
        from ..utils.flops import count_inference_flops
        self.set_model_params(w)
        return count_inference_flops(self.model, getattr(self.args, "dataset", "ABCD"))

    def count_communication_params(self, update_to_server):
        """Number of non-zero entries (``model_trainer.py:49-53``); accepts a state_dict or a slab."""
        if isinstance(update_to_server, ParamSlab):
            return ops.count_nonzero(update_to_server.flat)
        total = 0
        for name in update_to_server:
            total = total + torch.count_nonzero(update_to_server[name])
        return total


MaskLike = Union[None, ParamSlab, Dict[str, torch.Tensor]]


class SlabTrainer(ModelTrainer):
    """Local SGD on a slab.  One instance is shared by all simulated clients of a rank (as in the
    reference, ``sailentgrads_api.py:36-45``); per-client state is the slab it is attached to."""

    def __init__(self, model: nn.Module, args=None, logger=None, device=None, store=None, engine=None,
                 loss_kind: str = "bce", clip_norm: float = 10.0, staging: str = "auto"):
        """``staging``: how index batches reach the device -- ``client`` (each loader's volumes uploaded once per use, batches
        gathered on the device; the default on CUDA), ``stream`` (batch-by-batch copies), ``resident`` (whole store on the device);
        see :mod:`..runtime.staging`.  ``args.staging`` / ``--staging`` overrides ``auto``."""
        super().__init__(model, args)
        self.logger = logger
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        self.store = store
        self.engine = engine if engine is not None else TorchEngine(
            model, self.device, loss_kind=loss_kind, volumetric=store is not None,
            input_scale=getattr(store, "scale", 1.0), amp=bool(getattr(args, "amp", False)))
        self.layout = self.engine.layout
        self.clip_norm = clip_norm
        self.work = self.engine.new_slab()         # the slab API-compat calls operate on
        self.engine.attach(self.work)
        self.mom = ParamSlab(self.layout, device=self.device, params_only=True)
        self.masks: Optional[ParamSlab] = None
        self._stagers: Dict[int, BatchStager] = {}
        if staging == "auto":
            staging = getattr(args, "staging", None) or "auto"
        self.staging = "client" if staging == "auto" else staging
        self.double_sigmoid_eval = True
        self.step_counter = 0
        self.pending_losses = []
        self.lanes: list = []                  # sibling trainers (own engine, stagers, CUDA stream) for concurrent client slots
        self.stream = None                     # None: whatever stream is current

    # ------------------------------------------------------------------ slab plumbing
    @property
    def slab(self) -> ParamSlab:
        return self.engine.slab

    def attach(self, slab: ParamSlab) -> None:
        self.engine.attach(slab)

    def _stager(self, batch_size: int) -> BatchStager:
        st = self._stagers.get(batch_size)
        if st is None:
            st = make_stager(self.staging, self.store, self.device, batch_size)
            st.fixed_output = bool(getattr(self.engine, "use_graphs", False))
            self._stagers[batch_size] = st
        return st

    def add_lane(self, engine) -> "SlabTrainer":
        """A second client slot on this GPU: a sibling trainer with its own engine (activation buffers), stagers and CUDA stream.
        Orchestrators that see lanes train several clients at once, one per lane, interleaving their steps from one host thread
        (``FedAvgAPI.run_round``): the small layers of one client (grids far below 148 CTAs) run in the shadow of the other's
        large ones.  The reference trains its clients strictly one after another (``sailentgrads_api.py:126-138``)."""
        lane = SlabTrainer(self.model, self.args, self.logger, device=self.device, store=self.store, engine=engine, clip_norm=self.clip_norm,
                           staging=self.staging)
        lane.double_sigmoid_eval = self.double_sigmoid_eval
        lane.stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None
        self.lanes.append(lane)
        return lane

    def all_lanes(self) -> list:
        return [self] + self.lanes

    def set_staging(self, mode: str) -> None:
        """Switch the staging mode (``client`` | ``stream`` | ``resident``); existing device buffers are dropped."""
        for lane in self.lanes:
            lane.set_staging(mode)
        if mode != self.staging:
            if self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
            self._stagers.clear()
            self._prepared = None
            self.staging = mode

    @property
    def h2d_bytes(self) -> int:
        """Bytes copied host -> device by this trainer's stagers so far."""
        return sum(st.h2d_bytes for st in self._stagers.values()) + sum(l.h2d_bytes for l in self.lanes)

    def _cached(self, data) -> bool:
        return (self.staging != "stream" and self.store is not None and hasattr(data, "indices")
                and self._stager(data.batch_size).cacheable(data))

    def plan_data(self, loaders, kind: str = "train", group: bool = False) -> None:
        """Tell the stager which loaders come next (in order) so their volumes are uploaded ahead of use, behind the compute
        stream: the round's clients, then its test splits (``group=True``: one slot for all of them), then the first clients of
        the next round.  A no-op for tensor loaders and in ``stream`` / ``resident`` mode."""
        loaders = [l for l in loaders if self._cached(l)]
        if not loaders:
            return
        by_bs: Dict[int, list] = {}
        for l in loaders:
            by_bs.setdefault(l.batch_size if not group else max(x.batch_size for x in loaders), []).append(l)
        for bs, ls in by_bs.items():
            self._stager(bs).want(ls, kind=kind, group=group)

    def _batches(self, data):
        """Uniform ``(x, y)`` stream: index loaders go through the volume store, tensor loaders pass through."""
        if self.store is not None and hasattr(data, "indices"):
            yield from self._stager(data.batch_size).iterate(data)
        else:
            for batch in data:
                yield batch[0], batch[1]

    def _epoch_stream(self, data, epochs: int):
        """(index-batch generator over ``epochs`` passes of ``data``, list that receives the epoch tag of every batch)."""
        tags = []

        def index_stream():
            for e in range(epochs):
                for b in data:
                    tags.append(e)
                    yield b
        return index_stream(), tags

    def _announce_ids(self, data) -> None:
        """Tell the engine which samples (ids in the volume store) the batch it is about to train on holds, in row order: engines
        that cache per-sample quantities (``B200Engine``: Gram matrices of the first layer's input patches) key them on these."""
        if hasattr(self.engine, "batch_ids") and self.store is not None and hasattr(data, "indices"):
            st = self._stagers.get(data.batch_size) if hasattr(self, "_stagers") else None
            self.engine.batch_ids = getattr(st, "last_ids", None) if st is not None else None
            self.engine.batch_ids_key = id(self.store)

    def _batches_epochs(self, data, epochs: int, lookahead_data=None):
        """``(epoch, x, y)`` over ``epochs`` passes of ``data``.  Client staging: the loader's volumes are uploaded once (ahead of
        time when :meth:`plan_data` announced them) and every epoch's batches are gathered on the device.  Stream staging: ONE
        staged stream over all epochs.  ``lookahead_data``: the loader the next call will train on; its upload (client mode) / first
        batch (stream mode) is started while this client's steps still run."""
        if self._cached(data):
            if lookahead_data is not None:
                self.plan_data([lookahead_data])
            yield from self._stager(data.batch_size).cached_batches(data, epochs)
        elif self.store is not None and hasattr(data, "indices"):
            prepared, self._prepared = getattr(self, "_prepared", None), None
            if prepared is not None and prepared[0] is data and prepared[1] == epochs:
                stream, tags = prepared[2], prepared[3]
            else:
                stream, tags = self._epoch_stream(data, epochs)
            la_stream = None
            if lookahead_data is not None and hasattr(lookahead_data, "indices") and lookahead_data.batch_size == data.batch_size:
                la_stream, la_tags = self._epoch_stream(lookahead_data, epochs)
                self._prepared = (lookahead_data, epochs, la_stream, la_tags)
            for i, (x, y) in enumerate(self._stager(data.batch_size).iterate(stream, lookahead=la_stream)):
                yield tags[i], x, y
        else:
            for e in range(epochs):
                for batch in data:
                    yield e, batch[0], batch[1]

    def test_multi(self, loaders, slab_sets):
        """Evaluate several models on the same test loaders with the data staged ONCE: ``slab_sets[p][j]`` is the slab of pass
        ``p`` for ``loaders[j]`` (pass 0 = the global model on every client, pass 1 = the personalised models; the reference stages
        each client's test split once per pass, ``sailentgrads_api.py:247-256``).  Returns ``results[p][j]`` metric dicts."""
        loaders = list(loaders)
        flat = self.test_jobs([(loaders[j], slabs[j]) for slabs in slab_sets for j in range(len(loaders))])
        n = len(loaders)
        return [flat[p * n:(p + 1) * n] for p in range(len(slab_sets))]

    def _eval_split(self, loaders) -> list:
        """Which lane evaluates which test loader: loader i of the group goes to lane i % L (all jobs of a loader run where its
        volumes are staged)."""
        L = len(self.all_lanes()) if len(loaders) > 1 else 1
        return [loaders[k::L] for k in range(L)]

    def plan_eval(self, loaders) -> None:
        """Announce the round's evaluation group (see :meth:`plan_data`); with lanes, every lane stages its share."""
        loaders = list(loaders)
        for lane, ls in zip(self.all_lanes(), self._eval_split(loaders)):
            lane.plan_data(ls, kind="eval", group=True)

    def _test_jobs_steps(self, jobs, loaders, acc):
        """Generator: evaluate ``jobs`` (loader, slab) on this trainer with ``loaders`` staged as one group, counters into ``acc``
        rows; yields after every enqueued batch."""
        bs = max(l.batch_size for l in loaders)
        st = self._stager(bs)
        ent = st.open(loaders, kind="eval")
        try:
            for j, (loader, slab) in enumerate(jobs):
                self.attach(slab)
                c = EvalCounters(self.device, acc[j])
                for b in loader:
                    x, y = st.batch_from(ent, loader, b[0])
                    self.engine.eval_batch(x, y, c, double_sigmoid=self.double_sigmoid_eval)
                    yield
        finally:
            st.close(ent)

    def test_jobs(self, jobs, group=None):
        """Evaluate ``[(loader, slab), ...]`` in order; every distinct loader's volumes are uploaded once (one slot for the whole
        group) however many jobs use it.  ``group``: the loader list announced with ``plan_eval`` (defaults to the distinct loaders
        of ``jobs`` in order).  With lanes the loaders (and their jobs) are spread over the lanes and run concurrently.  One
        device->host read returns all counters.  Returns one metrics dict per job."""
        jobs = list(jobs)
        if not jobs:
            return []
        loaders = list(group) if group is not None else list(dict((id(l), l) for l, _ in jobs).values())
        if not all(self._cached(l) for l in loaders):
            return self.test_many([(slab, loader) for loader, slab in jobs])
        lanes = self.all_lanes()
        split = self._eval_split(loaders)
        lane_of = {id(l): k for k, ls in enumerate(split) for l in ls}
        per_lane = [[] for _ in split]
        for j, (loader, slab) in enumerate(jobs):
            per_lane[lane_of[id(loader)]].append((j, loader, slab))
        main = torch.cuda.current_stream(self.device)
        accs, gens = [], []
        for k, items in enumerate(per_lane):
            lane = lanes[k]
            acc = torch.zeros((max(1, len(items)), 3), dtype=torch.float64, device=self.device)
            accs.append(acc)
            if lane.stream is not None:
                lane.stream.wait_stream(main)
            gens.append(lane._test_jobs_steps([(l, s) for _, l, s in items], split[k], acc) if items else None)
        live = True
        while live:
            live = False
            for k, g in enumerate(gens):
                if g is None:
                    continue
                live = True
                ctx = torch.cuda.stream(lanes[k].stream) if lanes[k].stream is not None else contextlib.nullcontext()
                with ctx:
                    try:
                        next(g)
                    except StopIteration:
                        gens[k] = None
        for lane in lanes[:len(split)]:
            if lane.stream is not None:
                main.wait_stream(lane.stream)
        host = xfer.to_host(torch.cat(accs)).tolist()          # ONE device->host read for all jobs' counters
        out = [None] * len(jobs)
        row = 0
        for k, items in enumerate(per_lane):
            for n, (j, _, _) in enumerate(items):
                out[j] = EvalCounters.as_result(*host[row + n])
            row += max(1, len(items))
        return out

    def test_many(self, jobs):
        """Evaluate ``[(slab, loader), ...]`` back to back as one staged stream (the copy of the next client's test
        batch overlaps the current forward pass).  Returns one metrics dict per job."""
        jobs = list(jobs)
        if not jobs:
            return []
        if not (self.store is not None and all(hasattr(l, "indices") for _, l in jobs)):
            out = []
            for slab, loader in jobs:
                self.attach(slab)
                out.append(self.test(loader))
            return out
        bs = max(l.batch_size for _, l in jobs)
        tags = []

        def index_stream():
            for j, (_, loader) in enumerate(jobs):
                for b in loader:
                    tags.append(j)
                    yield b
        counters = [EvalCounters(self.device) for _ in jobs]
        cur = -1
        for i, (x, y) in enumerate(self._stager(bs).iterate(index_stream())):
            j = tags[i]
            if j != cur:
                self.attach(jobs[j][0])
                cur = j
            self.engine.eval_batch(x, y, counters[j], double_sigmoid=self.double_sigmoid_eval)
        return [c.result() for c in counters]

    def mask_slab(self, masks: MaskLike) -> Optional[ParamSlab]:
        if masks is None or isinstance(masks, ParamSlab):
            return masks
        m = make_mask_slab(self.layout, self.device)
        for name, t in masks.items():
            if name in self.layout:
                m.view(name).copy_(t.to(self.device, torch.float32))
        return m

    # ------------------------------------------------------------- reference-style API
    def set_masks(self, masks: MaskLike):
        self.masks = self.mask_slab(masks)

    def get_model_params(self):
        self.engine.detach_sync()
        return self.slab.state_dict()

    def set_model_params(self, model_parameters):
        if isinstance(model_parameters, ParamSlab):
            self.slab.copy_(model_parameters)
        else:
            self.slab.load_state_dict(model_parameters)
        self.slab.sync_int_buffers_to(self.model)

    def get_trainable_params(self):
        return OrderedDict((n, self.slab.view(n)) for n in self.layout.param_names)

    def get_model_sps(self, slab: Optional[ParamSlab] = None) -> float:
        """Percentage of zero entries among the trainable parameters (``my_model_trainer.py:144-158``)."""
        s = slab if slab is not None else self.slab
        nz = int(ops.count_nonzero(s.params_flat).item())
        total = self.layout.numel_real(params_only=True)
        return 100.0 * (total - nz) / total

    # ----------------------------------------------------------------- sparsity helpers
    def calculate_sparsities(self, params=None, tabu=(), distribution: str = "ERK", sparse: float = 0.5) -> Dict[str, float]:
        """Per-tensor sparsity: ``uniform`` (1-dense_ratio everywhere) or Erdos-Renyi-Kernel with the
        densify-and-retry loop (``DisPFL/my_model_trainer.py:43-117``)."""
        shapes = {n: tuple(self.layout[n].shape) for n in (params if params is not None else self.layout.param_names)}
        if distribution == "uniform":
            return {n: (0.0 if n in tabu else 1.0 - float(self.args.dense_ratio)) for n in shapes}
        power = float(getattr(self.args, "erk_power_scale", 1.0))
        density = sparse
        dense_layers = set(n for n in shapes if n in tabu)
        while True:
            divisor, rhs, raw = 0.0, 0.0, {}
            for n, shp in shapes.items():
                n_param = float(np.prod(shp))
                if n in dense_layers:
                    rhs -= n_param * (1 - density)
                else:
                    rhs += n_param * density
                    raw[n] = (float(np.sum(shp)) / n_param) ** power
                    divisor += raw[n] * n_param
            eps = rhs / divisor
            mx = max(raw.values())
            if mx * eps > 1:
                dense_layers |= {n for n, v in raw.items() if v == mx}
            else:
                break
        return {n: (0.0 if n in dense_layers else 1.0 - eps * raw[n]) for n in shapes}

    def init_masks(self, params=None, sparsities: Dict[str, float] = None, generator: Optional[torch.Generator] = None) -> ParamSlab:
        """Random per-tensor masks at the given sparsities (``DisPFL/my_model_trainer.py:31-41``)."""
        m = ParamSlab(self.layout, device=self.device, params_only=True)
        for n in (params if params is not None else self.layout.param_names):
            e = self.layout[n]
            dense = int((1 - sparsities[n]) * e.numel)
            if dense > 0:
                perm = torch.randperm(e.numel, generator=generator)[:dense].to(self.device)
                m.view(n).view(-1)[perm] = 1
        return m

    # -------------------------------------------------------------------------- train
    def train(self, train_data, device=None, args=None, round_idx: int = 0, masks: MaskLike = None,
              prox_ref: Optional[ParamSlab] = None, lamda: float = 0.0, epochs: Optional[int] = None,
              apply_mask: Optional[bool] = None, grad_mask: bool = False, lookahead_data=None):
        """``epochs`` of local SGD on the attached slab.  A fresh optimizer per call (momentum
        buffers reset), ``lr * lr_decay**round``, grad-clip 10, weight decay on every parameter and
        ``w *= mask`` after every step when ``snip_mask`` (``my_model_trainer.py:201-235``)."""
        gen = self.train_steps(train_data, device, args, round_idx, masks, prox_ref, lamda, epochs, apply_mask, grad_mask, lookahead_data)
        try:
            while True:
                next(gen)
        except StopIteration as stop:
            return stop.value

    def train_steps(self, train_data, device=None, args=None, round_idx: int = 0, masks: MaskLike = None,
                    prox_ref: Optional[ParamSlab] = None, lamda: float = 0.0, epochs: Optional[int] = None,
                    apply_mask: Optional[bool] = None, grad_mask: bool = False, lookahead_data=None):
        """:meth:`train` as a generator that yields after every SGD step has been enqueued, so that a caller can interleave the
        steps of several clients running on several lanes (CUDA streams).  The generator's return value is :meth:`train`'s."""
        args = args if args is not None else self.args
        mslab = self.mask_slab(masks) if masks is not None else self.masks
        use_mask = bool(getattr(args, "snip_mask", True)) if apply_mask is None else apply_mask
        mask_flat = mslab.flat if (mslab is not None and use_mask) else None
        if getattr(args, "client_optimizer", "sgd") != "sgd":
            raise ValueError("only client_optimizer='sgd' defines an optimizer (as in the reference)")
        hp = StepHParams(lr=args.lr * (args.lr_decay ** round_idx), momentum=float(getattr(args, "momentum", 0.0)),
                         wd=float(getattr(args, "wd", 0.0)), clip_norm=self.clip_norm)
        mom = self.mom.flat if hp.momentum != 0 else None
        first = True
        losses = []
        n_epochs = int(epochs if epochs is not None else args.epochs)

        def end_epoch(epoch):
            # the per-epoch loss stays on the device; it is read (one sync) after the client's last epoch, or at the end of the
            # round when args.loss_sync == "round" (the reference syncs on loss.item() every step, my_model_trainer.py:226)
            losses.append((self.id, epoch, self.engine.pop_mean_loss(sync=False)))

        # all epochs are one batch stream, so the host->device prefetch runs across epoch boundaries
        cur = 0
        for epoch, x, y in self._batches_epochs(train_data, n_epochs, lookahead_data):
            if epoch != cur:
                end_epoch(cur)
                cur = epoch
            self._announce_ids(train_data)
            if grad_mask:
                self.engine.train_step(x, y, hp, None, mom, first, grad_mask=mslab.flat if mslab is not None else None)
            else:
                self.engine.train_step(x, y, hp, mask_flat, mom, first)
            if prox_ref is not None and lamda != 0:
                ops.prox_pull(self.slab.params_flat, prox_ref.params_flat, hp.lr, lamda)
            first = False
            self.step_counter += 1
            yield
        if n_epochs > 0:
            end_epoch(cur)
        self.pending_losses.extend(losses)
        if getattr(args, "loss_sync", "client") != "round":
            return self.flush_losses()
        return losses

    def flush_losses(self):
        """Read the queued per-epoch losses (one host sync) and log them in the reference's format."""
        for lane in self.lanes:              # their losses were enqueued on their own streams
            if lane.pending_losses:
                if lane.stream is not None:
                    torch.cuda.current_stream(self.device).wait_stream(lane.stream)
                self.pending_losses.extend(lane.pending_losses)
                lane.pending_losses = []
        if not self.pending_losses:
            return []
        sums = torch.stack([(l.sum if isinstance(l, LazyMean) else l).detach().float().reshape(()) for _, _, l in self.pending_losses])
        vals = xfer.to_host(sums).tolist()
        vals = [v / l.steps if isinstance(l, LazyMean) else v for v, (_, _, l) in zip(vals, self.pending_losses)]
        out = []
        for (cid, epoch, _), v in zip(self.pending_losses, vals):
            out.append(v)
            if self.logger is not None:
                self.logger.info('Client Index = {}\tEpoch: {}\tLoss: {:.6f}'.format(cid, epoch, v))
        self.pending_losses = []
        return out

    # --------------------------------------------------------------------------- test
    def test(self, test_data, device=None, args=None):
        c = EvalCounters(self.device)
        for x, y in self._batches(test_data):
            self.engine.eval_batch(x, y, c, double_sigmoid=self.double_sigmoid_eval)
        return c.result()

    # --------------------------------------------------------------- gradient screening
    def screen_gradients(self, train_data, device=None) -> ParamSlab:
        """Dense gradient of one batch in eval mode (``DisPFL/my_model_trainer.py:166-189``)."""
        src = self.engine
        for x, y in self._batches(train_data):
            try:
                self.engine.compute_grads(x, y, train_mode=False)
            except NotImplementedError:
                # the sm_100a engines implement batch-statistics BatchNorm backward only; this once-per-round eval-mode
                # gradient goes through the PyTorch engine on the same slab
                if getattr(self, "_screen_engine", None) is None:
                    self._screen_engine = TorchEngine(self.model, self.device, loss_kind=getattr(self.engine, "loss_kind", "bce"),
                                                      volumetric=self.store is not None, input_scale=getattr(self.store, "scale", 1.0))
                src = self._screen_engine
                src.attach(self.slab)
                src.compute_grads(x, y, train_mode=False)
            break
        g = ParamSlab(self.layout, device=self.device, params_only=True)
        g.flat.copy_(src.grad.flat)
        return g

    def saliency_batch(self, x, y, scores: ParamSlab, scale: float = 1.0) -> None:
        """One SNIP evaluation: ``scores += scale*|g*w|`` with BN in batch-stat mode and dropout on,
        as the reference's deep-copied model is (``snip.py:38-66``); weights are not modified."""
        P = self.layout.param_numel
        self.engine.detach_sync()
        saved = self.slab.flat[P:].clone()           # BN running stats must not move (the reference scores a deep copy)
        fused = getattr(self.engine, "saliency_step", None)
        done = bool(fused(x, y, scores.flat, scale)) if fused is not None else False      # |dW*w| from the wgrad reductions themselves
        if not done:
            self.engine.compute_grads(x, y, train_mode=True)
        self.slab.flat[P:].copy_(saved)
        self.slab.sync_int_buffers_to(self.model)
        if not done:
            ops.saliency_accumulate(scores.flat, self.engine.grad.flat, self.slab.params_flat, scale)
