"""Shared machinery of the FL orchestrators: client set-up, deterministic sampling, device-side
aggregation, evaluation of all clients, statistics, phase timing and checkpoint/resume.

Reference orchestrators: ``fedml_api/standalone/{fedavg/fedavg_api.py:12-196,
sailentgrads/sailentgrads_api.py:18-346}``.  Departures: the global model, every
client's personal model and the mask are device-resident slabs; clients are placed on ranks
(one process per GPU) and train there; aggregation is a flat kernel (+ fused cross-rank collective).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .. import ops
from ..parallel.aggregation import Aggregator
from ..parallel.context import DistContext
from ..runtime.slab import ParamSlab
from ..utils.checkpoint import load_checkpoint, save_checkpoint
from ..utils.timing import PhaseTimer
from .client import Client
from .trainer import SlabTrainer


class FedAPIBase:
    NAME = "fed"
    PERSONAL_IN_CHECKPOINT = False     # algorithms whose result IS the per-client models set this
    BALANCE_EVAL = False               # spread the round's evaluation jobs over ranks by load (FedAvg family)

    def __init__(self, dataset, device, args, model_trainer: SlabTrainer, logger, dist: Optional[DistContext] = None,
                 agg_mode: str = "auto"):
        self.logger = logger
        self.device = torch.device(device)
        self.args = args
        self.dist = dist if dist is not None else DistContext(device=self.device)
        [train_data_num, test_data_num, train_data_global, test_data_global,
         train_data_local_num_dict, train_data_local_dict, test_data_local_dict, class_counts] = dataset[:8]
        self.train_global = train_data_global
        self.test_global = test_data_global
        self.val_global = None
        self.val_data_local_dict = dataset[8] if len(dataset) > 8 else None
        self.train_data_num_in_total = train_data_num
        self.test_data_num_in_total = test_data_num
        self.train_data_local_num_dict = train_data_local_num_dict
        self.train_data_local_dict = train_data_local_dict
        self.test_data_local_dict = test_data_local_dict
        self.class_counts = class_counts
        self.model_trainer = model_trainer
        self.layout = model_trainer.layout
        if getattr(args, "placement", "lpt") == "lpt" and self.dist.is_distributed and hasattr(self.dist, "place_clients"):
            self.dist.place_clients([train_data_local_num_dict[c] for c in range(args.client_num_in_total)])
        self.client_list: List[Client] = []
        self._trained: set = set()               # clients that hold a personal model on their owner (the same on every rank)
        self._train_load: List[float] = [0.0] * self.dist.world_size
        self._eval_tmp: List[ParamSlab] = []
        self._setup_clients(train_data_local_num_dict, train_data_local_dict, test_data_local_dict, model_trainer)
        self.init_stat_info()
        self.w_global: ParamSlab = model_trainer.engine.new_slab()
        self.aggregator = Aggregator(self.dist, self.layout, self.device, agg_mode)
        self.aggregator.adopt_global(self.w_global)
        self.timer = PhaseTimer(self.device)
        self.start_round = 0
        self.mask: Optional[ParamSlab] = None
        self.w_init: Optional[ParamSlab] = None   # personalised model of a client that never trained
        self._resumed_extra: dict = {}
        self.metrics_sink = None                 # callable(**metrics): the entry points plug a JSONL writer in (one line per evaluation)

    # ------------------------------------------------------------------------- set-up
    def _setup_clients(self, train_data_local_num_dict, train_data_local_dict, test_data_local_dict, model_trainer):
        self.logger.info("############setup_clients (START)#############")
        for client_idx in range(self.args.client_num_in_total):
            c = Client(client_idx, train_data_local_dict[client_idx], test_data_local_dict[client_idx],
                       train_data_local_num_dict[client_idx], self.args, self.device, model_trainer, self.logger)
            self.client_list.append(c)
        self.logger.info("############setup_clients (END)#############")

    def init_stat_info(self):
        self.stat_info = {
            "label_num": self.class_counts, "sum_comm_params": 0, "sum_training_flops": 0, "avg_inference_flops": 0,
            "old_mask_test_acc": [], "new_mask_test_acc": [], "final_masks": [], "mask_dis_matrix": [],
            "global_test_acc": [], "person_test_acc": [], "global_test_loss": [], "person_test_loss": [],
            "round_time_s": [],
        }
        self._comm_params_dev = None

    # ----------------------------------------------------------------------- sampling
    def _client_sampling(self, round_idx, client_num_in_total, client_num_per_round):
        """All clients when ``per_round == total``; otherwise ``np.random.seed(round_idx)`` +
        ``choice`` without replacement, so every framework picks the same clients each round
        (``sailentgrads_api.py:152-160``)."""
        if client_num_in_total == client_num_per_round:
            client_indexes = [i for i in range(client_num_in_total)]
        else:
            num_clients = min(client_num_per_round, client_num_in_total)
            np.random.seed(round_idx)
            client_indexes = np.random.choice(range(client_num_in_total), num_clients, replace=False)
        self.logger.info("client_indexes = %s" % str(client_indexes))
        return client_indexes

    def peek_client_sampling(self, round_idx) -> np.ndarray:
        """The (sorted) clients round ``round_idx`` will sample under the stock rule, computed without side effects (numpy's global
        stream, which the sampler re-seeds, is put back; nothing is logged; the ``_client_sampling`` hook is not called).  Only a
        prefetch hint: a mispredicted upload is evicted by the stager."""
        total, per_round = self.args.client_num_in_total, self.args.client_num_per_round
        if total == per_round:
            return np.arange(total)
        state = np.random.get_state()
        try:
            np.random.seed(round_idx)
            return np.sort(np.random.choice(range(total), min(per_round, total), replace=False))
        finally:
            np.random.set_state(state)

    def local_clients(self, indexes: Sequence[int]) -> List[int]:
        return [int(c) for c in indexes if self.dist.owns(int(c))]

    # -------------------------------------------------------------------- aggregation
    def _aggregate(self, w_locals: Sequence[Tuple[int, ParamSlab]], training_num: Optional[int] = None,
                   mask: Optional[ParamSlab] = None, out: Optional[ParamSlab] = None) -> ParamSlab:
        """``w_global = sum_i (n_i / sum n) * w_i`` over all 39 state entries incl. BN buffers
        (``sailentgrads_api.py:212-227``).  ``w_locals`` are this rank's sampled clients in client
        order; ``training_num`` is the global sample total (defaults to the local one)."""
        if training_num is None:
            training_num = sum(n for n, _ in w_locals)
        out = out if out is not None else self.w_global
        pairs = [(n / training_num, s) for n, s in w_locals]
        with self.timer.phase("aggregate"):
            self.aggregator.aggregate(out, pairs, mask)
        return out

    # --------------------------------------------------------------------- evaluation
    def _metrics_arrays(self, client_ids: Sequence[int], results) -> Tuple[np.ndarray, np.ndarray]:
        n = self.args.client_num_in_total
        acc = np.zeros(n)
        loss = np.zeros(n)
        for cid, m in zip(client_ids, results):
            tot = max(m["test_total"], 1.0)
            acc[cid] = m["test_correct"] / tot
            loss[cid] = m["test_loss"] / tot
        return acc, loss

    def _evaluate_clients(self, slabs: Dict[int, ParamSlab], client_ids: Sequence[int]) -> Tuple[np.ndarray, np.ndarray]:
        """Per-client (accuracy, loss/sample) for the locally owned ``client_ids``; zeros elsewhere."""
        results = self.model_trainer.test_many([(slabs[cid], self.client_list[cid].local_test_data) for cid in client_ids])
        return self._metrics_arrays(client_ids, results)

    def _eval_ids(self) -> List[int]:
        return [0] if getattr(self.args, "ci", 0) == 1 else list(range(self.args.client_num_in_total))

    def _eval_plan(self, ids: Optional[Sequence[int]] = None) -> List[List[Tuple[int, str]]]:
        """Per rank, the ``(client, 'g' | 'p')`` evaluation jobs of a full evaluation (:mod:`..parallel.placement`)."""
        from ..parallel.placement import plan_eval_jobs
        ids = list(ids) if ids is not None else self._eval_ids()
        W = self.dist.world_size
        # Off by default: the global-model half of the jobs can only start after the aggregation, i.e. after the most loaded rank
        # has finished training, so moving jobs to lighter ranks does not shorten the round unless training itself is uneven
        # enough (measured at 8 GPUs, 21 equal clients: 24.0 ms balanced vs 20.6 ms on the owners).  ``--balance_eval 1`` turns it on.
        balance = bool(self.BALANCE_EVAL and self.dist.is_distributed and getattr(self.args, "balance_eval", False))
        return plan_eval_jobs(ids, {c: self.dist.owner(c) for c in ids}, {c: self.client_list[c].local_test_data.num_samples
                                                                          if hasattr(self.client_list[c].local_test_data, "num_samples")
                                                                          else len(self.client_list[c].local_test_data) for c in ids},
                              sorted(self._trained), self._train_load if balance else [0.0] * W, W, balance)

    def eval_loaders(self, ids: Optional[Sequence[int]] = None) -> list:
        """Test loaders this rank's evaluation jobs draw from, in job order without repeats (what :meth:`_test_on_all_clients`
        stages as one group)."""
        mine = self._eval_plan(ids)[self.dist.rank]
        seen, out = set(), []
        for c, _ in mine:
            if c not in seen:
                seen.add(c)
                out.append(self.client_list[c].local_test_data)
        return out

    def _note_training(self, client_indexes: Sequence[int]) -> None:
        """Bookkeeping every rank does alike at the start of a round: who will hold a personal model after it, and how much
        training each rank is about to do (the base load the evaluation jobs are balanced on top of)."""
        from ..parallel.placement import train_cost
        self._trained.update(int(c) for c in client_indexes)
        load = [0.0] * self.dist.world_size
        for c in client_indexes:
            load[self.dist.owner(int(c))] += train_cost(self.client_list[int(c)].get_sample_number(), getattr(self.args, "epochs", 1))
        self._train_load = load

    def _sync_trained(self) -> None:
        """Rebuild the every-rank-alike set of clients that hold a personal model from what the owners actually hold (after a resume)."""
        n = self.args.client_num_in_total
        t = torch.zeros(n, dtype=torch.float32, device=self.device)
        for c in self.client_list:
            if c.slab is not None and self.dist.owns(c.client_idx):
                t[c.client_idx] = 1
        self.dist.all_reduce_sum(t)
        self._trained = set(int(i) for i in torch.nonzero(t > 0).flatten().tolist())

    def _fetch_personal(self, plan, w_per_mdls: Dict[int, ParamSlab]) -> Dict[int, ParamSlab]:
        """Personal models this rank evaluates for other owners: point-to-point copies owner -> evaluator (all ranks take part;
        the plan is the same everywhere).  Returns client -> received slab."""
        import torch.distributed as dist
        ops_, got = [], {}
        k = 0
        for r, jobs in enumerate(plan):
            for c, kind in jobs:
                if kind != "p" or c not in self._trained or self.dist.owner(c) == r:
                    continue
                if self.dist.rank == self.dist.owner(c):
                    ops_.append(dist.P2POp(dist.isend, w_per_mdls[c].flat, r))
                elif self.dist.rank == r:
                    while len(self._eval_tmp) <= k:
                        self._eval_tmp.append(self.model_trainer.engine.new_slab())
                    got[c] = self._eval_tmp[k]
                    k += 1
                    ops_.append(dist.P2POp(dist.irecv, got[c].flat, self.dist.owner(c)))
        if ops_:
            for req in dist.batch_isend_irecv(ops_):
                req.wait()
        return got

    def _reduce_metric(self, *arrays: np.ndarray) -> List[np.ndarray]:
        if not self.dist.is_distributed:
            return list(arrays)
        t = torch.tensor(np.stack(arrays), dtype=torch.float64, device=self.device)
        self.dist.all_reduce_sum(t)
        from ..runtime import xfer
        return [a for a in xfer.to_host(t).numpy()]

    def _test_on_all_clients(self, w_global: ParamSlab, w_per_mdls: Optional[Dict[int, ParamSlab]], round_idx):
        """Global model and personalised models on every client's test split; metrics are
        client-uniform means (``sailentgrads_api.py:231-285``).  ``--ci 1`` evaluates only the first
        client (and, unlike the reference, averages over what was evaluated)."""
        self.logger.info("################global_test_on_all_clients : {}".format(round_idx))
        ids = self._eval_ids()
        with self.timer.phase("eval"):
            # every test split is staged once per rank however many models it is evaluated with (the reference stages it once per
            # model); the jobs of all clients are spread over the ranks by load when BALANCE_EVAL
            plan = self._eval_plan(ids)
            mine = plan[self.dist.rank]
            if w_per_mdls is None:
                mine = [j for j in mine if j[1] == "g"]
            foreign = self._fetch_personal(plan, w_per_mdls) if (w_per_mdls is not None and self.dist.is_distributed) else {}

            def model_of(c, kind):
                if kind == "g":
                    return w_global
                return foreign.get(c) or w_per_mdls.get(c) or self.w_init or w_global
            res = self.model_trainer.test_jobs([(self.client_list[c].local_test_data, model_of(c, kind)) for c, kind in mine],
                                               group=self.eval_loaders(ids) if w_per_mdls is not None else None)
            g_acc, g_loss = self._metrics_arrays([c for c, k in mine if k == "g"], [m for (c, k), m in zip(mine, res) if k == "g"])
            if w_per_mdls is not None:
                p_acc, p_loss = self._metrics_arrays([c for c, k in mine if k == "p"], [m for (c, k), m in zip(mine, res) if k == "p"])
            g_acc, g_loss, p_acc, p_loss = self._reduce_metric(g_acc, g_loss, *((p_acc, p_loss) if w_per_mdls is not None else (g_acc * 0, g_loss * 0)))
            if w_per_mdls is None:
                p_acc, p_loss = g_acc.copy(), g_loss.copy()
        denom = len(ids)
        stats = {'global_test_acc': g_acc.sum() / denom, 'global_test_loss': g_loss.sum() / denom}
        self.stat_info["global_test_acc"].append(stats['global_test_acc'])
        self.stat_info["global_test_loss"].append(stats['global_test_loss'])
        self.logger.info(stats)
        pstats = {'person_test_acc': p_acc.sum() / denom, 'person_test_loss': p_loss.sum() / denom}
        self.stat_info["person_test_acc"].append(pstats['person_test_acc'])
        self.stat_info["person_test_loss"].append(pstats['person_test_loss'])
        self.logger.info(pstats)
        self._emit(round=int(round_idx), **{k: float(v) for k, v in {**stats, **pstats}.items()})
        return stats, pstats

    def _emit(self, **metrics) -> None:
        if self.metrics_sink is not None and self.dist.is_main:
            self.metrics_sink(event="eval", algorithm=self.NAME, **metrics)

    # ------------------------------------------------------------------------- stats
    def _add_comm_params(self, n) -> None:
        if torch.is_tensor(n):
            self._comm_params_dev = n.clone() if self._comm_params_dev is None else self._comm_params_dev + n
        else:
            self.stat_info["sum_comm_params"] += n

    def finalize_stats(self) -> dict:
        if self._comm_params_dev is not None:
            t = self._comm_params_dev.to(torch.float64)
            self.dist.all_reduce_sum(t)
            self.stat_info["sum_comm_params"] += int(t.item())
            self._comm_params_dev = None
        self._sync_flops()
        self.stat_info["phase_time_s"] = self.timer.summary()
        return self.stat_info

    def get_model_sps_for_weight(self, slab: ParamSlab) -> float:
        """Sparsity (%) over *all* state entries of a model (``sailentgrads_api.py:68-83``)."""
        nz = int(ops.count_nonzero(slab.flat).item())
        total = self.layout.numel_real()
        return 100.0 * (total - nz) / total

    # ---- helpers the reference orchestrator carries without calling them (``sailentgrads_api.py:165-209,319-331``)
    def _benefit_choose(self, round_idx, cur_clnt, client_num_in_total, client_num_per_round, dist_local=None, total_dist=None,
                        cs=False, active_ths_rnd=None):
        """Neighbour choice for decentralised variants: everyone, seeded-stream random peers, the ring neighbours or all active clients."""
        from .decentralized import choose_neighbors
        if client_num_in_total == client_num_per_round:
            return [i for i in range(client_num_in_total)]
        return choose_neighbors(cs, cur_clnt, client_num_in_total, client_num_per_round, active_ths_rnd)

    def average_weights(self, weights_list: Sequence) -> ParamSlab:
        """Plain (unweighted) mean of models given as slabs or state_dicts -- the sanity-check aggregate."""
        slabs = []
        for w in weights_list:
            if isinstance(w, ParamSlab):
                slabs.append(w)
            else:
                s = self.model_trainer.engine.new_slab()
                s.load_state_dict(w)
                slabs.append(s)
        out = self.model_trainer.engine.new_slab()
        ops.weighted_aggregate(out.flat, [s.flat for s in slabs], [1.0 / len(slabs)] * len(slabs))
        return out

    def record_avg_inference_flops(self, w_global: ParamSlab, mask_pers: Optional[Dict[int, ParamSlab]] = None) -> float:
        """Mean inference FLOPs over clients of the (optionally per-client masked) global model; zero weights are not counted."""
        flops = []
        for cid in range(self.args.client_num_in_total):
            if mask_pers is None:
                w = w_global
            else:
                w = w_global.clone()
                n = mask_pers[cid].flat.numel()
                w.flat[:n] *= mask_pers[cid].flat
            flops.append(self.model_trainer.count_inference_flops(w.state_dict()))
            if mask_pers is None:
                flops = flops * self.args.client_num_in_total      # identical for every client
                break
        self.stat_info["avg_inference_flops"] = sum(flops) / len(flops)
        return self.stat_info["avg_inference_flops"]

    # ------------------------------------------------------------------- checkpointing
    def personal_slabs(self) -> Dict[int, ParamSlab]:
        return {c.client_idx: c.slab for c in self.client_list if c.slab is not None}

    def _loader_states(self) -> Dict[int, torch.Tensor]:
        """Shuffle-generator state of every client's training loader, from the rank that iterates it (all ranks take part)."""
        mine = {c.client_idx: c.local_training_data.state() for c in self.client_list
                if self.dist.owns(c.client_idx) and hasattr(c.local_training_data, "state")}
        if not self.dist.is_distributed:
            return mine
        import torch.distributed as dist
        parts = [None] * self.dist.world_size
        dist.all_gather_object(parts, mine)
        return {k: v for part in parts for k, v in part.items()}

    def _gather_personal_state(self) -> Dict[int, dict]:
        """Every client's personal model as a CPU state_dict on rank 0, fetched from the rank that owns (= trained) it; all ranks
        take part.  Rank 0's own view is not enough: the clients of other ranks exist there only as untrained clones."""
        self._sync_trained()
        out: Dict[int, dict] = {}
        tmp = None
        for c in sorted(self._trained):
            slab = self.client_list[c].slab
            if self.dist.is_distributed:
                owner = self.dist.owner(c)
                if self.dist.rank != owner:
                    if tmp is None:
                        tmp = self.model_trainer.engine.new_slab()
                    slab = tmp
                self.dist.broadcast(slab.flat, src=owner)
            if self.dist.is_main:
                out[c] = {k: v.detach().cpu().clone() for k, v in slab.state_dict().items()}
        return out

    def _sync_flops(self) -> None:
        """``sum_training_flops`` is accumulated per rank; fold the ranks' shares since the last sync into one global value."""
        if not self.dist.is_distributed:
            return
        base = getattr(self, "_flops_base", 0)
        t = torch.tensor([float(self.stat_info["sum_training_flops"] - base)], dtype=torch.float64, device=self.device)
        self.dist.all_reduce_sum(t)
        delta = t.item()
        self.stat_info["sum_training_flops"] = base + (int(round(delta)) if float(delta).is_integer() else delta)
        self._flops_base = self.stat_info["sum_training_flops"]

    def save(self, path: str, round_idx: int) -> None:
        loaders = self._loader_states()
        keep_personal = self.PERSONAL_IN_CHECKPOINT or getattr(self.args, "save_personal", False)
        personal = self._gather_personal_state() if keep_personal else None
        self._sync_flops()
        if self.dist.is_main:
            extra = dict(self.extra_state())
            if self.w_init is not None:
                extra["w_init"] = {k: v.cpu() for k, v in self.w_init.state_dict().items()}
            save_checkpoint(path, round_idx, self.w_global, self.mask, self.args, self.stat_info,
                            personal=personal, loaders=loaders, extra=extra)

    def extra_state(self) -> dict:
        """Algorithm-specific state a resumed run needs beyond the global model, the mask and the personal models (CPU tensors /
        plain python).  Counterpart: :meth:`load_extra_state`."""
        return {}

    def load_extra_state(self, state: dict) -> None:
        self._resumed_extra = state

    def resume(self, path: str) -> int:
        """Continue after the checkpointed round.  On one rank the continuation is bit-identical to an uninterrupted run: the
        loaders' shuffle streams and the python / numpy / torch RNG states come back too (with several ranks only rank 0's global
        RNG state is in the file, so it is not restored; the loader streams are)."""
        ck = load_checkpoint(path, map_location=self.device, restore_rng=not self.dist.is_distributed)
        for cid, st in (ck.get("loaders") or {}).items():
            loader = self.client_list[int(cid)].local_training_data
            if hasattr(loader, "set_state"):
                loader.set_state(st)
        self.w_global.load_state_dict(ck["state_dict"])
        if ck.get("mask") is not None:
            self.mask = self.model_trainer.mask_slab(ck["mask"])
        for k, v in ck.get("stat_info", {}).items():
            self.stat_info[k] = v
        self._flops_base = self.stat_info.get("sum_training_flops", 0)
        for cid, sd in (ck.get("personal") or {}).items():
            c = self.client_list[int(cid)]
            c.ensure_slab(self.w_global).load_state_dict(sd)
        extra = dict(ck.get("extra") or {})
        if "w_init" in extra:
            self.w_init = self.model_trainer.engine.new_slab()
            self.w_init.load_state_dict(extra.pop("w_init"))
        self.load_extra_state(extra)
        self._sync_trained()
        self.start_round = int(ck["round"]) + 1
        self.logger.info("resumed from %s at round %d" % (path, self.start_round))
        return self.start_round

    def _maybe_checkpoint(self, round_idx: int) -> None:
        every = int(getattr(self.args, "checkpoint_every", 0) or 0)
        path = getattr(self.args, "checkpoint_path", None)
        if path and every > 0 and (round_idx + 1) % every == 0:
            self.save(path, round_idx)
