"""FedAvg (reference: ``fedml_api/standalone/fedavg/fedavg_api.py:12-196``)."""
from __future__ import annotations

import contextlib
import time

import numpy as np
import torch

from .base import FedAPIBase


class FedAvgAPI(FedAPIBase):
    NAME = "fedavg"
    BALANCE_EVAL = True

    def run_round(self, round_idx: int, mask=None, evaluate: bool = True):
        """sample -> local training of the owned sampled clients -> weighted aggregation -> eval."""
        args = self.args
        self.logger.info("################Communication round : {}".format(round_idx))
        client_indexes = np.sort(self._client_sampling(round_idx, args.client_num_in_total, args.client_num_per_round))
        self.logger.info("client_indexes = " + str(client_indexes))
        training_num = sum(self.client_list[int(c)].get_sample_number() for c in client_indexes)
        w_locals = []
        will_eval = evaluate and round_idx % max(1, int(getattr(args, "frequency_of_the_test", 1))) == 0
        tr = self.model_trainer
        self._note_training(client_indexes)
        with self.timer.phase("train"):
            mine = list(self.local_clients(client_indexes))
            # the whole round's data plan is known up front: this round's clients in order, then every test split (one group, staged
            # once for both evaluations), then the first clients of the next round -- uploads run behind the compute stream
            n_lanes = len(tr.all_lanes()) if (hasattr(tr, "all_lanes") and len(mine) > 1) else 1
            for k, lane in enumerate(tr.all_lanes()[:n_lanes] if n_lanes > 1 else [tr]):      # lane k trains clients mine[k::n_lanes]
                lane.plan_data([self.client_list[c].local_training_data for c in mine[k::n_lanes]])
            if will_eval:
                tr.plan_eval(self.eval_loaders()) if hasattr(tr, "plan_eval") else tr.plan_data(self.eval_loaders(), kind="eval", group=True)
            if round_idx + 1 < int(getattr(args, "comm_round", 0)):
                nxt_round = self.peek_client_sampling(round_idx + 1)
                nxt_mine = self.local_clients(nxt_round)
                n_nxt = n_lanes if len(nxt_mine) > 1 else 1
                for k, lane in enumerate(tr.all_lanes()[:n_nxt] if n_nxt > 1 else [tr]):
                    lane.plan_data([self.client_list[c].local_training_data for c in nxt_mine[k::n_nxt][:2]])
            if self._comm_params_dev is None and self.device.type == "cuda":
                self._comm_params_dev = torch.zeros((), dtype=torch.int64, device=self.device)
            lanes = tr.all_lanes() if hasattr(tr, "all_lanes") else [tr]
            if len(lanes) > 1 and len(mine) > 1:
                results = self._train_on_lanes(lanes, mine, round_idx, mask)
            else:
                results = {}
                for pos, cur_clnt in enumerate(mine):
                    self.logger.info('@@@@@@@@@@@@@@@@ Training Client CM({}): {}'.format(round_idx, cur_clnt))
                    # stream staging: the next client's first batch is copied while this one finishes
                    nxt = self.client_list[mine[pos + 1]].local_training_data if pos + 1 < len(mine) else None
                    results[cur_clnt] = self.client_list[cur_clnt].train(self.w_global, round_idx, mask, lookahead_data=nxt,
                                                                         comm_out=self._comm_params_dev)
            for cur_clnt in mine:          # client order: the aggregation adds in this order
                client = self.client_list[cur_clnt]
                slab, training_flops, num_comm_params = results[cur_clnt]
                self._defend(slab)
                w_locals.append((client.get_sample_number(), slab))
                self.stat_info["sum_training_flops"] += training_flops
                if num_comm_params is not self._comm_params_dev:
                    self._add_comm_params(num_comm_params)
        # the shared mask rides along in the aggregation kernel (a no-op on already-masked client weights, as in the reference
        # where the multiply at sailentgrads_api.py:109 is commented out; it keeps the global model exactly sparse)
        self._aggregate(w_locals, training_num, mask=mask)
        self.model_trainer.flush_losses()
        if will_eval:
            self._test_on_all_clients(self.w_global, self.personal_slabs(), round_idx)
        self._maybe_checkpoint(round_idx)

    def _train_on_lanes(self, lanes, mine, round_idx, mask) -> dict:
        """Concurrent client slots: lane k (its own engine and CUDA stream) trains clients ``mine[k::L]`` one after another, and the
        host interleaves the lanes step by step, so the GPU always holds kernels of ``L`` clients.  Every client's result is what the
        sequential schedule gives (its steps are the same kernels on its own slab; the dropout stream is tied to (round, client))."""
        dev = self.device
        main = torch.cuda.current_stream(dev) if dev.type == "cuda" else None
        L = len(lanes)
        queues = [list(mine[k::L]) for k in range(L)]
        gens, cur = [None] * L, [None] * L
        results = {}
        for lane in lanes:
            if lane.stream is not None:
                lane.stream.wait_stream(main)              # the global model / mask written on the main stream
        live = L
        while live:
            live = 0
            for k, lane in enumerate(lanes):
                if gens[k] is None and queues[k]:
                    cur[k] = queues[k].pop(0)
                    self.logger.info('@@@@@@@@@@@@@@@@ Training Client CM({}): {}'.format(round_idx, cur[k]))
                    gens[k] = self.client_list[cur[k]].train_steps(self.w_global, round_idx, mask, trainer=lane, comm_out=self._comm_params_dev)
                if gens[k] is None:
                    continue
                live += 1
                ctx = torch.cuda.stream(lane.stream) if lane.stream is not None else contextlib.nullcontext()
                with ctx:
                    try:
                        next(gens[k])
                    except StopIteration as stop:
                        results[cur[k]] = stop.value
                        gens[k] = None
        for lane in lanes:
            if lane.stream is not None:
                main.wait_stream(lane.stream)
        return results

    def _defend(self, slab) -> None:
        """Optional robust-aggregation defence on a client's upload (``--defense_type``; the reference ships the routines without a
        caller, ``fedml_core/robustness/robust_aggregation.py:32-55``): clip the update's norm to ``--norm_bound`` around the global
        model, and for ``weak_dp`` also add N(0, ``--stddev``^2) noise to the parameters."""
        kind = getattr(self.args, "defense_type", None)
        if not kind or kind == "none":
            return
        if getattr(self, "_robust", None) is None:
            from ..utils.robust import RobustAggregator
            self._robust = RobustAggregator(self.args)
        self._robust.norm_diff_clipping(slab, self.w_global)
        if kind == "weak_dp":
            P = slab.layout.param_numel
            slab.flat[:P] = self._robust.add_noise(slab.flat[:P])
        elif kind != "norm_diff_clipping":
            raise ValueError(f"unknown defense_type {kind!r} (none | norm_diff_clipping | weak_dp)")

    def train(self):
        args = self.args
        if self.w_init is None:
            self.w_init = self.w_global.clone()
        for round_idx in range(self.start_round, args.comm_round):
            t0 = time.time()
            self.run_round(round_idx)
            self.stat_info["round_time_s"].append(time.time() - t0)
        if getattr(args, "final_finetune", True):
            self.finetune_round()
        return self.finalize_stats()

    def finetune_round(self):
        """Final round: *every* client trains once from the global model with ``round=-1``
        (lr * decay^-1) and the personalised accuracy is re-evaluated (``fedavg_api.py:79-88``)."""
        self.logger.info("################ final fine-tune round")
        self._note_training(range(self.args.client_num_in_total))
        with self.timer.phase("train"):
            for cid in self.local_clients(range(self.args.client_num_in_total)):
                self.client_list[cid].train(self.w_global, -1, None)
        self._test_on_all_clients(self.w_global, self.personal_slabs(), -1)
