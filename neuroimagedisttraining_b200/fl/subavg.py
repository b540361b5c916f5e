"""Sub-FedAvg: iterative magnitude pruning per client with mask-count-weighted aggregation
(reference: ``fedml_api/standalone/subavg/subavg_api.py:43-139``, ``subavg/client.py:36-63``,
``subavg/my_model_trainer.py:30-82``, ``subavg/prune_func.py:9-90``)."""
from __future__ import annotations

import time
from typing import Tuple

import numpy as np
import torch

from .. import ops
from ..runtime.slab import ParamSlab, make_mask_slab
from .local import LocalAPI


def fake_prune(each_prune_ratio: float, weights: ParamSlab, mask: ParamSlab) -> ParamSlab:
    """New mask with the smallest ``each_prune_ratio`` of the *alive* weights of every prunable tensor removed
    (percentile with linear interpolation, as ``np.percentile``; biases / norm layers are never pruned)."""
    new = mask.clone()
    for name in weights.layout.param_names:
        e = weights.layout[name]
        if "weight" not in name or "bn" in name or len(e.shape) < 2:
            continue
        w, m = weights.view(name), mask.view(name)
        alive = (w * m)
        alive = alive[alive != 0].abs()
        if alive.numel() == 0:
            continue
        srt = torch.sort(alive.flatten())[0]
        pos = each_prune_ratio * (srt.numel() - 1)
        lo, hi = int(np.floor(pos)), int(np.ceil(pos))
        thr = srt[lo] + (srt[hi] - srt[lo]) * (pos - lo)
        new.view(name).copy_(torch.where(w.abs() < thr, torch.zeros_like(m), m))
    return new


def real_prune(weights: ParamSlab, mask: ParamSlab) -> ParamSlab:
    """Apply a mask to the trainable region (in place)."""
    weights.params_flat.mul_(mask.flat)
    return weights


def dist_masks(m1: ParamSlab, m2: ParamSlab) -> float:
    """Mean over tensors of the normalised Hamming distance."""
    d = []
    for name in m1.layout.param_names:
        a, b = m1.view(name), m2.view(name)
        d.append(float(((a != 0) ^ (b != 0)).float().mean()))
    return float(np.mean(d))


def print_pruning(weights: ParamSlab) -> Tuple[float, int]:
    nz = int(ops.count_nonzero(weights.flat).item())
    return nz / weights.layout.numel_real(), nz


class SubAvgAPI(LocalAPI):
    NAME = "subavg"

    def extra_state(self) -> dict:
        return {"subavg_masks": [m.flat.detach().cpu().clone() for m in self.mask_pers]} if hasattr(self, "mask_pers") else {}

    def _client_train(self, client, mask: ParamSlab, round_idx: int):
        args, tr = self.args, self.model_trainer
        slab = client.ensure_slab(self.w_global)
        slab.copy_(self.w_global)
        comm = tr.count_communication_params(self.w_global)
        real_prune(slab, mask)
        dense, _ = print_pruning(slab)
        tr.attach(slab)
        tr.set_id(client.client_idx)
        # gradient masking (not weight masking); m1 after the first epoch, m2 after the last
        tr.train(client.local_training_data, self.device, args, round_idx, mask, epochs=1, grad_mask=True)
        m1 = fake_prune(args.each_prune_ratio, slab, mask)
        if args.epochs > 1:
            tr.train(client.local_training_data, self.device, args, round_idx, mask, epochs=args.epochs - 1, grad_mask=True)
        m2 = fake_prune(args.each_prune_ratio, slab, mask)
        tr.engine.detach_sync()
        final = mask
        if dist_masks(m1, m2) > args.dist_thresh and dense > args.dense_ratio:
            trial = slab.clone()
            real_prune(trial, m2)
            tr.attach(trial)
            m = tr.test(client.local_training_data, self.device, args)
            if m["test_correct"] / max(m["test_total"], 1) > args.acc_thresh:
                slab.copy_(trial)
                final = m2
        comm = comm + tr.count_communication_params(slab)
        return final, slab, args.epochs * client.get_sample_number(), comm

    def _aggregate_masked(self, w_locals):
        """Per element: mean over the clients whose mask keeps it; elements nobody keeps retain the server value."""
        P = self.layout.param_numel
        cnt = torch.zeros(P, device=self.device)
        acc = torch.zeros_like(self.w_global.flat)
        for mask, slab in w_locals:
            cnt += mask.flat
            acc += slab.flat
        n_models = torch.tensor([float(len(w_locals))], device=self.device)
        with self.timer.phase("aggregate"):
            if self.dist.is_distributed:                 # sums of the ranks' sampled clients
                for t in (cnt, acc, n_models):
                    self.dist.all_reduce_sum(t)
            keep = cnt > 0
            self.w_global.flat[:P] = torch.where(keep, acc[:P] / cnt.clamp(min=1.0), self.w_global.flat[:P])
            self.w_global.flat[P:] = acc[P:] / n_models.clamp(min=1.0)
        return self.w_global

    def train(self):
        args = self.args
        n = args.client_num_in_total
        if self.w_init is None:
            self.w_init = self.w_global.clone()
        mask_pers = [make_mask_slab(self.layout, self.device) for _ in range(n)]
        if self._resumed_extra.get("subavg_masks") is not None:
            for m, f in zip(mask_pers, self._resumed_extra["subavg_masks"]):
                m.flat.copy_(f.to(m.flat.device))
        self.mask_pers = mask_pers
        for round_idx in range(self.start_round, args.comm_round):
            t0 = time.time()
            self.logger.info("################Communication round : {}".format(round_idx))
            self.record_mask_diffrence(mask_pers)
            client_indexes = self._client_sampling(round_idx, n, args.client_num_per_round)
            w_locals, next_masks = [], []
            with self.timer.phase("train"):
                for cid in self.local_clients(client_indexes):
                    new_mask, slab, flops, comm = self._client_train(self.client_list[cid], mask_pers[cid], round_idx)
                    w_locals.append((mask_pers[cid].clone(), slab))
                    next_masks.append((cid, new_mask))
                    self.stat_info["sum_training_flops"] += flops
                    self._add_comm_params(comm)
            self._aggregate_masked(w_locals)
            if round_idx == args.comm_round - 1 or round_idx % max(1, int(getattr(args, "frequency_of_the_test", 1))) == 0:
                self._local_test_masked(mask_pers, round_idx)
            for cid, m in next_masks:
                mask_pers[cid] = m
            if self.dist.is_distributed:                 # every rank keeps all masks (mask statistics, checkpoints): owners publish
                for cid in client_indexes:
                    self.dist.broadcast(mask_pers[int(cid)].flat, src=self.dist.owner(int(cid)))
            self.stat_info["round_time_s"].append(time.time() - t0)
            self._maybe_checkpoint(round_idx)
        self.record_information()
        return self.finalize_stats()

    def record_mask_diffrence(self, mask_pers):
        n = len(mask_pers)
        if n < 2:
            return
        d = [int(ops.hamming_distance(mask_pers[0].flat, mask_pers[j].flat)) for j in range(1, min(n, 5))]
        self.logger.info("mask difference to client 0: " + str(d))

    def _local_test_masked(self, mask_pers, round_idx):
        n = len(self.client_list)
        accs, losses = np.zeros(n), np.zeros(n)
        with self.timer.phase("eval"):
            for cid in self.local_clients(range(n)):
                w = self.w_global.clone()
                real_prune(w, mask_pers[cid])
                m = self.client_list[cid].local_test(w, True)
                accs[cid] = m["test_correct"] / max(m["test_total"], 1)
                losses[cid] = m["test_loss"] / max(m["test_total"], 1)
            accs, losses = self._reduce_metric(accs, losses)
        stats = {'person_test_acc': float(np.mean(accs)), 'person_test_loss': float(np.mean(losses))}
        self.stat_info["person_test_acc"].append(stats["person_test_acc"])
        self.stat_info.setdefault("person_test_loss", []).append(stats["person_test_loss"])
        self.stat_info.setdefault("test_acc", []).append(stats["person_test_acc"])        # the reference's key (subavg_api.py)
        self.eval_masks = [m.clone() for m in mask_pers]                                    # masks the evaluation used
        self.logger.info(stats)
        self._emit(round=int(round_idx), **stats)
        return stats
