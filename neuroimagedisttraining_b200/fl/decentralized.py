"""Decentralised algorithms: D-PSGD, DisPFL and FedFomo share neighbour selection and per-client models.

* :class:`DPSGDAPI`  -- each client averages its neighbours' models (ring / random / full), then trains; the
  reported global model is the plain mean (``fedml_api/standalone/dpsgd/dpsgd_api.py:41-103,116-178``).
* :class:`DisPFLAPI` -- personalised sparse models: ERK/uniform random masks, magnitude prune + gradient/random
  regrow each round with a cosine-annealed rate, client drop-out (``--active``), mask Hamming statistics
  (``fedml_api/standalone/DisPFL/dispfl_api.py:46-240``, ``DisPFL/client.py:32-99``).  The reference has the neighbour
  aggregation commented out (``dispfl_api.py:138-142``); ``args.neighbor_aggregation`` turns the intended
  mask-count-weighted aggregation (``_aggregate_func``, ``:222-240``) on.
* :class:`FedFomoAPI` -- first-order model optimisation: neighbour weights proportional to the validation-loss gain
  per unit of parameter distance (``fedml_api/standalone/fedfomo/fedfomo_api.py:53-217``).
"""
from __future__ import annotations

import math
import random
import time
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .. import ops
from ..runtime.slab import ParamSlab
from ..utils.slim import cosine_annealing
from .local import LocalAPI


def choose_neighbors(cs: str, cur: int, total: int, per_round: int, active: Optional[np.ndarray] = None,
                     seed: Optional[int] = None) -> np.ndarray:
    """``_benefit_choose`` (``dpsgd_api.py:116-139``): everyone when ``per_round == total``; else random peers
    (never ``cur``), the two ring neighbours, or all active clients.  ``seed``: D-PSGD re-seeds numpy with
    ``round + client`` before drawing (``dpsgd_api.py:122-123``); DisPFL draws from the running stream (``dispfl_api.py:203-208``)."""
    if total == per_round:
        return np.arange(total)
    if cs == "random":
        k = min(per_round, total)
        if seed is not None:
            np.random.seed(seed)
        idx = np.random.choice(range(total), k, replace=False)
        while cur in idx:
            idx = np.random.choice(range(total), k, replace=False)
        return idx
    if cs == "ring":
        return np.asarray([(cur - 1 + total) % total, (cur + 1) % total])
    if cs == "full":
        act = np.arange(total) if active is None else np.where(np.asarray(active) == 1)[0]
        return act[act != cur]
    raise ValueError(f"unknown neighbour selection {cs!r}")


class _PerClientModels(LocalAPI):
    """Base for algorithms that keep one model per client and mix them.  Decentralised mixing reads other clients'
    models, so these run single-rank (the reference simulates them on one GPU as well)."""

    MULTI_RANK = False      # DisPFL sets this: its clients train independently and the owners publish their results once per round

    def _init_models(self) -> List[ParamSlab]:
        if self.dist.is_distributed and not self.MULTI_RANK:
            raise NotImplementedError(f"{self.NAME} mixes arbitrary clients' models and runs on one rank")
        if self.w_init is None:
            self.w_init = self.w_global.clone()
        for c in self.client_list:
            if c.slab is None:                      # a resumed run brings the personal models along
                c.slab = self.w_global.clone()
        return [c.slab for c in self.client_list]

    def _slabs_to_state(self, slabs: Sequence[ParamSlab]) -> list:
        return [s.flat.detach().cpu().clone() for s in slabs]

    def _slabs_from_state(self, flats, like: ParamSlab) -> List[ParamSlab]:
        out = []
        for f in flats:
            s = like.zeros_like()
            s.flat.copy_(f.to(s.flat.device))
            out.append(s)
        return out

    def _mean_of(self, slabs: Sequence[ParamSlab], out: ParamSlab) -> ParamSlab:
        w = 1.0 / len(slabs)
        ops.weighted_aggregate(out.flat, [s.flat for s in slabs], [w] * len(slabs))
        return out

    def _test_global_and_personal(self, round_idx):
        return self._test_on_all_clients(self.w_global, self.personal_slabs(), round_idx)


class DPSGDAPI(_PerClientModels):
    """Several ranks: a client's neighbour average and local training run on its owner (the neighbour draw is re-seeded per round
    and client, so ranks agree on it).  On GPUs the owners keep their clients' models in a peer-mapped arena and the mixing kernel
    pulls just the neighbours it needs over NVLink (:mod:`..parallel.neighbors`, W11); on CPU / gloo every rank keeps all models and
    the owners broadcast theirs at the end of the round."""
    NAME = "dpsgd"
    MULTI_RANK = True

    def train(self):
        args = self.args
        models = self._init_models()
        n = args.client_num_in_total
        mixed = self.w_global.clone()
        topology = self._mixing_matrix(n)
        # W11: on several GPUs the owners keep their clients' models in a peer-mapped arena and the mixing kernel pulls the
        # neighbours it needs over NVLink; otherwise (one rank, CPU / gloo) every rank holds all models and owners broadcast
        from ..parallel.neighbors import make_arena
        arena = make_arena(self.dist, self.layout.total_numel, n, self.device) if getattr(args, "peer_arena", True) else None
        self.arena = arena
        last = [m.clone() for m in models] if arena is None else None
        if arena is not None:
            for cid in self.local_clients(range(n)):
                arena.publish(cid, models[cid].flat, (self.start_round - 1) & 1)
            arena.align()

        def mix(out, ids, weights, round_idx):
            if arena is not None:
                arena.average(out.flat, [int(j) for j in ids], weights, (round_idx - 1) & 1)
            else:
                ops.weighted_aggregate(out.flat, [last[int(j)].flat for j in ids], weights)
        for round_idx in range(self.start_round, args.comm_round):
            t0 = time.time()
            self.logger.info("################Communication round : {}".format(round_idx))
            if arena is None:
                for i in range(n):
                    last[i].copy_(models[i])
            with self.timer.phase("train"):
                for cid in self.local_clients(range(n)):
                    self.logger.info('@@@@@@@@@@@@@@@@ Training Client CM({}): {}'.format(round_idx, cid))
                    if topology is not None:                                   # row of a mixing matrix instead of a plain mean
                        row = topology[cid]
                        nei = np.nonzero(row)[0]
                        self.logger.info("client_indexes = " + str(nei))
                        mix(mixed, nei, [float(row[j]) for j in nei], round_idx)
                    else:
                        nei = choose_neighbors(getattr(args, "cs", "ring"), cid, n, args.client_num_per_round, seed=round_idx + cid)
                        if n != args.client_num_per_round:
                            nei = np.append(nei, cid)
                        nei = np.sort(nei)
                        self.logger.info("client_indexes = " + str(nei))
                        mix(mixed, nei, [1.0 / len(nei)] * len(nei), round_idx)      # neighbour averaging (K: _aggregate_func)
                    _, flops, comm = self.client_list[cid].train(mixed, round_idx, None)
                    self.stat_info["sum_training_flops"] += flops
                    self._add_comm_params(comm)
                    if arena is not None:
                        arena.publish(cid, models[cid].flat, round_idx & 1)
            if arena is not None:
                arena.align()
                arena.average(self.w_global.flat, list(range(n)), [1.0 / n] * n, round_idx & 1)      # _avg_aggregate, in client order
            else:
                if self.dist.is_distributed:
                    for cid in range(n):
                        self.dist.broadcast(models[cid].flat, src=self.dist.owner(cid))
                self._mean_of(models, self.w_global)                           # _avg_aggregate
            self._test_global_and_personal(round_idx)
            if round_idx % 100 == 99:
                self._finetune(round_idx)
            self.stat_info["round_time_s"].append(time.time() - t0)
            self._maybe_checkpoint(round_idx)
        return self.finalize_stats()

    def _mixing_matrix(self, n: int):
        """``--cs topology_sym`` / ``topology_asym``: gossip with the row-stochastic mixing matrix of the FedML topology managers
        (ring plus ``client_num_per_round`` extra links; ``fedml_core/distributed/topology/*.py``, which the reference never wires
        up).  None for the reference's own ring / random / full neighbour rules."""
        cs = str(getattr(self.args, "cs", "ring"))
        if not cs.startswith("topology"):
            return None
        from ..utils.topology import AsymmetricTopologyManager, SymmetricTopologyManager
        k = max(2, int(self.args.client_num_per_round))
        k -= k % 2                                                             # ring lattices need an even neighbour count
        if cs == "topology_asym":
            tm = AsymmetricTopologyManager(n, k, max(1, k // 2), seed=int(getattr(self.args, "seed", 0)))
        else:
            tm = SymmetricTopologyManager(n, k)
        tm.generate_topology()
        return np.asarray(tm.topology, dtype=np.float64)

    def _finetune(self, round_idx):
        """Every 100 rounds: one extra training pass from the averaged model, evaluated but not kept."""
        self.logger.info("################Communication round Fine Tune Round After CM({})".format(round_idx))
        mine = [c for c in self.client_list if self.dist.owns(c.client_idx)]
        keep = {c.client_idx: c.slab.clone() for c in mine}
        for c in mine:
            c.train(self.w_global, -1, None)
        self._test_on_all_clients(self.w_global, self.personal_slabs(), -1)
        for c in mine:
            c.slab.copy_(keep[c.client_idx])


class DisPFLAPI(_PerClientModels):
    """Several ranks: every rank keeps all clients' models and masks (tens of MB) and replays the cheap bookkeeping of all clients
    -- so the numpy streams that pick active clients and neighbours stay aligned with the single-process run -- but trains,
    screens and prunes only the clients it owns; at the end of a round every owner broadcasts its clients' new model and mask.
    Identical to one process when masks regrow by gradient and the model has no dropout (those draw from torch's per-rank stream)."""
    NAME = "dispfl"
    MULTI_RANK = True

    def extra_state(self) -> dict:
        if not hasattr(self, "masks_per_client"):
            return {}
        return {"dispfl_masks": self._slabs_to_state(self.masks_per_client), "dispfl_shared": self._slabs_to_state(self._shared_masks),
                "dispfl_dist": self._dist_locals.tolist(), "dispfl_w_spa": list(self._w_spa)}

    # ------------------------------------------------------------------ mask evolution
    def fire_mask(self, mask: ParamSlab, weights: ParamSlab, round_idx: int):
        """Magnitude prune per tensor with the cosine-annealed drop ratio (``DisPFL/client.py:71-82``)."""
        drop = cosine_annealing(self.args, round_idx)
        new, removed = mask.clone(), {}
        if mask.flat.is_cuda:
            # one host sync for all per-tensor counts, then one radix-select kernel per tensor (csrc/mask_select.cu)
            names = list(self.layout.param_names)
            nnz = torch.stack([mask.view(n).sum() for n in names]).tolist()
            for name, cnt in zip(names, nnz):
                k = math.ceil(drop * int(cnt))
                ops.mask_select_k(weights.view(name), new.view(name), k, "fire")
                removed[name] = k
            return new, removed
        for name in self.layout.param_names:
            nm, k = ops.reference.fire_mask(weights.view(name), mask.view(name), drop)
            new.view(name).copy_(nm)
            removed[name] = k
        return new, removed

    def regrow_mask(self, mask: ParamSlab, removed: Dict[str, int], grad: Optional[ParamSlab]):
        new = mask.clone()
        for name in self.layout.param_names:
            g = grad.view(name) if grad is not None else None
            if g is not None and new.flat.is_cuda:
                ops.mask_select_k(g, new.view(name), removed[name], "regrow")
            else:
                new.view(name).copy_(ops.reference.regrow_mask(mask.view(name), removed[name], g))
        return new

    def _aggregate_func(self, cid: int, nei: Sequence[int], last: List[ParamSlab], masks: List[ParamSlab], shared: List[ParamSlab]):
        """Mask-count-weighted neighbour average over the parameters, then the client's own mask
        (``dispfl_api.py:222-240``).  Returns (masked personal start, dense consensus)."""
        P = self.layout.param_numel
        dense = last[cid].clone()
        cnt = torch.zeros(P, device=self.device)
        acc = torch.zeros(P, device=self.device)
        for j in nei:
            cnt += shared[int(j)].flat
            acc += last[int(j)].flat[:P]
        dense.flat[:P] = torch.where(cnt > 0, acc / cnt.clamp(min=1.0), torch.zeros_like(acc))
        start = dense.clone()
        start.flat[:P] *= masks[cid].flat
        return start, dense

    def train(self):
        args = self.args
        tr = self.model_trainer
        n = args.client_num_in_total
        models = self._init_models()
        names = self.layout.param_names
        # initial masks: one shared random mask, or one per client, optionally with different sparsities
        w_spa = [args.dense_ratio] * n

        def make(sparse):
            sp = tr.calculate_sparsities(names, distribution="uniform" if getattr(args, "uniform", False) else "ERK", sparse=sparse)
            return tr.init_masks(names, sp)
        P = self.layout.param_numel
        if self._resumed_extra.get("dispfl_masks") is not None:     # continue with the evolved masks (no RNG draw for fresh ones)
            st = self._resumed_extra
            like = ParamSlab(self.layout, device=self.device, params_only=True)
            masks = self._slabs_from_state(st["dispfl_masks"], like)
            shared = self._slabs_from_state(st["dispfl_shared"], like)
            dist_locals, w_spa = np.asarray(st["dispfl_dist"]), list(st["dispfl_w_spa"])
        else:
            if not getattr(args, "different_initial", False):
                base = make(args.dense_ratio)
                masks = [base.clone() for _ in range(n)]
            elif not getattr(args, "diff_spa", False):
                masks = [make(args.dense_ratio) for _ in range(n)]
            else:
                levels = [0.2, 0.4, 0.6, 0.8, 1.0]
                masks = []
                for i in range(n):
                    masks.append(make(levels[i % 5]))
                    w_spa[i] = levels[i % 5]
            shared = [m.clone() for m in masks]
            dist_locals = np.zeros((n, n))
            for i in range(n):          # (not on resume: a model keeps its just-pruned weights until its next SGD step)
                models[i].flat[:P] *= masks[i].flat
        self.masks_per_client, self._shared_masks, self._dist_locals, self._w_spa = masks, shared, dist_locals, w_spa
        for round_idx in range(self.start_round, args.comm_round):
            t0 = time.time()
            self.logger.info("################Communication round : {}".format(round_idx))
            active = np.random.choice([0, 1], size=n, p=[1.0 - args.active, args.active])
            last = [m.clone() for m in models]
            shared_last = [m.clone() for m in shared]
            before, after = [], []
            with self.timer.phase("train"):
                for cid in range(n):
                    if active[cid] == 0:
                        self.logger.info('@@@@@@@@@@@@@@@@ Client Drop this round CM({}) with spasity {}: {}'.format(round_idx, w_spa[cid], cid))
                    self.logger.info('@@@@@@@@@@@@@@@@ Training Client CM({}) with spasity {}: {}'.format(round_idx, w_spa[cid], cid))
                    dist_locals[cid][cid] = float(ops.hamming_distance(shared_last[cid].flat, masks[cid].flat))
                    self.logger.info("local mask changes: {} / {}".format(dist_locals[cid][cid], self.layout.numel_real(params_only=True)))
                    nei = np.array([], dtype=int) if active[cid] == 0 else choose_neighbors("random" if n != args.client_num_per_round else "full",
                                                                                            cid, n, args.client_num_per_round, active)
                    if n != args.client_num_per_round:
                        nei = np.append(nei, cid)
                    nei = np.sort(nei).astype(int)
                    for j in nei:
                        if j != cid:
                            dist_locals[cid][j] = float(ops.hamming_distance(masks[cid].flat, shared_last[j].flat))
                    self.logger.info("choose client_indexes: {}, accoring to {}".format(str(nei), args.cs))
                    if getattr(args, "neighbor_aggregation", False) and active[cid] == 1:
                        start, _ = self._aggregate_func(cid, nei, last, masks, shared_last)
                    else:
                        start = last[cid]
                    shared[cid].copy_(masks[cid])
                    client = self.client_list[cid]
                    if not self.dist.owns(cid):
                        continue                                  # its owner trains it; the result arrives with the round's broadcast
                    before.append((cid, client.local_test(start, True)))
                    # local training with the personal mask applied after every step
                    _, flops, comm = client.train(start, round_idx, masks[cid], apply_mask=True)
                    m_after = client.local_test(client.slab, True)
                    after.append((cid, m_after))
                    self.logger.info("test acc on this client after {} / {} : {:.2f}".format(m_after['test_correct'], m_after['test_total'], m_after['test_acc']))
                    if not getattr(args, "static", False):
                        grad = None
                        if not getattr(args, "dis_gradient_check", False):
                            tr.attach(client.slab)
                            grad = tr.screen_gradients(client.local_training_data, self.device)
                        fired, removed = self.fire_mask(masks[cid], client.slab, round_idx)
                        masks[cid] = self.regrow_mask(fired, removed, grad)
                    self.stat_info["sum_training_flops"] += flops
                    self._add_comm_params(comm)
            if self.dist.is_distributed:                        # owners publish what they trained / pruned this round
                for cid in range(n):
                    self.dist.broadcast(models[cid].flat, src=self.dist.owner(cid))
                    self.dist.broadcast(masks[cid].flat, src=self.dist.owner(cid))
            self._log_round(after, round_idx, "old_mask_test_acc")
            self._log_round(before, round_idx, "new_mask_test_acc")
            self.stat_info["round_time_s"].append(time.time() - t0)
            self._maybe_checkpoint(round_idx)
        for i in range(n):
            self.stat_info["mask_dis_matrix"].append([int(ops.hamming_distance(masks[i].flat, masks[j].flat)) for j in range(n)])
        if getattr(args, "save_masks", False):
            self.stat_info["final_masks"] = [{name: m.view(name).bool().cpu() for name in names} for m in masks]
        return self.finalize_stats()

    def _log_round(self, results, round_idx, key):
        """``results``: (client, metrics) pairs of the clients this rank evaluated; means are over all clients of all ranks."""
        n = self.args.client_num_in_total
        acc, loss, seen = np.zeros(n), np.zeros(n), np.zeros(n)
        for cid, r in results:
            acc[cid] = r["test_correct"] / max(r["test_total"], 1)
            loss[cid] = r["test_loss"] / max(r["test_total"], 1)
            seen[cid] = 1.0
        acc, loss, seen = self._reduce_metric(acc, loss, seen)
        denom = max(float(seen.sum()), 1.0)
        acc, loss = float(acc.sum() / denom), float(loss.sum() / denom)
        self.logger.info({'test_acc': acc, 'test_loss': loss})
        self._emit(round=int(round_idx), series=key, test_acc=acc, test_loss=loss)
        self.stat_info[key].append(acc)
        if key == "old_mask_test_acc":
            self.stat_info["person_test_acc"].append(acc)


class FedFomoAPI(_PerClientModels):
    """Several ranks: all ranks replay every client's neighbour draw (its RNG consumption does not depend on per-client state), the
    owner trains, weighs and mixes, then broadcasts the client's model and its rows of the weight / preference matrices."""
    NAME = "fedfomo"
    MULTI_RANK = True

    def extra_state(self) -> dict:
        if not hasattr(self, "_fomo_state"):
            return {}
        return {"fomo_weights": self._fomo_state[0].tolist(), "fomo_p_choose": self._fomo_state[1].tolist()}

    def _val_loss(self, cid: int, slab: ParamSlab) -> float:
        client = self.client_list[cid]
        data = self.val_data_local_dict[cid] if self.val_data_local_dict is not None else client.local_test_data
        self.model_trainer.attach(slab)
        m = self.model_trainer.test(data, self.device, self.args)
        return m["test_loss"]

    def _benefit_choose(self, cur, total, per_round, p_choose):
        if total == per_round:
            return np.arange(total)
        k = min(per_round, total)
        p_choose[cur] = 0
        if random.random() >= 0.5:
            return np.argsort(p_choose)[-k:]
        idx = np.random.choice(range(total), k, replace=False)
        while cur in idx:
            idx = np.random.choice(range(total), k, replace=False)
        return idx

    def train(self):
        args = self.args
        n = args.client_num_in_total
        models = self._init_models()
        weights_locals = np.full((n, n), 1.0 / n)
        p_choose = np.ones((n, n))
        if self._resumed_extra.get("fomo_weights") is not None:
            weights_locals, p_choose = np.asarray(self._resumed_extra["fomo_weights"]), np.asarray(self._resumed_extra["fomo_p_choose"])
        self._fomo_state = (weights_locals, p_choose)
        trained = self.w_global.clone()
        for round_idx in range(self.start_round, args.comm_round):
            t0 = time.time()
            self.logger.info("################Communication round : {}".format(round_idx))
            last = [m.clone() for m in models]
            after_train, after_agg = [], []
            with self.timer.phase("train"):
                for cid in range(n):
                    client = self.client_list[cid]
                    if not self.dist.owns(cid):              # keep the python / numpy streams in step with the owner's draw
                        self._benefit_choose(cid, n, args.client_num_per_round, p_choose[cid].copy())
                        continue
                    self.logger.info('@@@@@@@@@@@@@@@@ Training Client CM({}): {}'.format(round_idx, cid))
                    _, flops, comm = client.train(last[cid], round_idx, None)
                    trained.copy_(client.slab)
                    after_train.append((cid, client.local_test(trained, True)))
                    nei = self._benefit_choose(cid, n, args.client_num_per_round, p_choose[cid])
                    if n != args.client_num_per_round:
                        nei = np.append(nei, cid)
                    nei = np.sort(nei).astype(int)
                    self.logger.info("client_indexes = " + str(nei))
                    # first-order weights: (validation-loss gain) / ||model difference||
                    loss_cur = self._val_loss(cid, last[cid])
                    for j in nei:
                        cand = trained if j == cid else last[j]
                        gain = loss_cur - self._val_loss(cid, cand)
                        dist = float(torch.sqrt(ops.squared_distance(cand.flat, last[cid].flat)))
                        weights_locals[cid][j] = 0.0 if dist == 0 else gain / dist
                    self.logger.info("Computing Agg weights: " + str([weights_locals[cid][j] for j in nei]))
                    p_choose[cid] = p_choose[cid] + weights_locals[cid]
                    wpos = np.maximum(weights_locals[cid], 0)
                    total_w = float(wpos[nei].sum())
                    if total_w == 0.0:
                        client.slab.copy_(last[cid])
                    else:
                        # w = w_old + sum_j (w_j - w_old) * weight_j / total  ==  weighted aggregate with the residual on w_old
                        srcs, coefs = [last[cid].flat], [1.0 - sum(wpos[j] for j in nei) / total_w]
                        for j in nei:
                            srcs.append((trained if j == cid else last[j]).flat)
                            coefs.append(wpos[j] / total_w)
                        ops.weighted_aggregate(client.slab.flat, srcs, coefs)
                    m = client.local_test(client.slab, True)
                    self.logger.info("test acc on this client after agg {} / {} : {:.2f}".format(m['test_correct'], m['test_total'], m['test_acc']))
                    after_agg.append((cid, m))
                    self.stat_info["sum_training_flops"] += flops
                    self._add_comm_params(comm)
            if self.dist.is_distributed:
                for cid in range(n):
                    src = self.dist.owner(cid)
                    self.dist.broadcast(models[cid].flat, src=src)
                    rows = torch.tensor(np.stack([weights_locals[cid], p_choose[cid]]), dtype=torch.float64, device=self.device)
                    self.dist.broadcast(rows, src=src)
                    weights_locals[cid], p_choose[cid] = rows[0].cpu().numpy(), rows[1].cpu().numpy()
            for key, res in (("old_mask_test_acc", after_train), ("new_mask_test_acc", after_agg)):
                accs = np.zeros(n)
                for cid, r in res:
                    accs[cid] = r["test_correct"] / max(r["test_total"], 1)
                acc = float(self._reduce_metric(accs)[0].sum() / n)
                self.stat_info[key].append(acc)
                self.logger.info({'test_acc': acc})
                self._emit(round=int(round_idx), series=key, test_acc=acc)
            self.stat_info["person_test_acc"].append(self.stat_info["new_mask_test_acc"][-1])
            self.stat_info["round_time_s"].append(time.time() - t0)
            self._maybe_checkpoint(round_idx)
        self.record_information()
        return self.finalize_stats()
