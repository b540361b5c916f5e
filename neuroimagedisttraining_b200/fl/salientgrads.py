"""SalientGrads: sparse FL with one saliency-derived global mask
(reference: ``fedml_api/standalone/sailentgrads/sailentgrads_api.py:18-346``, ``snip.py``).

Before round 0 every client scores ``|dL/dm| = |g*w|`` on ``itersnip_iteration`` mini-batches, the
scores are averaged over *all* clients (unweighted), a global top-k over the Conv/Linear weights gives
one shared binary mask, and FedAvg then runs with ``w *= mask`` fused into every SGD step.
"""
from __future__ import annotations

import time


from .. import ops
from ..runtime.slab import ParamSlab, make_mask_slab
from .fedavg import FedAvgAPI


class SalientGradsAPI(FedAvgAPI):
    NAME = "salientgrads"

    def generate_global_mask_snip(self) -> ParamSlab:
        """Scores of every client (not just the sampled subset) -> mean -> global top-k mask
        (``sailentgrads_api.py:47-66``).  Cross-rank: each rank sums its clients' scores; the
        reduction over ranks, the top-k threshold and the mask emission run on the device."""
        args = self.args
        n_clients = args.client_num_in_total
        scores = ParamSlab(self.layout, device=self.device, params_only=True)
        tmp = ParamSlab(self.layout, device=self.device, params_only=True)
        with self.timer.phase("saliency"):
            for cid in self.local_clients(range(n_clients)):
                tmp.flat.zero_()
                self.client_list[cid].generate_sailency_scores_from_each_client(
                    itersnip_iteration=int(args.itersnip_iteration), stratified_sampling=bool(args.stratified_sampling),
                    w_global=self.w_global, out=tmp)
                scores.flat.add_(tmp.flat, alpha=1.0 / n_clients)
            self.logger.info("@@@@@Aggregate the local masks@@@@@@@@@@@@@@@@@@@")
            mask = ParamSlab(self.layout, device=self.device, params_only=True)
            fused = getattr(self.aggregator, "_fused", None)
            if fused is not None:
                # W2: NVLink all-reduce of the scores + global top-k + mask emission in one kernel per rank
                if not hasattr(self, "_prunable_flags"):
                    self._prunable_flags = ops.prunable_flags(self.layout, self.device)
                _, info = fused.saliency_topk_mask(scores.flat, self.layout, float(args.dense_ratio), mask.flat, self._prunable_flags)
            else:
                self.dist.all_reduce_sum(scores.flat)
                _, info = ops.global_topk_mask(scores.flat, self.layout, float(args.dense_ratio), mask.flat)
        self.mask_info = info
        self.scores = scores
        return mask

    def train(self):
        args = self.args
        if self.w_init is None:
            self.w_init = self.w_global.clone()
        if self.mask is None:
            self.mask = self.generate_global_mask_snip()
            if not getattr(args, "snip_mask", True):
                # dense ablation of the reference: the mask is generated, then replaced by ones
                self.mask = make_mask_slab(self.layout, self.device)
        for round_idx in range(self.start_round, args.comm_round):
            t0 = time.time()
            self.run_round(round_idx, mask=self.mask)
            self.stat_info["round_time_s"].append(time.time() - t0)
        self._test_on_all_clients(self.w_global, self.personal_slabs(), -1)
        return self.finalize_stats()

    def mask_state_dict(self):
        return {n: self.mask.view(n).clone() for n in self.layout.param_names} if self.mask is not None else None


SailentGradsAPI = SalientGradsAPI  # the reference's spelling
