"""Ditto: FedAvg global model + a personalised model per client pulled towards it,
``w -= lr*lamda*(w - w_global)`` after every SGD step (reference: ``fedml_api/standalone/ditto/ditto_api.py:40-78``,
``ditto/my_model_trainer.py:38-68``, ``ditto/client.py:32-50``)."""
from __future__ import annotations

import time

import numpy as np

from .local import LocalAPI


class DittoAPI(LocalAPI):
    NAME = "ditto"

    def train(self):
        args = self.args
        if self.w_init is None:
            self.w_init = self.w_global.clone()
        lamda = float(getattr(args, "lamda", 0.5))
        local_epochs = int(getattr(args, "local_epochs", 3))
        for round_idx in range(self.start_round, args.comm_round):
            t0 = time.time()
            self.logger.info("################Communication round : {}".format(round_idx))
            client_indexes = np.sort(self._client_sampling(round_idx, args.client_num_in_total, args.client_num_per_round))
            training_num = sum(self.client_list[int(c)].get_sample_number() for c in client_indexes)
            w_locals = []
            with self.timer.phase("train"):
                for cid in self.local_clients(client_indexes):
                    client = self.client_list[cid]
                    self.logger.info('@@@@@@@@@@@@@@@@ Training Client CM({}): {}'.format(round_idx, cid))
                    if not hasattr(client, "global_copy"):
                        client.global_copy = self.w_global.clone()
                        if client.slab is None:                   # (a resumed run already has the personal model)
                            client.slab = self.w_init.clone()     # personal model starts at the initial weights
                    # (1) the client's contribution to the global model: plain local training from w_global
                    tr = self.model_trainer
                    client.global_copy.copy_(self.w_global)
                    tr.attach(client.global_copy)
                    tr.set_id(cid)
                    comm = tr.count_communication_params(self.w_global)
                    tr.train(client.local_training_data, self.device, args, round_idx, None)
                    comm = comm + tr.count_communication_params(client.global_copy)
                    # (2) the personalised model: local_epochs of SGD with the proximal pull towards w_global
                    tr.attach(client.slab)
                    tr.train(client.local_training_data, self.device, args, round_idx, None, prox_ref=self.w_global, lamda=lamda,
                             epochs=local_epochs)
                    tr.engine.detach_sync()
                    w_locals.append((client.get_sample_number(), client.global_copy))
                    self.stat_info["sum_training_flops"] += (args.epochs + local_epochs) * client.get_sample_number()
                    self._add_comm_params(comm)
            self._aggregate(w_locals, training_num)
            if round_idx == args.comm_round - 1 or round_idx % max(1, int(getattr(args, "frequency_of_the_test", 1))) == 0:
                for c in self.client_list:
                    if c.slab is None:
                        c.slab = self.w_init.clone()
                self._local_test_on_all_clients(round_idx)
            self.stat_info["round_time_s"].append(time.time() - t0)
            self._maybe_checkpoint(round_idx)
        self.record_information()
        return self.finalize_stats()
