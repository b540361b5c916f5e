"""Local-only baseline: every client trains alone, no communication
(reference: ``fedml_api/standalone/local/local_api.py:51-84``; results pickled as in ``:196-199``)."""
from __future__ import annotations

import os
import pickle
import time


from .base import FedAPIBase


class LocalAPI(FedAPIBase):
    NAME = "local"
    PERSONAL_IN_CHECKPOINT = True

    def train(self):
        args = self.args
        if self.w_init is None:
            self.w_init = self.w_global.clone()
        for c in self.client_list:
            c.ensure_slab(self.w_global)
        for round_idx in range(self.start_round, args.comm_round):
            t0 = time.time()
            self.logger.info("################Communication round : {}".format(round_idx))
            with self.timer.phase("train"):
                for cid in self.local_clients(range(args.client_num_in_total)):
                    client = self.client_list[cid]
                    self.logger.info('@@@@@@@@@@@@@@@@ Training Client CM({}): {}'.format(round_idx, cid))
                    _, flops, comm = client.train(client.slab.clone(), round_idx, None)
                    self.stat_info["sum_training_flops"] += flops
            if round_idx % max(1, int(getattr(args, "frequency_of_the_test", 1))) == 0 or round_idx == args.comm_round - 1:
                self._local_test_on_all_clients(round_idx)
            self.stat_info["round_time_s"].append(time.time() - t0)
            self._maybe_checkpoint(round_idx)
        self.record_information()
        return self.finalize_stats()

    def _local_test_on_all_clients(self, round_idx):
        self.logger.info("################local_test_on_all_clients : {}".format(round_idx))
        ids = list(range(self.args.client_num_in_total))
        mine = self.local_clients(ids)
        with self.timer.phase("eval"):
            acc, loss = self._evaluate_clients({c: self.client_list[c].slab for c in mine}, mine)
            acc, loss = self._reduce_metric(acc, loss)
        stats = {'person_test_acc': acc.sum() / len(ids), 'person_test_loss': loss.sum() / len(ids)}
        self.stat_info["person_test_acc"].append(stats["person_test_acc"])
        self.stat_info["person_test_loss"].append(stats["person_test_loss"])
        # the key the reference's Local / Ditto / D-PSGD / FedFomo / Sub-FedAvg APIs record (e.g. ditto_api.py:149-150)
        self.stat_info.setdefault("test_acc", []).append(stats["person_test_acc"])
        self.logger.info(stats)
        self._emit(round=int(round_idx), **{k: float(v) for k, v in stats.items()})
        return stats

    def record_information(self):
        """Pickle ``stat_info`` to ``results/<dataset>/<identity>`` (the reference crashes when the directory is
        missing, ``subavg/error2937452.err``; here it is created)."""
        ident = getattr(self.args, "identity", None)
        if not ident or not self.dist.is_main:
            return
        path = os.path.join(getattr(self.args, "results_dir", "results"), str(getattr(self.args, "dataset", "data")))
        os.makedirs(path, exist_ok=True)
        with open(os.path.join(path, ident), "wb") as f:
            pickle.dump({k: v for k, v in self.stat_info.items() if k != "phase_time_s"}, f)
