"""TurboAggregate-style secure aggregation on top of FedAvg.

The reference ships only stubs for this algorithm (``fedml_api/standalone/turboaggregate/TA_trainer.py:38-72`` -- the protocol
body is ``pass``, the client constructor does not even match its caller -- plus the field arithmetic in ``mpc_function.py``).
Here the protocol is functional: the sampled clients are arranged in the vanilla ring of groups (``TA_topology_vanilla``,
``TA_trainer.py:87-97``); each client quantises its sample-weighted update into a prime field, splits it into additive
shares (one per member of the next group) and every group forwards the running masked aggregate, so no party ever sees an
individual update; the server reconstructs only the sum.  Optional client drop-out (``set_dropout``, ``TA_client.py:25-26``)
is tolerated by re-sharing within the surviving members.  The result equals plain FedAvg up to the quantisation step.
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np
import torch

from ..utils import mpc
from .fedavg import FedAvgAPI


class TurboAggregateAPI(FedAvgAPI):
    NAME = "turboaggregate"

    def TA_topology_vanilla(self, client_indexes: Sequence[int], group_size: int) -> List[List[int]]:
        ids = [int(c) for c in client_indexes]
        return [ids[i:i + group_size] for i in range(0, len(ids), group_size)]

    def _secure_sum(self, updates: List[np.ndarray], groups: List[List[int]], order: List[int], dropped: Sequence[int], p: int,
                    rng: np.random.RandomState) -> np.ndarray:
        """Ring of groups: group l receives, per member, a share of everything aggregated so far, adds its own members'
        shares and passes the result on.  Returns the field-sum of the surviving clients' quantised updates."""
        d = updates[0].shape[0]
        pos = {c: i for i, c in enumerate(order)}
        carried = None                                   # shares held by the members of the current group
        for gi, group in enumerate(groups):
            alive = [c for c in group if c not in dropped]
            if not alive:
                continue
            nxt = groups[gi + 1] if gi + 1 < len(groups) else [-1]        # -1 = the server
            nxt_alive = [c for c in nxt if c not in dropped] or [-1]
            n_out = len(nxt_alive)
            outgoing = np.zeros((n_out, d), dtype=np.int64)
            for c in alive:
                masks = mpc.Gen_Additive_SS(d, n_out, p, rng)             # sums to 0: hides the individual update
                shares = masks.copy()
                shares[0] = (shares[0] + updates[pos[c]]) % p
                outgoing = (outgoing + shares) % p
            if carried is not None:                                       # forward what earlier groups aggregated
                total_prev = np.mod(carried.sum(axis=0), p)
                resplit = mpc.Gen_Additive_SS(d, n_out, p, rng)
                resplit[0] = (resplit[0] + total_prev) % p
                outgoing = (outgoing + resplit) % p
            carried = outgoing
        return np.mod(carried.sum(axis=0), p) if carried is not None else np.zeros(d, dtype=np.int64)

    def _aggregate(self, w_locals, training_num=None, mask=None, out=None):
        args = self.args
        if training_num is None:
            training_num = sum(n for n, _ in w_locals)
        p = int(getattr(args, "ta_prime", mpc.DEFAULT_PRIME))
        scale = int(getattr(args, "ta_scale", 1 << 16))
        group_size = int(getattr(args, "ta_group_size", 2))
        rng = np.random.RandomState(int(getattr(args, "seed", 0)) + 7)
        order = list(range(len(w_locals)))
        updates = [mpc.quantize((s.flat * (n / training_num)).detach().cpu().numpy(), scale, p) for n, s in w_locals]
        groups = self.TA_topology_vanilla(order, group_size)
        dropped = set(getattr(self, "dropped_clients", ()))
        with self.timer.phase("aggregate"):
            if updates:
                total = self._secure_sum(updates, groups, order, dropped, p, rng)
            else:                                        # this rank owns none of the sampled clients
                total = np.zeros(self.w_global.flat.numel(), dtype=np.int64)
            if self.dist.is_distributed:
                # the ranks' masked partial aggregates are field elements too: the ring continues across ranks as one more sum mod p
                t = torch.from_numpy(total).to(self.device)
                self.dist.all_reduce_sum(t)
                total = np.mod(t.cpu().numpy(), p)
        out = out if out is not None else self.w_global
        res = torch.from_numpy(mpc.dequantize(total, scale, p)).to(self.device, torch.float32)
        if mask is not None:
            res[:mask.flat.numel()] *= mask.flat
        out.flat.copy_(res)
        return out
