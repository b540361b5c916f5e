"""Where work runs: clients on ranks, evaluation jobs on ranks.

The reference is one process; its clients train and are evaluated one after another (``sailentgrads_api.py:126-138,231-285``).
Here a rank owns a client (its personal model lives there), but an *evaluation* is just (model, test split): the global model is
on every rank, a personal model is a 10 MB peer copy away, and every rank can stage any test split.  So the 2 x clients evaluation
jobs of a round are spread by load on top of what each rank already spent training -- with 21 clients on 8 GPUs the ranks that
trained 2 clients instead of 3 take the evaluations, and the round ends when the 3-client ranks finish training.

Every function here is a pure function of values all ranks hold, so all ranks compute the same plan without talking.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

TRAIN_COST_PER_SAMPLE = 3.0      # forward + backward ~ 3 forward passes
EVAL_COST_PER_SAMPLE = 1.4       # small evaluation batches run below the training batch's efficiency (measured 0.35 ms vs 6.1 ms)
FOREIGN_PENALTY = 0.15           # relative extra cost of evaluating a personal model away from its owner (the slab copy)


def lpt_assign(costs: Sequence[float], world: int) -> List[int]:
    """Longest-processing-time placement: items by decreasing cost (ties by index), each to the least loaded rank (ties to the
    lowest rank).  Equal costs give round-robin."""
    load = [0.0] * world
    owner = [0] * len(costs)
    for i in sorted(range(len(costs)), key=lambda j: (-float(costs[j]), j)):
        r = min(range(world), key=lambda q: (load[q], q))
        owner[i] = r
        load[r] += float(costs[i])
    return owner


def train_cost(n_samples: int, epochs: int) -> float:
    return TRAIN_COST_PER_SAMPLE * float(n_samples) * max(1, int(epochs))


def eval_cost(n_samples: int) -> float:
    return EVAL_COST_PER_SAMPLE * float(n_samples)


def plan_eval_jobs(client_ids: Sequence[int], owners: Dict[int, int], test_sizes: Dict[int, int], personal: Sequence[int],
                   train_load: Sequence[float], world: int, balance: bool = True) -> List[List[Tuple[int, str]]]:
    """Per rank, the ``(client, 'g' | 'p')`` evaluations it runs: 'g' = global model, 'p' = the client's personal model (only for
    clients in ``personal``; the others' personalised entry is evaluated with the shared fall-back model, as a 'p' job without a
    transfer).  ``balance=False`` keeps every job on the client's owner."""
    plan: List[List[Tuple[int, str]]] = [[] for _ in range(world)]
    if not balance or world == 1:
        for c in client_ids:
            plan[owners[c]].extend([(c, "g"), (c, "p")])
        return plan
    load = [float(x) for x in train_load]
    has = set(int(c) for c in personal)
    jobs = [(c, "p") for c in client_ids] + [(c, "g") for c in client_ids]
    for c, kind in sorted(jobs, key=lambda j: (-eval_cost(test_sizes[j[0]]), j[1] != "p", j[0])):
        cost = eval_cost(test_sizes[c])

        def total(r):
            foreign = kind == "p" and c in has and r != owners[c]
            return load[r] + cost * (1.0 + (FOREIGN_PENALTY if foreign else 0.0))
        r = min(range(world), key=lambda q: (total(q), q != owners[c], q))
        load[r] = total(r)
        plan[r].append((c, kind))
    for r in range(world):      # global-model jobs first (one weight re-pack), then personal ones, each in client order
        plan[r].sort(key=lambda j: (j[1] != "g", j[0]))
    return plan
