"""Concurrent client slots (two engines on two CUDA streams, steps interleaved by the host) against the sequential schedule:
every client's slab, the aggregated global model and the evaluation metrics must be bit-identical -- with dropout on, which is
only true because the dropout stream is tied to (round, client) instead of to the engine's step count."""
import logging
import types

import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(lanes: int):
    from neuroimagedisttraining_b200.data import SyntheticABCD, load_partition_data_abcd
    from neuroimagedisttraining_b200.fl import SalientGradsAPI, SlabTrainer
    from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
    from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
    dev = torch.device("cuda:0")
    store = SyntheticABCD(5 * 10, n_sites=5, seed=0, pool=2, pin=True, site_sizes=[10] * 5)
    args = types.SimpleNamespace(client_num_in_total=5, client_num_per_round=5, comm_round=2, epochs=2, lr=0.01, lr_decay=0.998, momentum=0.0,
                                 wd=5e-4, client_optimizer="sgd", batch_size=4, dense_ratio=0.5, itersnip_iteration=1, stratified_sampling=False,
                                 snip_mask=True, ci=0, frequency_of_the_test=1, dataset="ABCD", loss_sync="round")
    logger = logging.getLogger("lanes"); logger.setLevel(logging.ERROR)
    torch.manual_seed(1)
    model = AlexNet3D_Dropout(1, p_drop=0.5)
    tr = SlabTrainer(model, args, logger, device=dev, store=store, engine=B200Engine(model, dev, max_batch=4, input_scale=store.scale))
    for _ in range(lanes - 1):
        tr.add_lane(B200Engine(model, dev, max_batch=4, input_scale=store.scale))
    api = SalientGradsAPI(load_partition_data_abcd(store, client_number=5, batch_size=4), dev, args, tr, logger)
    api.w_init = api.w_global.clone()
    api.mask = api.generate_global_mask_snip()
    for r in range(2):
        api.run_round(r, mask=api.mask)
    torch.cuda.synchronize()
    st = api.finalize_stats()
    return ([c.slab.flat.clone() for c in api.client_list], api.w_global.flat.clone(), st["global_test_loss"], st["person_test_loss"],
            st["sum_comm_params"])


def test_two_client_lanes_equal_the_sequential_schedule_bit_for_bit():
    seq = _run(1)
    par = _run(2)
    for c, (a, b) in enumerate(zip(seq[0], par[0])):
        assert torch.equal(a, b), f"client {c}: max diff {(a - b).abs().max().item()}"
    assert torch.equal(seq[1], par[1])
    assert seq[2] == par[2] and seq[3] == par[3] and seq[4] == par[4]
