"""End-to-end parity against the unmodified reference on CPU: same
synthetic data, same initial weights, dropout p=0, unshuffled loaders -> the SalientGrads mask must
be identical and the aggregated global model must match to fp32 round-off."""
import numpy as np
import pytest
import torch

from baseline import ref_harness
from neuroimagedisttraining_b200.data import SyntheticABCD, site_partition
from neuroimagedisttraining_b200.data.loaders import IndexLoader
from neuroimagedisttraining_b200.fl import FedAvgAPI, SalientGradsAPI, SlabTrainer
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout

from helpers import make_args, quiet_logger

pytestmark = pytest.mark.skipif(ref_harness.reference_available() is not None, reason="reference not installed")


def _setup(n_clients=2, per_site=5, batch=2):
    store = SyntheticABCD(n_clients * per_site, n_sites=n_clients, seed=3, pool=2, site_sizes=[per_site] * n_clients)
    train, test, _ = site_partition(store.sites)
    return store, train, test


def _ours(algo, store, train, test, args, init_state):
    torch.manual_seed(0)
    model = AlexNet3D_Dropout(num_classes=1, p_drop=0.0)
    model.load_state_dict(init_state)
    tr = SlabTrainer(model, args, quiet_logger(), device="cpu", store=store)
    local_num = {c: len(t) for c, t in enumerate(train)}
    trl = {c: IndexLoader(t, store.labels[t], store.sites[t], args.batch_size, shuffle=False) for c, t in enumerate(train)}
    tel = {c: IndexLoader(t, store.labels[t], store.sites[t], args.batch_size, shuffle=False) for c, t in enumerate(test)}
    ds = [None, None, None, None, local_num, trl, tel, 2]
    api = (SalientGradsAPI if algo == "salientgrads" else FedAvgAPI)(ds, "cpu", args, tr, quiet_logger())
    return api


@pytest.mark.slow
def test_salientgrads_matches_reference_cpu():
    torch.set_num_threads(max(1, torch.get_num_threads()))
    store, train, test = _setup()
    args = make_args(client_num_in_total=2, batch_size=2, epochs=1, comm_round=1, dense_ratio=0.5)
    torch.manual_seed(123)
    init = AlexNet3D_Dropout(num_classes=1).state_dict()

    ref_api, ref_model, _ = ref_harness.build_reference_api("salientgrads", store, train, test, args, "cpu", shuffle=False,
                                                           model_state=init, p_drop=0.0)
    ref_mask = ref_api.generate_global_mask_snip()
    ours = _ours("salientgrads", store, train, test, args, init)
    mask = ours.generate_global_mask_snip()
    total = diff = 0
    for name, m in ref_mask.items():
        d = int((mask.view(name) != m).sum())
        diff += d
        total += m.numel()
    # identical up to ties / last-bit differences of the normalising sum (documented tolerance)
    assert diff <= max(2, total // 100000), f"mask differs at {diff}/{total} positions"
    assert abs(float(mask.flat.sum()) - sum(float(m.sum()) for m in ref_mask.values())) <= 2

    # one full round with the shared mask, then compare the aggregated global model
    ref_api.generate_global_mask_snip = lambda: ref_mask
    ref_model.load_state_dict(init)
    captured = {}
    orig_agg = ref_api._aggregate
    def agg(w_locals):
        out = orig_agg(w_locals)
        captured["w"] = out
        return out
    ref_api._aggregate = agg
    ref_api.train()
    ours.mask = mask
    ours.train()
    sd = ours.w_global.state_dict(cast_int=False)
    for k, v in captured["w"].items():
        torch.testing.assert_close(sd[k].float(), v.float(), rtol=2e-4, atol=2e-5, msg=lambda m: f"{k}: {m}")
    # metric keys and values
    assert abs(ours.stat_info["global_test_acc"][0] - ref_api.stat_info["global_test_acc"][0]) < 1e-6
    assert abs(ours.stat_info["person_test_acc"][0] - ref_api.stat_info["person_test_acc"][0]) < 1e-6


@pytest.mark.slow
def test_fedavg_matches_reference_cpu():
    store, train, test = _setup()
    args = make_args(client_num_in_total=2, batch_size=2, epochs=1, comm_round=1, lr=0.001)
    torch.manual_seed(7)
    init = AlexNet3D_Dropout(num_classes=1).state_dict()
    ref_api, ref_model, _ = ref_harness.build_reference_api("fedavg", store, train, test, args, "cpu", shuffle=False,
                                                           model_state=init, p_drop=0.0)
    captured = {}
    orig_agg = ref_api._aggregate
    def agg(w_locals):
        out = orig_agg(w_locals)
        captured.setdefault("w", {k: v.clone() for k, v in out.items()})
        return out
    ref_api._aggregate = agg
    # the reference's main_fedavg hard-requires CUDA; on CPU its trainer's model.to(device) is a no-op
    ref_api.train()
    ours = _ours("fedavg", store, train, test, args, init)
    args.final_finetune = False
    ours.run_round(0)
    sd = ours.w_global.state_dict(cast_int=False)
    for k, v in captured["w"].items():
        torch.testing.assert_close(sd[k].float(), v.float(), rtol=2e-4, atol=2e-5, msg=lambda m: f"{k}: {m}")


def test_dispfl_prune_regrow_match_reference_client_methods():
    """DisPFL mask evolution: the reference's ``Client.fire_mask`` / ``regrow_mask`` (DisPFL/client.py:71-99, full sorts) against
    the select-k op the GPU kernel restates (``ops.mask_select_k``) and the per-tensor helpers, on tie-free random tensors."""
    import logging
    import math
    import types
    from neuroimagedisttraining_b200 import ops
    from neuroimagedisttraining_b200.utils.slim import cosine_annealing
    ref_harness.import_reference()
    from fedml_api.standalone.DisPFL.client import Client as RefClient

    args = types.SimpleNamespace(anneal_factor=0.5, comm_round=10, dis_gradient_check=False, static=False)
    log = logging.getLogger("ref-dispfl"); log.setLevel(logging.ERROR)
    rc = RefClient(0, None, None, 5, args, "cpu", None, log)
    g = torch.Generator().manual_seed(7)
    shapes = {"features.0.weight": (8, 1, 3, 3, 3), "features.4.weight": (16, 8, 3, 3, 3), "classifier.1.weight": (16, 32)}
    w = {n: torch.randn(*s, generator=g) for n, s in shapes.items()}
    grad = {n: torch.randn(*s, generator=g) for n, s in shapes.items()}
    masks = {n: (torch.rand(*s, generator=g) > 0.5).float() for n, s in shapes.items()}
    for rnd in (0, 3, 9):
        ref_fired, ref_removed = rc.fire_mask(masks, w, rnd)
        ref_grown = rc.regrow_mask(ref_fired, ref_removed, grad)
        drop = cosine_annealing(args, rnd)
        for n in shapes:
            k = math.ceil(drop * int(masks[n].sum()))
            assert k == ref_removed[n]
            fired = ops.mask_select_k(w[n], masks[n].clone(), k, "fire")
            assert torch.equal(fired, ref_fired[n]), (n, rnd)
            fired2, k2 = ops.reference.fire_mask(w[n], masks[n], drop)
            assert k2 == k and torch.equal(fired2, ref_fired[n])
            grown = ops.mask_select_k(grad[n], fired.clone(), k, "regrow")
            assert torch.equal(grown, ref_grown[n]), (n, rnd)
            assert torch.equal(ops.reference.regrow_mask(fired, k, grad[n]), ref_grown[n])
            assert int(grown.sum()) == int(masks[n].sum())          # density preserved


@pytest.mark.parametrize("active,static", [(1.0, False), (0.7, False), (1.0, True)])
def test_dispfl_matches_reference_cpu(active, static):
    """DisPFL end to end through the reference's unmodified ``dispflAPI`` / ``Client`` / ``MyModelTrainer``: ERK random masks from the
    same torch RNG stream, masked local SGD, eval-mode gradient screening, cosine-annealed prune / regrow, seeded client drop-out,
    both accuracy series and the final mask Hamming matrix.  The model is the small 3-D CNN at 32^3 (the reference trainer takes any
    module; its only ABCD-specific part is the HDF5 fetch, replaced by the synthetic source as in the other parity tests)."""
    import copy
    from neuroimagedisttraining_b200.fl import DisPFLAPI
    from neuroimagedisttraining_b200.models import Tiny3DCNN
    ref_harness.import_reference()
    from fedml_api.standalone.DisPFL import my_model_trainer as dp_trainer
    from fedml_api.standalone.DisPFL.dispfl_api import dispflAPI as RefAPI

    n, per_site = 3, 10
    store = SyntheticABCD(n * per_site, shape=(32, 32, 32), n_sites=n, seed=5, pool=3, site_sizes=[per_site] * n)
    train, test, _ = site_partition(store.sites)
    args = make_args(client_num_in_total=n, batch_size=4, epochs=1, comm_round=3, lr=0.02, dense_ratio=0.5, anneal_factor=0.3, cs="random",
                     active=active, static=static, uniform=False, different_initial=True, diff_spa=False, dis_gradient_check=False,
                     save_masks=True, record_mask_diff=False)
    torch.manual_seed(11)
    init = copy.deepcopy(Tiny3DCNN(1, in_shape=(32, 32, 32)).state_dict())

    source = ref_harness.SyntheticSource(store, "cpu")
    dp_trainer.MyModelTrainer.load_data_chunks_batch = lambda self, idx: source.batch(idx)
    rmodel = Tiny3DCNN(1, in_shape=(32, 32, 32))
    rmodel.load_state_dict(init)
    ds = ref_harness.build_dataset(store, train, test, args.batch_size, shuffle=False)
    rapi = RefAPI(ds, torch.device("cpu"), args, dp_trainer.MyModelTrainer(rmodel, args, quiet_logger()), quiet_logger())
    finals = {}
    for c in rapi.client_list:
        orig = c.train
        c.train = (lambda o, cid: (lambda w, m, r: (lambda out: (finals.__setitem__(cid, copy.deepcopy(out[1])), out)[1])(o(w, m, r))))(orig, c.client_idx)
    torch.manual_seed(3); np.random.seed(3)
    rapi.train()

    model = Tiny3DCNN(1, in_shape=(32, 32, 32))
    model.load_state_dict(init)
    tr = SlabTrainer(model, args, quiet_logger(), device="cpu", store=store)
    trl = {c: IndexLoader(t, store.labels[t], store.sites[t], args.batch_size, shuffle=False) for c, t in enumerate(train)}
    tel = {c: IndexLoader(t, store.labels[t], store.sites[t], args.batch_size, shuffle=False) for c, t in enumerate(test)}
    api = DisPFLAPI([None, None, None, None, {c: len(t) for c, t in enumerate(train)}, trl, tel, 2], "cpu", args, tr, quiet_logger())
    torch.manual_seed(3); np.random.seed(3)
    api.train()

    for c in range(n):
        sd = api.client_list[c].slab.state_dict(cast_int=False)
        for k, v in finals[c].items():
            torch.testing.assert_close(sd[k].float(), v.float(), rtol=2e-4, atol=2e-5, msg=lambda m: f"client {c} {k}: {m}")
        for name, m in rapi.stat_info["final_masks"][c].items():
            assert torch.equal(api.stat_info["final_masks"][c][name], m), (c, name)
    assert api.stat_info["mask_dis_matrix"] == [[int(v) for v in row] for row in rapi.stat_info["mask_dis_matrix"]]
    if not static:
        assert any(v > 0 for row in api.stat_info["mask_dis_matrix"] for v in row), "masks must have evolved"
    for key in ("old_mask_test_acc", "new_mask_test_acc"):
        assert len(api.stat_info[key]) == len(rapi.stat_info[key]) == 3
        for a, b in zip(api.stat_info[key], rapi.stat_info[key]):
            assert abs(float(a) - float(b)) < 1e-6, key


@pytest.mark.slow
def test_salientgrads_two_rounds_with_partial_participation_matches_reference_cpu():
    """Two rounds, 3 site clients of which 2 are sampled per round (``np.random.seed(round)`` sampler), lr decay by round, the
    personalised models of clients that did / did not train, and both accuracy series incl. the closing ``round = -1`` evaluation."""
    store, train, test = _setup(n_clients=3, per_site=5)
    args = make_args(client_num_in_total=3, frac=0.67, batch_size=2, epochs=1, comm_round=2, dense_ratio=0.3, lr=0.02)
    assert args.client_num_per_round == 2
    torch.manual_seed(77)
    init = AlexNet3D_Dropout(num_classes=1).state_dict()
    ref_api, ref_model, _ = ref_harness.build_reference_api("salientgrads", store, train, test, args, "cpu", shuffle=False,
                                                           model_state=init, p_drop=0.0)
    captured = []
    orig_agg, orig_sample = ref_api._aggregate, ref_api._client_sampling
    ref_api._aggregate = lambda w: (captured.append({k: v.clone() for k, v in orig_agg(w).items()}), captured[-1])[1]
    sampled = []
    ref_api._client_sampling = lambda r, n, k: (sampled.append(sorted(int(c) for c in orig_sample(r, n, k))), orig_sample(r, n, k))[1]
    ref_api.train()

    # Round by round, re-synchronising our global model to the reference's after each comparison: a batch-2 BatchNorm network
    # amplifies the 1e-6 round-off of one round a hundredfold in the next (ReLU / max-pool decisions flip), which would hide what
    # is pinned here -- the per-round semantics (sampling, lr decay, masking, aggregation weights, personal models).
    ours = _ours("salientgrads", store, train, test, args, init)
    ours_sampled = []
    orig_ours = ours._client_sampling
    ours._client_sampling = lambda r, n, k: (ours_sampled.append(sorted(int(c) for c in orig_ours(r, n, k))), orig_ours(r, n, k))[1]
    ours.mask = ours.generate_global_mask_snip()
    ours.w_init = ours.w_global.clone()
    for r in range(2):
        ours.run_round(r, mask=ours.mask)
        sd = ours.w_global.state_dict(cast_int=False)
        for k, v in captured[r].items():
            torch.testing.assert_close(sd[k].float(), v.float(), rtol=2e-4, atol=2e-5, msg=lambda m: f"round {r} {k}: {m}")
        ours.w_global.load_state_dict(captured[r])
    ours._test_on_all_clients(ours.w_global, ours.personal_slabs(), -1)
    assert ours_sampled == sampled and len(sampled) == 2 and all(len(s) == 2 for s in sampled)
    for key in ("global_test_acc", "person_test_acc"):
        assert len(ours.stat_info[key]) == len(ref_api.stat_info[key]) == 3
        for a, b in zip(ours.stat_info[key], ref_api.stat_info[key]):
            assert abs(float(a) - float(b)) < 1e-6, key


@pytest.mark.slow
def test_fedavg_final_finetune_round_matches_reference_cpu():
    """FedAvg's closing fine-tune (``fedavg_api.py:79-88``): every client trains once more from the final global model with
    ``round = -1`` (lr * decay^-1) and is evaluated with that personal model.  Our global model is set to the reference's before the
    fine-tune so that only this step's semantics are compared (see the two-round SalientGrads test on error amplification)."""
    import copy
    store, train, test = _setup(n_clients=2, per_site=5)
    args = make_args(client_num_in_total=2, batch_size=2, epochs=1, comm_round=1, lr=0.01, lr_decay=0.9)
    torch.manual_seed(5)
    init = AlexNet3D_Dropout(num_classes=1).state_dict()
    ref_api, _, _ = ref_harness.build_reference_api("fedavg", store, train, test, args, "cpu", shuffle=False, model_state=init, p_drop=0.0)
    cap = {}
    orig_test = ref_api._test_on_all_clients
    ref_api._test_on_all_clients = lambda wg, wp, r: (cap.__setitem__(r, (copy.deepcopy(wg), copy.deepcopy(wp))), orig_test(wg, wp, r))[1]
    ref_api.train()
    assert sorted(cap) == [-1, 0]
    ours = _ours("fedavg", store, train, test, args, init)
    args.final_finetune = False
    ours.train()
    ours.w_global.load_state_dict(cap[-1][0])
    ours.finetune_round()
    for c in range(2):
        sd = ours.client_list[c].slab.state_dict(cast_int=False)
        for k, v in cap[-1][1][c].items():
            torch.testing.assert_close(sd[k].float(), v.float(), rtol=2e-4, atol=2e-5, msg=lambda m: f"personal[{c}] {k}: {m}")
    assert len(ours.stat_info["person_test_acc"]) == len(ref_api.stat_info["person_test_acc"]) == 2
    for key in ("global_test_acc", "person_test_acc"):
        for a, b in zip(ours.stat_info[key], ref_api.stat_info[key]):
            assert abs(float(a) - float(b)) < 1e-6, key
