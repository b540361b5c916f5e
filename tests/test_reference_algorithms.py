"""End-to-end parity of further algorithms against the UNMODIFIED reference APIs on CPU: same initial weights, same tensors,
unshuffled loaders, a dropout-free CIFAR-shape CNN.  (SalientGrads / FedAvg on the 3-D model: tests/test_reference_parity.py.)"""
import copy
import logging
import types

import numpy as np
import pytest
import torch
from torch.utils.data import DataLoader, TensorDataset

from baseline import ref_harness

pytestmark = pytest.mark.skipif(ref_harness.reference_available() is not None, reason="reference not installed")


def _log():
    log = logging.getLogger("ref-algo")
    log.setLevel(logging.ERROR)
    return log


def _image_dataset(n_clients=3, n_train=16, n_test=8, batch=8, seed=0):
    """The FedML 8-tuple over fixed random CIFAR-shape tensors; identical loaders are handed to both implementations."""
    g = torch.Generator().manual_seed(seed)
    trl, tel, num = {}, {}, {}
    for c in range(n_clients):
        x, y = torch.randn(n_train, 3, 32, 32, generator=g), torch.randint(0, 10, (n_train,), generator=g)
        xt, yt = torch.randn(n_test, 3, 32, 32, generator=g), torch.randint(0, 10, (n_test,), generator=g)
        trl[c] = DataLoader(TensorDataset(x, y), batch_size=batch, shuffle=False)
        tel[c] = DataLoader(TensorDataset(xt, yt), batch_size=batch, shuffle=False)
        num[c] = n_train
    return [n_clients * n_train, n_clients * n_test, None, None, num, trl, tel, 10]


def _args(**kw):
    d = dict(model="cnn_cifar10", dataset="cifar10", batch_size=8, client_optimizer="sgd", lr=0.05, lr_decay=0.998, wd=5e-4, momentum=0.0,
             epochs=1, local_epochs=2, lamda=0.5, client_num_in_total=3, client_num_per_round=3, frac=1.0, comm_round=2,
             frequency_of_the_test=1, ci=0, seed=0, gpu=0, type="epoch")
    d.update(kw)
    return types.SimpleNamespace(**d)


def _model(seed=4):
    from neuroimagedisttraining_b200.models import create_model
    torch.manual_seed(seed)
    return create_model("cnn_cifar10", 10)


def _assert_state_close(slab, state, rtol=2e-4, atol=2e-5, what=""):
    sd = slab.state_dict(cast_int=False)
    for k, v in state.items():
        torch.testing.assert_close(sd[k].float(), v.float().cpu(), rtol=rtol, atol=atol, msg=lambda m: f"{what} {k}: {m}")


def test_ditto_matches_reference_cpu():
    ref_harness.import_reference()
    from fedml_api.standalone.ditto.ditto_api import DittoAPI as RefAPI
    from fedml_api.standalone.ditto.my_model_trainer import MyModelTrainer as RefTrainer
    from neuroimagedisttraining_b200.fl import DittoAPI, SlabTrainer
    args, ds = _args(), _image_dataset()
    init = copy.deepcopy(_model().state_dict())

    rmodel = _model()
    rmodel.load_state_dict(init)
    rapi = RefAPI(ds, torch.device("cpu"), args, RefTrainer(rmodel, args, _log()), _log())
    cap = {}
    orig_agg, orig_test = rapi._aggregate, rapi._local_test_on_all_clients
    rapi._aggregate = lambda w: cap.__setitem__("g", orig_agg(w)) or cap["g"]
    rapi._local_test_on_all_clients = lambda w_pers, r: (cap.__setitem__("p", copy.deepcopy(w_pers)), orig_test(w_pers, r))[1]
    rapi.train()

    model = _model()
    model.load_state_dict(init)
    tr = SlabTrainer(model, args, _log(), device="cpu", loss_kind="ce")
    api = DittoAPI(ds, "cpu", args, tr, _log())
    api.train()
    _assert_state_close(api.w_global, cap["g"], what="global")
    for c in range(args.client_num_in_total):
        _assert_state_close(api.client_list[c].slab, cap["p"][c], what=f"personal[{c}]")
    assert len(api.stat_info["test_acc"]) == len(rapi.stat_info["test_acc"]) == 2
    for a, b in zip(api.stat_info["test_acc"], rapi.stat_info["test_acc"]):
        assert abs(float(a) - float(b)) < 1e-6


@pytest.mark.parametrize("cs,per_round", [("ring", 2), ("random", 2), ("full", 5)])
def test_dpsgd_matches_reference_cpu(cs, per_round):
    """Neighbour choice (ring / seeded random / everyone), neighbour averaging, local training and the plain-mean global model."""
    ref_harness.import_reference()
    from fedml_api.standalone.dpsgd.dpsgd_api import DPSGDAPI as RefAPI
    from fedml_api.standalone.dpsgd.my_model_trainer import MyModelTrainer as RefTrainer
    from neuroimagedisttraining_b200.fl import DPSGDAPI, SlabTrainer
    n = 5
    args, ds = _args(client_num_in_total=n, client_num_per_round=per_round, cs=cs, comm_round=2), _image_dataset(n_clients=n)
    init = copy.deepcopy(_model().state_dict())

    rmodel = _model()
    rmodel.load_state_dict(init)
    rapi = RefAPI(ds, torch.device("cpu"), args, RefTrainer(rmodel, args), _log())
    cap = {}
    orig = rapi._test_on_all_clients
    rapi._test_on_all_clients = lambda wg, wp, r: (cap.update(g=copy.deepcopy(wg), p=copy.deepcopy(wp)), orig(wg, wp, r))[1]
    rapi.train()

    model = _model()
    model.load_state_dict(init)
    api = DPSGDAPI(ds, "cpu", args, SlabTrainer(model, args, _log(), device="cpu", loss_kind="ce"), _log())
    api.train()
    _assert_state_close(api.w_global, cap["g"], what="global")
    for c in range(n):
        _assert_state_close(api.client_list[c].slab, cap["p"][c], what=f"personal[{c}]")
    for key in ("global_test_acc", "person_test_acc"):
        assert len(api.stat_info[key]) == len(rapi.stat_info[key]) == 2
        for a, b in zip(api.stat_info[key], rapi.stat_info[key]):
            assert abs(float(a) - float(b)) < 1e-6, key


def test_local_only_matches_reference_cpu():
    """No communication: every client keeps training its own model (the reference's test bookkeeping only works when all clients
    train every round, local_api.py:105-121, which is also what ours does)."""
    ref_harness.import_reference()
    from fedml_api.standalone.local.local_api import LocalAPI as RefAPI
    from fedml_api.standalone.local.my_model_trainer import MyModelTrainer as RefTrainer
    from neuroimagedisttraining_b200.fl import LocalAPI, SlabTrainer
    args, ds = _args(comm_round=3), _image_dataset()
    init = copy.deepcopy(_model().state_dict())
    rmodel = _model()
    rmodel.load_state_dict(init)
    rtrainer = RefTrainer(rmodel, args, _log())
    rapi = RefAPI(ds, torch.device("cpu"), args, rtrainer, _log())
    rapi.record_information = lambda *a, **k: None          # the reference pickles into ../../results (crashes without the directory)
    finals = {}
    for c in rapi.client_list:
        orig = c.train
        c.train = (lambda o, cid: (lambda w, r: (lambda out: (finals.__setitem__(cid, copy.deepcopy(out[0])), out)[1])(o(w, r))))(orig, c.client_idx)
    rapi.train()

    model = _model()
    model.load_state_dict(init)
    api = LocalAPI(ds, "cpu", args, SlabTrainer(model, args, _log(), device="cpu", loss_kind="ce"), _log())
    api.record_information = lambda *a, **k: None
    api.train()
    for c in range(args.client_num_in_total):
        _assert_state_close(api.client_list[c].slab, finals[c], what=f"personal[{c}]")
    assert len(api.stat_info["test_acc"]) == len(rapi.stat_info["test_acc"]) == 3
    for a, b in zip(api.stat_info["test_acc"], rapi.stat_info["test_acc"]):
        assert abs(float(a) - float(b)) < 1e-6


def test_subavg_matches_reference_cpu():
    """Sub-FedAvg: masked local training, percentile pruning after the first / last epoch, accept-if-accurate rule, mask-count-weighted
    aggregation.  Thresholds are set so that clients really prune (dense > dense_ratio, acc > acc_thresh = 0)."""
    ref_harness.import_reference()
    from fedml_api.standalone.subavg.subavg_api import SubAvgAPI as RefAPI
    from fedml_api.standalone.subavg.my_model_trainer import MyModelTrainer as RefTrainer
    from neuroimagedisttraining_b200.fl import SlabTrainer, SubAvgAPI
    args = _args(comm_round=3, epochs=2, each_prune_ratio=0.2, dense_ratio=0.3, dist_thresh=1e-4, acc_thresh=-1.0)
    ds = _image_dataset()
    init = copy.deepcopy(_model().state_dict())
    rmodel = _model()
    rmodel.load_state_dict(init)
    rapi = RefAPI(ds, torch.device("cpu"), args, RefTrainer(rmodel, args, _log()), _log())
    rapi.record_information = lambda *a, **k: None
    rapi.record_avg_inference_flops = lambda *a, **k: None
    cap = {}
    orig_agg = rapi._aggregate
    rapi._aggregate = lambda ws, wl: cap.__setitem__("g", copy.deepcopy(orig_agg(ws, wl))) or cap["g"]
    orig_test = rapi._local_test_on_all_clients
    rapi._local_test_on_all_clients = lambda wg, mp, r: (cap.__setitem__("masks", copy.deepcopy(mp)), orig_test(wg, mp, r))[1]
    rapi.train()

    model = _model()
    model.load_state_dict(init)
    api = SubAvgAPI(ds, "cpu", args, SlabTrainer(model, args, _log(), device="cpu", loss_kind="ce"), _log())
    api.record_information = lambda *a, **k: None
    api.train()
    _assert_state_close(api.w_global, cap["g"], what="global")
    pruned = 0
    for c in range(args.client_num_in_total):          # masks as they were when the last evaluation ran
        for name, m in cap["masks"][c].items():
            ours = api.eval_masks[c].view(name) if hasattr(api, "eval_masks") else None
            if ours is not None:
                assert torch.equal(ours, m.float()), (c, name)
            pruned += int((m == 0).sum())
    assert pruned > 0, "the configuration must exercise pruning"
    for a, b in zip(api.stat_info["test_acc"], rapi.stat_info["test_acc"]):
        assert abs(float(a) - float(b)) < 1e-6


@pytest.mark.parametrize("per_round", [4, 2])
def test_fedfomo_matches_reference_cpu(per_round):
    """FedFomo: first-order weights (validation-loss gain / model distance), positive-part normalised neighbour mix, the benefit-driven
    neighbour choice (python + numpy RNG streams, re-seeded identically for both runs) and both accuracy series."""
    import random
    ref_harness.import_reference()
    from fedml_api.standalone.fedfomo.fedfomo_api import FEDFOMOAPI as RefAPI
    from fedml_api.standalone.fedfomo.my_model_trainer import MyModelTrainer as RefTrainer
    from neuroimagedisttraining_b200.fl import FedFomoAPI, SlabTrainer
    n = 4
    args = _args(client_num_in_total=n, client_num_per_round=per_round, comm_round=3)
    ds = _image_dataset(n_clients=n)
    g = torch.Generator().manual_seed(123)
    val = {c: DataLoader(TensorDataset(torch.randn(8, 3, 32, 32, generator=g), torch.randint(0, 10, (8,), generator=g)), batch_size=8)
           for c in range(n)}
    init = copy.deepcopy(_model().state_dict())

    rmodel = _model()
    rmodel.load_state_dict(init)
    rapi = RefAPI(ds[:6] + [val] + ds[6:], torch.device("cpu"), args, RefTrainer(rmodel, args, _log()), _log())
    rapi.record_information = lambda *a, **k: None
    finals = {}
    for c in rapi.client_list:      # the last local_test of a client in a round sees its post-aggregation personal model
        orig = c.local_test
        c.local_test = (lambda o, cid: (lambda w, b=True: (finals.__setitem__(cid, copy.deepcopy(w)), o(w, b))[1]))(orig, c.client_idx)
    random.seed(7); np.random.seed(7)
    rapi.train()

    model = _model()
    model.load_state_dict(init)
    api = FedFomoAPI(ds + [val], "cpu", args, SlabTrainer(model, args, _log(), device="cpu", loss_kind="ce"), _log())
    api.record_information = lambda *a, **k: None
    random.seed(7); np.random.seed(7)
    api.train()
    for c in range(n):
        _assert_state_close(api.client_list[c].slab, finals[c], rtol=1e-3, atol=1e-4, what=f"personal[{c}]")
    for key in ("old_mask_test_acc", "new_mask_test_acc"):
        assert len(api.stat_info[key]) == len(rapi.stat_info[key]) == 3
        for a, b in zip(api.stat_info[key], rapi.stat_info[key]):
            assert abs(float(a) - float(b)) < 1e-6, key
