"""Every FL algorithm end-to-end on CPU with tiny synthetic data (plumbing + invariants)."""
import numpy as np
import pytest
import torch

from neuroimagedisttraining_b200.data import SyntheticABCD, load_partition_data_abcd, load_partition_data_images
from neuroimagedisttraining_b200.fl import (DisPFLAPI, DittoAPI, DPSGDAPI, FedAvgAPI, FedFomoAPI, LocalAPI, SalientGradsAPI, SlabTrainer,
                                            SubAvgAPI, TurboAggregateAPI, choose_neighbors, fake_prune)
from neuroimagedisttraining_b200.models import create_model
from neuroimagedisttraining_b200.runtime.slab import make_mask_slab
from neuroimagedisttraining_b200.utils import mpc

from helpers import make_args, quiet_logger


def _vol_setup(n_clients=4, **kw):
    store = SyntheticABCD(n_clients * 10, (32, 32, 32), n_sites=n_clients, seed=1, site_sizes=[10] * n_clients)
    args = make_args(client_num_in_total=n_clients, batch_size=8, epochs=1, comm_round=2, lr=0.01, **kw)
    ds = load_partition_data_abcd(store, client_number=n_clients, batch_size=8)
    torch.manual_seed(0)
    model = create_model("tiny3DCNN", 1, (32, 32, 32))
    tr = SlabTrainer(model, args, quiet_logger(), device="cpu", store=store)
    return ds, args, tr


def _img_setup(n_clients=4, val=False, **kw):
    args = make_args(client_num_in_total=n_clients, batch_size=16, epochs=2, comm_round=2, lr=0.05, dataset="cifar10", **kw)
    ds = load_partition_data_images("cifar10", None, "dir", 0.5, n_clients, 16, synthetic_size=(240, 80), val_split=0.1 if val else 0.0)
    torch.manual_seed(0)
    model = create_model("lenet5_cifar", 10)
    tr = SlabTrainer(model, args, quiet_logger(), device="cpu", loss_kind="ce")
    return ds, args, tr


def test_fedavg_finetune_round_and_sampling():
    ds, args, tr = _vol_setup(frac=0.5)
    args.client_num_per_round = 2
    api = FedAvgAPI(ds, "cpu", args, tr, quiet_logger())
    st = api.train()
    # 2 rounds + the final fine-tune round are evaluated
    assert len(st["global_test_acc"]) == 3 and len(st["person_test_acc"]) == 3
    # deterministic sampling: np.random.seed(round) + choice without replacement
    np.random.seed(1)
    expect = np.random.choice(range(4), 2, replace=False)
    assert list(api._client_sampling(1, 4, 2)) == list(expect)
    assert list(api._client_sampling(0, 4, 4)) == [0, 1, 2, 3]


def test_salientgrads_mask_density_and_masked_weights():
    ds, args, tr = _vol_setup(dense_ratio=0.3, itersnip_iteration=2)
    api = SalientGradsAPI(ds, "cpu", args, tr, quiet_logger())
    api.train()
    lay = api.layout
    kept = sum(float(api.mask.view(n).sum()) for n in lay.prunable_names)
    assert kept == int(lay.prunable_numel() * 0.3)
    # aggregated weights are zero wherever the mask is zero
    for n in lay.prunable_names:
        assert float((api.w_global.view(n)[api.mask.view(n) == 0]).abs().max()) == 0.0
    # snip_mask=False: mask generated then replaced by ones (dense ablation of the reference)
    ds, args, tr = _vol_setup(snip_mask=False)
    api = SalientGradsAPI(ds, "cpu", args, tr, quiet_logger())
    api.train()
    assert float(api.mask.flat.sum()) == lay.numel_real(params_only=True)


def test_orchestrator_helper_methods_of_the_reference_api():
    """``_benefit_choose`` / ``average_weights`` / ``record_avg_inference_flops``: carried (uncalled) by the reference's orchestrator
    (``sailentgrads_api.py:165-209,319-331``) and therefore part of the surface a user may call."""
    ds, args, tr = _vol_setup(dense_ratio=0.5)
    api = SalientGradsAPI(ds, "cpu", args, tr, quiet_logger())
    assert api._benefit_choose(0, 1, 5, 5) == [0, 1, 2, 3, 4]
    assert list(api._benefit_choose(0, 0, 5, 2, cs="ring")) == [4, 1]
    active = np.array([1, 0, 1, 1, 1])
    assert list(api._benefit_choose(0, 2, 5, 2, cs="full", active_ths_rnd=active)) == [0, 3, 4]
    np.random.seed(0)
    assert 3 not in api._benefit_choose(0, 3, 5, 2, cs="random")
    a, b = api.w_global.clone(), api.w_global.clone()
    b.flat.mul_(3.0)
    torch.testing.assert_close(api.average_weights([a, b.state_dict()]).flat, 2.0 * a.flat)
    mask = api.generate_global_mask_snip()
    dense = api.record_avg_inference_flops(api.w_global)
    sparse = api.record_avg_inference_flops(api.w_global, {c: mask for c in range(args.client_num_in_total)})
    assert 0 < sparse < dense and api.stat_info["avg_inference_flops"] == sparse


def test_salientgrads_stratified_and_ci_flag():
    ds, args, tr = _vol_setup(stratified_sampling=True, ci=1)
    api = SalientGradsAPI(ds, "cpu", args, tr, quiet_logger())
    st = api.train()
    assert len(st["global_test_acc"]) == 3 and all(0.0 <= a <= 1.0 for a in st["global_test_acc"])


@pytest.mark.parametrize("API", [LocalAPI, DittoAPI, DPSGDAPI, FedFomoAPI, SubAvgAPI, TurboAggregateAPI])
def test_image_algorithms_run(API):
    extra = dict(lamda=0.5, local_epochs=1, cs="ring", each_prune_ratio=0.1, dist_thresh=1e-4, acc_thresh=0.0, dense_ratio=0.5)
    ds, args, tr = _img_setup(val=(API is FedFomoAPI), **extra)
    args.final_finetune = False
    api = API(ds, "cpu", args, tr, quiet_logger())
    st = api.train()
    assert len(st["person_test_acc"]) >= 1 or len(st["global_test_acc"]) >= 1
    assert torch.isfinite(api.w_global.flat).all()


def test_turboaggregate_equals_fedavg_up_to_quantisation():
    ds, args, tr = _img_setup()
    args.final_finetune = False
    args.comm_round = 1
    a = FedAvgAPI(ds, "cpu", args, tr, quiet_logger())
    w0 = a.w_global.clone()
    a.train()
    ds, args2, tr2 = _img_setup()
    args2.final_finetune = False
    args2.comm_round = 1
    b = TurboAggregateAPI(ds, "cpu", args2, tr2, quiet_logger())
    b.w_global.copy_(w0)
    b.train()
    assert (a.w_global.flat - b.w_global.flat).abs().max().item() < 4 * 4 / (1 << 16)


def test_dispfl_masks_keep_density_and_weights_stay_masked():
    ds, args, tr = _vol_setup(dense_ratio=0.5, active=1.0, cs="random", anneal_factor=0.5, frac=0.5)
    args.client_num_per_round = 2
    api = DisPFLAPI(ds, "cpu", args, tr, quiet_logger())
    st = api.train()
    lay = api.layout
    for cid, m in enumerate(api.masks_per_client):
        w = api.client_list[cid].slab
        for n in lay.prunable_names:
            mv, wv = m.view(n), w.view(n)
            assert set(mv.unique().tolist()) <= {0.0, 1.0}
    assert len(st["mask_dis_matrix"]) == 4 and st["mask_dis_matrix"][0][0] == 0
    assert len(st["old_mask_test_acc"]) == 2
    # prune + regrow keeps the number of active weights of a tensor constant
    name = lay.prunable_names[0]
    m0 = api.masks_per_client[0]
    fired, removed = api.fire_mask(m0, api.client_list[0].slab, 0)
    regrown = api.regrow_mask(fired, removed, None)
    assert float(regrown.view(name).sum()) == float(m0.view(name).sum())
    assert float(fired.view(name).sum()) == float(m0.view(name).sum()) - removed[name]


def test_dispfl_neighbor_aggregation_flag():
    ds, args, tr = _vol_setup(dense_ratio=0.5, active=0.75, cs="random", anneal_factor=0.5, frac=0.5, static=True)
    args.client_num_per_round = 2
    args.neighbor_aggregation = True
    np.random.seed(0)
    api = DisPFLAPI(ds, "cpu", args, tr, quiet_logger())
    api.train()
    assert torch.isfinite(api.client_list[0].slab.flat).all()


def test_choose_neighbors_topologies():
    np.random.seed(0)
    assert list(choose_neighbors("ring", 0, 5, 2)) == [4, 1]
    assert 3 not in choose_neighbors("random", 3, 6, 2)
    assert list(choose_neighbors("full", 1, 4, 2, np.array([1, 1, 0, 1]))) == [0, 3]
    assert list(choose_neighbors("ring", 2, 4, 4)) == [0, 1, 2, 3]


def test_fake_prune_removes_requested_fraction():
    _, args, tr = _img_setup()
    slab = tr.engine.new_slab()
    mask = make_mask_slab(tr.layout, "cpu")
    m2 = fake_prune(0.2, slab, mask)
    for n in tr.layout.prunable_names:
        frac = 1.0 - float(m2.view(n).mean())
        assert abs(frac - 0.2) < 0.02, (n, frac)
    for n in tr.layout.param_names:
        if not tr.layout[n].prunable:
            assert float(m2.view(n).min()) == 1.0


def test_mpc_primitives():
    p = mpc.DEFAULT_PRIME
    rng = np.random.RandomState(0)
    for a in (1, 2, 12345, p - 1):
        assert a * mpc.modular_inv(a, p) % p == 1
    X = rng.randint(0, 1000, size=(4, 6))
    shares = mpc.BGW_encoding(X, N=5, T=2, p=p, rng=rng)
    assert np.array_equal(mpc.BGW_decoding(shares[[0, 2, 4]], [0, 2, 4], p)[0], X)
    enc = mpc.LCC_encoding(X, N=6, K=2, T=1, p=p, rng=rng)
    dec = mpc.LCC_decoding(enc[[0, 1, 2]], 1, 6, 2, 1, [0, 1, 2], p)
    assert np.array_equal(dec.reshape(X.shape), X)
    ss = mpc.Gen_Additive_SS(10, 4, p, rng)
    assert np.all(np.mod(ss.sum(0), p) == 0)
    x = rng.randn(50)
    assert np.allclose(mpc.dequantize(mpc.quantize(x, 1 << 16, p), 1 << 16, p), x, atol=1 / (1 << 16))
    sk_a, sk_b, g = 1234567, 7654321, 5
    assert mpc.my_key_agreement(sk_a, mpc.my_pk_gen(sk_b, p, g), p, g) == mpc.my_key_agreement(sk_b, mpc.my_pk_gen(sk_a, p, g), p, g)


def test_checkpoint_resume_roundtrip(tmp_path):
    ds, args, tr = _vol_setup()
    args.checkpoint_path = str(tmp_path / "ck.pt")
    args.checkpoint_every = 1
    api = SalientGradsAPI(ds, "cpu", args, tr, quiet_logger())
    api.train()
    ds, args2, tr2 = _vol_setup()
    api2 = SalientGradsAPI(ds, "cpu", args2, tr2, quiet_logger())
    start = api2.resume(args.checkpoint_path)
    assert start == 2
    assert torch.equal(api2.w_global.flat, api.w_global.flat)
    assert torch.equal(api2.mask.flat, api.mask.flat)
    ck = torch.load(args.checkpoint_path, weights_only=False)
    ref_keys = list(create_model("tiny3DCNN", 1, (32, 32, 32)).state_dict().keys())
    assert list(ck["state_dict"].keys()) == ref_keys and ck["state_dict"]["features.1.num_batches_tracked"].dtype == torch.int64


def test_resumed_run_is_bit_identical_to_an_uninterrupted_one(tmp_path):
    """2 rounds + checkpoint + resume for 2 more == 4 rounds in one go: global model, mask and accuracy series, with dropout active
    (global torch RNG) and shuffled loaders (private generators) -- both streams are part of the checkpoint."""
    from neuroimagedisttraining_b200.models import Tiny3DCNN

    store = SyntheticABCD(40, (32, 32, 32), n_sites=4, seed=1, site_sizes=[10] * 4)

    def build(comm_round, ck=None):
        args = make_args(client_num_in_total=4, batch_size=4, epochs=1, comm_round=comm_round, lr=0.01, dense_ratio=0.5)
        ds = load_partition_data_abcd(store, client_number=4, batch_size=4)
        torch.manual_seed(5)
        model = Tiny3DCNN(1, in_shape=(32, 32, 32), p_drop=0.3)
        tr = SlabTrainer(model, args, quiet_logger(), device="cpu", store=store)
        args.checkpoint_every, args.checkpoint_path, args.save_personal, args.final_finetune = 1, ck, True, False
        return SalientGradsAPI(ds, "cpu", args, tr, quiet_logger())

    torch.manual_seed(9); np.random.seed(9)
    full = build(4, str(tmp_path / "full.pt"))
    full.train()
    torch.manual_seed(9); np.random.seed(9)
    first = build(2, str(tmp_path / "part.pt"))
    first.train()
    torch.manual_seed(12345); np.random.seed(777)        # a fresh process would start from unrelated RNG states
    second = build(4, str(tmp_path / "part.pt"))
    assert second.resume(str(tmp_path / "part.pt")) == 2
    second.train()
    assert torch.equal(second.mask.flat, full.mask.flat)
    assert torch.equal(second.w_global.flat, full.w_global.flat)
    assert second.stat_info["global_test_acc"] == full.stat_info["global_test_acc"]
    assert second.stat_info["person_test_acc"] == full.stat_info["person_test_acc"]


@pytest.mark.parametrize("API,extra", [(LocalAPI, {}), (DittoAPI, dict(lamda=0.5, local_epochs=1)), (DPSGDAPI, dict(cs="random")),
                                       (FedFomoAPI, {}), (SubAvgAPI, dict(each_prune_ratio=0.2, dist_thresh=1e-4, acc_thresh=-1.0, dense_ratio=0.3)),
                                       (DisPFLAPI, dict(dense_ratio=0.5, different_initial=True, cs="random", active=0.8))])
def test_per_client_algorithms_resume_bit_identically(tmp_path, API, extra):
    """Personal models, the initial model, per-client masks (DisPFL, Sub-FedAvg) and FedFomo's mixing weights travel in the
    checkpoint: 2 rounds + resume + 2 rounds equals 4 rounds, bit for bit, for every algorithm that keeps per-client state."""
    import random

    def build(comm_round, ck):
        ds, args, tr = _img_setup(n_clients=4, val=(API is FedFomoAPI), frac=0.5, **extra)
        args.comm_round, args.client_num_per_round, args.epochs = comm_round, 2, 1
        args.checkpoint_every, args.checkpoint_path, args.final_finetune = 1, ck, False
        api = API(ds, "cpu", args, tr, quiet_logger())
        api.record_information = lambda *a, **k: None
        return api

    def seed(v):
        random.seed(v); np.random.seed(v); torch.manual_seed(v)

    seed(3)
    full = build(4, str(tmp_path / "full.pt"))
    full.train()
    class Killed(Exception):
        pass

    seed(3)
    first = build(4, str(tmp_path / "part.pt"))          # same job (schedules depend on comm_round), killed after round 1's checkpoint
    orig = first._maybe_checkpoint

    def checkpoint_then_die(r):
        orig(r)
        if r == 1:
            raise Killed
    first._maybe_checkpoint = checkpoint_then_die
    with pytest.raises(Killed):
        first.train()
    seed(999)
    second = build(4, str(tmp_path / "part.pt"))
    assert second.resume(str(tmp_path / "part.pt")) == 2
    second.train()
    assert torch.equal(second.w_global.flat, full.w_global.flat)
    for a, b in zip(second.client_list, full.client_list):
        assert (a.slab is None) == (b.slab is None)
        if a.slab is not None:
            assert torch.equal(a.slab.flat, b.slab.flat), a.client_idx
    for key in ("test_acc", "person_test_acc", "old_mask_test_acc", "new_mask_test_acc", "global_test_acc"):
        assert [float(v) for v in second.stat_info.get(key, [])] == [float(v) for v in full.stat_info.get(key, [])], key
    if API is DisPFLAPI:
        assert all(torch.equal(x.flat, y.flat) for x, y in zip(second.masks_per_client, full.masks_per_client))
    if API is SubAvgAPI:
        assert all(torch.equal(x.flat, y.flat) for x, y in zip(second.mask_pers, full.mask_pers))


def test_fedavg_with_norm_clipping_and_weak_dp_defences():
    """``--defense_type``: every upload is pulled back to within ``norm_bound`` of the global model before it is averaged."""
    ds, args, tr = _vol_setup()
    args.final_finetune, args.comm_round = False, 1
    plain = FedAvgAPI(ds, "cpu", args, tr, quiet_logger())
    start = plain.w_global.clone()
    plain.train()
    P = plain.layout.param_numel
    moved = float((plain.w_global.flat[:P] - start.flat[:P]).norm())
    ds, args, tr = _vol_setup()
    args.final_finetune, args.comm_round, args.defense_type, args.norm_bound = False, 1, "norm_diff_clipping", moved / 4
    clipped = FedAvgAPI(ds, "cpu", args, tr, quiet_logger())
    clipped.train()
    for c in clipped.client_list:
        assert float((c.slab.flat[:P] - start.flat[:P]).norm()) <= moved / 4 * (1 + 1e-4)
    assert float((clipped.w_global.flat[:P] - start.flat[:P]).norm()) <= moved / 4 * (1 + 1e-4) < moved
    ds, args, tr = _vol_setup()
    args.final_finetune, args.comm_round, args.defense_type, args.norm_bound, args.stddev = False, 1, "weak_dp", 100.0, 0.01
    torch.manual_seed(1)
    noisy = FedAvgAPI(ds, "cpu", args, tr, quiet_logger())
    noisy.train()
    d = noisy.w_global.flat[:P] - plain.w_global.flat[:P]
    real = d[plain.w_global.flat[:P] != 0]
    assert 0.001 < float(real.std()) < 0.01              # mean of 4 clients' N(0, 0.01^2) noise: 0.005
    with pytest.raises(ValueError, match="defense_type"):
        args.defense_type = "bogus"
        FedAvgAPI(ds, "cpu", args, tr, quiet_logger()).train()


@pytest.mark.parametrize("cs", ["topology_sym", "topology_asym"])
def test_dpsgd_gossip_with_topology_manager_mixing_matrix(cs):
    """D-PSGD whose neighbour step is a row of the FedML topology managers' mixing matrix: one round from identical models leaves the
    plain-mean global model where row-stochastic mixing of identical models must leave it, and models stay finite / distinct."""
    ds, args, tr = _img_setup(n_clients=6, cs=cs, frac=0.5)
    args.client_num_per_round, args.comm_round = 2, 2
    api = DPSGDAPI(ds, "cpu", args, tr, quiet_logger())
    W = api._mixing_matrix(6)
    assert W.shape == (6, 6) and np.allclose(W.sum(1), 1.0) and (np.diag(W) > 0).all()
    if cs == "topology_sym":
        assert np.allclose(W, W.T)
    st = api.train()
    assert len(st["global_test_acc"]) == 2 and torch.isfinite(api.w_global.flat).all()
    assert not torch.equal(api.client_list[0].slab.flat, api.client_list[3].slab.flat)
    mean = sum(c.slab.flat for c in api.client_list) / 6
    torch.testing.assert_close(api.w_global.flat, mean, rtol=1e-5, atol=1e-6)
