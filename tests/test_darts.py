"""DARTS / GDAS package (reference: ``fedml_api/model/cv/darts/*.py``, dead code there): search networks, genotype networks for CIFAR
and ImageNet resolutions, helpers, training loop, DOT export; the GDAS genotype rule is checked against the reference class."""
import numpy as np
import pytest
import torch

from baseline import ref_harness
from neuroimagedisttraining_b200.models import darts
from neuroimagedisttraining_b200.models.darts import utils as U, visualize as V


def _loader(n=2, bs=4, hw=16, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(bs, 3, hw, hw, generator=g), torch.randint(0, 10, (bs,), generator=g)) for _ in range(n)]


def test_gdas_samples_one_op_per_edge_and_learns_alphas():
    torch.manual_seed(0)
    net = darts.NetworkGumbelSoftmax(C=4, layers=3, tau=5.0)
    x, y = _loader(1)[0]
    loss = net._loss(x, y)
    loss.backward()
    assert net(x).shape == (4, 10)
    for a in net.arch_parameters():                       # straight-through estimator reaches every alpha row
        assert a.grad is not None and (a.grad.abs().sum(dim=1) > 0).all()
    w = torch.nn.functional.gumbel_softmax(net.alphas_normal, net.tau, hard=True)
    assert torch.equal(w.sum(dim=1), torch.ones(w.shape[0])) and ((w == 0) | (w == 1)).all()
    g, n_conv, r_conv = net.genotype_with_counts()
    assert len(g.normal) == len(g.reduce) == 8 and 0 <= n_conv <= 8 and 0 <= r_conv <= 8
    clone = net.new()
    assert all(torch.equal(a, b) for a, b in zip(clone.arch_parameters(), net.arch_parameters())) and clone.get_tau() == net.get_tau()
    model, geno = darts.search(_loader(), _loader(seed=1), C=4, layers=3, epochs=2, max_batches=1, arch="gdas", tau_max=8.0, tau_min=2.0)
    assert model.get_tau() == 2.0 and all(op != "none" for op, _ in geno.normal + geno.reduce)


def test_genotype_networks_train_and_infer():
    torch.manual_seed(0)
    model, hist = darts.train(_loader(hw=32), _loader(hw=32, seed=1), genotype=darts.FedNAS_V1, C=4, layers=3, auxiliary=True, epochs=2,
                              max_batches=1)
    assert len(hist) == 2 and hist[1]["lr"] < hist[0]["lr"] and model.drop_path_prob > 0
    model.train()
    out, aux = model(torch.randn(2, 3, 32, 32))
    assert out.shape == aux.shape == (2, 10)
    model.eval()
    assert model(torch.randn(2, 3, 32, 32)).shape == (2, 10)
    inet = darts.NetworkImageNet(C=8, num_classes=7, layers=3, auxiliary=True)
    inet.train()
    out, aux = inet(torch.randn(2, 3, 224, 224))
    assert out.shape == aux.shape == (2, 7)


def test_helpers(tmp_path):
    m = U.AvgrageMeter()
    m.update(1.0, 2); m.update(4.0, 1)
    assert m.avg == 2.0 and m.cnt == 3
    logits = torch.tensor([[0.1, 0.9, 0.0], [0.8, 0.1, 0.1], [0.2, 0.3, 0.5]])
    top1, top2 = U.accuracy(logits, torch.tensor([1, 1, 2]), topk=(1, 2))
    assert abs(float(top1) - 200 / 3) < 1e-4 and abs(float(top2) - 100.0) < 1e-4
    np.random.seed(0)
    img = torch.ones(3, 16, 16)
    cut = U.Cutout(8)(img)
    assert 0 < int((cut == 0).sum()) <= 3 * 64 and torch.equal(img, torch.ones(3, 16, 16))       # input left untouched
    torch.manual_seed(0)
    x = torch.ones(64, 2, 3, 3)
    d = U.drop_path(x, 0.5)
    per_sample = d.flatten(1)
    assert ((per_sample == 0).all(1) | (per_sample == 2).all(1)).all() and U.drop_path(x, 0.0) is x
    net = darts.NetworkCIFAR(C=4, layers=3, auxiliary=True)
    assert U.count_parameters_in_MB(net) < sum(p.numel() for p in net.parameters()) / 1e6
    U.create_exp_dir(str(tmp_path / "exp"), scripts_to_save=[__file__])
    U.save_checkpoint({"epoch": 1}, True, str(tmp_path / "exp"))
    assert (tmp_path / "exp" / "model_best.pth.tar").exists() and (tmp_path / "exp" / "scripts" / "test_darts.py").exists()
    U.save(net, str(tmp_path / "w.pt"))
    other = darts.NetworkCIFAR(C=4, layers=3, auxiliary=True)
    U.load(other, str(tmp_path / "w.pt"))
    assert all(torch.equal(a, b) for a, b in zip(net.state_dict().values(), other.state_dict().values()))


def test_genotype_dot_export(tmp_path):
    dot = V.to_dot(darts.DARTS_V2.normal, "normal")
    assert dot.startswith("digraph normal {") and dot.count("->") == 8 + 4 and '"c_{k-2}" -> "0" [label="sep_conv_3x3"]' in dot
    assert "dil_conv_3x3" in V.describe(darts.DARTS_V2)
    path = V.plot(darts.DARTS_V2.reduce, str(tmp_path / "reduce"))
    assert open(path).read().count("max_pool_3x3") == 5


@pytest.mark.skipif(ref_harness.reference_available() is not None, reason="reference not installed")
def test_genotype_rule_matches_the_reference_search_networks():
    """Same alphas -> same discrete architecture as the reference's softmax (DARTS) and Gumbel-softmax (GDAS) networks."""
    ref_harness.import_reference()
    from fedml_api.model.cv.darts.model_search import Network as RefNet
    from fedml_api.model.cv.darts.model_search_gdas import Network_GumbelSoftmax as RefGdas
    torch.manual_seed(3)
    ours, gd = darts.SearchNetwork(C=4, layers=3), darts.NetworkGumbelSoftmax(C=4, layers=3)
    ref = RefNet(4, 10, 3, torch.nn.CrossEntropyLoss(), "cpu")
    rgd = RefGdas(4, 10, 3, torch.nn.CrossEntropyLoss(), "cpu")
    a_n, a_r = torch.randn(14, 8), torch.randn(14, 8)
    for net in (ours, gd, ref, rgd):
        net.alphas_normal.data.copy_(a_n); net.alphas_reduce.data.copy_(a_r)
    want = ref.genotype()
    want = want[0] if isinstance(want, tuple) else want
    for got in (ours.genotype(), gd.genotype()):
        assert list(got.normal) == list(want.normal) and list(got.reduce) == list(want.reduce)
        assert list(got.normal_concat) == list(want.normal_concat)
    g, nc, rc = gd.genotype_with_counts()
    rg, rnc, rrc = rgd.genotype()
    assert list(g.normal) == list(rg.normal) and (nc, rc) == (rnc, rrc)
    sd = {k: v for k, v in ref.state_dict().items()}
    assert set(ours.state_dict()) == set(sd)                                 # same parameter names as the reference's search network
