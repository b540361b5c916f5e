"""Smaller components: partitioners, utilities, experimental models, DARTS, SyncBN, preprocessing, FLOP counter."""
import numpy as np
import pytest
import torch

from neuroimagedisttraining_b200.data import (SyntheticABCD, load_partition_data_abcd_rescale, partition_labels, rescale_partition, site_partition)
from neuroimagedisttraining_b200.data import preprocess as P
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout, create_model, experimental as E, MODEL_NAMES
from neuroimagedisttraining_b200.models.darts import DARTS_V2, NetworkCIFAR, search
from neuroimagedisttraining_b200.runtime.slab import ParamSlab
from neuroimagedisttraining_b200.utils import (AsymmetricTopologyManager, RobustAggregator, SymmetricTopologyManager, count_inference_flops,
                                               count_training_flops, cosine_annealing, hamming_distance, model_difference)


def test_site_partition_matches_reference_rule():
    sites = np.repeat(np.arange(3), [10, 7, 5])
    tr, te, ids = site_partition(sites)
    for s, (a, b) in enumerate(zip(tr, te)):
        idx = np.where(sites == s)[0]
        np.random.seed(42)
        np.random.shuffle(idx)
        n_test = int(len(idx) * 0.2)
        assert list(a) == list(idx[:len(idx) - n_test]) and list(b) == list(idx[len(idx) - n_test:])
    tr2, te2 = rescale_partition(tr, te, 4)
    assert all(len(x) == sum(map(len, tr)) // 4 for x in tr2) and len(tr2) == 4


@pytest.mark.parametrize("method,alpha", [("homo", 0), ("hetero", 0.5), ("dir", 0.3), ("n_cls", 3), ("my_part", 2)])
def test_label_partitioners_cover_disjointly(method, alpha):
    y = np.random.RandomState(0).randint(0, 10, size=1000)
    parts = partition_labels(y, method, 8, alpha, seed=1)
    allidx = np.concatenate([parts[c] for c in range(8)])
    assert allidx.min() >= 0 and allidx.max() < 1000
    if method in ("n_cls", "my_part"):
        # as in the reference, every client gets its full quota: an exhausted class pool is refilled, so samples repeat
        assert all(len(parts[c]) == 125 for c in range(8))
    else:
        assert len(np.unique(allidx)) == len(allidx) and len(allidx) >= 900
    if method == "n_cls":
        assert all(len(np.unique(y[parts[c]])) <= 3 for c in range(8))


def test_rescale_loader_84_clients():
    store = SyntheticABCD(21 * 24, (8, 8, 8), n_sites=21, seed=0, site_sizes=[24] * 21)
    ds = load_partition_data_abcd_rescale(store, client_number=84, batch_size=4)
    assert len(ds[5]) == 84 and all(v == ds[4][0] for v in ds[4].values())


def test_flop_counter_handles_conv3d_and_sparsity():
    m = AlexNet3D_Dropout(1)
    dense = count_inference_flops(m, "ABCD")
    assert abs(dense / 1e9 - 7.47) < 0.05                     # 7.45 GFLOP forward per sample
    assert count_training_flops(m, "ABCD") == 3 * dense
    with torch.no_grad():
        m.features[4].weight.mul_((torch.rand_like(m.features[4].weight) > 0.5).float())
    assert count_inference_flops(m, "ABCD") < dense
    assert count_inference_flops(create_model("cnn_cifar10", 10), "cifar10") > 0


def test_slim_utils_and_robust_aggregation():
    args = type("A", (), {"anneal_factor": 0.5, "comm_round": 10})()
    assert cosine_annealing(args, 0) == 0.5 and abs(cosine_annealing(args, 10)) < 1e-9
    a, b = {"w": torch.tensor([1., 0., 1.])}, {"w": torch.tensor([1., 1., 0.])}
    d, tot = hamming_distance(a, b)
    assert int(d) == 2 and tot == 3
    assert float(model_difference(a, b)) == 2.0
    ra = RobustAggregator(type("A", (), {"norm_bound": 1.0, "stddev": 0.1, "defense_type": "norm_diff_clipping"})())
    g = {"w": torch.zeros(4), "bn.running_mean": torch.zeros(2)}
    l = {"w": torch.full((4,), 2.0), "bn.running_mean": torch.ones(2)}
    c = ra.norm_diff_clipping(l, g)
    assert abs(float(c["w"].norm()) - 1.0) < 1e-5 and torch.equal(c["bn.running_mean"], l["bn.running_mean"])
    m = create_model("tiny3DCNN", 1, (32, 32, 32))
    s1, s2 = ParamSlab.from_module(m, bind=False), ParamSlab.from_module(m, bind=False)
    s1.params_flat.add_(1.0)
    ra.norm_diff_clipping(s1, s2)
    Pn = s1.layout.param_numel
    assert abs(float((s1.flat[:Pn] - s2.flat[:Pn]).norm()) - 1.0) < 1e-3
    assert ra.add_noise(torch.zeros(1000)).std() > 0.05


def test_topology_managers():
    t = SymmetricTopologyManager(8, 4)
    t.generate_topology()
    assert np.allclose(t.topology.sum(1), 1) and np.allclose(t.topology, t.topology.T)
    assert t.get_in_neighbor_idx_list(0) == [1, 2, 6, 7] == t.get_out_neighbor_idx_list(0)
    a = AsymmetricTopologyManager(8, 2, 3, seed=0)
    a.generate_topology()
    assert np.allclose(a.topology.sum(1), 1) and not np.allclose(a.topology, a.topology.T)


def test_model_registry_builds_everything():
    for name in MODEL_NAMES:
        shape = (32, 32, 32) if name == "tiny3DCNN" else (121, 145, 121)
        classes = 1 if "3D" in name else (100 if name == "cnn_cifar100" else 10)
        m = create_model(name, classes, shape)
        assert sum(p.numel() for p in m.parameters()) > 0
    r = create_model("3DResNet", 1)
    with torch.no_grad():
        out = r.eval()(torch.randn(1, 1, 121, 145, 121))
    assert out.shape == (1, 1)      # FC sized from the input (the reference's hard-coded 9216 would fail here)


def test_experimental_models_and_darts():
    x = torch.randn(2, 3, 32, 32)
    for f in (E.resnet20, E.resnet29_ip, E.cnn_cifar10_meta, E.ResNet20):
        assert f()(x).shape == (2, 10)
    assert E.ResNet20()(x, [0] * 9).shape == (2, 10)
    mp = E.MetaPruningResNet20(10)                       # hyper-network generated weights, widths picked per call
    assert mp(x).shape == (2, 10) and len(E.META_CHANNEL_SCALES) == 31 and mp.num_blocks == 6
    y = mp(x, [30, 10, 10, 20, 20, 5, 5], [3, 7, 11, 15, 19, 23])
    y.sum().backward()
    assert y.shape == (2, 10) and all(p.grad is not None for p in mp.parameters())
    assert not any(k.endswith("conv1.weight") for k in mp.state_dict())      # no block owns a convolution weight
    ref = torch.nn.GroupNorm(2, 8)
    xx = torch.randn(2, 8, 5, 5)
    torch.testing.assert_close(E.GroupNorm2d(8, 4, affine=True)(xx), ref(xx), rtol=1e-4, atol=1e-5)
    net = NetworkCIFAR(C=8, layers=5, genotype=DARTS_V2)
    assert net(x).shape == (2, 10)
    loader = [(torch.randn(4, 3, 16, 16), torch.randint(0, 10, (4,))) for _ in range(2)]
    for unrolled in (False, True):
        model, geno = search(loader, loader, C=4, layers=3, epochs=1, unrolled=unrolled, max_batches=1)
        assert len(geno.normal) == 8 and all(op != "none" for op, _ in geno.normal)


def test_preprocessing_pipeline(tmp_path):
    rng = np.random.RandomState(0)
    vols = [np.clip(rng.rand(6, 7, 6) + (0.5 if i % 2 else 0.0), 0, None).astype(np.float32) for i in range(6)]
    for v in vols:
        v[0] = 0.0                                       # background plane
    mask = P.group_mean_mask(vols, 0.2)
    assert not mask[0].any() and mask[1:].all()
    q = P.minmax_quantize_u8(vols[0], mask)
    assert q.dtype == np.uint8 and q[0].max() == 0 and q.max() == 255 and q[mask].min() == 0
    store = P.build_store(vols, [i % 2 for i in range(6)], ["s2", "s1", "s2", "s1", "s3", "s3"])
    assert list(store.sites) == [1, 0, 1, 0, 2, 2] and store.volumes.shape == (6, 6, 7, 6)
    P.save_npz(store, str(tmp_path / "abcd.npz"))
    back = P.load_npz(str(tmp_path / "abcd.npz"))
    assert torch.equal(back.volumes, store.volumes) and list(back.labels) == list(store.labels)


def test_executor_plan_ctypes_mirror_matches_the_c_structs():
    """The ctypes mirror of nidt_plan_t / nidt_block_t (ops/engine_b200.py) must have the C layout (the library loads without a GPU)."""
    import ctypes as C
    from neuroimagedisttraining_b200.ops import native
    from neuroimagedisttraining_b200.ops.engine_b200 import _BlockC, _PlanC
    lib = native.load()
    for fn in (lib.nidt_plan_sizeof, lib.nidt_block_sizeof, lib.nidt_plan_offsetof):
        fn.restype = C.c_int
    lib.nidt_plan_offsetof.argtypes = [C.c_int]
    assert C.sizeof(_PlanC) == lib.nidt_plan_sizeof()
    assert C.sizeof(_BlockC) == lib.nidt_block_sizeof()
    for field, attr in ((0, _PlanC.blocks), (1, _PlanC.off_fc1_w), (2, _PlanC.head_s), (3, _PlanC.p_drop), (4, _PlanC.loss_acc),
                        (5, _PlanC.norm_scratch), (10, _BlockC.off_w), (11, _BlockC.y), (12, _BlockC.wf), (13, _BlockC.stats_tiles),
                        (14, _BlockC.box), (15, _BlockC.dbox)):
        assert attr.offset == lib.nidt_plan_offsetof(field), field


def test_native_choose_box_matches_python_and_partial_batches_fit_the_buffers():
    """The executor picks M-tile boxes for ragged last batches itself: same choice as the Python search, and never more
    tiles than the full batch (the BN-statistics buffers are sized for the full batch)."""
    import ctypes as C
    from neuroimagedisttraining_b200.ops import native
    from neuroimagedisttraining_b200.ops.engine_b200 import choose_box
    lib = native.load()
    lib.nidt_choose_box.argtypes = [C.c_int] * 4 + [C.POINTER(C.c_int)]
    lib.nidt_choose_box.restype = C.c_int
    out = (C.c_int * 4)()
    tiles = lambda W, H, D, N, b: -(-W // b[0]) * -(-H // b[1]) * -(-D // b[2]) * -(-N // b[3])
    for (W, H, D) in ((17, 21, 17), (19, 23, 19), (5, 7, 5), (32, 38, 32), (16, 19, 16), (8, 10, 8), (3, 4, 3), (1, 2, 1)):
        for full in (16, 8, 4):
            t_full = tiles(W, H, D, full, choose_box(W, H, D, full))
            for n in range(1, full + 1):
                assert lib.nidt_choose_box(W, H, D, n, out) == 0
                assert tuple(out) == choose_box(W, H, D, n), (W, H, D, n)
                assert tiles(W, H, D, n, tuple(out)) <= t_full, (W, H, D, n, full)


def test_halo_kernel_shape_support_query_runs_without_a_gpu():
    from neuroimagedisttraining_b200.ops import volumetric as V
    lib = V._lib()
    assert lib.nidt_conv3d_halo_supported(8, 32, 38, 32, 64, 64, 1, 64) == 1       # ResNet layer1
    assert lib.nidt_conv3d_halo_supported(16, 17, 21, 17, 64, 128, 0, 128) == 1    # AlexNet conv2
    assert lib.nidt_conv3d_halo_supported(3, 4, 5, 4, 256, 256, 1, 128) == 0       # four Cin blocks of planes exceed shared memory
    assert lib.nidt_conv3d_halo_supported(2, 8, 8, 200, 64, 64, 1, 64) == 0        # row pitch above 128
    assert lib.nidt_conv3d_halo_supported(2, 8, 8, 8, 32, 64, 1, 64) == 0          # Cin not a multiple of 64


def test_batch_augment_crops_flips_and_cuts_reproducibly():
    import itertools
    from neuroimagedisttraining_b200.data.loaders import BatchAugment, TensorBatchLoader, load_partition_data_images
    x = torch.arange(3 * 2 * 6 * 6).float().view(3, 2, 6, 6) + 1
    y = BatchAugment(2, True, 0)(x, torch.Generator().manual_seed(0))
    for i in range(3):                                   # each output is a shifted, maybe mirrored window of the zero-padded input
        xp = torch.nn.functional.pad(x[i], (2, 2, 2, 2))
        wins = [xp[:, dy:dy + 6, dx:dx + 6] for dy, dx in itertools.product(range(5), range(5))]
        assert any(torch.equal(w, y[i]) or torch.equal(w.flip(2), y[i]) for w in wins)
    holes = (BatchAugment(0, False, 4)(torch.ones(4, 2, 8, 8), torch.Generator().manual_seed(1)) == 0).sum(dim=(1, 2, 3))
    assert ((holes > 0) & (holes <= 2 * 16)).all()
    a = [b[0] for b in TensorBatchLoader(x, torch.arange(3), 2, True, seed=5, augment=BatchAugment())]
    b = [b[0] for b in TensorBatchLoader(x, torch.arange(3), 2, True, seed=5, augment=BatchAugment())]
    assert all(torch.equal(p, q) for p, q in zip(a, b))
    ds = load_partition_data_images("cifar10", None, "homo", 0, 2, 8, synthetic_size=(64, 16), augment=True, cutout=8)
    xb, yb = next(iter(ds[5][0]))
    assert xb.shape == (8, 3, 32, 32) and ds[5][0].augment is not None and ds[6][0].augment is None


def test_nifti_reader_roundtrip_scaling_and_byte_order(tmp_path):
    rng = np.random.RandomState(0)
    vol = rng.rand(7, 9, 5).astype(np.float32)
    for name in ("a.nii", "a.nii.gz"):
        P.save_nifti(vol, str(tmp_path / name))
        assert np.array_equal(P.load_nifti(str(tmp_path / name)), vol)
    # big-endian int16 file with an intensity scaling and a trailing singleton dimension, written by hand
    q = rng.randint(-500, 500, size=(4, 3, 2, 1)).astype(">i2")
    hdr = bytearray(352)
    hdr[0:4] = np.array(348, ">i4").tobytes()
    hdr[40:56] = np.array([4, 4, 3, 2, 1, 1, 1, 1], ">i2").tobytes()
    hdr[70:72] = np.array(4, ">i2").tobytes()
    hdr[72:74] = np.array(16, ">i2").tobytes()
    hdr[108:120] = np.array([352, 0.5, 10.0], ">f4").tobytes()
    hdr[344:348] = b"n+1\x00"
    (tmp_path / "be.nii").write_bytes(bytes(hdr) + q.tobytes(order="F"))
    got = P.load_nifti(str(tmp_path / "be.nii"))
    assert got.shape == (4, 3, 2) and np.allclose(got, q[..., 0].astype(np.float32) * 0.5 + 10.0)
    with pytest.raises(ValueError, match="NIfTI"):
        (tmp_path / "bad.nii").write_bytes(b"\x00" * 400)
        P.load_nifti(str(tmp_path / "bad.nii"))
    paths = []
    for i in range(4):
        v = np.clip(rng.rand(6, 7, 6) + 0.3 * (i % 2), 0, None).astype(np.float32)
        v[0] = 0
        P.save_nifti(v, str(tmp_path / f"s{i}.nii.gz"))
        paths.append(str(tmp_path / f"s{i}.nii.gz"))
    store = P.build_store_from_nifti(paths, [0, 1, 0, 1], ["b", "a", "b", "a"])
    assert store.volumes.shape == (4, 6, 7, 6) and store.volumes.dtype == torch.uint8 and list(store.sites) == [1, 0, 1, 0]


def test_hdf5_volume_store_against_a_stand_in_h5py(monkeypatch):
    """h5py is not in this image; a dict-backed stand-in with h5py's indexing rules (increasing unique indices only) exercises the
    reader: volumes in one file, labels / sites in another (the reference's two files), rows returned in the requested order."""
    import sys
    import types
    from neuroimagedisttraining_b200.data import HDF5VolumeStore
    from neuroimagedisttraining_b200.data.loaders import load_partition_data_abcd

    class DS:
        def __init__(self, a):
            self.a, self.shape, self.dtype = a, a.shape, a.dtype

        def __getitem__(self, k):
            if isinstance(k, tuple) and k == ():
                return self.a
            k = np.asarray(k)
            if k.ndim == 1 and (np.diff(k) <= 0).any():
                raise TypeError("Indexing elements must be in increasing order")
            return self.a[k]

    rng = np.random.RandomState(0)
    files = {"x.h5": {"X": DS(rng.randint(0, 256, size=(12, 4, 5, 3)).astype(np.uint8))},
             "l.h5": {"y": DS((rng.rand(12) > 0.5).astype(np.float64)), "site": DS(np.repeat(np.arange(3), 4))}}
    monkeypatch.setitem(sys.modules, "h5py", types.SimpleNamespace(File=lambda path, mode="r": files[path]))
    store = HDF5VolumeStore("x.h5", "l.h5")
    assert store.shape == (4, 5, 3) and store.labels.dtype == np.float32 and list(store.sites[:5]) == [0, 0, 0, 0, 1]
    x, y = store.fetch([7, 2, 2, 11])
    assert torch.equal(x, torch.from_numpy(files["x.h5"]["X"].a[[7, 2, 2, 11]])) and list(y) == list(store.labels[[7, 2, 2, 11]])
    out = torch.zeros(6, 4, 5, 3, dtype=torch.uint8)
    x2, _ = store.fetch([5, 0], out=out)
    assert x2.data_ptr() == out.data_ptr() and torch.equal(out[1], torch.from_numpy(files["x.h5"]["X"].a[0]))
    ds = load_partition_data_abcd(store, batch_size=2)
    assert len(ds[5]) == 3 and sum(ds[4].values()) == 3 * 4 - 0 - 0       # 4 per site, int(4 * 0.2) = 0 test samples each
    files["f.h5"] = {"X": DS(rng.rand(3, 2, 2, 2).astype(np.float32)), "y": DS(np.zeros(3))}      # float volumes are quantised on the fly
    fs = HDF5VolumeStore("f.h5")
    xf, _ = fs.fetch([2, 0])
    assert xf.dtype == torch.uint8 and torch.equal(xf[0], torch.from_numpy(np.rint(files["f.h5"]["X"].a[2] * 255).astype(np.uint8)))


def test_c_abi_header_matches_the_exported_symbols():
    """``include/nidt_b200.h`` declares exactly the entry points the library exports (and the library exports nothing else)."""
    import glob
    import os
    import re
    import subprocess
    from neuroimagedisttraining_b200.ops import native
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = set(re.findall(r"\b(nidt_\w+)\s*\(", open(os.path.join(root, "include", "nidt_b200.h")).read()))
    defined = set()
    for f in glob.glob(os.path.join(root, "neuroimagedisttraining_b200", "csrc", "*.cu")):
        defined |= set(m.group(1) for m in re.finditer(r"NIDT_API\s+[\w\s\*]+?\b(nidt_\w+)\s*\(", open(f).read()))
    assert header == defined and len(header) > 50
    if native.available():
        out = subprocess.run(["nm", "-D", "--defined-only", str(native.lib_path())], capture_output=True, text=True).stdout
        exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
        assert exported == header


def test_tiny_imagenet_folder_reader_and_loader(tmp_path):
    """The Tiny-ImageNet folder layout (train/<wnid>/images/*.JPEG, val/images + val_annotations.txt) is decoded once, cached, and
    feeds the 200-class loader path (``tiny_imagenet/datasets.py:20-145``); 3 classes x 4 images of 64x64 stand in for the dataset."""
    from PIL import Image
    from neuroimagedisttraining_b200.data.loaders import load_partition_data_images
    from neuroimagedisttraining_b200.data.tiny_imagenet import TinyImageNet
    rng = np.random.RandomState(0)
    root = tmp_path / "tiny"
    wnids = ["n01", "n02", "n03"]
    ann = []
    for ci, w in enumerate(wnids):
        (root / "train" / w / "images").mkdir(parents=True)
        for i in range(4):
            Image.fromarray(np.full((64, 64, 3), 40 * ci + i, np.uint8)).save(root / "train" / w / "images" / f"{w}_{i}.JPEG", quality=100)
    (root / "val" / "images").mkdir(parents=True)
    for i in range(6):
        Image.fromarray(rng.randint(0, 255, (64, 64, 3)).astype(np.uint8)).save(root / "val" / "images" / f"val_{i}.JPEG")
        ann.append(f"val_{i}.JPEG\t{wnids[i % 3]}\t0\t0\t10\t10")
    (root / "val" / "val_annotations.txt").write_text("\n".join(ann) + "\n")
    tr = TinyImageNet(str(root), True)
    assert len(tr) == 12 and sorted(set(tr.targets.tolist())) == [0, 1, 2] and tr.data.shape == (12, 64, 64, 3)
    assert (root / "train_cache.pkl").exists()
    x, y = tr.tensors()
    assert x.shape == (12, 3, 64, 64) and x.dtype == torch.float32 and y.dtype == torch.int64
    again = TinyImageNet(str(root), True, dataidxs=np.array([0, 5, 11]))                     # from the cache, truncated
    assert len(again) == 3 and np.array_equal(again.data, tr.data[[0, 5, 11]])
    va = TinyImageNet(str(root), False)
    assert va.targets.tolist() == [0, 1, 2, 0, 1, 2]
    ds = load_partition_data_images("tiny", str(root), "homo", 0, 2, 4)
    assert ds[0] == 12 and ds[1] == 6 and sum(ds[4].values()) == 12
    xb, yb = next(iter(ds[5][0]))
    assert xb.shape[1:] == (3, 64, 64)


def test_cifar_pickle_reader_and_no_silent_fallback(tmp_path):
    """CIFAR-10 / 100 from the python-pickle layout; a given but wrong ``data_dir`` raises instead of switching to synthetic data."""
    import pickle
    from neuroimagedisttraining_b200.data.loaders import load_partition_data_images
    rng = np.random.RandomState(0)
    d10 = tmp_path / "c10" / "cifar-10-batches-py"
    d10.mkdir(parents=True)
    rows = {}
    for name in [f"data_batch_{i}" for i in range(1, 6)] + ["test_batch"]:
        rows[name] = (rng.randint(0, 256, (20, 3072)).astype(np.uint8), rng.randint(0, 10, 20).tolist())
        with open(d10 / name, "wb") as f:
            pickle.dump({b"data": rows[name][0], b"labels": rows[name][1]}, f)
    ds = load_partition_data_images("cifar10", str(tmp_path / "c10"), "homo", 0, 4, 8)
    assert ds[0] == 100 and ds[1] == 20 and sum(ds[4].values()) == 100
    x0 = ds[2].x[0]                                       # first training image, normalised with the reference's statistics
    mean, std = torch.tensor([0.49139968, 0.48215827, 0.44653124]).view(3, 1, 1), torch.tensor([0.24703233, 0.24348505, 0.26158768]).view(3, 1, 1)
    want = (torch.from_numpy(rows["data_batch_1"][0][0].reshape(3, 32, 32)).float() / 255 - mean) / std
    torch.testing.assert_close(x0, want)
    d100 = tmp_path / "cifar-100-python"
    d100.mkdir()
    for name, n in (("train", 60), ("test", 30)):
        with open(d100 / name, "wb") as f:
            pickle.dump({b"data": rng.randint(0, 256, (n, 3072)).astype(np.uint8), b"fine_labels": rng.randint(0, 100, n).tolist()}, f)
    ds = load_partition_data_images("cifar100", str(d100), "homo", 0, 3, 8)          # the dataset folder itself also works
    assert ds[0] == 60 and ds[1] == 30
    with pytest.raises(FileNotFoundError, match="python-pickle layout"):
        load_partition_data_images("cifar10", str(tmp_path / "nowhere"), "homo", 0, 2, 8)


def test_phase_timer_and_cpu_batch_stager():
    import time
    from neuroimagedisttraining_b200.data.loaders import IndexLoader
    from neuroimagedisttraining_b200.runtime.staging import BatchStager
    from neuroimagedisttraining_b200.utils.timing import PhaseTimer
    t = PhaseTimer("cpu")
    for _ in range(3):
        with t.phase("train"):
            time.sleep(0.01)
    with pytest.raises(RuntimeError):
        with t.phase("eval"):                              # a phase that raises is still accounted for
            raise RuntimeError("boom")
    s = t.summary()
    assert s["train"]["calls"] == 3 and s["train"]["total_s"] >= 0.03 and s["eval"]["calls"] == 1
    store = SyntheticABCD(12, (6, 7, 5), n_sites=2, seed=0, site_sizes=[6, 6])
    stager = BatchStager(store, torch.device("cpu"), batch_size=4)
    idx = np.array([9, 2, 7, 0, 11, 5])
    loader = IndexLoader(idx, store.labels[idx], store.sites[idx], 4, shuffle=False)
    got = list(stager.iterate(loader))
    assert [len(y) for _, y in got] == [4, 2]
    first = np.sort(idx[:4])                               # volumes come in sorted-index order, labels re-read to match (as the reference)
    want_x, want_y = store.fetch(first)
    assert torch.equal(got[0][0], want_x) and torch.equal(got[0][1], want_y)


def test_roofline_script_on_the_committed_step_profile():
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("roofline", os.path.join(root, "scripts", "roofline.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    L = mod.layer_table(16)
    assert [r["M"] for r in L] == [16 * 247151, 16 * 6069, 16 * 175, 16 * 175, 16 * 175] and [r["K"] for r in L] == [125, 1728, 3456, 5184, 5184]
    text = mod.main([os.path.join(root, "profiles", "profile_step_v18.txt")])
    assert "| `conv3d_fprop_tc[1]` |" in text and "| math |" in text and "| memory |" in text


def test_syncbn_single_process_matches_batchnorm_with_cma_momentum_and_without_affine():
    """momentum=None (cumulative moving average) and affine=False go through convert_sync_batchnorm unchanged and must behave like
    torch.nn.BatchNorm3d (one process: the cross-rank reduction is the identity)."""
    import copy
    from neuroimagedisttraining_b200.parallel.syncbn import convert_sync_batchnorm
    torch.manual_seed(0)
    for kw in (dict(momentum=None), dict(affine=False), dict(momentum=None, affine=False), dict()):
        ref = torch.nn.BatchNorm3d(5, **kw)
        mod = convert_sync_batchnorm(copy.deepcopy(ref))
        for _ in range(3):
            x = torch.randn(4, 5, 3, 3, 3, requires_grad=True)
            x2 = x.detach().clone().requires_grad_(True)
            y, y2 = ref(x), mod(x2)
            y.square().sum().backward(); y2.square().sum().backward()
            torch.testing.assert_close(y2, y, rtol=1e-5, atol=1e-6)
            torch.testing.assert_close(x2.grad, x.grad, rtol=1e-4, atol=1e-5)
        torch.testing.assert_close(mod.running_mean, ref.running_mean, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(mod.running_var, ref.running_var, rtol=1e-5, atol=1e-6)
