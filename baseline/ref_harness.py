"""Harness that runs the UNMODIFIED reference (installed under ``baseline/_ref``) offline.

Nothing of this repo's models / kernels / engine is on that path.  What the harness adds
(SURVEY Appendix C): stub modules for ``h5py`` / ``pudb`` (imported unconditionally by the
reference, absent here), replacement of the two HDF5 loaders that hard-code ``/data/users2/...``
paths with a synthetic in-memory source, and construction of the FedML 8-tuple.
"""
from __future__ import annotations

import logging
import sys
import types
from pathlib import Path
from typing import Optional

import numpy as np
import torch
from torch.utils.data import DataLoader, TensorDataset

REF_ROOT = Path(__file__).resolve().parent / "_ref"


def reference_available() -> Optional[str]:
    """None if importable, else a one-line reason."""
    if not (REF_ROOT / "fedml_api").is_dir():
        return f"reference not installed under {REF_ROOT} (see DESIGN.md: pip install --target baseline/_ref)"
    return None


def import_reference():
    why = reference_available()
    if why:
        raise ImportError(why)
    for name in ("h5py", "pudb"):
        if name not in sys.modules:
            try:
                __import__(name)
            except ImportError:
                sys.modules[name] = types.ModuleType(name)
    if str(REF_ROOT) not in sys.path:
        sys.path.insert(0, str(REF_ROOT))
    from fedml_api.model.cv.salient_models import AlexNet3D_Dropout  # noqa: F401
    from fedml_api.standalone.sailentgrads import my_model_trainer as sg_trainer, client as sg_client, sailentgrads_api as sg_api
    from fedml_api.standalone.fedavg import my_model_trainer as fa_trainer, client as fa_client, fedavg_api as fa_api
    return types.SimpleNamespace(AlexNet3D_Dropout=AlexNet3D_Dropout, sg_trainer=sg_trainer, sg_client=sg_client, sg_api=sg_api,
                                 fa_trainer=fa_trainer, fa_client=fa_client, fa_api=fa_api)


class SyntheticSource:
    """fp32 volumes on the reference's device, by sorted index -- what its HDF5 read would return.
    Volumes are the same 8-bit synthetic scans our arm uses, normalised by ``store.scale`` on the
    host, then copied H2D as fp32 (the reference's own ``torch.tensor(np, float32, device='cuda')``)."""

    def __init__(self, store, device):
        self.store = store
        self.device = torch.device(device)
        self.h2d_bytes = 0

    def batch(self, indexes: torch.Tensor):
        idx = np.sort(indexes.numpy().astype(int))
        x_u8, y = self.store.fetch(idx)
        xb = (x_u8.numpy().astype(np.float32) * np.float32(self.store.scale))
        X = torch.tensor(xb, dtype=torch.float32, device=self.device)
        Y = torch.tensor(y.numpy(), dtype=torch.float32, device=self.device)
        self.h2d_bytes += X.numel() * 4 + Y.numel() * 4
        return X, Y


def patch_loaders(ref, source: SyntheticSource):
    def load_data_chunks_batch(self, indexes):
        return source.batch(indexes)

    def load_data_chunks(self, train_mini_batch):
        X, Y = source.batch(train_mini_batch[0])
        train_mini_batch[0] = X
        train_mini_batch[1] = Y
        return train_mini_batch

    ref.sg_trainer.MyModelTrainer.load_data_chunks_batch = load_data_chunks_batch
    ref.fa_trainer.MyModelTrainer.load_data_chunks_batch = load_data_chunks_batch
    ref.sg_client.Client.load_data_chunks = load_data_chunks


def build_dataset(store, train_idx, test_idx, batch_size: int, shuffle: bool = True):
    local_num, train_local, test_local = {}, {}, {}
    for cid, (tr, te) in enumerate(zip(train_idx, test_idx)):
        def ds(idx):
            return TensorDataset(torch.as_tensor(idx, dtype=torch.float32), torch.as_tensor(store.labels[idx], dtype=torch.float32),
                                 torch.as_tensor(store.sites[idx], dtype=torch.float32))
        train_local[cid] = DataLoader(ds(tr), batch_size=batch_size, shuffle=shuffle)
        test_local[cid] = DataLoader(ds(te), batch_size=batch_size, shuffle=False)
        local_num[cid] = len(tr)
    return [None, None, None, None, local_num, train_local, test_local, 2]


def make_args(**kw):
    d = dict(model="3DCNN", dataset="ABCD", batch_size=16, client_optimizer="sgd", lr=0.01, lr_decay=0.998, wd=5e-4, epochs=2,
             client_num_in_total=4, frac=1.0, momentum=0.0, comm_round=1, frequency_of_the_test=1, gpu=0, ci=0, dense_ratio=0.5,
             anneal_factor=0.5, seed=1024, cs="v0", itersnip_iteration=1, stratified_sampling=False, active=1.0,
             snip_mask=True, erk_power_scale=1.0)
    d.update(kw)
    d.setdefault("client_num_per_round", int(d["client_num_in_total"] * d["frac"]))
    return types.SimpleNamespace(**d)


def quiet_logger(name="ref"):
    lg = logging.getLogger(name)
    lg.setLevel(logging.WARNING)
    return lg


def build_reference_api(algo: str, store, train_idx, test_idx, args, device, logger=None, shuffle=True, model_state=None, p_drop=None):
    """Returns (api, model, source).  ``algo``: 'salientgrads' | 'fedavg'."""
    ref = import_reference()
    source = SyntheticSource(store, device)
    patch_loaders(ref, source)
    logger = logger or quiet_logger()
    dataset = build_dataset(store, train_idx, test_idx, args.batch_size, shuffle)
    model = ref.AlexNet3D_Dropout(num_classes=1)
    if p_drop is not None:
        for m in model.modules():
            if isinstance(m, torch.nn.Dropout):
                m.p = p_drop
    if model_state is not None:
        model.load_state_dict(model_state)
    if algo == "salientgrads":
        model = model.to(device)   # main_sailentgrads.py:276
        trainer = ref.sg_trainer.MyModelTrainer(model, args, logger)
        api = ref.sg_api.SailentGradsAPI(dataset, torch.device(device), args, trainer, logger)
    elif algo == "fedavg":
        trainer = ref.fa_trainer.MyModelTrainer(model, args, logger)
        api = ref.fa_api.FedAvgAPI(dataset, torch.device(device), args, trainer, logger)
    else:
        raise ValueError(algo)
    return api, model, source
