"""The stagers tell the engine which samples every training batch holds (``engine.batch_ids``, row order) -- what the sm_100a engine
keys its per-sample Gram-matrix cache on.  Checked on the CPU path with a recording engine."""
import logging

import numpy as np
import torch

from helpers import make_args
from neuroimagedisttraining_b200.data import SyntheticABCD, load_partition_data_abcd
from neuroimagedisttraining_b200.fl import SlabTrainer
from neuroimagedisttraining_b200.fl.engine import TorchEngine
from neuroimagedisttraining_b200.models import create_model


class _Recorder(TorchEngine):
    def __init__(self, *a, store=None, **k):
        super().__init__(*a, **k)
        self.batch_ids = None
        self.batch_ids_key = None
        self.seen = []
        self._store = store

    def train_step(self, x, y, *a, **k):
        ids, self.batch_ids = self.batch_ids, None
        assert ids is not None and len(ids) == x.shape[0] and self.batch_ids_key == id(self._store)
        want, _ = self._store.fetch(ids)
        assert torch.equal(want.to(x.device), x if x.dtype == torch.uint8 else want.to(x.device))     # row i of the batch IS sample ids[i]
        self.seen.append(np.asarray(ids).copy())
        return super().train_step(x, y, *a, **k)


def test_every_training_batch_is_announced_with_its_sample_ids():
    store = SyntheticABCD(24, (32, 32, 32), n_sites=3, seed=1, site_sizes=[8, 8, 8])
    args = make_args(client_num_in_total=3, batch_size=4, epochs=2, comm_round=1, lr=0.01, model="tiny3DCNN")
    log = logging.getLogger("t"); log.setLevel(logging.ERROR)
    ds = load_partition_data_abcd(store, client_number=3, batch_size=4)
    torch.manual_seed(0)
    model = create_model("tiny3DCNN", 1, (32, 32, 32))
    eng = _Recorder(model, "cpu", volumetric=True, input_scale=store.scale, store=store)
    tr = SlabTrainer(model, args, log, device=torch.device("cpu"), store=store, engine=eng)
    train_loader = ds[5][0]
    tr.set_id(0)
    tr.train(train_loader, torch.device("cpu"), args, round_idx=0)          # on the trainer's own work slab
    per_epoch = len(train_loader)
    assert len(eng.seen) == 2 * per_epoch
    want = set(int(i) for i in train_loader.indices)
    for e in range(2):
        got = np.concatenate(eng.seen[e * per_epoch:(e + 1) * per_epoch])
        assert set(int(i) for i in got) == want and len(got) == len(want)       # every sample of the client once per epoch
