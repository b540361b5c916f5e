import logging
import types



def quiet_logger(name="test"):
    lg = logging.getLogger(name)
    lg.setLevel(logging.ERROR)
    return lg


def make_args(**kw):
    d = dict(model="3DCNN", dataset="ABCD", batch_size=16, client_optimizer="sgd", lr=0.01, lr_decay=0.998, wd=5e-4, epochs=2,
             client_num_in_total=4, frac=1.0, momentum=0.0, comm_round=1, frequency_of_the_test=1, gpu=0, ci=0, dense_ratio=0.5,
             anneal_factor=0.5, seed=1024, cs="v0", itersnip_iteration=1, stratified_sampling=False, active=1.0,
             snip_mask=True, erk_power_scale=1.0)
    d.update(kw)
    d.setdefault("client_num_per_round", int(d["client_num_in_total"] * d["frac"]))
    return types.SimpleNamespace(**d)
