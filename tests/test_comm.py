"""Message-passing runtime: in-process and gRPC transports, and FedAvg over messages == the slab aggregation."""
import threading
import time

import pytest
import torch

from neuroimagedisttraining_b200.comm import (FedAvgClientManager, FedAvgServerManager, GrpcCommManager, InProcessCommManager, Message,
                                              MqttCommManager)


def test_message_roundtrip():
    m = Message("t", 1, 2)
    m.add_params("w", {"a": torch.arange(6).float().view(2, 3)})
    m2 = Message.from_bytes(m.to_bytes())
    assert m2.get_type() == "t" and m2.get_sender_id() == 1 and m2.get_receiver_id() == 2
    assert torch.equal(m2.get("w")["a"], m.get("w")["a"])
    assert '"a": [[0.0, 1.0, 2.0]' in m.to_json()


def _run_fedavg(make_comm, n_clients=3, rounds=2):
    torch.manual_seed(0)
    state = {"w": torch.randn(50), "b": torch.randn(5)}
    ns = [3, 5, 9][:n_clients]

    def train_fn(i):
        def fn(sd):
            return ns[i], {k: v + (i + 1) * 0.1 for k, v in sd.items()}
        return fn
    size = n_clients + 1
    server = FedAvgServerManager(state, rounds, comm=make_comm(0, size), rank=0, size=size)
    clients = [FedAvgClientManager(train_fn(i), comm=make_comm(i + 1, size), rank=i + 1, size=size) for i in range(n_clients)]
    threads = [threading.Thread(target=m.run, daemon=True) for m in [server] + clients]
    for t in threads:
        t.start()
    time.sleep(0.2)
    server.start()
    assert server.done.wait(30)
    for t in threads:
        t.join(10)
    # expected: every round adds sum_i n_i/N * (i+1)*0.1
    delta = sum(n * (i + 1) * 0.1 for i, n in enumerate(ns)) / sum(ns)
    for k in state:
        torch.testing.assert_close(server.state[k], state[k] + rounds * delta, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(clients[0].final_state[k], server.state[k])


def test_fedavg_over_inprocess_messages():
    _run_fedavg(lambda r, s: InProcessCommManager(r, s, network="t1"))


def test_fedavg_over_grpc_loopback():
    pytest.importorskip("grpc")
    _run_fedavg(lambda r, s: GrpcCommManager("127.0.0.1", 52310, r, s), n_clients=2, rounds=1)


def test_fedavg_over_mqtt_with_the_builtin_broker():
    """MQTT transport end to end: the in-tree broker on a free loopback port, one topic per participant (``fedml<id>``), model-sized
    payloads (multi-byte remaining-length field) and the same FedAvg result as the other transports."""
    from neuroimagedisttraining_b200.comm import MiniMqttBroker
    broker = MiniMqttBroker("127.0.0.1", 0).start()
    try:
        _run_fedavg(lambda r, s: MqttCommManager("127.0.0.1", broker.port, r, s - 1), n_clients=3, rounds=2)
        assert broker.published >= 2 * 3 * 2
    finally:
        broker.stop()


def test_mqtt_wire_format_and_fanout():
    import queue
    from neuroimagedisttraining_b200.comm import MiniMqttBroker, MiniMqttClient
    from neuroimagedisttraining_b200.comm import mqtt_mini as M
    assert M._encode_len(0) == b"\x00" and M._encode_len(127) == b"\x7f" and M._encode_len(128) == b"\x80\x01" and M._encode_len(16384) == b"\x80\x80\x01"
    assert M._publish_packet("a/b", b"hi") == b"\x30\x07\x00\x03a/bhi"            # fixed header, topic length, topic, payload
    broker = MiniMqttBroker("127.0.0.1", 0).start()
    try:
        got = [queue.Queue(), queue.Queue()]
        subs = [MiniMqttClient(f"s{i}", on_message=lambda t, b, i=i: got[i].put((t, b)), keepalive=0).connect("127.0.0.1", broker.port) for i in range(2)]
        for c in subs:
            c.subscribe("room")
        subs[1].subscribe("only1")
        pub = MiniMqttClient("p", keepalive=0).connect("127.0.0.1", broker.port)
        big = bytes(range(256)) * 1200                                             # 307 200 bytes -> 3-byte length field
        pub.publish("room", big)
        pub.publish("only1", b"x")
        pub.publish("nobody", b"dropped")
        assert got[0].get(timeout=5) == ("room", big) and got[1].get(timeout=5) == ("room", big)
        assert got[1].get(timeout=5) == ("only1", b"x") and got[0].empty()
        for c in subs + [pub]:
            c.disconnect()
    finally:
        broker.stop()


def test_mqtt_paho_request_reports_missing_dependency():
    try:
        import paho.mqtt.client  # noqa: F401
        pytest.skip("paho-mqtt is installed")
    except ImportError:
        with pytest.raises(ImportError, match="paho-mqtt"):
            MqttCommManager("127.0.0.1", 1883, 0, 1, impl="paho")


def test_salientgrads_over_messages_matches_a_sequential_oracle():
    """Saliency phase (clients upload |g*w|, server emits the global top-k mask) + masked FedAvg rounds over the message runtime."""
    import copy
    import torch.nn as nn
    from neuroimagedisttraining_b200.comm import SalientGradsClientManager, SalientGradsServerManager
    from neuroimagedisttraining_b200.ops import reference
    torch.manual_seed(0)
    net = nn.Sequential(nn.Linear(10, 8), nn.ReLU(), nn.Linear(8, 1))
    init = copy.deepcopy(net.state_dict())
    data = [(torch.randn(12 + 4 * i, 10), torch.randn(12 + 4 * i, 1)) for i in range(3)]

    def score_fn(i):
        def fn(sd):
            m = copy.deepcopy(net); m.load_state_dict(sd)
            nn.functional.mse_loss(m(data[i][0]), data[i][1]).backward()
            return {k: (p.grad * p).abs().detach() for k, p in m.named_parameters() if p.dim() > 1}
        return fn

    def train_fn(i):
        def fn(sd, mask):
            m = copy.deepcopy(net); m.load_state_dict(sd)
            opt = torch.optim.SGD(m.parameters(), lr=0.05)
            for _ in range(2):
                opt.zero_grad(); nn.functional.mse_loss(m(data[i][0]), data[i][1]).backward(); opt.step()
                with torch.no_grad():
                    for k, p in m.named_parameters():
                        if k in mask:
                            p.mul_(mask[k])
            return len(data[i][0]), {k: v.detach().clone() for k, v in m.state_dict().items()}
        return fn

    size, rounds, dense = 4, 2, 0.5
    server = SalientGradsServerManager(init, rounds, dense, comm=InProcessCommManager(0, size, network="sg"), rank=0, size=size)
    clients = [SalientGradsClientManager(score_fn(i), train_fn(i), comm=InProcessCommManager(i + 1, size, network="sg"), rank=i + 1, size=size)
               for i in range(3)]
    threads = [threading.Thread(target=m.run, daemon=True) for m in [server] + clients]
    for t in threads:
        t.start()
    time.sleep(0.2)
    server.start()
    assert server.done.wait(30)
    for t in threads:
        t.join(10)
    # oracle: the same protocol, sequentially
    scores = [score_fn(i)(init) for i in range(3)]
    names = list(scores[0])
    mean = torch.cat([sum(s[k] for s in scores).flatten() / 3 for k in names])
    mask_flat = torch.empty_like(mean)
    reference.global_topk_mask(mean, [(0, mean.numel())], dense, mask_flat)
    off, mask = 0, {}
    for k in names:
        mask[k] = mask_flat[off:off + scores[0][k].numel()].view_as(scores[0][k]); off += scores[0][k].numel()
    state = init
    for _ in range(rounds):
        ups = [train_fn(i)(state, mask) for i in range(3)]
        total = sum(n for n, _ in ups)
        state = {k: sum(w[k] * (n / total) for n, w in ups) for k in state}
    for k in names:
        assert torch.equal(server.mask[k], mask[k])
        assert float((server.state[k] * (1 - mask[k])).abs().sum()) == 0.0
    kept = sum(float(m.sum()) for m in server.mask.values())
    assert kept >= int(mean.numel() * dense)
    for k in state:
        torch.testing.assert_close(server.state[k], state[k], rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(clients[1].final_state[k], state[k], rtol=1e-5, atol=1e-6)
