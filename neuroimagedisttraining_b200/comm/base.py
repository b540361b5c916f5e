"""Observer pattern + communication-manager interface (reference: ``fedml_core/distributed/communication/observer.py``,
``base_com_manager.py``) and three transports:

* :class:`InProcessCommManager` -- queues between threads of one process (tests, single-host simulation);
* :class:`TorchDistCommManager` -- ``torch.distributed`` point-to-point (gloo / NCCL); replaces the reference's
  mpi4py send/receive threads with 0.3 s / 3 ms polling (``mpi/com_manager.py:13-98``);
* :class:`GrpcCommManager` -- unary RPC with raw-bytes payloads on port ``base_port + rank``
  (``gRPC/grpc_comm_manager.py:20-106``; the reference's JSON-string payload and broken imports are not reproduced).

MQTT (``mqtt/mqtt_comm_manager.py``): :class:`MqttCommManager` runs on ``paho-mqtt`` when installed, else on the built-in
MQTT 3.1.1 client of :mod:`.mqtt_mini`, which also provides a small broker (the image has neither paho nor a broker).
"""
from __future__ import annotations

import queue
import threading
from abc import ABC, abstractmethod
from concurrent import futures
from typing import Dict, List, Optional

from .message import Message


class Observer(ABC):
    @abstractmethod
    def receive_message(self, msg_type, msg_params) -> None:
        ...


class BaseCommunicationManager(ABC):
    def __init__(self):
        self._observers: List[Observer] = []
        self._running = False

    def add_observer(self, observer: Observer):
        self._observers.append(observer)

    def remove_observer(self, observer: Observer):
        self._observers.remove(observer)

    def _notify(self, msg: Message):
        for o in list(self._observers):
            o.receive_message(msg.get_type(), msg)

    @abstractmethod
    def send_message(self, msg: Message):
        ...

    @abstractmethod
    def _poll(self, timeout: float) -> Optional[Message]:
        ...

    def handle_receive_message(self):
        """Blocking receive loop; returns after :meth:`stop_receive_message`."""
        self._running = True
        while self._running:
            msg = self._poll(0.05)
            if msg is not None:
                self._notify(msg)

    def stop_receive_message(self):
        self._running = False


class InProcessCommManager(BaseCommunicationManager):
    """All ranks live in one process; a shared dict of queues is the network."""

    _networks: Dict[str, Dict[int, "queue.Queue[bytes]"]] = {}
    _lock = threading.Lock()

    def __init__(self, rank: int, size: int, network: str = "default"):
        super().__init__()
        self.rank, self.size = rank, size
        with self._lock:
            net = self._networks.setdefault(network, {})
            for r in range(size):
                net.setdefault(r, queue.Queue())
        self.net = net

    def send_message(self, msg: Message):
        self.net[int(msg.get_receiver_id())].put(msg.to_bytes())

    def _poll(self, timeout):
        try:
            return Message.from_bytes(self.net[self.rank].get(timeout=timeout))
        except queue.Empty:
            return None


class TorchDistCommManager(BaseCommunicationManager):
    """Point-to-point messages over an initialised ``torch.distributed`` group: an 8-byte length header, then the payload."""

    def __init__(self, rank: int, size: int, device=None):
        super().__init__()
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank, self.size = rank, size
        self.device = device or torch.device("cpu")
        self._q: "queue.Queue[Message]" = queue.Queue()
        self._recv_thread = None

    def send_message(self, msg: Message):
        data = msg.to_bytes()
        t, d, dst = self.torch, self.dist, int(msg.get_receiver_id())
        d.send(t.tensor([len(data)], dtype=t.int64, device=self.device), dst)
        d.send(t.frombuffer(bytearray(data), dtype=t.uint8).to(self.device), dst)

    def _recv_loop(self):
        t, d = self.torch, self.dist
        while self._running:
            n = t.zeros(1, dtype=t.int64, device=self.device)
            src = d.recv(n)
            buf = t.empty(int(n.item()), dtype=t.uint8, device=self.device)
            d.recv(buf, src=src)
            self._q.put(Message.from_bytes(bytes(buf.cpu().numpy().tobytes())))

    def handle_receive_message(self):
        self._running = True
        self._recv_thread = threading.Thread(target=self._recv_loop, daemon=True)
        self._recv_thread.start()
        super().handle_receive_message()

    def _poll(self, timeout):
        try:
            return self._q.get(timeout=timeout)
        except queue.Empty:
            return None


class GrpcCommManager(BaseCommunicationManager):
    SERVICE, METHOD = "nidt.Comm", "SendMessage"

    def __init__(self, host: str, base_port: int, rank: int, size: int, ip_table: Optional[Dict[int, str]] = None, max_mb: int = 1024):
        super().__init__()
        import grpc
        self.grpc = grpc
        self.rank, self.size, self.base_port = rank, size, base_port
        self.ip_table = ip_table or {r: host for r in range(size)}
        self._q: "queue.Queue[bytes]" = queue.Queue()
        opts = [("grpc.max_send_message_length", max_mb << 20), ("grpc.max_receive_message_length", max_mb << 20)]
        self._opts = opts
        handler = grpc.method_handlers_generic_handler(self.SERVICE, {
            self.METHOD: grpc.unary_unary_rpc_method_handler(self._on_rpc, request_deserializer=None, response_serializer=None)})
        self.server = grpc.server(futures.ThreadPoolExecutor(max_workers=4), options=opts)
        self.server.add_generic_rpc_handlers((handler,))
        self.server.add_insecure_port(f"{host}:{base_port + rank}")
        self.server.start()

    def _on_rpc(self, request: bytes, context) -> bytes:
        self._q.put(request)
        return b"ok"

    def send_message(self, msg: Message):
        dst = int(msg.get_receiver_id())
        with self.grpc.insecure_channel(f"{self.ip_table[dst]}:{self.base_port + dst}", options=self._opts) as ch:
            ch.unary_unary(f"/{self.SERVICE}/{self.METHOD}", request_serializer=None, response_deserializer=None)(msg.to_bytes(), timeout=60)

    def _poll(self, timeout):
        try:
            return Message.from_bytes(self._q.get(timeout=timeout))
        except queue.Empty:
            return None

    def stop_receive_message(self):
        super().stop_receive_message()
        self.server.stop(0)


class MqttCommManager(BaseCommunicationManager):
    """Pub/sub transport: every participant subscribes to ``<topic><its id>`` and publishes to the receiver's topic
    (``mqtt_comm_manager.py:99-120``).  ``impl="auto"`` uses ``paho-mqtt`` when it is installed and otherwise the built-in MQTT 3.1.1
    client (:mod:`.mqtt_mini`, same wire protocol); a broker must be reachable -- :class:`.mqtt_mini.MiniMqttBroker` is one."""

    def __init__(self, host: str, port: int, client_id: int, client_num: int, topic: str = "fedml", impl: str = "auto"):
        super().__init__()
        self.client_id, self.client_num, self.topic = client_id, client_num, topic
        self._q: "queue.Queue[bytes]" = queue.Queue()
        mqtt = None
        if impl in ("auto", "paho"):
            try:
                import paho.mqtt.client as mqtt  # type: ignore
            except ImportError as e:
                if impl == "paho":
                    raise ImportError("impl='paho' needs the `paho-mqtt` package, which does not ship with this image; "
                                      "impl='auto' falls back to the built-in client") from e
        if mqtt is not None:
            self.client = mqtt.Client(client_id=str(client_id))
            self.client.on_message = lambda c, u, m: self._q.put(m.payload)
            self.client.connect(host, port)
            self.client.subscribe(f"{topic}{client_id}")
            self.client.loop_start()
            self._publish = lambda t, b: self.client.publish(t, payload=b)
            self._close = lambda: (self.client.loop_stop(), self.client.disconnect())
        else:
            from .mqtt_mini import MiniMqttClient
            self.client = MiniMqttClient(f"{topic}-{client_id}", on_message=lambda t, b: self._q.put(b)).connect(host, port)
            self.client.subscribe(f"{topic}{client_id}")
            self._publish = self.client.publish
            self._close = self.client.disconnect

    def send_message(self, msg: Message):
        self._publish(f"{self.topic}{int(msg.get_receiver_id())}", msg.to_bytes())

    def _poll(self, timeout):
        try:
            return Message.from_bytes(self._q.get(timeout=timeout))
        except queue.Empty:
            return None

    def stop_receive_message(self):
        super().stop_receive_message()
        self._close()
