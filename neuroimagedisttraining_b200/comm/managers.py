"""Client / server managers: register handlers per message type, run the receive loop, send
(reference: ``fedml_core/distributed/client/client_manager.py:13-73``, ``server/server_manager.py:13-68``).  Plus a complete
message-passing FedAvg (server + clients) built on them, which the reference's vendored runtime never had a caller for."""
from __future__ import annotations

import logging
import threading
from typing import Callable, Dict, Optional

import torch

from .base import BaseCommunicationManager, GrpcCommManager, InProcessCommManager, MqttCommManager, Observer, TorchDistCommManager
from .message import Message


def make_comm_manager(backend: str, rank: int, size: int, **kw) -> BaseCommunicationManager:
    backend = backend.upper()
    if backend in ("INPROC", "INPROCESS", "LOCAL"):
        return InProcessCommManager(rank, size, kw.get("network", "default"))
    if backend in ("MPI", "TORCH", "TORCHDIST", "GLOO", "NCCL"):
        return TorchDistCommManager(rank, size, kw.get("device"))
    if backend == "GRPC":
        return GrpcCommManager(kw.get("host", "127.0.0.1"), int(kw.get("base_port", 50000)), rank, size, kw.get("ip_table"))
    if backend == "MQTT":
        return MqttCommManager(kw.get("host", "127.0.0.1"), int(kw.get("port", 1883)), rank, size - 1, kw.get("topic", "fedml"), kw.get("impl", "auto"))
    raise ValueError(f"unknown communication backend {backend!r}")


class _Manager(Observer):
    def __init__(self, args=None, comm: Optional[BaseCommunicationManager] = None, rank: int = 0, size: int = 0, backend: str = "INPROC", **kw):
        self.args, self.rank, self.size, self.backend = args, rank, size, backend
        self.com_manager = comm if comm is not None else make_comm_manager(backend, rank, size, **kw)
        self.com_manager.add_observer(self)
        self.message_handler_dict: Dict[object, Callable[[Message], None]] = {}

    def register_message_receive_handlers(self) -> None:
        """Subclasses register their handlers here."""

    def register_message_receive_handler(self, msg_type, handler_callback_func) -> None:
        self.message_handler_dict[msg_type] = handler_callback_func

    def run(self) -> None:
        self.register_message_receive_handlers()
        self.com_manager.handle_receive_message()

    def get_sender_id(self) -> int:
        return self.rank

    def receive_message(self, msg_type, msg_params) -> None:
        handler = self.message_handler_dict.get(msg_type)
        if handler is None:
            logging.warning("rank %d: no handler for message type %r", self.rank, msg_type)
            return
        handler(msg_params)

    def send_message(self, message: Message) -> None:
        self.com_manager.send_message(message)

    def finish(self) -> None:
        self.com_manager.stop_receive_message()


class ClientManager(_Manager):
    pass


class ServerManager(_Manager):
    pass


# ------------------------------------------------------------------------- message-passing FedAvg
MSG_INIT, MSG_SYNC, MSG_UPLOAD, MSG_FINISH = "s2c_init", "s2c_sync", "c2s_upload", "s2c_finish"


class FedAvgServerManager(ServerManager):
    """Rank 0: broadcasts the global state_dict, collects (n_i, w_i) from every client, averages, repeats."""

    def __init__(self, state_dict, rounds: int, **kw):
        super().__init__(**kw)
        self.state = {k: v.clone() for k, v in state_dict.items()}
        self.rounds, self.round = rounds, 0
        self.uploads = {}
        self.done = threading.Event()

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_UPLOAD, self._on_upload)

    def start(self):
        for c in range(1, self.size):
            m = Message(MSG_INIT, 0, c)
            m.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, self.state)
            self.send_message(m)

    def _on_upload(self, msg: Message):
        self.uploads[msg.get_sender_id()] = (msg.get("num_samples"), msg.get(Message.MSG_ARG_KEY_MODEL_PARAMS))
        if len(self.uploads) < self.size - 1:
            return
        total = float(sum(n for n, _ in self.uploads.values()))
        ordered = [self.uploads[c] for c in sorted(self.uploads)]
        for k in self.state:
            acc = ordered[0][1][k].float() * (ordered[0][0] / total)
            for n, w in ordered[1:]:
                acc = acc + w[k].float() * (n / total)
            self.state[k] = acc.to(self.state[k].dtype) if self.state[k].is_floating_point() else acc
        self.uploads = {}
        self.round += 1
        last = self.round >= self.rounds
        for c in range(1, self.size):
            m = Message(MSG_FINISH if last else MSG_SYNC, 0, c)
            m.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, self.state)
            self.send_message(m)
        if last:
            self.done.set()
            self.finish()


class FedAvgClientManager(ClientManager):
    """Rank >= 1: on every model it receives, runs ``train_fn(state_dict) -> (num_samples, new_state_dict)`` and uploads."""

    def __init__(self, train_fn, **kw):
        super().__init__(**kw)
        self.train_fn = train_fn
        self.final_state = None

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_INIT, self._on_model)
        self.register_message_receive_handler(MSG_SYNC, self._on_model)
        self.register_message_receive_handler(MSG_FINISH, self._on_finish)

    def _on_model(self, msg: Message):
        n, state = self.train_fn(msg.get(Message.MSG_ARG_KEY_MODEL_PARAMS))
        up = Message(MSG_UPLOAD, self.rank, 0)
        up.add_params("num_samples", n)
        up.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, state)
        self.send_message(up)

    def _on_finish(self, msg: Message):
        self.final_state = msg.get(Message.MSG_ARG_KEY_MODEL_PARAMS)
        self.finish()


# ---------------------------------------------------------------------- message-passing SalientGrads
MSG_SCORE_REQUEST, MSG_SCORES, MSG_MASK = "s2c_score_request", "c2s_scores", "s2c_mask"


class SalientGradsServerManager(FedAvgServerManager):
    """SalientGrads over messages (``sailentgrads_api.py:47-66,86-147`` as a protocol): first every client scores the initial model
    and uploads ``|dL/dw * w|`` per prunable tensor; the server averages the scores unweighted, keeps the global top
    ``dense_ratio`` fraction (ties kept, as ``snip.get_mask_from_grads``) and sends the mask; then FedAvg rounds run, with the
    aggregated model multiplied by the mask before it goes out again."""

    def __init__(self, state_dict, rounds: int, dense_ratio: float, prunable: Optional[Callable[[str], bool]] = None, **kw):
        super().__init__(state_dict, rounds, **kw)
        self.dense_ratio = float(dense_ratio)
        self.prunable = prunable or (lambda k: k.endswith("weight") and self.state[k].dim() > 1)
        self.scores: Dict[int, dict] = {}
        self.mask: Optional[dict] = None

    def register_message_receive_handlers(self):
        super().register_message_receive_handlers()
        self.register_message_receive_handler(MSG_SCORES, self._on_scores)

    def start(self):
        for c in range(1, self.size):
            m = Message(MSG_SCORE_REQUEST, 0, c)
            m.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, self.state)
            self.send_message(m)

    def _on_scores(self, msg: Message):
        self.scores[msg.get_sender_id()] = msg.get("scores")
        if len(self.scores) < self.size - 1:
            return
        names = [k for k in self.state if self.prunable(k)]
        mean = {k: sum(self.scores[c][k].float() for c in sorted(self.scores)) / len(self.scores) for k in names}
        flat = torch.cat([mean[k].flatten() for k in names])
        norm = flat.sum()
        flat = flat / norm
        k = int(flat.numel() * self.dense_ratio)
        thr = torch.topk(flat, k, sorted=True)[0][-1] if k > 0 else torch.tensor(float("inf"))
        self.mask = {n: ((mean[n] / norm) >= thr).float() for n in names}
        for c in range(1, self.size):
            m = Message(MSG_MASK, 0, c)
            m.add_params("mask", self.mask)
            m.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, self.state)
            self.send_message(m)

    def _on_upload(self, msg: Message):
        before = self.round
        super()._on_upload(msg)
        if self.round != before and self.mask is not None:
            # every upload is already masked (clients multiply after each step), so the average is too up to round-off; pin the zeros
            for k, m in self.mask.items():
                self.state[k] = self.state[k] * m


class SalientGradsClientManager(FedAvgClientManager):
    """``score_fn(state_dict) -> {name: |g*w| tensor}`` for the saliency phase; ``train_fn(state_dict, mask) -> (n, new_state_dict)``
    afterwards (the client applies the mask after every step, as the reference's trainer does)."""

    def __init__(self, score_fn, train_fn, **kw):
        super().__init__(train_fn, **kw)
        self.score_fn = score_fn
        self.mask = None

    def register_message_receive_handlers(self):
        super().register_message_receive_handlers()
        self.register_message_receive_handler(MSG_SCORE_REQUEST, self._on_score_request)
        self.register_message_receive_handler(MSG_MASK, self._on_mask)

    def _on_score_request(self, msg: Message):
        up = Message(MSG_SCORES, self.rank, 0)
        up.add_params("scores", self.score_fn(msg.get(Message.MSG_ARG_KEY_MODEL_PARAMS)))
        self.send_message(up)

    def _on_mask(self, msg: Message):
        self.mask = msg.get("mask")
        self._on_model(msg)

    def _on_model(self, msg: Message):
        n, state = self.train_fn(msg.get(Message.MSG_ARG_KEY_MODEL_PARAMS), self.mask)
        up = Message(MSG_UPLOAD, self.rank, 0)
        up.add_params("num_samples", n)
        up.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, state)
        self.send_message(up)
