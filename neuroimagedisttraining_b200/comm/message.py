"""Typed key/value message exchanged between FL managers (reference: ``fedml_core/distributed/communication/message.py:5-74``).
The reference serialises to a JSON *string* (tensors as nested lists, gRPC payload cap 100 MB) or pickles the whole object
for MPI; here the wire format is ``torch.save`` bytes, so tensors travel as raw buffers."""
from __future__ import annotations

import io
import json
from typing import Any, Dict

import torch


class Message:
    MSG_ARG_KEY_OPERATION = "operation"
    MSG_ARG_KEY_TYPE = "msg_type"
    MSG_ARG_KEY_SENDER = "sender"
    MSG_ARG_KEY_RECEIVER = "receiver"
    MSG_OPERATION_SEND = "send"
    MSG_OPERATION_RECEIVE = "receive"
    MSG_OPERATION_BROADCAST = "broadcast"
    MSG_OPERATION_REDUCE = "reduce"
    MSG_ARG_KEY_MODEL_PARAMS = "model_params"

    def __init__(self, type: Any = 0, sender_id: int = 0, receiver_id: int = 0):  # noqa: A002
        self.type = type
        self.sender_id = sender_id
        self.receiver_id = receiver_id
        self.msg_params: Dict[str, Any] = {Message.MSG_ARG_KEY_TYPE: type, Message.MSG_ARG_KEY_SENDER: sender_id,
                                           Message.MSG_ARG_KEY_RECEIVER: receiver_id}

    def init(self, msg_params: Dict[str, Any]) -> None:
        self.msg_params = msg_params
        self.type = msg_params.get(Message.MSG_ARG_KEY_TYPE)
        self.sender_id = msg_params.get(Message.MSG_ARG_KEY_SENDER)
        self.receiver_id = msg_params.get(Message.MSG_ARG_KEY_RECEIVER)

    def init_from_json_string(self, json_string: str) -> None:
        self.init(json.loads(json_string))

    def get_sender_id(self):
        return self.sender_id

    def get_receiver_id(self):
        return self.receiver_id

    def add_params(self, key: str, value: Any) -> None:
        self.msg_params[key] = value

    add = add_params

    def get_params(self) -> Dict[str, Any]:
        return self.msg_params

    def get(self, key: str) -> Any:
        return self.msg_params.get(key)

    def get_type(self):
        return self.msg_params[Message.MSG_ARG_KEY_TYPE]

    def to_string(self) -> Dict[str, Any]:
        return self.msg_params

    def to_json(self) -> str:
        def conv(v):
            if torch.is_tensor(v):
                return v.detach().cpu().tolist()
            if isinstance(v, dict):
                return {k: conv(x) for k, x in v.items()}
            return v
        return json.dumps({k: conv(v) for k, v in self.msg_params.items()})

    def get_content(self) -> str:
        return f"{self.msg_params.get(Message.MSG_ARG_KEY_TYPE)}: {self.sender_id} -> {self.receiver_id} ({list(self.msg_params)})"

    # ------------------------------------------------------------------ wire format
    def to_bytes(self) -> bytes:
        buf = io.BytesIO()
        torch.save(self.msg_params, buf)
        return buf.getvalue()

    @classmethod
    def from_bytes(cls, data: bytes) -> "Message":
        m = cls()
        m.init(torch.load(io.BytesIO(data), map_location="cpu", weights_only=False))
        return m
