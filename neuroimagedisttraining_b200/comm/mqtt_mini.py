"""A small MQTT 3.1.1 client and broker over plain TCP sockets (QoS 0 publish / subscribe, keep-alive pings).

The reference's MQTT transport (``fedml_core/distributed/communication/mqtt/mqtt_comm_manager.py:14-126``) needs the ``paho-mqtt``
package and a broker listening on ``0.0.0.0:1883`` (``client_manager.py:23-26``); neither exists in this image.  This module speaks
the same wire protocol (so :class:`MiniMqttClient` can talk to mosquitto and paho clients can talk to :class:`MiniMqttBroker`), and
covers exactly what the FL runtime uses: CONNECT / CONNACK, SUBSCRIBE / SUBACK on literal topics, PUBLISH at QoS 0, PINGREQ /
PINGRESP, DISCONNECT.  No retained messages, wildcards, wills, sessions or QoS 1/2.
"""
from __future__ import annotations

import socket
import struct
import threading
from typing import Callable, Dict, Optional, Set, Tuple

CONNECT, CONNACK, PUBLISH, SUBSCRIBE, SUBACK, UNSUBSCRIBE, UNSUBACK, PINGREQ, PINGRESP, DISCONNECT = 1, 2, 3, 8, 9, 10, 11, 12, 13, 14


def _encode_len(n: int) -> bytes:
    out = bytearray()
    while True:
        d, n = n % 128, n // 128
        out.append(d | (0x80 if n else 0))
        if not n:
            return bytes(out)


def _utf8(s: str) -> bytes:
    b = s.encode()
    return struct.pack("!H", len(b)) + b


def _packet(ptype: int, flags: int, body: bytes) -> bytes:
    return bytes([(ptype << 4) | flags]) + _encode_len(len(body)) + body


def _recv_exact(sock: socket.socket, n: int) -> Optional[bytes]:
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(min(n - len(buf), 1 << 20))
        if not chunk:
            return None
        buf += chunk
    return bytes(buf)


def read_packet(sock: socket.socket) -> Optional[Tuple[int, int, bytes]]:
    """(type, flags, body) or None when the peer closed."""
    h = _recv_exact(sock, 1)
    if h is None:
        return None
    mult, n = 1, 0
    for _ in range(4):
        b = _recv_exact(sock, 1)
        if b is None:
            return None
        n += (b[0] & 0x7F) * mult
        if not b[0] & 0x80:
            break
        mult *= 128
    else:
        raise ValueError("malformed MQTT remaining-length field")
    body = _recv_exact(sock, n) if n else b""
    if body is None:
        return None
    return h[0] >> 4, h[0] & 0x0F, body


def _publish_packet(topic: str, payload: bytes) -> bytes:
    return _packet(PUBLISH, 0, _utf8(topic) + payload)


def _parse_publish(flags: int, body: bytes) -> Tuple[str, bytes]:
    (tl,) = struct.unpack("!H", body[:2])
    topic = body[2:2 + tl].decode()
    off = 2 + tl + (2 if (flags >> 1) & 3 else 0)        # packet id present for QoS > 0
    return topic, body[off:]


# ---------------------------------------------------------------------------------------------------- broker
class MiniMqttBroker:
    """Threaded broker: one reader thread per connection, fan-out under a lock.  ``port=0`` picks a free port (``.port``)."""

    def __init__(self, host: str = "127.0.0.1", port: int = 1883):
        self._srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self._srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self._srv.bind((host, port))
        self._srv.listen(64)
        self.host, self.port = host, self._srv.getsockname()[1]
        self._subs: Dict[str, Set[socket.socket]] = {}
        self._locks: Dict[socket.socket, threading.Lock] = {}
        self._mu = threading.Lock()
        self._running = False
        self.published = 0

    def start(self) -> "MiniMqttBroker":
        self._running = True
        threading.Thread(target=self._accept_loop, daemon=True, name="mqtt-broker").start()
        return self

    def stop(self) -> None:
        self._running = False
        try:
            self._srv.close()
        except OSError:
            pass
        with self._mu:
            conns = list(self._locks)
        for c in conns:
            try:
                c.close()
            except OSError:
                pass

    def _accept_loop(self):
        while self._running:
            try:
                conn, _ = self._srv.accept()
            except OSError:
                return
            conn.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            with self._mu:
                self._locks[conn] = threading.Lock()
            threading.Thread(target=self._serve, args=(conn,), daemon=True).start()

    def _send(self, conn: socket.socket, data: bytes) -> None:
        lock = self._locks.get(conn)
        if lock is None:
            return
        try:
            with lock:
                conn.sendall(data)
        except OSError:
            self._drop(conn)

    def _drop(self, conn: socket.socket) -> None:
        with self._mu:
            self._locks.pop(conn, None)
            for s in self._subs.values():
                s.discard(conn)
        try:
            conn.close()
        except OSError:
            pass

    def _serve(self, conn: socket.socket):
        try:
            while self._running:
                pkt = read_packet(conn)
                if pkt is None:
                    break
                ptype, flags, body = pkt
                if ptype == CONNECT:
                    self._send(conn, _packet(CONNACK, 0, b"\x00\x00"))
                elif ptype == SUBSCRIBE:
                    pid, off, granted = body[:2], 2, bytearray()
                    while off < len(body):
                        (tl,) = struct.unpack("!H", body[off:off + 2])
                        topic = body[off + 2:off + 2 + tl].decode()
                        off += 2 + tl + 1                      # + requested QoS byte
                        with self._mu:
                            self._subs.setdefault(topic, set()).add(conn)
                        granted.append(0)
                    self._send(conn, _packet(SUBACK, 0, pid + bytes(granted)))
                elif ptype == UNSUBSCRIBE:
                    pid, off = body[:2], 2
                    while off < len(body):
                        (tl,) = struct.unpack("!H", body[off:off + 2])
                        with self._mu:
                            self._subs.get(body[off + 2:off + 2 + tl].decode(), set()).discard(conn)
                        off += 2 + tl
                    self._send(conn, _packet(UNSUBACK, 0, pid))
                elif ptype == PUBLISH:
                    topic, payload = _parse_publish(flags, body)
                    with self._mu:
                        targets = list(self._subs.get(topic, ()))
                        self.published += 1
                    out = _publish_packet(topic, payload)
                    for t in targets:
                        self._send(t, out)
                elif ptype == PINGREQ:
                    self._send(conn, _packet(PINGRESP, 0, b""))
                elif ptype == DISCONNECT:
                    break
        except (OSError, ValueError):
            pass
        finally:
            self._drop(conn)


# ---------------------------------------------------------------------------------------------------- client
class MiniMqttClient:
    """``connect`` -> ``subscribe`` / ``publish``; ``on_message(topic, payload)`` runs on the reader thread."""

    def __init__(self, client_id: str, on_message: Optional[Callable[[str, bytes], None]] = None, keepalive: int = 60):
        self.client_id, self.on_message, self.keepalive = client_id, on_message, keepalive
        self._sock: Optional[socket.socket] = None
        self._wlock = threading.Lock()
        self._acks: Dict[int, threading.Event] = {}
        self._pid = 0
        self._connack = threading.Event()
        self._running = False

    def connect(self, host: str, port: int = 1883, timeout: float = 10.0) -> "MiniMqttClient":
        self._sock = socket.create_connection((host, port), timeout=timeout)
        self._sock.settimeout(None)
        self._sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        body = _utf8("MQTT") + bytes([4, 0x02]) + struct.pack("!H", self.keepalive) + _utf8(self.client_id)   # level 4, clean session
        self._running = True
        threading.Thread(target=self._reader, daemon=True, name=f"mqtt-{self.client_id}").start()
        self._write(_packet(CONNECT, 0, body))
        if not self._connack.wait(timeout):
            raise ConnectionError(f"no CONNACK from {host}:{port}")
        if self.keepalive:
            threading.Thread(target=self._pinger, daemon=True).start()
        return self

    def _write(self, data: bytes) -> None:
        with self._wlock:
            self._sock.sendall(data)

    def _pinger(self):
        ev = threading.Event()
        while self._running and not ev.wait(self.keepalive / 2):
            try:
                self._write(_packet(PINGREQ, 0, b""))
            except OSError:
                return

    def _reader(self):
        try:
            while self._running:
                pkt = read_packet(self._sock)
                if pkt is None:
                    break
                ptype, flags, body = pkt
                if ptype == CONNACK:
                    if body[1] != 0:
                        raise ConnectionError(f"broker refused the connection (code {body[1]})")
                    self._connack.set()
                elif ptype in (SUBACK, UNSUBACK):
                    ev = self._acks.pop(struct.unpack("!H", body[:2])[0], None)
                    if ev:
                        ev.set()
                elif ptype == PUBLISH and self.on_message is not None:
                    self.on_message(*_parse_publish(flags, body))
        except (OSError, ValueError, ConnectionError):
            pass
        finally:
            self._running = False

    def subscribe(self, topic: str, timeout: float = 10.0) -> None:
        self._pid = self._pid % 65535 + 1
        ev = self._acks[self._pid] = threading.Event()
        self._write(_packet(SUBSCRIBE, 0x02, struct.pack("!H", self._pid) + _utf8(topic) + b"\x00"))
        if not ev.wait(timeout):
            raise TimeoutError(f"no SUBACK for {topic!r}")

    def publish(self, topic: str, payload: bytes) -> None:
        self._write(_publish_packet(topic, payload))

    def disconnect(self) -> None:
        self._running = False
        try:
            self._write(_packet(DISCONNECT, 0, b""))
            self._sock.close()
        except OSError:
            pass
