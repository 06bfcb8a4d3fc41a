"""A small topic broker + client speaking a length-prefixed JSON protocol over TCP.

The reference's wide-area backends need an external MQTT broker (paho client to ``mqtt.fedml.ai`` /
``0.0.0.0:1883``; communication/mqtt/mqtt_comm_manager.py:47-124) — unreachable offline and paho is not
installable here.  This module supplies the pub/sub semantics those backends rely on (topics, per-client
subscriptions, retained last-will published when a connection drops) so the MQTT, MQTT_S3 and MQTT_S3_MNN
managers and the MLOps status channel run and are tested locally.  If ``paho-mqtt`` is installed and a real
broker is configured, :class:`BrokerClient` is a drop-in for ``paho.mqtt.client.Client`` usage in those
managers (same connect / subscribe / publish / on_message shape)."""
import json
import socket
import struct
import threading
from typing import Callable, Dict, Optional, Set


def _send_frame(sock, obj, payload: bytes = b""):
    head = json.dumps(obj).encode()
    sock.sendall(struct.pack("<II", len(head), len(payload)) + head + payload)


def _recv_exact(sock, n):
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(n - len(buf))
        if not chunk:
            return None
        buf += chunk
    return bytes(buf)


def _recv_frame(sock):
    head = _recv_exact(sock, 8)
    if head is None:
        return None, None
    hl, pl = struct.unpack("<II", head)
    h = _recv_exact(sock, hl)
    p = _recv_exact(sock, pl) if pl else b""
    if h is None or p is None:
        return None, None
    return json.loads(h), p


class LocalBroker:
    """Threaded broker: ``start()`` returns the bound port; ``stop()`` closes everything."""

    def __init__(self, host="127.0.0.1", port=0):
        self._srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self._srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self._srv.bind((host, port))
        self._srv.listen(128)
        self.host, self.port = self._srv.getsockname()
        self._subs: Dict[str, Set[socket.socket]] = {}
        self._wills: Dict[socket.socket, tuple] = {}
        self._lock = threading.Lock()
        self._threads = []

    def start(self):
        t = threading.Thread(target=self._accept, daemon=True)
        t.start()
        self._threads.append(t)
        return self.port

    def _accept(self):
        while True:
            try:
                conn, _ = self._srv.accept()
            except OSError:
                return
            conn.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            threading.Thread(target=self._serve, args=(conn,), daemon=True).start()

    def _publish(self, topic, payload, exclude=None):
        with self._lock:
            targets = list(self._subs.get(topic, ()))
        for s in targets:
            try:
                _send_frame(s, {"op": "msg", "topic": topic}, payload)
            except OSError:
                pass

    def _serve(self, conn):
        try:
            while True:
                h, p = _recv_frame(conn)
                if h is None:
                    break
                op = h.get("op")
                if op == "sub":
                    with self._lock:
                        self._subs.setdefault(h["topic"], set()).add(conn)
                    _send_frame(conn, {"op": "suback", "topic": h["topic"]})
                elif op == "unsub":
                    with self._lock:
                        self._subs.get(h["topic"], set()).discard(conn)
                elif op == "pub":
                    self._publish(h["topic"], p)
                elif op == "will":
                    self._wills[conn] = (h["topic"], p)
                elif op == "bye":
                    self._wills.pop(conn, None)
                    break
        finally:
            with self._lock:
                for subs in self._subs.values():
                    subs.discard(conn)
            will = self._wills.pop(conn, None)
            if will:
                self._publish(will[0], will[1])
            try:
                conn.close()
            except OSError:
                pass

    def stop(self):
        try:
            self._srv.close()
        except OSError:
            pass


class BrokerClient:
    def __init__(self, host="127.0.0.1", port=1883, client_id="", will_topic=None, will_payload=None):
        self.client_id = client_id
        self.on_message: Optional[Callable[[str, bytes], None]] = None
        self.on_connect: Optional[Callable[[], None]] = None
        self._sock = socket.create_connection((host, int(port)), timeout=30)
        self._sock.settimeout(None)
        self._sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        self._wlock = threading.Lock()
        self._suback = {}
        if will_topic:
            self._send({"op": "will", "topic": will_topic},
                       will_payload if isinstance(will_payload, bytes) else json.dumps(will_payload).encode())
        self._alive = True
        self._thread = threading.Thread(target=self._loop, daemon=True)
        self._thread.start()

    def _send(self, head, payload=b""):
        with self._wlock:
            _send_frame(self._sock, head, payload)

    def _loop(self):
        while self._alive:
            try:
                h, p = _recv_frame(self._sock)
            except OSError:
                break
            if h is None:
                break
            if h["op"] == "suback":
                ev = self._suback.get(h["topic"])
                if ev:
                    ev.set()
            elif h["op"] == "msg" and self.on_message:
                try:
                    self.on_message(h["topic"], p)
                except Exception:  # noqa: BLE001 - a handler error must not kill the network thread
                    import logging, traceback
                    logging.error(traceback.format_exc())

    def subscribe(self, topic, wait=True):
        ev = self._suback[topic] = threading.Event()
        self._send({"op": "sub", "topic": topic})
        if wait:
            ev.wait(10)

    def publish(self, topic, payload):
        if isinstance(payload, str):
            payload = payload.encode()
        elif not isinstance(payload, (bytes, bytearray)):
            payload = json.dumps(payload).encode()
        self._send({"op": "pub", "topic": topic}, bytes(payload))

    def disconnect(self, graceful=True):
        self._alive = False
        try:
            if graceful:
                self._send({"op": "bye"})
            self._sock.close()
        except OSError:
            pass
