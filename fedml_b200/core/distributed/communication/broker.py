"""MQTT 3.1.1 over TCP: a small broker and a client, both speaking the standard wire protocol.

The reference's wide-area backends use a paho client against an external broker (``mqtt.fedml.ai`` / ``0.0.0.0:1883``;
communication/mqtt/mqtt_comm_manager.py:47-124).  Neither is reachable / installable offline, so this module carries
its own implementation of the protocol subset those backends need — CONNECT (with last will), CONNACK, PUBLISH
(QoS 0 sent, QoS 0/1 accepted), PUBACK, SUBSCRIBE / SUBACK (``+`` and ``#`` wildcards), UNSUBSCRIBE / UNSUBACK,
PINGREQ / PINGRESP, DISCONNECT — byte-compatible with any MQTT 3.1.1 peer: a paho or mosquitto client can connect to
:class:`LocalBroker`, and :class:`BrokerClient` can connect to a real broker.  :func:`make_client` returns a paho-backed
client with the same small API when ``paho-mqtt`` is importable and requested (``args.mqtt_client: "paho"``).
"""
import json
import logging
import socket
import struct
import threading
import traceback
from typing import Callable, Dict, List, Optional, Tuple

CONNECT, CONNACK, PUBLISH, PUBACK, SUBSCRIBE, SUBACK, UNSUBSCRIBE, UNSUBACK, PINGREQ, PINGRESP, DISCONNECT = \
    1, 2, 3, 4, 8, 9, 10, 11, 12, 13, 14
MAX_REMAINING = 268_435_455          # protocol limit of one packet (256 MiB - 1)


# ---------------------------------------------------------------- wire helpers
def _enc_len(n: int) -> bytes:
    if n > MAX_REMAINING:
        raise ValueError(f"MQTT packet of {n} bytes exceeds the protocol limit of {MAX_REMAINING}; use the MQTT_S3 backend "
                         "for large payloads")
    out = bytearray()
    while True:
        b = n % 128
        n //= 128
        out.append(b | (0x80 if n else 0))
        if not n:
            return bytes(out)


def _utf8(s: str) -> bytes:
    b = s.encode("utf-8")
    return struct.pack(">H", len(b)) + b


def _packet(ptype: int, flags: int, body: bytes) -> bytes:
    return bytes([(ptype << 4) | flags]) + _enc_len(len(body)) + body


def _recv_exact(sock, n):
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(min(n - len(buf), 1 << 20))
        if not chunk:
            return None
        buf += chunk
    return bytes(buf)


def _recv_packet(sock) -> Tuple[Optional[int], int, bytes]:
    h = _recv_exact(sock, 1)
    if h is None:
        return None, 0, b""
    mult, length = 1, 0
    while True:
        b = _recv_exact(sock, 1)
        if b is None:
            return None, 0, b""
        length += (b[0] & 0x7F) * mult
        if not b[0] & 0x80:
            break
        mult *= 128
        if mult > 128 ** 3:
            raise ValueError("malformed MQTT remaining length")
    body = _recv_exact(sock, length) if length else b""
    if body is None:
        return None, 0, b""
    return h[0] >> 4, h[0] & 0x0F, body


def _rd_utf8(buf: bytes, i: int) -> Tuple[str, int]:
    (n,) = struct.unpack_from(">H", buf, i)
    return buf[i + 2:i + 2 + n].decode("utf-8"), i + 2 + n


def topic_matches(filt: str, topic: str) -> bool:
    """MQTT topic filter match (``+`` one level, ``#`` the rest)."""
    f, t = filt.split("/"), topic.split("/")
    for i, part in enumerate(f):
        if part == "#":
            return True
        if i >= len(t) or (part != "+" and part != t[i]):
            return False
    return len(f) == len(t)


# ---------------------------------------------------------------- broker
class LocalBroker:
    """Threaded MQTT 3.1.1 broker: ``start()`` returns the bound port; ``stop()`` closes everything."""

    def __init__(self, host="127.0.0.1", port=0, certfile=None, keyfile=None):
        """``certfile`` / ``keyfile``: serve MQTT over TLS (the 8883 flavour) with this certificate chain and private key."""
        self._tls = None
        if certfile:
            import ssl
            self._tls = ssl.SSLContext(ssl.PROTOCOL_TLS_SERVER)
            self._tls.load_cert_chain(certfile, keyfile)
        self._srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self._srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self._srv.bind((host, port))
        self._srv.listen(128)
        self.host, self.port = self._srv.getsockname()
        self._subs: Dict[socket.socket, List[str]] = {}
        self._wlocks: Dict[socket.socket, threading.Lock] = {}
        self._lock = threading.Lock()

    def start(self):
        threading.Thread(target=self._accept, daemon=True).start()
        return self.port

    def _accept(self):
        while True:
            try:
                conn, _ = self._srv.accept()
            except OSError:
                return
            conn.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            if self._tls is not None:
                try:
                    conn = self._tls.wrap_socket(conn, server_side=True)
                except (OSError, ValueError):          # failed handshake (plain-text client, unknown CA): drop the connection
                    conn.close()
                    continue
            threading.Thread(target=self._serve, args=(conn,), daemon=True).start()

    def _send(self, conn, data: bytes):
        lock = self._wlocks.get(conn)
        if lock is None:
            return
        with lock:
            conn.sendall(data)

    def _publish(self, topic: str, payload: bytes):
        with self._lock:
            targets = [c for c, filters in self._subs.items() if any(topic_matches(f, topic) for f in filters)]
        pkt = _packet(PUBLISH, 0, _utf8(topic) + payload)
        for c in targets:
            try:
                self._send(c, pkt)
            except OSError:
                pass

    def _serve(self, conn):
        will = None
        clean_exit = False
        with self._lock:
            self._subs[conn] = []
            self._wlocks[conn] = threading.Lock()
        try:
            ptype, flags, body = _recv_packet(conn)
            if ptype != CONNECT:
                return
            proto, i = _rd_utf8(body, 0)
            level, cflags = body[i], body[i + 1]
            i += 4                                              # level, flags, keep-alive
            _cid, i = _rd_utf8(body, i)
            if cflags & 0x04:
                wt, i = _rd_utf8(body, i)
                (wn,) = struct.unpack_from(">H", body, i)
                will = (wt, body[i + 2:i + 2 + wn])
                i += 2 + wn
            rc = 0 if (proto == "MQTT" and level == 4) or (proto == "MQIsdp" and level == 3) else 1
            self._send(conn, _packet(CONNACK, 0, bytes([0, rc])))
            if rc:
                return
            while True:
                ptype, flags, body = _recv_packet(conn)
                if ptype is None:
                    break
                if ptype == PUBLISH:
                    qos = (flags >> 1) & 3
                    topic, i = _rd_utf8(body, 0)
                    if qos:
                        (pid,) = struct.unpack_from(">H", body, i)
                        i += 2
                        self._send(conn, _packet(PUBACK, 0, struct.pack(">H", pid)))
                    self._publish(topic, body[i:])
                elif ptype == SUBSCRIBE:
                    (pid,) = struct.unpack_from(">H", body, 0)
                    i, codes = 2, bytearray()
                    while i < len(body):
                        filt, i = _rd_utf8(body, i)
                        i += 1                                  # requested QoS: granted 0
                        with self._lock:
                            if filt not in self._subs[conn]:
                                self._subs[conn].append(filt)
                        codes.append(0)
                    self._send(conn, _packet(SUBACK, 0, struct.pack(">H", pid) + bytes(codes)))
                elif ptype == UNSUBSCRIBE:
                    (pid,) = struct.unpack_from(">H", body, 0)
                    i = 2
                    while i < len(body):
                        filt, i = _rd_utf8(body, i)
                        with self._lock:
                            if filt in self._subs[conn]:
                                self._subs[conn].remove(filt)
                    self._send(conn, _packet(UNSUBACK, 0, struct.pack(">H", pid)))
                elif ptype == PINGREQ:
                    self._send(conn, _packet(PINGRESP, 0, b""))
                elif ptype == DISCONNECT:
                    clean_exit = True
                    break
        except (OSError, ValueError, struct.error, IndexError):
            pass
        finally:
            with self._lock:
                self._subs.pop(conn, None)
                self._wlocks.pop(conn, None)
            if will and not clean_exit:                         # last will: published when the connection drops unannounced
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


# ---------------------------------------------------------------- client
class BrokerClient:
    """Minimal MQTT 3.1.1 client.  ``on_message(topic, payload_bytes)`` runs on the network thread."""

    def __init__(self, host="127.0.0.1", port=1883, client_id="", will_topic=None, will_payload=None, keepalive=0,
                 username=None, password=None, tls=False, ca_certs=None, tls_insecure=False):
        """``tls`` (or a ``ca_certs`` bundle path): MQTT over TLS; the server certificate is verified against ``ca_certs`` (the
        system store when None) unless ``tls_insecure``."""
        self.client_id = client_id
        self.on_message: Optional[Callable[[str, bytes], None]] = None
        self.on_connect: Optional[Callable[[], None]] = None
        self._sock = socket.create_connection((host, int(port)), timeout=30)
        self._sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        if tls or ca_certs:
            import ssl
            ctx = ssl.create_default_context(cafile=ca_certs)
            if tls_insecure:
                ctx.check_hostname, ctx.verify_mode = False, ssl.CERT_NONE
            self._sock = ctx.wrap_socket(self._sock, server_hostname=host)
        self._wlock = threading.Lock()
        self._acks: Dict[int, threading.Event] = {}
        self._pid = 0
        flags = 0x02                                            # clean session
        payload = _utf8(client_id or "")
        if will_topic:
            wp = will_payload if isinstance(will_payload, (bytes, bytearray)) else json.dumps(will_payload).encode()
            flags |= 0x04
            payload += _utf8(will_topic) + struct.pack(">H", len(wp)) + bytes(wp)
        if username is not None:
            flags |= 0x80
            payload += _utf8(username)
            if password is not None:
                flags |= 0x40
                payload += _utf8(password)
        self._sock.sendall(_packet(CONNECT, 0, _utf8("MQTT") + bytes([4, flags]) + struct.pack(">H", int(keepalive)) + payload))
        ptype, _, body = _recv_packet(self._sock)
        if ptype != CONNACK or len(body) < 2 or body[1] != 0:
            raise ConnectionError(f"MQTT broker refused the connection (CONNACK {body!r})")
        self._sock.settimeout(None)
        self._alive = True
        self._thread = threading.Thread(target=self._loop, daemon=True)
        self._thread.start()
        if keepalive:
            threading.Thread(target=self._ping, args=(max(1, int(keepalive) // 2),), daemon=True).start()

    def _send(self, data: bytes):
        with self._wlock:
            self._sock.sendall(data)

    def _ping(self, every):
        import time
        while self._alive:
            time.sleep(every)
            try:
                self._send(_packet(PINGREQ, 0, b""))
            except OSError:
                return

    def _next_pid(self) -> int:
        self._pid = self._pid % 65535 + 1
        return self._pid

    def _loop(self):
        while self._alive:
            try:
                ptype, flags, body = _recv_packet(self._sock)
            except (OSError, ValueError):
                break
            if ptype is None:
                break
            if ptype == PUBLISH:
                qos = (flags >> 1) & 3
                topic, i = _rd_utf8(body, 0)
                if qos:
                    (pid,) = struct.unpack_from(">H", body, i)
                    i += 2
                    try:
                        self._send(_packet(PUBACK, 0, struct.pack(">H", pid)))
                    except OSError:
                        pass
                if self.on_message:
                    try:
                        self.on_message(topic, body[i:])
                    except Exception:  # noqa: BLE001 - a handler error must not kill the network thread
                        logging.error(traceback.format_exc())
            elif ptype in (SUBACK, UNSUBACK, PUBACK):
                (pid,) = struct.unpack_from(">H", body, 0)
                ev = self._acks.pop(pid, None)
                if ev:
                    ev.set()

    def subscribe(self, topic, wait=True):
        pid = self._next_pid()
        ev = self._acks[pid] = threading.Event()
        self._send(_packet(SUBSCRIBE, 0x2, struct.pack(">H", pid) + _utf8(topic) + b"\x00"))
        if wait:
            ev.wait(10)

    def unsubscribe(self, topic):
        pid = self._next_pid()
        self._send(_packet(UNSUBSCRIBE, 0x2, struct.pack(">H", pid) + _utf8(topic)))

    def publish(self, topic, payload):
        if isinstance(payload, str):
            payload = payload.encode()
        elif not isinstance(payload, (bytes, bytearray)):
            payload = json.dumps(payload).encode()
        self._send(_packet(PUBLISH, 0, _utf8(topic) + bytes(payload)))

    def disconnect(self, graceful=True):
        self._alive = False
        try:
            if graceful:
                self._send(_packet(DISCONNECT, 0, b""))
            self._sock.close()
        except OSError:
            pass


class PahoClient:
    """The same small API on top of ``paho.mqtt.client`` (reference: mqtt_comm_manager.py:47-124), for deployments that
    have paho and a production broker."""

    def __init__(self, host="127.0.0.1", port=1883, client_id="", will_topic=None, will_payload=None, keepalive=180,
                 username=None, password=None):
        import paho.mqtt.client as mqtt
        self.on_message: Optional[Callable[[str, bytes], None]] = None
        try:
            self._c = mqtt.Client(client_id=client_id, clean_session=True)
        except TypeError:                                       # paho >= 2.0 wants the callback API version first
            self._c = mqtt.Client(mqtt.CallbackAPIVersion.VERSION1, client_id=client_id, clean_session=True)
        if username is not None:
            self._c.username_pw_set(username, password)
        if will_topic:
            wp = will_payload if isinstance(will_payload, (bytes, bytearray, str)) else json.dumps(will_payload)
            self._c.will_set(will_topic, wp, qos=0, retain=False)
        self._c.on_message = lambda _c, _u, m: self.on_message and self.on_message(m.topic, bytes(m.payload))
        self._c.connect(host, int(port), keepalive)
        self._c.loop_start()

    def subscribe(self, topic, wait=True):
        self._c.subscribe(topic, qos=0)

    def unsubscribe(self, topic):
        self._c.unsubscribe(topic)

    def publish(self, topic, payload):
        if not isinstance(payload, (bytes, bytearray, str)):
            payload = json.dumps(payload)
        self._c.publish(topic, payload, qos=0)

    def disconnect(self, graceful=True):
        self._c.loop_stop()
        if graceful:
            self._c.disconnect()


def make_client(host, port, client_id="", will_topic=None, will_payload=None, kind="auto", **kw):
    """``kind``: "native" (this module's client), "paho" (must be importable) or "auto" (native: it speaks the same
    protocol and has no dependency)."""
    if str(kind).lower() == "paho":
        return PahoClient(host, port, client_id, will_topic, will_payload, **kw)
    return BrokerClient(host, port, client_id, will_topic, will_payload, **kw)
