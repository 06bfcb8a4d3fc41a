"""gRPC transport (reference: communication/grpc/*.py, proto grpc_comm_manager.proto:3-16).

Same topology as the reference — every node serves on ``base_port + rank`` (8888 + rank by default), a unary
``sendMessage`` RPC carries one message, receiver ids map to IPs through a CSV ``receiver_id,ip`` — but
without generated stubs (generic method handlers with identity (de)serialisers, so no protoc step) and with
the raw-tensor ``Message.to_bytes`` codec instead of pickle.  Channels are cached per receiver instead of
being re-opened for every message; max message size 1000 MiB as in the reference.

Wire formats.  Every node serves BOTH:
  * ``/fedml_b200.gRPCCommManager/sendMessage`` — request = ``Message.to_bytes()`` (raw tensors, no pickle), the default
    between fedml_b200 nodes;
  * ``/gRPCCommManager/sendMessage`` — the reference's service (proto grpc_comm_manager.proto:3-16): request =
    protobuf ``CommRequest{int32 client_id = 1; bytes message = 2}`` whose ``message`` is ``pickle.dumps(Message)``
    (grpc_comm_manager.py:64-88), answer = an empty ``CommResponse``.  A reference client can therefore talk to this
    server.  ``wire_format="reference"`` (args.grpc_wire_format) makes this node SEND that format too, so it can talk to a
    reference server.  Unpickling network bytes executes code chosen by the peer — exactly what the reference does on
    every message — so the reference service only accepts pickles when ``wire_format == "reference"`` or
    ``allow_pickle`` is set; otherwise its ``message`` must be the raw codec."""
import pickle
import csv
import queue
import threading
from concurrent import futures
from typing import Dict, List

from .base_com_manager import BaseCommunicationManager
from .message import Message
from .observer import Observer

_STOP = object()
_METHOD = "/fedml_b200.gRPCCommManager/sendMessage"
_REF_METHOD = "/gRPCCommManager/sendMessage"
_REF_MESSAGE_CLASS = ("fedml.core.distributed.communication.message", "Message")


def _varint(n: int) -> bytes:
    n &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        out.append(b | (0x80 if n else 0))
        if not n:
            return bytes(out)


def encode_comm_request(client_id: int, message: bytes) -> bytes:
    """protobuf wire encoding of CommRequest / CommResponse {int32 client_id = 1; bytes message = 2} (proto3: zero / empty
    fields are omitted)."""
    out = b""
    if client_id:
        out += b"\x08" + _varint(int(client_id))
    if message:
        out += b"\x12" + _varint(len(message)) + message
    return out


def decode_comm_request(buf: bytes):
    """-> (client_id, message bytes); unknown fields are skipped."""
    i, n, cid, msg = 0, len(buf), 0, b""

    def rd_varint(i):
        shift = val = 0
        while True:
            b = buf[i]; i += 1
            val |= (b & 0x7F) << shift
            shift += 7
            if not b & 0x80:
                return val, i
    while i < n:
        key, i = rd_varint(i)
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, i = rd_varint(i)
            if field == 1:
                cid = v - (1 << 64) if v >> 63 else v
        elif wt == 2:
            ln, i = rd_varint(i)
            if field == 2:
                msg = bytes(buf[i:i + ln])
            i += ln
        elif wt == 5:
            i += 4
        elif wt == 1:
            i += 8
        else:
            raise ValueError(f"unsupported protobuf wire type {wt}")
    return cid, msg


class CommRequest:
    """The reference's protobuf message (proto/grpc_comm_manager.proto:3-9) without generated code: same field names, same wire
    bytes (``SerializeToString`` / ``FromString``), so code that builds requests by hand keeps working."""

    def __init__(self, client_id: int = 0, message: bytes = b""):
        self.client_id, self.message = int(client_id), bytes(message)

    def SerializeToString(self) -> bytes:      # noqa: N802  (protobuf API)
        return encode_comm_request(self.client_id, self.message)

    @classmethod
    def FromString(cls, data: bytes):          # noqa: N802
        return cls(*decode_comm_request(data))

    def __eq__(self, other):
        return isinstance(other, CommRequest) and (self.client_id, self.message) == (other.client_id, other.message)

    def __repr__(self):
        return f"{type(self).__name__}(client_id={self.client_id}, message=<{len(self.message)} bytes>)"


class CommResponse(CommRequest):
    """Same two fields (grpc_comm_manager.proto:11-14)."""


def pickle_like_reference(msg: Message) -> bytes:
    """``pickle.dumps`` of the message as the reference's class ``fedml.core...message.Message`` (same attributes: type,
    sender_id, receiver_id, msg_params), whatever module this class lives in: protocol-2 stream = GLOBAL class, NEWOBJ,
    state dict, BUILD."""
    state = {"type": msg.type, "sender_id": msg.sender_id, "receiver_id": msg.receiver_id, "msg_params": msg.get_params()}
    body = pickle.dumps(state, protocol=2)
    assert body[:2] == b"\x80\x02" and body[-1:] == b"."
    mod, name = _REF_MESSAGE_CLASS
    return b"\x80\x02c" + mod.encode() + b"\n" + name.encode() + b"\n)\x81" + body[2:-1] + b"b."


class _MessageUnpickler(pickle.Unpickler):
    """Maps the reference's Message class onto ours; everything else resolves as usual (tensors, numpy arrays, ...)."""

    def find_class(self, module, name):
        if (module, name) == _REF_MESSAGE_CLASS or (module.endswith("communication.message") and name == "Message"):
            return Message
        return super().find_class(module, name)


def unpickle_message(data: bytes) -> Message:
    import io
    obj = _MessageUnpickler(io.BytesIO(data)).load()
    if not isinstance(obj, Message):
        raise TypeError(f"expected a Message, got {type(obj)}")
    obj._sync_header()
    return obj
_OPTS = [("grpc.max_send_message_length", 1000 * 1024 * 1024), ("grpc.max_receive_message_length", 1000 * 1024 * 1024)]


def read_ip_config(path) -> Dict[int, str]:
    table = {}
    with open(path, newline="") as fh:
        rd = csv.reader(fh)
        next(rd, None)                 # header: receiver_id,ip
        for row in rd:
            if len(row) >= 2:
                table[int(row[0])] = row[1].strip()
    return table


class GRPCCommManager(BaseCommunicationManager):
    def __init__(self, host, port, ip_config_path=None, topic="fedml", client_id=0, client_num=0, base_port=8888,
                 announce_ready=False, wire_format="fedml_b200", allow_pickle=False):
        import grpc
        self._grpc = grpc
        self.wire_format = str(wire_format or "fedml_b200").lower()
        self.allow_pickle = bool(allow_pickle) or self.wire_format == "reference"
        self.rank, self.size, self.base_port = int(client_id), int(client_num) + 1, int(base_port)
        self.ip_table = read_ip_config(ip_config_path) if ip_config_path else {}
        self._observers: List[Observer] = []
        self._inbox: "queue.Queue" = queue.Queue()
        self._channels = {}
        self._lock = threading.Lock()
        self._running = False

        def handle(request: bytes, context):
            self._inbox.put(Message.from_bytes(request))
            return b"ok"

        def handle_reference(request: bytes, context):
            sender, payload = decode_comm_request(request)
            if payload[:4] == b"FB2M":                    # raw codec inside the reference envelope
                self._inbox.put(Message.from_bytes(payload))
            elif self.allow_pickle:
                self._inbox.put(unpickle_message(payload))
            else:
                context.abort(grpc.StatusCode.PERMISSION_DENIED,
                              "pickled messages are refused unless grpc_wire_format is 'reference' (or grpc_allow_pickle)")
            return encode_comm_request(0, b"")          # empty CommResponse, as the reference answers

        handler = grpc.method_handlers_generic_handler(
            "fedml_b200.gRPCCommManager",
            {"sendMessage": grpc.unary_unary_rpc_method_handler(handle, request_deserializer=None,
                                                                 response_serializer=None)})
        ref_handler = grpc.method_handlers_generic_handler(
            "gRPCCommManager",
            {"sendMessage": grpc.unary_unary_rpc_method_handler(handle_reference, request_deserializer=None,
                                                                 response_serializer=None),
             "handleReceiveMessage": grpc.unary_unary_rpc_method_handler(handle_reference, request_deserializer=None,
                                                                          response_serializer=None)})
        self._server = grpc.server(futures.ThreadPoolExecutor(max_workers=max(4, self.size)), options=_OPTS)
        self._server.add_generic_rpc_handlers((handler, ref_handler))
        self._server.add_insecure_port(f"{host}:{int(port)}")
        self._server.start()
        if announce_ready:
            self._inbox.put(Message(0, self.rank, self.rank))

    def _stub(self, receiver):
        with self._lock:
            ch = self._channels.get(receiver)
            if ch is None:
                ip = self.ip_table.get(receiver, "127.0.0.1")
                ch = self._grpc.insecure_channel(f"{ip}:{self.base_port + receiver}", options=_OPTS)
                self._channels[receiver] = ch
            method = _REF_METHOD if self.wire_format == "reference" else _METHOD
            return ch.unary_unary(method, request_serializer=None, response_deserializer=None)

    def send_message(self, msg: Message):
        if self.wire_format == "reference":
            req = encode_comm_request(self.rank, pickle_like_reference(msg))
        else:
            req = msg.to_bytes()
        self._stub(int(msg.get_receiver_id()))(req, wait_for_ready=True, timeout=1800)

    def add_observer(self, observer: Observer):
        self._observers.append(observer)

    def remove_observer(self, observer: Observer):
        self._observers.remove(observer)

    def handle_receive_message(self):
        self._running = True
        while self._running:
            msg = self._inbox.get()
            if msg is _STOP:
                break
            for obs in list(self._observers):
                obs.receive_message(msg.get_type(), msg.get_params())

    def stop_receive_message(self):
        self._running = False
        self._inbox.put(_STOP)
        self._server.stop(0.5)
        with self._lock:
            for ch in self._channels.values():
                ch.close()
            self._channels.clear()
