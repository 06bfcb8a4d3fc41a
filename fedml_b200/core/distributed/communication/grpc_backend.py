"""gRPC transport (reference: communication/grpc/*.py, proto grpc_comm_manager.proto:3-16).

Same topology as the reference — every node serves on ``base_port + rank`` (8888 + rank by default), a unary
``sendMessage`` RPC carries one message, receiver ids map to IPs through a CSV ``receiver_id,ip`` — but
without generated stubs (generic method handlers with identity (de)serialisers, so no protoc step) and with
the raw-tensor ``Message.to_bytes`` codec instead of pickle.  Channels are cached per receiver instead of
being re-opened for every message; max message size 1000 MiB as in the reference."""
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
                 announce_ready=False):
        import grpc
        self._grpc = grpc
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

        handler = grpc.method_handlers_generic_handler(
            "fedml_b200.gRPCCommManager",
            {"sendMessage": grpc.unary_unary_rpc_method_handler(handle, request_deserializer=None,
                                                                 response_serializer=None)})
        self._server = grpc.server(futures.ThreadPoolExecutor(max_workers=max(4, self.size)), options=_OPTS)
        self._server.add_generic_rpc_handlers((handler,))
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
            return ch.unary_unary(_METHOD, request_serializer=None, response_deserializer=None)

    def send_message(self, msg: Message):
        self._stub(int(msg.get_receiver_id()))(msg.to_bytes(), wait_for_ready=True, timeout=1800)

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
