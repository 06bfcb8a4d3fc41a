"""Socket transport: the multi-process stand-in for the reference's mpi4py backend.

Reference (communication/mpi/*.py): two daemon threads per process around blocking ``comm.send/recv`` of
pickled dicts plus a main loop polling every 0.3 s.  Here every rank listens on
``base_port + rank``; frames are ``<u64 length><Message.to_bytes()>`` (raw tensor buffers, no pickle); the
receive loop blocks on a queue fed by reader threads, so there is no polling latency floor."""
import logging
import queue
import socket
import struct
import threading
import time
from typing import Dict, List

from .base_com_manager import BaseCommunicationManager
from .message import Message
from .observer import Observer

_STOP = object()


def _recv_exact(sock, n):
    """``n`` bytes into one preallocated buffer (no per-chunk concatenation, no final copy)."""
    buf = bytearray(n)
    view, got = memoryview(buf), 0
    while got < n:
        k = sock.recv_into(view[got:], min(4 << 20, n - got))
        if not k:
            return None
        got += k
    return buf


def _send_buffers(sock, buffers):
    """Length prefix + buffers: small pieces are coalesced into one write, large ones (tensor payloads) go to the kernel as they
    are - the message is never assembled in one Python bytes object."""
    total = sum(b.nbytes if isinstance(b, memoryview) else len(b) for b in buffers)
    pending = bytearray(struct.pack("<Q", total))
    for b in buffers:
        n = b.nbytes if isinstance(b, memoryview) else len(b)
        if n < 65536:
            pending += b
            continue
        if pending:
            sock.sendall(pending)
            pending = bytearray()
        sock.sendall(b)
    if pending:
        sock.sendall(pending)


class TcpCommManager(BaseCommunicationManager):
    def __init__(self, rank, size, host_table: Dict[int, str] = None, base_port=50000, announce_ready=False):
        self.rank, self.size, self.base_port = int(rank), int(size), int(base_port)
        self.hosts = host_table or {}
        self._observers: List[Observer] = []
        self._inbox: "queue.Queue" = queue.Queue()
        self._out: Dict[int, socket.socket] = {}
        self._out_lock = threading.Lock()
        self._running = False
        self._srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self._srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self._srv.bind(("0.0.0.0", self.base_port + self.rank))
        self._srv.listen(64)
        self._accept_thread = threading.Thread(target=self._accept_loop, daemon=True)
        self._accept_thread.start()
        if announce_ready:
            self._inbox.put(Message(0, self.rank, self.rank))

    def _accept_loop(self):
        while True:
            try:
                conn, _ = self._srv.accept()
            except OSError:
                return
            threading.Thread(target=self._reader, args=(conn,), daemon=True).start()

    def _reader(self, conn):
        with conn:
            while True:
                head = _recv_exact(conn, 8)
                if head is None:
                    return
                (n,) = struct.unpack("<Q", head)
                body = _recv_exact(conn, n)
                if body is None:
                    return
                self._inbox.put(Message.from_bytes(body))

    def _connect(self, receiver):
        host = self.hosts.get(receiver, "127.0.0.1")
        deadline = time.time() + 60
        while True:
            try:
                s = socket.create_connection((host, self.base_port + receiver), timeout=30)
                s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                return s
            except OSError:
                if time.time() > deadline:
                    raise
                time.sleep(0.05)

    def send_message(self, msg: Message):
        buffers = msg.to_buffers()
        r = int(msg.get_receiver_id())
        with self._out_lock:
            s = self._out.get(r)
            if s is None:
                s = self._out[r] = self._connect(r)
            _send_buffers(s, buffers)

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
        self._running = False

    def stop_receive_message(self):
        self._running = False
        self._inbox.put(_STOP)
        try:
            self._srv.close()
        except OSError:
            pass
        with self._out_lock:
            for s in self._out.values():
                try:
                    s.close()
                except OSError:
                    pass
            self._out.clear()
