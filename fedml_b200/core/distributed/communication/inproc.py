"""In-process transport: every "rank" is a thread, messages are queue hand-offs (tensors by reference).

The reference has no fake transport: its only way to run a distributed algorithm without a
cluster is mpirun on localhost.  This backend makes the manager/handler protocol testable in one process and
is what ``SimulatorMPI`` uses by default on a single machine."""
import queue
import threading
from typing import Dict, List, Tuple

from .base_com_manager import BaseCommunicationManager
from .message import Message
from .observer import Observer

_STOP = object()


class InProcHub:
    _lock = threading.Lock()
    _queues: Dict[Tuple[str, int], "queue.Queue"] = {}

    @classmethod
    def queue_for(cls, run_id, rank) -> "queue.Queue":
        with cls._lock:
            return cls._queues.setdefault((str(run_id), int(rank)), queue.Queue())

    @classmethod
    def reset(cls, run_id=None):
        with cls._lock:
            for k in [k for k in cls._queues if run_id is None or k[0] == str(run_id)]:
                del cls._queues[k]


class InProcCommManager(BaseCommunicationManager):
    def __init__(self, run_id, rank, size, announce_ready=False):
        self.run_id, self.rank, self.size = str(run_id), int(rank), int(size)
        self._observers: List[Observer] = []
        self._inbox = InProcHub.queue_for(self.run_id, self.rank)
        self._running = False
        if announce_ready:      # cross-silo handshake: synthetic "connection ready" message (type 0)
            self._inbox.put(Message(0, self.rank, self.rank))

    def send_message(self, msg: Message):
        InProcHub.queue_for(self.run_id, msg.get_receiver_id()).put(msg)

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
