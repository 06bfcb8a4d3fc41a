"""What a communication manager delivers to: the ``Observer`` contract plus two ready-made observers.

Every back-end (MPI queues, MQTT, gRPC, TRPC, the in-process loopback) ends in the same call,
``observer.receive_message(msg_type, msg_params)`` (reference: communication/observer.py:4-7), made from the back-end's
receive thread.  ``FedMLCommManager`` is the observer used by the algorithm managers; the two below cover the other uses
in this tree:

* ``CallbackObserver`` - wraps a plain function, for tools and tests that only want to look at the traffic;
* ``MailboxObserver`` - a thread-safe mailbox with a blocking ``get(msg_type, timeout)``, for request/response style code
  (the cross-device round trip test, the broker self-test) that wants to wait for one specific message type.
"""
from __future__ import annotations

import threading
from abc import ABC, abstractmethod
from collections import defaultdict, deque
from typing import Callable, Optional


class Observer(ABC):
    """Receives one decoded message.  Called on the back-end's receive thread: do not block in it."""

    @abstractmethod
    def receive_message(self, msg_type, msg_params) -> None:
        ...


class CallbackObserver(Observer):
    def __init__(self, fn: Callable[[object, object], None]):
        self._fn = fn

    def receive_message(self, msg_type, msg_params) -> None:
        self._fn(msg_type, msg_params)


class MailboxObserver(Observer):
    """Stores what arrives, per message type, in arrival order."""

    def __init__(self):
        self._boxes = defaultdict(deque)
        self._cv = threading.Condition()
        self.received = 0

    def receive_message(self, msg_type, msg_params) -> None:
        with self._cv:
            self._boxes[str(msg_type)].append(msg_params)
            self.received += 1
            self._cv.notify_all()

    def get(self, msg_type, timeout: Optional[float] = None):
        """Oldest undelivered message of ``msg_type``; waits up to ``timeout`` seconds, then raises ``TimeoutError``."""
        key = str(msg_type)
        with self._cv:
            if not self._cv.wait_for(lambda: self._boxes[key], timeout):
                raise TimeoutError(f"no message of type {key} within {timeout}s")
            return self._boxes[key].popleft()

    def pending(self, msg_type=None) -> int:
        with self._cv:
            if msg_type is None:
                return sum(len(b) for b in self._boxes.values())
            return len(self._boxes[str(msg_type)])
