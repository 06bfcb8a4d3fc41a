"""Transport contract shared by every backend (INPROC, TCP / MPI, gRPC, TRPC, MQTT, MQTT_S3, MQTT_S3_MNN).

The five abstract methods are the reference's public interface (communication/base_com_manager.py:7-26), so managers written
against the reference plug in unchanged.  On top of that contract this class carries what every backend here needs anyway:
observer bookkeeping with a thread-safe snapshot, a ``notify`` helper, and context-manager support so a test or a script can
write ``with create_comm_manager(...) as cm:`` and be sure the receive loop is stopped."""
import threading
from abc import ABC, abstractmethod
from typing import List

from .message import Message
from .observer import Observer


class BaseCommunicationManager(ABC):
    # ---- the reference's contract -------------------------------------------------------------------------------------
    @abstractmethod
    def send_message(self, msg: Message):
        """Deliver ``msg`` to ``msg.get_receiver_id()``; must be callable from any thread."""

    @abstractmethod
    def add_observer(self, observer: Observer):
        """Register a receiver of ``receive_message(msg_type, msg_params)`` callbacks."""

    @abstractmethod
    def remove_observer(self, observer: Observer):
        """Undo :meth:`add_observer`."""

    @abstractmethod
    def handle_receive_message(self):
        """Blocking receive loop: returns after :meth:`stop_receive_message`."""

    @abstractmethod
    def stop_receive_message(self):
        """Make :meth:`handle_receive_message` return and release sockets / threads."""

    # ---- shared helpers -------------------------------------------------------------------------------------------------
    _obs_lock = threading.Lock()

    def observers_snapshot(self) -> List[Observer]:
        """Copy of the observer list that is safe to iterate while another thread registers observers."""
        with self._obs_lock:
            return list(getattr(self, "_observers", None) or getattr(self, "observers", None) or [])

    def notify(self, msg: Message):
        """Hand a received message to every observer (``msg_type`` is the message's type field)."""
        for ob in self.observers_snapshot():
            ob.receive_message(msg.get_type(), msg)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        try:
            self.stop_receive_message()
        except Exception:           # a backend that never started its loop has nothing to stop
            pass
        return False
