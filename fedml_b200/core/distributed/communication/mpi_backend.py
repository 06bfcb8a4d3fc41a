"""MPI transport through mpi4py, used when mpi4py is importable and the job was started by ``mpirun``
(reference: communication/mpi/com_manager.py, mpi_send_thread.py, mpi_receive_thread.py — two daemon threads per
rank around ``comm.send`` / ``comm.recv`` of pickled Message objects and a polling observer loop).

Here: one receive thread blocks in ``comm.recv(source=ANY_SOURCE)`` and feeds a queue, sends go out from the caller's
thread under a lock (mpi4py needs MPI_THREAD_MULTIPLE for concurrent calls: it is requested at import by default; with a
lower level the lock plus ``iprobe`` polling keeps all MPI calls serialised).  The payload is the dict of message
parameters, pickled by mpi4py exactly as the reference's ``comm.send(msg, dest=...)`` does, so mixed jobs interoperate.
Without mpi4py the same star topology runs over sockets (tcp.py)."""
import queue
import threading
import time
from typing import List

from .base_com_manager import BaseCommunicationManager
from .message import Message
from .observer import Observer

_STOP = object()
_TAG = 0


def mpi_available() -> bool:
    try:
        import mpi4py  # noqa: F401
        return True
    except Exception:  # noqa: BLE001
        return False


class MpiCommManager(BaseCommunicationManager):
    def __init__(self, comm=None, rank=None, size=None, node_type="client", announce_ready=False):
        from mpi4py import MPI
        self._MPI = MPI
        self.comm = comm if comm is not None else MPI.COMM_WORLD
        self.rank = self.comm.Get_rank() if rank is None else int(rank)
        self.size = self.comm.Get_size() if size is None else int(size)
        self.node_type = node_type
        self._observers: List[Observer] = []
        self._inbox: "queue.Queue" = queue.Queue()
        self._lock = threading.Lock()
        self._running = True
        self._multiple = getattr(MPI, "Query_thread", lambda: 3)() >= getattr(MPI, "THREAD_MULTIPLE", 3)
        self._rx = threading.Thread(target=self._receive_loop, daemon=True)
        self._rx.start()
        if announce_ready:
            self._inbox.put(Message(0, self.rank, self.rank))

    def _receive_loop(self):
        MPI = self._MPI
        while self._running:
            if self._multiple:
                status = MPI.Status()
                obj = self.comm.recv(source=MPI.ANY_SOURCE, tag=MPI.ANY_TAG, status=status)
            else:
                with self._lock:
                    ready = self.comm.iprobe(source=MPI.ANY_SOURCE, tag=MPI.ANY_TAG)
                    obj = self.comm.recv(source=MPI.ANY_SOURCE, tag=MPI.ANY_TAG) if ready else None
                if obj is None:
                    time.sleep(0.0005)
                    continue
            if obj is _STOP or (isinstance(obj, str) and obj == "__fb_stop__"):
                break
            if isinstance(obj, Message):
                msg = obj
            else:
                msg = Message()
                msg.init(obj)
            self._inbox.put(msg)

    def send_message(self, msg: Message):
        dest = int(msg.get_receiver_id())
        with self._lock:
            self.comm.send(msg.get_params(), dest=dest, tag=_TAG)

    def add_observer(self, observer: Observer):
        self._observers.append(observer)

    def remove_observer(self, observer: Observer):
        self._observers.remove(observer)

    def handle_receive_message(self):
        while True:
            msg = self._inbox.get()
            if msg is _STOP:
                break
            for obs in list(self._observers):
                obs.receive_message(msg.get_type(), msg.get_params())

    def stop_receive_message(self):
        self._running = False
        self._inbox.put(_STOP)
        try:                                   # wake the receive thread of this rank
            with self._lock:
                self.comm.send("__fb_stop__", dest=self.rank, tag=_TAG)
        except Exception:  # noqa: BLE001
            pass
