"""TorchRPC (TensorPipe) transport (reference: communication/trpc/*.py).

``rpc.init_rpc("worker{rank}")`` with ``num_worker_threads=16, rpc_timeout=1800`` and the master address/port
from a CSV (``master_ip,master_port``); a message is delivered with ``rpc_sync("worker{r}", deliver, (msg,))``
so tensors travel through TensorPipe without pickling their storage; ``enable_cuda_rpc`` adds device maps so
CUDA tensors stay on the GPU end to end."""
import csv
import os
import queue
from typing import List

from .base_com_manager import BaseCommunicationManager
from .message import Message
from .observer import Observer

_STOP = object()
_INBOX: "queue.Queue" = queue.Queue()


def _deliver(params: dict):
    m = Message()
    m.init(params)
    _INBOX.put(m)
    return True


def read_master_config(path):
    with open(path, newline="") as fh:
        rd = csv.reader(fh)
        next(rd, None)
        row = next(rd)
        return row[0].strip(), str(row[1]).strip()


class TRPCCommManager(BaseCommunicationManager):
    def __init__(self, trpc_master_config_path, process_id=0, world_size=1, args=None, announce_ready=False):
        import torch
        import torch.distributed.rpc as rpc
        self._rpc = rpc
        self.rank, self.size = int(process_id), int(world_size)
        addr, port = ("127.0.0.1", "29600")
        if trpc_master_config_path and os.path.exists(trpc_master_config_path):
            addr, port = read_master_config(trpc_master_config_path)
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = addr, port
        opts = rpc.TensorPipeRpcBackendOptions(num_worker_threads=16, rpc_timeout=1800,
                                               init_method=f"tcp://{addr}:{port}", _transports=["uv"])
        if args is not None and getattr(args, "enable_cuda_rpc", False) and torch.cuda.is_available():
            for r in range(self.size):
                if r != self.rank:
                    opts.set_device_map(f"worker{r}", {0: 0})
        rpc.init_rpc(f"worker{self.rank}", backend=rpc.BackendType.TENSORPIPE, rank=self.rank, world_size=self.size,
                     rpc_backend_options=opts)
        self._observers: List[Observer] = []
        self._running = False
        if announce_ready:
            _INBOX.put(Message(0, self.rank, self.rank))

    def send_message(self, msg: Message):
        self._rpc.rpc_sync(f"worker{int(msg.get_receiver_id())}", _deliver, args=(msg.get_params(),))

    def add_observer(self, observer: Observer):
        self._observers.append(observer)

    def remove_observer(self, observer: Observer):
        self._observers.remove(observer)

    def handle_receive_message(self):
        self._running = True
        while self._running:
            msg = _INBOX.get()
            if msg is _STOP:
                break
            for obs in list(self._observers):
                obs.receive_message(msg.get_type(), msg.get_params())

    def stop_receive_message(self):
        self._running = False
        _INBOX.put(_STOP)
        try:
            self._rpc.shutdown(graceful=False)
        except Exception:  # noqa: BLE001
            pass
