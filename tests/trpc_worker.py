"""Two-process TorchRPC round trip (run by tests/test_comm_cpu.py): rank 1 sends a tensor message to rank 0, rank 0 answers."""
import os
import sys
import threading

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from fedml_b200.core.distributed.communication.message import Message  # noqa: E402
from fedml_b200.core.distributed.communication.trpc_backend import TRPCCommManager  # noqa: E402

rank, cfg = int(sys.argv[1]), sys.argv[2]
mgr = TRPCCommManager(cfg, rank, 2)
done = threading.Event()
got = []


class Obs:
    def receive_message(self, t, p):
        got.append((t, p))
        if rank == 0:
            back = Message(8, 0, 1)
            back.add_params("echo", p["w"] * 2)
            mgr.send_message(back)
        done.set()


mgr.add_observer(Obs())
th = threading.Thread(target=mgr.handle_receive_message, daemon=True)
th.start()
if rank == 1:
    m = Message(7, 1, 0)
    m.add_params("w", torch.arange(5.0))
    mgr.send_message(m)
assert done.wait(60), "no message"
if rank == 1:
    assert got[0][0] == 8 and torch.equal(got[0][1]["echo"], torch.arange(5.0) * 2)
else:
    assert got[0][0] == 7 and torch.equal(got[0][1]["w"], torch.arange(5.0))
import time
time.sleep(0.5)
print(f"TRPC_OK rank {rank}", flush=True)
os._exit(0)
