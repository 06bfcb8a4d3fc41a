"""Round-trip latency of the message transports with tensor payloads - the micro-benchmark of the reference's
``test/grpc_benchmark`` (gRPC vs TorchRPC, ``identity`` / ``heavy`` remote functions, NUM_RPC = 10 requests in flight, cpu / cuda
tensors, CUDA-event timing: common.py:7-74, grpc_client.py:53-109), extended to every transport of this tree.

Two endpoints in one process (rank 1 = caller, rank 0 = callee) talk through ``create_comm_manager``; the callee applies the remote
function to the tensor and sends it back.  One measurement = NUM_RPC requests issued back to back, time until the last reply
(wall clock; with ``--cuda`` the tensors live on the GPU, the callee's work is enqueued on the current stream and the interval is
closed by CUDA events like in the reference).  Reported: median over ``--repeat`` measurements, per request.

    python tools/bench_comm.py                       # INPROC TCP GRPC GRPC_REF MQTT MQTT_S3, 4 KB .. 16 MB, cpu
    python tools/bench_comm.py --backends GRPC TCP --sizes 1048576 --cuda --markdown profiles/r2/comm_latency.md
"""
import argparse
import os
import statistics
import sys
import tempfile
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from fedml_b200.arguments import Arguments  # noqa: E402
from fedml_b200.core.distributed.communication import Message, Observer, create_comm_manager  # noqa: E402

NUM_RPC = 10
MSG_CALL, MSG_REPLY = 101, 102


def identity(x):
    return x


def heavy(x):
    for _ in range(100):
        x *= 2.0
        x /= 2.0
    return x


class Callee(Observer):
    def __init__(self, mgr, fn, cuda):
        self.mgr, self.fn, self.cuda = mgr, fn, cuda

    def receive_message(self, msg_type, msg):
        if int(msg_type) != MSG_CALL:
            return
        x = torch.as_tensor(msg[Message.MSG_ARG_KEY_MODEL_PARAMS]["x"])
        if self.cuda:
            x = x.cuda(non_blocking=True)
        y = self.fn(x.clone() if not x.is_floating_point() else x.float().clone())
        out = Message(MSG_REPLY, 0, 1)
        out.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, {"x": y.cpu() if self.cuda else y})
        out.add_params("seq", msg["seq"])
        self.mgr.send_message(out)


class Caller(Observer):
    def __init__(self):
        self.pending, self.done = 0, threading.Event()
        self.lock = threading.Lock()

    def receive_message(self, msg_type, msg):
        if int(msg_type) != MSG_REPLY:
            return
        with self.lock:
            self.pending -= 1
            if self.pending == 0:
                self.done.set()


def free_port():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def run_backend(name, sizes, fn, repeat, cuda, broker_port, store_dir):
    backend, extra = name, {}
    if name == "GRPC_REF":                                   # the reference's wire format (protobuf + pickled Message)
        backend, extra = "GRPC", {"grpc_wire_format": "reference", "grpc_allow_pickle": True}
    if name == "MQTT_BIN":                                   # MQTT carrying the binary tensor codec instead of JSON lists
        backend, extra = "MQTT", {"mqtt_binary": True}
    args = Arguments.from_dict(dict(run_id=f"bench-{name}-{fn.__name__}", tcp_base_port=free_port(), grpc_base_port=free_port(),
                                    mqtt_port=broker_port, object_store_dir=store_dir, client_id_list="[1]", **extra))
    if backend == "INPROC":
        from fedml_b200.core.distributed.communication.inproc import InProcHub
        InProcHub.reset(args.run_id)
    callee_mgr = create_comm_manager(args, 0, 2, backend)
    caller_mgr = create_comm_manager(args, 1, 2, backend)
    caller = Caller()
    callee_mgr.add_observer(Callee(callee_mgr, fn, cuda))
    caller_mgr.add_observer(caller)
    for m in (callee_mgr, caller_mgr):
        threading.Thread(target=m.handle_receive_message, daemon=True).start()
    time.sleep(0.3)
    rows = []
    try:
        for n_bytes in sizes:
            x = torch.ones(max(1, n_bytes // 4), dtype=torch.float32)
            samples = []
            for it in range(repeat + 1):                     # first one warms connections up
                caller.pending, _ = NUM_RPC, caller.done.clear()
                if cuda:
                    tik, tok = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    tik.record()
                t0 = time.perf_counter()
                for k in range(NUM_RPC):
                    m = Message(MSG_CALL, 1, 0)
                    m.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, {"x": x})      # where models travel: MQTT_S3 moves this key through the store
                    m.add_params("seq", k)
                    caller_mgr.send_message(m)
                if not caller.done.wait(120):
                    raise TimeoutError(f"{name}: no reply within 120 s at {n_bytes} bytes")
                if cuda:
                    tok.record()
                    tok.synchronize()
                    dt = tik.elapsed_time(tok) / 1e3
                else:
                    dt = time.perf_counter() - t0
                if it:
                    samples.append(dt / NUM_RPC)
            med = statistics.median(samples)
            rows.append((name, fn.__name__, n_bytes, med * 1e3, 2 * n_bytes / med / 1e6))
            print(f"{name:9s} {fn.__name__:8s} {n_bytes:>10d} B  {med * 1e3:9.3f} ms / request  {2 * n_bytes / med / 1e6:10.1f} MB/s", flush=True)
    finally:
        for m in (callee_mgr, caller_mgr):
            try:
                m.stop_receive_message()
            except Exception:  # noqa: BLE001
                pass
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backends", nargs="+", default=["INPROC", "TCP", "GRPC", "GRPC_REF", "MQTT", "MQTT_BIN", "MQTT_S3"])
    ap.add_argument("--sizes", nargs="+", type=int, default=[4096, 262144, 4194304, 16777216])
    ap.add_argument("--functions", nargs="+", default=["identity", "heavy"])
    ap.add_argument("--repeat", type=int, default=5)
    ap.add_argument("--cuda", action="store_true")
    ap.add_argument("--markdown", default="")
    a = ap.parse_args()
    if a.cuda and not torch.cuda.is_available():
        raise SystemExit("--cuda needs a GPU")
    from fedml_b200.core.distributed.communication.broker import LocalBroker
    broker = LocalBroker()
    port = broker.start()
    rows = []
    with tempfile.TemporaryDirectory() as store:
        for fn_name in a.functions:
            fn = {"identity": identity, "heavy": heavy}[fn_name]
            for b in a.backends:
                try:
                    rows += run_backend(b, a.sizes, fn, a.repeat, a.cuda, port, store)
                except Exception as exc:  # noqa: BLE001 - keep the table going
                    print(f"{b}: failed: {exc!r}", flush=True)
    broker.stop()
    if a.markdown:
        os.makedirs(os.path.dirname(os.path.abspath(a.markdown)) or ".", exist_ok=True)
        with open(a.markdown, "w") as fh:
            fh.write(f"# Transport round-trip latency ({'cuda' if a.cuda else 'cpu'} tensors, {NUM_RPC} requests in flight, median of "
                     f"{a.repeat}, both endpoints in one process on localhost)\n\n"
                     "`python tools/bench_comm.py`: the caller sends a float32 tensor, the callee applies the function and returns it.\n"
                     "GRPC_REF = the reference's wire format (protobuf `CommRequest` + pickled `Message`); GRPC = the binary tensor codec; MQTT = JSON lists (the reference's format), MQTT_BIN = `mqtt_binary: true`.\n\n"
                     "| transport | function | payload | ms / request | MB/s (both directions) |\n|---|---|---:|---:|---:|\n")
            for name, fn_name, n, ms, mbs in rows:
                fh.write(f"| {name} | {fn_name} | {n} | {ms:.3f} | {mbs:.1f} |\n")
        print("wrote", a.markdown)
    sys.stdout.flush()
    os._exit(0)          # transports keep daemon threads (gRPC pools, socket readers); skip interpreter teardown


if __name__ == "__main__":
    main()
