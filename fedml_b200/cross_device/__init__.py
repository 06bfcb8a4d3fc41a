"""``fedml.cross_device`` — Beehive server (reference: cross_device/**, 1241 LoC).

The Python server aggregates model *files* uploaded by phone clients: read each upload as
``{layer_idx: tensor}``, sample-weighted average (server_mnn/fedml_aggregator.py:67-94), write back, evaluate
on the server's test loader, publish the new global file through MQTT_S3_MNN; it also pushes a
``start_train`` JSON to the device agents (fedml_server_manager.py:60-157)."""
import json
import logging
import os
import time

import torch

from ..core import aggregation as A
from ..core.distributed.communication.message import Message
from ..core.distributed.managers import ServerManager
from ..simulation.mp.fedavg import MyMessage
from ..simulation.sp.fedavg_api import client_sampling
from .mnn_file import load_into_module, read_mnn_as_tensor_dict, read_mnn_like, write_mnn_like, write_tensor_dict_to_mnn


class FedMLAggregatorMNN:
    def __init__(self, test_dataloader, worker_num, device, args, model):
        self.test_global, self.worker_num, self.device, self.args, self.model = test_dataloader, worker_num, device, args, model
        self.model_dict, self.sample_num_dict = {}, {}
        self.flag = {i: False for i in range(worker_num)}
        self.history = []
        self.global_model_file_path = getattr(args, "global_model_file_path", "./model_file_cache/global_model.mnn")
        if not os.path.exists(self.global_model_file_path):
            write_mnn_like(self.global_model_file_path, model.state_dict())

    def get_global_model_params(self):
        return self.global_model_file_path

    def add_local_trained_result(self, index, model_file, sample_num):
        self.model_dict[index], self.sample_num_dict[index], self.flag[index] = model_file, sample_num, True

    def check_whether_all_receive(self):
        if all(self.flag.values()):
            self.flag = {i: False for i in self.flag}
            return True
        return False

    def aggregate(self):
        t0 = time.time()
        locals_ = [(self.sample_num_dict[i], read_mnn_as_tensor_dict(self.model_dict[i])) for i in sorted(self.model_dict)]
        avg = A.fedavg_aggregate(locals_)
        write_tensor_dict_to_mnn(self.global_model_file_path, avg)
        self.model_dict.clear(); self.sample_num_dict.clear()
        logging.info("aggregate time cost: %.3f", time.time() - t0)
        return self.global_model_file_path

    def client_sampling(self, round_idx, total, per_round):
        return client_sampling(round_idx, total, per_round)

    def test_on_server_for_all_clients(self, round_idx):
        # the reference evaluates the global .mnn file with the MNN runtime (fedml_aggregator.py:169-228); here the graph is
        # executed by mnn_format.run_graph when it consists of supported ops, else its parameters are loaded into the
        # equivalent torch module
        from . import mnn_format as MF
        graph_ops = None
        try:
            graph_ops = MF.describe(self.global_model_file_path)
            convs = [o for o in graph_ops if o["type"] == MF.OP_CONV]
            if not any(o["type"] not in (MF.OP_TRAINABLE_PARAM, MF.OP_INPUT) for o in graph_ops):
                graph_ops = None                       # parameters-only file: nothing to execute
        except (MF.MnnFormatError, Exception):  # noqa: BLE001  (legacy container, unreadable file)
            graph_ops = None
        if graph_ops is None:
            load_into_module(self.global_model_file_path, self.model)
            self.model.to(self.device).eval()
        self.eval_engine = "mnn_graph" if graph_ops is not None else "torch_module"
        correct = loss = total = 0.0
        with torch.no_grad():
            for x, y in self.test_global:
                if graph_ops is not None:
                    xin = x.float()
                    if convs:                           # NCHW input of the first convolution
                        cin = convs[0]["conv"]["in"]
                        side = int(round((xin[0].numel() / cin) ** 0.5))
                        xin = xin.reshape(len(xin), cin, side, side)
                    try:
                        probs = MF.run_graph(self.global_model_file_path, xin)
                    except MF.MnnFormatError:
                        graph_ops, self.eval_engine = None, "torch_module"
                        load_into_module(self.global_model_file_path, self.model)
                        self.model.to(self.device).eval()
                if graph_ops is not None:
                    out = probs.clamp_min(1e-12).log()       # the graph ends in a softmax
                    loss += float(torch.nn.functional.nll_loss(out, y, reduction="sum"))
                else:
                    out = self.model(x.to(self.device)).cpu()
                    loss += float(torch.nn.functional.cross_entropy(out, y, reduction="sum"))
                correct += float((out.argmax(1) == y).sum()); total += len(y)
        stats = {"round": round_idx, "test_acc": correct / max(total, 1), "test_loss": loss / max(total, 1)}
        self.history.append(stats)
        logging.info("%s", stats)
        return stats


class FedMLServerManagerMNN(ServerManager):
    announce_ready = True

    def __init__(self, args, aggregator, comm=None, rank=0, size=0, backend="MQTT_S3_MNN"):
        super().__init__(args, comm, rank, size, backend)
        self.aggregator, self.round_num, self.round_idx = aggregator, int(args.comm_round), 0
        self.started = False

    def start_train_payload(self):
        """JSON pushed to device agents on ``flserver_agent/<edge>/start_train`` (fedml_server_manager.py:60-110)."""
        ids = self.args.client_ids() if hasattr(self.args, "client_ids") else []
        return {"edges": [{"id": e, "device_id": str(e), "os_type": "Android"} for e in ids], "runId": getattr(self.args, "run_id", 0),
                "run_config": {"configName": "fedml_b200", "parameters": {k: v for k, v in vars(self.args).items()
                                                                          if isinstance(v, (int, float, str, bool))}},
                "timestamp": str(int(time.time()))}

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MyMessage.MSG_TYPE_CONNECTION_IS_READY, self.handle_connection_ready)
        self.register_message_receive_handler(MyMessage.MSG_TYPE_C2S_SEND_MODEL_TO_SERVER, self.handle_message_receive_model)

    def handle_connection_ready(self, msg_params):
        if not self.started:
            self.started = True
            self.send_init_msg()

    def send_init_msg(self):
        idx = self.aggregator.client_sampling(self.round_idx, int(self.args.client_num_in_total), int(self.args.client_num_per_round))
        for pid in range(1, self.size):
            self._send(MyMessage.MSG_TYPE_S2C_INIT_CONFIG, pid, idx[(pid - 1) % len(idx)])

    def _send(self, mtype, pid, client_index):
        m = Message(mtype, 0, pid)
        m.add_params(MyMessage.MSG_ARG_KEY_MODEL_PARAMS, self.aggregator.get_global_model_params())
        m.add_params(MyMessage.MSG_ARG_KEY_CLIENT_INDEX, str(client_index))
        self.send_message(m)

    def handle_message_receive_model(self, msg_params):
        sender = int(msg_params.get(MyMessage.MSG_ARG_KEY_SENDER))
        self.aggregator.add_local_trained_result(sender - 1, msg_params.get(MyMessage.MSG_ARG_KEY_MODEL_PARAMS),
                                                 msg_params.get(MyMessage.MSG_ARG_KEY_NUM_SAMPLES))
        if self.aggregator.check_whether_all_receive():
            self.aggregator.aggregate()
            self.aggregator.test_on_server_for_all_clients(self.round_idx)
            self.round_idx += 1
            if self.round_idx == self.round_num:
                for pid in range(1, self.size):
                    self.send_message(Message(MyMessage.MSG_TYPE_S2C_FINISH, 0, pid))
                self.finish()
                return
            idx = self.aggregator.client_sampling(self.round_idx, int(self.args.client_num_in_total),
                                                  int(self.args.client_num_per_round))
            for pid in range(1, self.size):
                self._send(MyMessage.MSG_TYPE_S2C_SYNC_MODEL_TO_CLIENT, pid, idx[(pid - 1) % len(idx)])


class ServerMNN:
    def __init__(self, args, device, test_dataloader, model, model_trainer=None):
        self.args, self.device, self.model = args, device, model
        if isinstance(test_dataloader, (list, tuple)) and len(test_dataloader) >= 8 and isinstance(test_dataloader[0], int):
            test_dataloader = test_dataloader[3]
        n = int(getattr(args, "client_num_per_round", 1))
        self.aggregator = FedMLAggregatorMNN(test_dataloader, n, device, args, model)
        self.manager = FedMLServerManagerMNN(args, self.aggregator, None, 0, n + 1, getattr(args, "backend", "MQTT_S3_MNN"))

    def run(self):
        self.manager.run()
        return self.aggregator


__all__ = ["ServerMNN", "FedMLAggregatorMNN", "FedMLServerManagerMNN"]
