"""Worker-per-process FedAvg family over the message managers (the reference's "MPI simulator").

Protocol (reference: simulation/mpi_p2p_mp/fedavg/*.py): rank 0 = aggregator + server
manager; ranks 1..W = trainer + client manager; ``worker_num`` processes stand in for the clients of a round:
each round the server samples K client indices and tells worker j to *become* client ``client_indexes[j-1]``;
the worker swaps its data, trains, uploads ``(state_dict, n_i)``; the server aggregates once all W results
are in.  Message types 1..4 and payload keys are the reference's (message_define.py:7-30).

One implementation serves FedAvg / FedOpt / FedProx / FedAvg_robust / FedSeg-style metric upload: the
aggregator takes the same algorithm hooks as the single-process simulator."""
import logging
import time
from collections import OrderedDict

import numpy as np
import torch

from ...core import aggregation as A
from ...core.distributed.communication.message import Message
from ...core.distributed.managers import ClientManager, ServerManager
from ...core.robustness import RobustAggregator
from ...utils.metrics import MetricsSink
from ..sp.fedavg_api import client_sampling
from ..trainers import create_model_trainer


class MyMessage:
    MSG_TYPE_CONNECTION_IS_READY = 0
    MSG_TYPE_S2C_INIT_CONFIG = 1
    MSG_TYPE_S2C_SYNC_MODEL_TO_CLIENT = 2
    MSG_TYPE_C2S_SEND_MODEL_TO_SERVER = 3
    MSG_TYPE_C2S_SEND_STATS_TO_SERVER = 4
    MSG_TYPE_C2S_CLIENT_STATUS = 5
    MSG_TYPE_S2C_FINISH = 6

    MSG_ARG_KEY_TYPE = "msg_type"
    MSG_ARG_KEY_SENDER = "sender"
    MSG_ARG_KEY_RECEIVER = "receiver"
    MSG_ARG_KEY_NUM_SAMPLES = "num_samples"
    MSG_ARG_KEY_MODEL_PARAMS = "model_params"
    MSG_ARG_KEY_MODEL_PARAMS_URL = "model_params_url"
    MSG_ARG_KEY_CLIENT_INDEX = "client_idx"
    MSG_ARG_KEY_CLIENT_STATUS = "client_status"
    MSG_ARG_KEY_CLIENT_OS = "client_os"
    MSG_ARG_KEY_TRAIN_CORRECT = "train_correct"
    MSG_ARG_KEY_TRAIN_ERROR = "train_error"
    MSG_ARG_KEY_TRAIN_NUM = "train_num_sample"
    MSG_ARG_KEY_TEST_CORRECT = "test_correct"
    MSG_ARG_KEY_TEST_ERROR = "test_error"
    MSG_ARG_KEY_TEST_NUM = "test_num_sample"


def transform_tensor_to_list(model_params):
    return {k: v.detach().cpu().numpy().tolist() for k, v in model_params.items()}


def transform_list_to_tensor(model_params_list):
    return OrderedDict((k, torch.as_tensor(np.asarray(v))) for k, v in model_params_list.items())


class FedAVGAggregator:
    def __init__(self, train_global, test_global, all_train_data_num, train_data_local_dict, test_data_local_dict,
                 train_data_local_num_dict, worker_num, device, args, model_trainer, server_aggregator=None):
        self.trainer, self.args, self.device = model_trainer, args, device
        self.train_global, self.test_global, self.all_train_data_num = train_global, test_global, all_train_data_num
        self.train_data_local_dict, self.test_data_local_dict = train_data_local_dict, test_data_local_dict
        self.train_data_local_num_dict = train_data_local_num_dict
        self.worker_num = worker_num
        self.model_dict, self.sample_num_dict = {}, {}
        self.flag_client_model_uploaded_dict = {i: False for i in range(worker_num)}
        self.server_aggregator = server_aggregator
        self.metrics = MetricsSink(args)
        self.history = []
        alg = str(getattr(args, "federated_optimizer", "FedAvg"))
        self.server_opt = None
        if alg.lower() in ("fedopt", "fedadam", "fedyogi", "fedadagrad"):
            self.server_opt = A.ServerOptimizer(getattr(args, "server_optimizer", "sgd"),
                                                lr=float(getattr(args, "server_lr", 1.0)),
                                                momentum=float(getattr(args, "server_momentum", 0.0)))
            self._param_names = [n for n, _ in self.trainer.model.named_parameters()]
        self.robust = RobustAggregator(args) if getattr(args, "defense_type", None) else None

    def get_global_model_params(self):
        return self.trainer.get_model_params()

    def set_global_model_params(self, model_parameters):
        self.trainer.set_model_params(model_parameters)

    def add_local_trained_result(self, index, model_params, sample_num):
        self.model_dict[index] = model_params
        self.sample_num_dict[index] = sample_num
        self.flag_client_model_uploaded_dict[index] = True

    def check_whether_all_receive(self):
        if not all(self.flag_client_model_uploaded_dict.get(i, False) for i in range(self.worker_num)):
            return False
        for i in range(self.worker_num):
            self.flag_client_model_uploaded_dict[i] = False
        return True

    def received_indices(self):
        return [i for i in range(self.worker_num) if i in self.model_dict]

    def aggregate(self, indices=None):
        """Sample-weighted mean over the received models.  ``indices`` restricts it to a subset — used by the
        round deadline (partial aggregation with re-normalised weights)."""
        t0 = time.time()
        idx = list(range(self.worker_num)) if indices is None else list(indices)
        w_locals = []
        for i in idx:
            mp = self.model_dict[i]
            if int(getattr(self.args, "is_mobile", 0)) == 1 and not torch.is_tensor(next(iter(mp.values()))):
                mp = transform_list_to_tensor(mp)
            w_locals.append((self.sample_num_dict[i], mp))
        w_global = self.get_global_model_params()
        if self.server_aggregator is not None:
            agg = self.server_aggregator.on_after_aggregation(
                self.server_aggregator.aggregate(self.server_aggregator.on_before_aggregation(w_locals)))
        elif self.robust is not None:
            agg = self.robust.defend(w_locals, w_global)
        else:
            agg = A.fedavg_aggregate(w_locals)
        if self.server_opt is not None:
            agg = self.server_opt.step(w_global, agg, self._param_names)
        agg = A.cast_like(agg, w_global)
        self.set_global_model_params(agg)
        self.model_dict.clear(); self.sample_num_dict.clear()
        logging.info("aggregate time cost: %.3f", time.time() - t0)
        return agg

    def client_sampling(self, round_idx, client_num_in_total, client_num_per_round):
        return client_sampling(round_idx, client_num_in_total, client_num_per_round)

    def test_on_server_for_all_clients(self, round_idx):
        if self.trainer.test_on_the_server(self.train_data_local_dict, self.test_data_local_dict, self.device, self.args):
            return
        freq = int(getattr(self.args, "frequency_of_the_test", 1) or 1)
        if not (round_idx % freq == 0 or round_idx == int(self.args.comm_round) - 1):
            return
        agg = [0.0, 0.0, 0.0]
        n_total = int(self.args.client_num_in_total)
        for cid in range(n_total):
            m = self.trainer.test(self.train_data_local_dict[cid], self.device, self.args)
            agg[0] += m["test_correct"]; agg[1] += m["test_loss"]; agg[2] += m["test_total"]
            if int(getattr(self.args, "ci", 0)) == 1:
                break
        tm = self.trainer.test(self.test_global, self.device, self.args)
        stats = {"round": round_idx, "train_acc": agg[0] / max(agg[2], 1), "train_loss": agg[1] / max(agg[2], 1),
                 "test_acc": tm["test_correct"] / max(tm["test_total"], 1), "test_loss": tm["test_loss"] / max(tm["test_total"], 1)}
        self.history.append(stats)
        self.metrics.log({"Train/Acc": stats["train_acc"], "Train/Loss": stats["train_loss"], "Test/Acc": stats["test_acc"],
                          "Test/Loss": stats["test_loss"], "round": round_idx})


class FedAVGTrainer:
    def __init__(self, client_index, train_data_local_dict, train_data_local_num_dict, test_data_local_dict,
                 train_data_num, device, args, model_trainer):
        self.trainer, self.client_index = model_trainer, client_index
        self.train_data_local_dict, self.train_data_local_num_dict = train_data_local_dict, train_data_local_num_dict
        self.test_data_local_dict, self.all_train_data_num = test_data_local_dict, train_data_num
        self.device, self.args = device, args
        self.train_local = self.local_sample_number = self.test_local = None

    def update_model(self, weights):
        if not torch.is_tensor(next(iter(weights.values()))):
            weights = transform_list_to_tensor(weights)
        self.trainer.set_model_params(weights)

    def update_dataset(self, client_index):
        self.client_index = client_index
        self.train_local = self.train_data_local_dict[client_index]
        self.local_sample_number = self.train_data_local_num_dict[client_index]
        self.test_local = self.test_data_local_dict[client_index]
        self.trainer.set_id(client_index)

    def train(self, round_idx=None):
        self.args.round_idx = round_idx
        self.trainer.train(self.train_local, self.device, self.args)
        weights = self.trainer.get_model_params()
        if int(getattr(self.args, "is_mobile", 0)) == 1:
            weights = transform_tensor_to_list(weights)
        return weights, self.local_sample_number

    def test(self):
        tr = self.trainer.test(self.train_local, self.device, self.args)
        te = self.trainer.test(self.test_local, self.device, self.args)
        return (tr["test_correct"], tr["test_loss"], tr["test_total"], te["test_correct"], te["test_loss"], te["test_total"])


class FedAVGServerManager(ServerManager):
    def __init__(self, args, aggregator, comm=None, rank=0, size=0, backend="MPI", is_preprocessed=False,
                 preprocessed_client_lists=None):
        super().__init__(args, comm, rank, size, backend)
        self.aggregator, self.round_num, self.round_idx = aggregator, int(args.comm_round), 0
        self.is_preprocessed, self.preprocessed_client_lists = is_preprocessed, preprocessed_client_lists
        self.deadline_s = float(getattr(args, "round_deadline_s", 0.0) or 0.0)
        self._round_started = time.time()
        self._timer = None

    def send_init_msg(self):
        client_indexes = self.aggregator.client_sampling(self.round_idx, int(self.args.client_num_in_total),
                                                         int(self.args.client_num_per_round))
        params = self.aggregator.get_global_model_params()
        if int(getattr(self.args, "is_mobile", 0)) == 1:
            params = transform_tensor_to_list(params)
        for pid in range(1, self.size):
            self._send_model(MyMessage.MSG_TYPE_S2C_INIT_CONFIG, pid, params, client_indexes[(pid - 1) % len(client_indexes)])
        self._arm_deadline()

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MyMessage.MSG_TYPE_C2S_SEND_MODEL_TO_SERVER,
                                              self.handle_message_receive_model_from_client)

    def _arm_deadline(self):
        """Per-round deadline: when it fires with at least one result in, aggregate what arrived
        (re-normalised weights) instead of hanging forever on a dead client (reference has no timeout)."""
        if self.deadline_s <= 0:
            return
        import threading
        if self._timer is not None:
            self._timer.cancel()
        rnd = self.round_idx
        self._timer = threading.Timer(self.deadline_s, lambda: self._on_deadline(rnd))
        self._timer.daemon = True
        self._timer.start()

    def _on_deadline(self, rnd):
        if rnd != self.round_idx:
            return
        got = self.aggregator.received_indices()
        if got:
            logging.warning("round %d deadline: aggregating %d of %d results", rnd, len(got), self.aggregator.worker_num)
            self._finish_round(got)
        else:
            self._arm_deadline()        # nothing in yet: keep waiting in deadline-sized steps

    def handle_message_receive_model_from_client(self, msg_params):
        sender = msg_params.get(MyMessage.MSG_ARG_KEY_SENDER)
        self.aggregator.add_local_trained_result(sender - 1, msg_params.get(MyMessage.MSG_ARG_KEY_MODEL_PARAMS),
                                                 msg_params.get(MyMessage.MSG_ARG_KEY_NUM_SAMPLES))
        if self.aggregator.check_whether_all_receive():
            self._finish_round(None)

    def _finish_round(self, indices):
        params = self.aggregator.aggregate(indices)
        for i in range(self.aggregator.worker_num):
            self.aggregator.flag_client_model_uploaded_dict[i] = False
        self.aggregator.test_on_server_for_all_clients(self.round_idx)
        self.round_idx += 1
        if self.round_idx == self.round_num:
            for pid in range(1, self.size):
                self.send_message(Message(MyMessage.MSG_TYPE_S2C_FINISH, self.get_sender_id(), pid))
            if self._timer is not None:
                self._timer.cancel()
            self.finish()
            return
        if self.is_preprocessed:
            client_indexes = [self.round_idx] * int(self.args.client_num_per_round) if self.preprocessed_client_lists is None \
                else self.preprocessed_client_lists[self.round_idx]
        else:
            client_indexes = self.aggregator.client_sampling(self.round_idx, int(self.args.client_num_in_total),
                                                             int(self.args.client_num_per_round))
        if int(getattr(self.args, "is_mobile", 0)) == 1:
            params = transform_tensor_to_list(params)
        for pid in range(1, self.size):
            self._send_model(MyMessage.MSG_TYPE_S2C_SYNC_MODEL_TO_CLIENT, pid, params,
                             client_indexes[(pid - 1) % len(client_indexes)])
        self._arm_deadline()

    def _send_model(self, mtype, receive_id, params, client_index):
        m = Message(mtype, self.get_sender_id(), receive_id)
        m.add_params(MyMessage.MSG_ARG_KEY_MODEL_PARAMS, params)
        m.add_params(MyMessage.MSG_ARG_KEY_CLIENT_INDEX, str(client_index))
        self.send_message(m)

    # reference method names
    def send_message_init_config(self, receive_id, global_model_params, client_index):
        self._send_model(MyMessage.MSG_TYPE_S2C_INIT_CONFIG, receive_id, global_model_params, client_index)

    def send_message_sync_model_to_client(self, receive_id, global_model_params, client_index):
        self._send_model(MyMessage.MSG_TYPE_S2C_SYNC_MODEL_TO_CLIENT, receive_id, global_model_params, client_index)


class FedAVGClientManager(ClientManager):
    def __init__(self, args, trainer, comm=None, rank=0, size=0, backend="MPI"):
        super().__init__(args, comm, rank, size, backend)
        self.trainer, self.num_rounds, self.round_idx = trainer, int(args.comm_round), 0

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MyMessage.MSG_TYPE_S2C_INIT_CONFIG, self.handle_message_init)
        self.register_message_receive_handler(MyMessage.MSG_TYPE_S2C_SYNC_MODEL_TO_CLIENT,
                                              self.handle_message_receive_model_from_server)
        self.register_message_receive_handler(MyMessage.MSG_TYPE_S2C_FINISH, lambda p: self.finish())

    def handle_message_init(self, msg_params):
        self.round_idx = 0
        self._train_on(msg_params)

    def handle_message_receive_model_from_server(self, msg_params):
        self.round_idx += 1
        self._train_on(msg_params)

    def _train_on(self, msg_params):
        self.trainer.update_model(msg_params.get(MyMessage.MSG_ARG_KEY_MODEL_PARAMS))
        self.trainer.update_dataset(int(msg_params.get(MyMessage.MSG_ARG_KEY_CLIENT_INDEX)))
        weights, n = self.trainer.train(self.round_idx)
        self.send_model_to_server(0, weights, n)

    def send_model_to_server(self, receive_id, weights, local_sample_num):
        m = Message(MyMessage.MSG_TYPE_C2S_SEND_MODEL_TO_SERVER, self.get_sender_id(), receive_id)
        m.add_params(MyMessage.MSG_ARG_KEY_MODEL_PARAMS, weights)
        m.add_params(MyMessage.MSG_ARG_KEY_NUM_SAMPLES, local_sample_num)
        self.send_message(m)


def FedML_init():  # noqa: N802
    """``(comm, process_id, worker_number)`` of this process (reference FedAvgAPI.py:12-16).  Under ``mpirun`` with mpi4py:
    ``MPI.COMM_WORLD`` and its rank / size.  Otherwise the job was started by ``torchrun`` (or any launcher that exports ``RANK`` /
    ``WORLD_SIZE``): ``comm`` is None and the managers connect the same star over TCP (``args.host_table``, ``args.tcp_base_port``)."""
    import os
    from ...core.distributed.communication.mpi_backend import mpi_available
    if mpi_available():
        from mpi4py import MPI
        comm = MPI.COMM_WORLD
        if comm.Get_size() > 1 or "RANK" not in os.environ:
            return comm, comm.Get_rank(), comm.Get_size()
    rank = int(os.environ.get("RANK", os.environ.get("OMPI_COMM_WORLD_RANK", os.environ.get("PMI_RANK", 0))))
    size = int(os.environ.get("WORLD_SIZE", os.environ.get("OMPI_COMM_WORLD_SIZE", os.environ.get("PMI_SIZE", 1))))
    return None, rank, size


def init_server(args, device, comm, rank, size, model, train_data_num, train_data_global, test_data_global, train_data_local_dict,
                test_data_local_dict, train_data_local_num_dict, model_trainer=None, preprocessed_sampling_lists=None, backend=None,
                server_aggregator=None):
    """Builds the aggregator + server manager of rank 0, sends the first global model and serves until the last round."""
    trainer = model_trainer or create_model_trainer(model, args)
    trainer.set_id(-1)
    agg = FedAVGAggregator(train_data_global, test_data_global, train_data_num, train_data_local_dict, test_data_local_dict,
                           train_data_local_num_dict, size - 1, device, args, trainer, server_aggregator)
    mgr = FedAVGServerManager(args, agg, comm, rank, size, backend or getattr(args, "mp_transport", "INPROC"),
                              is_preprocessed=preprocessed_sampling_lists is not None,
                              preprocessed_client_lists=preprocessed_sampling_lists)
    mgr.send_init_msg()
    mgr.run()
    return agg


def init_client(args, device, comm, process_id, size, model, train_data_num, train_data_local_num_dict, train_data_local_dict,
                test_data_local_dict, model_trainer=None, backend=None):
    """Builds the local trainer + client manager of a worker rank and serves until the server says finish."""
    client_index = process_id - 1
    trainer = model_trainer or create_model_trainer(model, args)
    trainer.set_id(client_index)
    tr = FedAVGTrainer(client_index, train_data_local_dict, train_data_local_num_dict, test_data_local_dict, train_data_num, device, args,
                       trainer)
    FedAVGClientManager(args, tr, comm, process_id, size, backend or getattr(args, "mp_transport", "INPROC")).run()
    return tr


def FedML_FedAvg_distributed(args, process_id, worker_number, comm, device, dataset, model, model_trainer=None,  # noqa: N802
                             preprocessed_sampling_lists=None, backend=None, server_aggregator=None):
    """Reference entry point (FedAvgAPI.py:19): builds the server (rank 0) or a client (rank > 0) and runs it."""
    (train_num, test_num, train_global, test_global, local_num, train_local, test_local, class_num) = list(dataset)[:8]
    if process_id == 0:
        return init_server(args, device, comm, process_id, worker_number, model, train_num, train_global, test_global, train_local,
                           test_local, local_num, model_trainer, preprocessed_sampling_lists, backend, server_aggregator)
    return init_client(args, device, comm, process_id, worker_number, model, train_num, local_num, train_local, test_local, model_trainer,
                       backend)


FedML_FedOpt_distributed = FedML_FedAvg_distributed          # noqa: N816  server optimizer chosen by args
FedML_FedProx_distributed = FedML_FedAvg_distributed         # noqa: N816  proximal term lives in the trainer


def FedML_FedAvgRobust_distributed(*a, **kw):                # noqa: N802  backdoor protocol: see fedavg_robust.py
    from .fedavg_robust import FedML_FedAvgRobust_distributed as impl
    return impl(*a, **kw)
