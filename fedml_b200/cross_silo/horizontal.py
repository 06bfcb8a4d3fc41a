"""Cross-silo (Octopus) horizontal FL: server + client over any WAN-style backend.

Behaviour of the reference (cross_silo/horizontal/*.py): same aggregator/trainer as the
worker-per-process FedAvg, plus real edge ids from ``client_id_list``; handshake — the transport emits a
synthetic message 0 "connection ready" → each client reports ``C2S_CLIENT_STATUS {ONLINE, os}`` → the server
waits until every id is online → ``send_init_msg``; separate *client selection* (which edges train) and *data
silo selection* (which partition each gets), both seeded by the round index; MLOps events around ``train``,
``aggregate``, ``comm_c2s``, ``server.wait``.  Added: a per-round deadline with partial aggregation and
re-normalised weights (the reference hangs forever on a dead client)."""
import logging
import platform
import time

import numpy as np

from ..core.distributed.communication.message import Message
from ..core.distributed.managers import ClientManager, ServerManager
from ..core.mlops import MLOpsMetrics, MLOpsProfilerEvent
from ..simulation.mp.fedavg import FedAVGAggregator, FedAVGTrainer, MyMessage
from ..simulation.trainers import create_model_trainer


class FedMLAggregator(FedAVGAggregator):
    def data_silo_selection(self, round_idx, client_num_in_total, client_num_per_round):
        assert client_num_in_total >= client_num_per_round
        if client_num_in_total == client_num_per_round:
            return list(range(client_num_in_total))
        np.random.seed(round_idx)
        return np.random.choice(range(client_num_in_total), client_num_per_round, replace=False).tolist()

    def client_selection(self, round_idx, client_id_list_in_total, client_num_per_round):
        if client_num_per_round == len(client_id_list_in_total):
            return list(client_id_list_in_total)
        np.random.seed(round_idx)
        return np.random.choice(client_id_list_in_total, client_num_per_round, replace=False).tolist()


class FedMLServerManager(ServerManager):
    announce_ready = True

    def __init__(self, args, aggregator, comm=None, rank=0, client_num=0, backend="INPROC"):
        super().__init__(args, comm, rank, client_num + 1, backend)
        self.aggregator, self.round_num, self.round_idx = aggregator, int(args.comm_round), 0
        ids = args.client_ids() if hasattr(args, "client_ids") else []
        self.client_real_ids = ids or list(range(1, client_num + 1))
        self.client_online_mapping = {}
        self.is_initialized = False
        self.client_id_list_in_this_round = None
        self.data_silo_index_list = None
        self.global_model_file_path = getattr(args, "global_model_file_path", None)
        self.events = MLOpsProfilerEvent.get_instance(args)
        self.mlops = MLOpsMetrics.get_instance(args)
        self.deadline_s = float(getattr(args, "round_deadline_s", 0.0) or 0.0)
        self._timer = None

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MyMessage.MSG_TYPE_CONNECTION_IS_READY, self.handle_message_connection_ready)
        self.register_message_receive_handler(MyMessage.MSG_TYPE_C2S_CLIENT_STATUS, self.handle_message_client_status_update)
        self.register_message_receive_handler(MyMessage.MSG_TYPE_C2S_SEND_MODEL_TO_SERVER,
                                              self.handle_message_receive_model_from_client)

    def handle_message_connection_ready(self, msg_params):
        self.mlops.report_server_training_status(getattr(self.args, "run_id", 0), "RUNNING")

    def handle_message_client_status_update(self, msg_params):
        if msg_params.get(MyMessage.MSG_ARG_KEY_CLIENT_STATUS) == "ONLINE":
            self.client_online_mapping[str(msg_params.get(MyMessage.MSG_ARG_KEY_SENDER))] = True
        all_online = all(self.client_online_mapping.get(str(r), False) for r in range(1, self.size))
        logging.info("online clients: %s", sorted(self.client_online_mapping))
        if all_online and not self.is_initialized:
            self.is_initialized = True
            self.send_init_msg()

    def _select(self):
        self.client_id_list_in_this_round = self.aggregator.client_selection(
            self.round_idx, self.client_real_ids, int(self.args.client_num_per_round))
        self.data_silo_index_list = self.aggregator.data_silo_selection(
            self.round_idx, int(self.args.client_num_in_total), len(self.client_id_list_in_this_round))

    def _rank_of(self, edge_id):
        return self.client_real_ids.index(edge_id) + 1

    def send_init_msg(self):
        self._select()
        params = self.aggregator.get_global_model_params()
        self.events.log_event_started("server.wait", str(self.round_idx))
        for i, edge in enumerate(self.client_id_list_in_this_round):
            self._send(MyMessage.MSG_TYPE_S2C_INIT_CONFIG, self._rank_of(edge), params, self.data_silo_index_list[i])
        self._arm_deadline()

    def _send(self, mtype, receive_id, params, client_index):
        m = Message(mtype, self.get_sender_id(), receive_id)
        m.add_params(MyMessage.MSG_ARG_KEY_MODEL_PARAMS, params)
        m.add_params(MyMessage.MSG_ARG_KEY_CLIENT_INDEX, str(client_index))
        self.send_message(m)

    def _arm_deadline(self):
        if self.deadline_s <= 0:
            return
        import threading
        if self._timer is not None:
            self._timer.cancel()
        rnd = self.round_idx
        self._timer = threading.Timer(self.deadline_s, lambda: self._deadline(rnd))
        self._timer.daemon = True
        self._timer.start()

    def _deadline(self, rnd):
        if rnd != self.round_idx:
            return
        if self.aggregator.received_indices():
            logging.warning("round %d deadline hit: partial aggregation", rnd)
            self._finish_round(self.aggregator.received_indices())
        else:
            self._arm_deadline()        # nothing in yet (slow first result): keep waiting in deadline-sized steps

    def handle_message_receive_model_from_client(self, msg_params):
        sender = int(msg_params.get(MyMessage.MSG_ARG_KEY_SENDER))
        self.events.log_event_ended("comm_c2s", str(self.round_idx), sender)
        slot = self.client_id_list_in_this_round.index(self.client_real_ids[sender - 1])
        self.aggregator.add_local_trained_result(slot, msg_params.get(MyMessage.MSG_ARG_KEY_MODEL_PARAMS),
                                                 msg_params.get(MyMessage.MSG_ARG_KEY_NUM_SAMPLES))
        n_round = len(self.client_id_list_in_this_round)
        if len(self.aggregator.received_indices()) == n_round:
            self._finish_round(list(range(n_round)))

    def _finish_round(self, indices):
        self.events.log_event_ended("server.wait", str(self.round_idx))
        self.events.log_event_started("aggregate", str(self.round_idx))
        params = self.aggregator.aggregate(indices)
        self.events.log_event_ended("aggregate", str(self.round_idx))
        self.aggregator.test_on_server_for_all_clients(self.round_idx)
        if self.aggregator.history:
            h = self.aggregator.history[-1]
            self.mlops.report_server_training_metric({"run_id": getattr(self.args, "run_id", 0), "round_idx": self.round_idx,
                                                      "timestamp": time.time(), "accuracy": h["test_acc"], "loss": h["test_loss"]})
        if self.global_model_file_path:
            from ..utils.checkpoint import save_state_dict
            save_state_dict(self.global_model_file_path, params)
        self.round_idx += 1
        if self.round_idx == self.round_num:
            for r in range(1, self.size):
                self.send_message(Message(MyMessage.MSG_TYPE_S2C_FINISH, self.get_sender_id(), r))
            self.mlops.report_server_training_status(getattr(self.args, "run_id", 0), "FINISHED")
            if self._timer is not None:
                self._timer.cancel()
            self.finish()
            return
        self._select()
        self.events.log_event_started("server.wait", str(self.round_idx))
        for i, edge in enumerate(self.client_id_list_in_this_round):
            self._send(MyMessage.MSG_TYPE_S2C_SYNC_MODEL_TO_CLIENT, self._rank_of(edge), params, self.data_silo_index_list[i])
        self._arm_deadline()


class FedMLClientManager(ClientManager):
    announce_ready = True

    def __init__(self, args, trainer, comm=None, rank=0, size=0, backend="INPROC"):
        super().__init__(args, comm, rank, size, backend)
        self.trainer, self.num_rounds, self.round_idx = trainer, int(args.comm_round), 0
        self.events = MLOpsProfilerEvent.get_instance(args)
        self.mlops = MLOpsMetrics.get_instance(args)
        self.has_sent_online_msg = False
        self._got_model = False

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MyMessage.MSG_TYPE_CONNECTION_IS_READY, self.handle_message_connection_ready)
        self.register_message_receive_handler(MyMessage.MSG_TYPE_S2C_INIT_CONFIG, self.handle_message_init)
        self.register_message_receive_handler(MyMessage.MSG_TYPE_S2C_SYNC_MODEL_TO_CLIENT,
                                              self.handle_message_receive_model_from_server)
        self.register_message_receive_handler(MyMessage.MSG_TYPE_S2C_FINISH, self.handle_message_finish)

    def handle_message_connection_ready(self, msg_params):
        if not self.has_sent_online_msg:
            self.has_sent_online_msg = True
            self.send_client_status(0)
            # pub/sub transports drop messages published before the server has subscribed: repeat the ONLINE
            # report until the first model arrives (the server de-duplicates by sender)
            import threading

            def again():
                while not self._got_model and self.com_manager_status != "STOPPED":
                    time.sleep(0.5)
                    if not self._got_model:
                        try:
                            self.send_client_status(0)
                        except Exception:  # noqa: BLE001 - transport already closed
                            return

            self._got_model = False
            threading.Thread(target=again, daemon=True).start()

    def send_client_status(self, receive_id, status="ONLINE"):
        m = Message(MyMessage.MSG_TYPE_C2S_CLIENT_STATUS, self.rank, receive_id)
        m.add_params(MyMessage.MSG_ARG_KEY_CLIENT_STATUS, status)
        m.add_params(MyMessage.MSG_ARG_KEY_CLIENT_OS, platform.system())
        self.send_message(m)

    def handle_message_init(self, msg_params):
        self.round_idx = 0
        self._got_model = True
        self.mlops.report_client_training_status(getattr(self.args, "edge_id", self.rank), "TRAINING")
        self._train_on(msg_params)

    def handle_message_receive_model_from_server(self, msg_params):
        self.round_idx += 1
        self._got_model = True
        self._train_on(msg_params)

    def handle_message_finish(self, msg_params):
        self.mlops.report_client_training_status(getattr(self.args, "edge_id", self.rank), "FINISHED")
        self.finish()

    def _train_on(self, msg_params):
        self.trainer.update_model(msg_params.get(MyMessage.MSG_ARG_KEY_MODEL_PARAMS))
        self.trainer.update_dataset(int(msg_params.get(MyMessage.MSG_ARG_KEY_CLIENT_INDEX)))
        self.events.log_event_started("train", str(self.round_idx))
        weights, n = self.trainer.train(self.round_idx)
        self.events.log_event_ended("train", str(self.round_idx))
        self.events.log_event_started("comm_c2s", str(self.round_idx))
        m = Message(MyMessage.MSG_TYPE_C2S_SEND_MODEL_TO_SERVER, self.rank, 0)
        m.add_params(MyMessage.MSG_ARG_KEY_MODEL_PARAMS, weights)
        m.add_params(MyMessage.MSG_ARG_KEY_NUM_SAMPLES, n)
        self.send_message(m)
        url = m.get(MyMessage.MSG_ARG_KEY_MODEL_PARAMS_URL)
        if url:
            self.mlops.report_client_model_info({"run_id": getattr(self.args, "run_id", 0), "edge_id": self.rank,
                                                 "round_idx": self.round_idx, "client_model_s3_address": url})


def FedML_Horizontal(args, client_rank, client_num, comm, device, dataset, model, model_trainer=None,  # noqa: N802
                     preprocessed_sampling_lists=None, server_aggregator=None):
    """Reference entry (cross_silo/horizontal/fedml_horizontal_api.py:10-120)."""
    (train_num, test_num, train_global, test_global, local_num, train_local, test_local, class_num) = list(dataset)[:8]
    backend = getattr(args, "backend", "INPROC")
    trainer = model_trainer or create_model_trainer(model, args)
    if client_rank == 0:
        agg = FedMLAggregator(train_global, test_global, train_num, train_local, test_local, local_num, client_num, device,
                              args, trainer, server_aggregator)
        mgr = FedMLServerManager(args, agg, comm, 0, client_num, backend)
        mgr.run()
        return agg
    tr = FedAVGTrainer(client_rank - 1, train_local, local_num, test_local, train_num, device, args, trainer)
    mgr = FedMLClientManager(args, tr, comm, client_rank, client_num + 1, backend)
    mgr.run()
    return tr
