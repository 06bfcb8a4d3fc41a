"""Algorithm templates over the managers (reference: mpi_p2p_mp/base_framework/*.py — clients send a number,
the server sums and broadcasts; mpi_p2p_mp/decentralized_framework/*.py — ring topology, neighbour exchange,
no server).  They are the smallest complete examples of the worker-oriented message-passing style."""
import logging

from ...core.distributed.communication.message import Message
from ...core.distributed.managers import ClientManager, ServerManager
from ...core.distributed.topology import SymmetricTopologyManager

MSG_S2C_INIT, MSG_S2C_INFO, MSG_C2S_INFO, MSG_FINISH, MSG_NEIGHBOR = 1, 2, 3, 6, 7


class BaseCentralManager(ServerManager):
    def __init__(self, args, comm, rank, size, backend):
        super().__init__(args, comm, rank, size, backend)
        self.round_num, self.round_idx, self.got, self.global_result = int(args.comm_round), 0, {}, 0
        self.history = []

    def send_init_msg(self):
        for r in range(1, self.size):
            m = Message(MSG_S2C_INIT, 0, r); m.add_params("global_result", self.global_result); self.send_message(m)

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_C2S_INFO, self.handle_client_info)

    def handle_client_info(self, p):
        self.got[p["sender"]] = p["client_info"]
        if len(self.got) == self.size - 1:
            self.global_result = sum(self.got.values())
            self.history.append(self.global_result)
            self.got = {}
            self.round_idx += 1
            last = self.round_idx == self.round_num
            for r in range(1, self.size):
                m = Message(MSG_FINISH if last else MSG_S2C_INFO, 0, r); m.add_params("global_result", self.global_result)
                self.send_message(m)
            if last:
                self.finish()


class BaseClientManager(ClientManager):
    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_S2C_INIT, self.handle)
        self.register_message_receive_handler(MSG_S2C_INFO, self.handle)
        self.register_message_receive_handler(MSG_FINISH, lambda p: self.finish())

    def handle(self, p):
        m = Message(MSG_C2S_INFO, self.rank, 0)
        m.add_params("client_info", float(p["global_result"]) / max(self.size - 1, 1) + self.rank)   # "local training"
        self.send_message(m)


def FedML_Base_distributed(args, process_id, worker_number, comm, device=None, dataset=None, model=None,  # noqa: N802
                           model_trainer=None, backend="INPROC", **_):
    if process_id == 0:
        mgr = BaseCentralManager(args, comm, 0, worker_number, backend)
        mgr.send_init_msg(); mgr.run()
        return mgr
    mgr = BaseClientManager(args, comm, process_id, worker_number, backend)
    mgr.run()
    return mgr


class DecentralizedWorkerManager(ClientManager):
    """Every rank is a worker: per round it sends its value to its out-neighbours and, once all in-neighbours
    reported, mixes with the topology weights (decentralized_framework/decentralized_worker_manager.py)."""

    def __init__(self, args, comm, rank, size, backend, topology):
        super().__init__(args, comm, rank, size, backend)
        self.topology, self.round_num, self.round_idx = topology, int(args.comm_round), 0
        self.value = float(rank)
        self.inbox = {}
        self.history = [self.value]

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_NEIGHBOR, self.handle_neighbor)

    def start(self):
        self._send_to_neighbors()

    def _send_to_neighbors(self):
        for nb in self.topology.get_out_neighbor_idx_list(self.rank):
            m = Message(MSG_NEIGHBOR, self.rank, nb); m.add_params("value", self.value); m.add_params("round", self.round_idx)
            self.send_message(m)

    def handle_neighbor(self, p):
        self.inbox.setdefault(p["round"], {})[p["sender"]] = p["value"]
        self._maybe_mix()

    def _maybe_mix(self):
        need = self.topology.get_in_neighbor_idx_list(self.rank)
        got = self.inbox.get(self.round_idx, {})
        while all(n in got for n in need):
            w = self.topology.get_in_neighbor_weights(self.rank)
            self.value = w[self.rank] * self.value + sum(w[n] * got[n] for n in need)
            self.history.append(self.value)
            self.inbox.pop(self.round_idx, None)
            self.round_idx += 1
            if self.round_idx == self.round_num:
                self.finish()
                return
            self._send_to_neighbors()
            got = self.inbox.get(self.round_idx, {})


def FedML_Decentralized_Demo_distributed(args, process_id, worker_number, comm, device=None, dataset=None,  # noqa: N802
                                         model=None, model_trainer=None, backend="INPROC", **_):
    tm = SymmetricTopologyManager(worker_number, 2)
    tm.generate_topology()
    mgr = DecentralizedWorkerManager(args, comm, process_id, worker_number, backend, tm)
    mgr.register_message_receive_handlers()
    mgr.start()
    mgr.run()
    return mgr
