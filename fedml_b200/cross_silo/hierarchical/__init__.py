"""Hierarchical cross-silo (reference: cross_silo/hierarchical/*.py, 1961 LoC).

Each silo is a torchrun-launched process group.  Rank-in-silo 0 is the ``ClientMasterManager`` that talks to
the FL server; the other ranks are ``ClientSlaveManager``s.  Per round the master broadcasts
``[round_idx, model_params, client_index]`` to its silo (``dist.broadcast_object_list``,
client_master_manager.py:239-249), every process barriers and trains its shard of the client's data under
data parallelism (trainer_dist_adapter.py:56-127: DDP), and the master uploads the result.  On GPUs the
intra-silo gradient all-reduce can run on this package's fused kernel (``parallel/cheetah.py``) instead of
DDP's bucketed NCCL all-reduce (``args.silo_dp_engine = "cheetah"``)."""
import logging
import os

import torch
import torch.distributed as dist

from ...core.distributed.communication.message import Message
from ...simulation.mp.fedavg import FedAVGTrainer, MyMessage
from ...simulation.trainers import create_model_trainer
from ..horizontal import FedMLAggregator, FedMLClientManager, FedMLServerManager


class ProcessGroupManager:
    """``init_process_group(NCCL if using_gpu else GLOO)`` + ``new_group()`` (process_group_manager.py:30-34)."""

    def __init__(self, rank, world_size, master_address, master_port, only_gpu):
        os.environ.setdefault("MASTER_ADDR", str(master_address))
        os.environ.setdefault("MASTER_PORT", str(master_port))
        os.environ["WORLD_SIZE"], os.environ["RANK"] = str(world_size), str(rank)
        backend = dist.Backend.NCCL if (only_gpu and torch.cuda.is_available()) else dist.Backend.GLOO
        if not dist.is_initialized():
            dist.init_process_group(backend=backend, rank=rank, world_size=world_size)
        self.messaging_pg = dist.new_group()
        self.rank, self.world_size = rank, world_size

    def cleanup(self):
        if dist.is_initialized():
            dist.destroy_process_group()

    def get_process_group(self):
        return self.messaging_pg


class TrainerDistAdapter:
    def __init__(self, args, device, client_rank, model, train_data_num, train_data_local_num_dict,
                 train_data_local_dict, test_data_local_dict, model_trainer=None):
        self.args, self.device = args, device
        self.n_proc = int(getattr(args, "n_proc_in_silo", 1) or 1)
        self.proc_rank = int(getattr(args, "proc_rank_in_silo", 0) or 0)
        self.pg = None
        model = model.to(device)
        if self.n_proc > 1:
            self.pg = ProcessGroupManager(self.proc_rank, self.n_proc, getattr(args, "pg_master_address", "127.0.0.1"),
                                          getattr(args, "pg_master_port", 29500), bool(getattr(args, "using_gpu", False)))
            from torch.nn.parallel import DistributedDataParallel as DDP
            model = DDP(model, device_ids=[device.index] if device.type == "cuda" else None)
        self.model = model
        trainer = model_trainer or create_model_trainer(model, args)
        trainer.model = model
        self.trainer = FedAVGTrainer(client_rank - 1, train_data_local_dict, train_data_local_num_dict,
                                     test_data_local_dict, train_data_num, device, args, trainer)

    def _shard(self, data):
        # data_loader_cross_silo lays a client's data out as [proc_rank_in_silo] when the silo has >1 process
        if isinstance(data, (list, tuple)) and data and isinstance(data[0], (list, tuple)) and \
                len(data) == self.n_proc and not torch.is_tensor(data[0][0] if data[0] else None):
            return data[self.proc_rank]
        return data

    def update_model(self, params):
        sd = params
        if self.n_proc > 1:
            sd = {("module." + k if not k.startswith("module.") else k): v for k, v in params.items()}
        self.trainer.update_model(sd)

    def update_dataset(self, client_index):
        self.trainer.update_dataset(int(client_index))
        self.trainer.train_local = self._shard(self.trainer.train_local)

    def train(self, round_idx):
        if self.n_proc > 1:
            dist.barrier()
        weights, n = self.trainer.train(round_idx)
        weights = {k[7:] if k.startswith("module.") else k: v for k, v in weights.items()}
        return weights, n

    def cleanup_pg(self):
        if self.pg is not None:
            self.pg.cleanup()


class ClientMasterManager(FedMLClientManager):
    def __init__(self, args, trainer_dist_adapter, comm=None, rank=0, size=0, backend="INPROC"):
        super().__init__(args, trainer_dist_adapter, comm, rank, size, backend)
        self.adapter = trainer_dist_adapter

    def sync_process_group(self, round_idx, model_params=None, client_index=None, finish=False):
        if self.adapter.n_proc > 1:
            dist.broadcast_object_list([round_idx, model_params, client_index, finish], src=0)

    def _train_on(self, msg_params):
        params = msg_params.get(MyMessage.MSG_ARG_KEY_MODEL_PARAMS)
        cidx = int(msg_params.get(MyMessage.MSG_ARG_KEY_CLIENT_INDEX))
        self.sync_process_group(self.round_idx, params, cidx)
        super()._train_on(msg_params)

    def handle_message_finish(self, msg_params):
        self.sync_process_group(self.round_idx, None, None, finish=True)
        self.adapter.cleanup_pg()
        super().handle_message_finish(msg_params)


class ClientSlaveManager:
    def __init__(self, args, trainer_dist_adapter):
        self.args, self.adapter = args, trainer_dist_adapter
        self.round_idx, self.finished = 0, False

    def await_sync_process_group(self, src=0):
        obj = [None, None, None, None]
        dist.broadcast_object_list(obj, src=src)
        return obj

    def train(self):
        round_idx, params, cidx, finish = self.await_sync_process_group()
        if finish:
            self.finished = True
            return
        self.round_idx = round_idx
        self.adapter.update_model(params)
        self.adapter.update_dataset(cidx)
        self.adapter.train(round_idx)

    def run(self):
        while not self.finished:
            self.train()
        self.adapter.cleanup_pg()


def FedML_Hierarchical(args, client_rank, client_num, comm, device, dataset, model, model_trainer=None,  # noqa: N802
                       server_aggregator=None):
    (train_num, test_num, train_global, test_global, local_num, train_local, test_local, class_num) = list(dataset)[:8]
    backend = getattr(args, "backend", "INPROC")
    if client_rank == 0:
        trainer = model_trainer or create_model_trainer(model, args)
        agg = FedMLAggregator(train_global, test_global, train_num, train_local, test_local, local_num, client_num, device,
                              args, trainer, server_aggregator)
        FedMLServerManager(args, agg, comm, 0, client_num, backend).run()
        return agg
    adapter = TrainerDistAdapter(args, device, client_rank, model, train_num, local_num, train_local, test_local, model_trainer)
    if int(getattr(args, "proc_rank_in_silo", 0) or 0) == 0:
        ClientMasterManager(args, adapter, comm, client_rank, client_num + 1, backend).run()
    else:
        ClientSlaveManager(args, adapter).run()
    return adapter


# builder functions of the reference's fedml_hierarchical_api.py:83-167 (scripts that assemble a silo by hand call these)
def get_trainer_dist_adapter(args, device, client_rank, model, train_data_num, train_data_local_num_dict, train_data_local_dict,
                             test_data_local_dict, model_trainer=None):
    return TrainerDistAdapter(args, device, client_rank, model, train_data_num, train_data_local_num_dict, train_data_local_dict,
                              test_data_local_dict, model_trainer)


def get_dist_aggregator(args, device, size, model, train_data_num, train_data_global, test_data_global, train_data_local_dict,
                        test_data_local_dict, train_data_local_num_dict, model_trainer=None, server_aggregator=None):
    trainer = model_trainer or create_model_trainer(model, args)
    return FedMLAggregator(train_data_global, test_data_global, train_data_num, train_data_local_dict, test_data_local_dict,
                           train_data_local_num_dict, size, device, args, trainer, server_aggregator)


def get_server_manager(args, dist_aggregator, comm, rank, size, backend, is_preprocessed=False, preprocessed_client_lists=None):
    return FedMLServerManager(args, dist_aggregator, comm, rank, size, backend)


def get_clinet_manager_master(args, trainer_dist_adapter, comm, client_rank, size, backend):      # (sic) the reference's spelling
    return ClientMasterManager(args, trainer_dist_adapter, comm, client_rank, size, backend)


def get_clinet_manager_salve(args, trainer_dist_adapter):                                        # (sic)
    return ClientSlaveManager(args, trainer_dist_adapter)


class Server:
    def __init__(self, args, device, dataset, model, model_trainer=None, server_aggregator=None):
        self.a = (args, device, dataset, model, model_trainer, server_aggregator)

    def run(self):
        args, device, dataset, model, mt, sa = self.a
        ids = args.client_ids() if hasattr(args, "client_ids") else []
        return FedML_Hierarchical(args, 0, len(ids) or int(args.client_num_per_round), None, device, dataset, model, mt, sa)


class Client:
    def __init__(self, args, device, dataset, model, model_trainer=None):
        self.a = (args, device, dataset, model, model_trainer)

    def run(self):
        args, device, dataset, model, mt = self.a
        ids = args.client_ids() if hasattr(args, "client_ids") else []
        return FedML_Hierarchical(args, int(args.rank), len(ids) or int(args.client_num_per_round), None, device, dataset,
                                  model, mt)


def build_torchrun_command(args, entry="torch_client.py", nodes=None):
    """``pdsh -w nodes torchrun --nnodes --nproc_per_node --rdzv_backend=c10d --rdzv_endpoint master:port
    --node_rank=%n entry`` (reference: dist_trainer_launcher.py:23-48); returned as a list for the CLI to run."""
    nodes = nodes or getattr(args, "node_addresses", ["127.0.0.1"])
    cmd = ["torchrun", f"--nnodes={int(getattr(args, 'n_node_in_silo', 1))}",
           f"--nproc_per_node={int(getattr(args, 'n_proc_per_node', 1))}", "--rdzv_backend=c10d",
           f"--rdzv_endpoint={getattr(args, 'master_address', '127.0.0.1')}:{getattr(args, 'launcher_rdzv_port', 29410)}",
           "--node_rank=%n", "--rdzv_id=hi_fl", entry, "--cf", str(getattr(args, "yaml_config_file", ""))]
    if len(nodes) > 1:
        return ["pdsh", "-w", ",".join(nodes)] + cmd
    return cmd
