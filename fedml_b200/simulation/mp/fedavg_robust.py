"""Robust FedAvg under a backdoor ("edge-case") attack, worker-per-process form.

Reference: simulation/mpi_p2p_mp/fedavg_robust/ — FedAvgRobustAggregator.py (forced sampling of the attacker every
``attack_freq`` rounds :192-194 / :277-296, defence by norm-difference clipping and weak DP :230-262, raw-task and
targeted-task evaluation :298-352) and FedAvgRobustTrainer.py (client 1 trains on the poisoned loader :32-37).
Poisoned / targeted-task sets come from ``data.load_poisoned_dataset_from_edge_case_examples`` (the reference's
edge_case_examples/data_loader.py:393; Southwest-airline / ARDIS images are not downloadable here, a trigger-patch
stand-in of the same interface is used).
"""
import logging

import numpy as np

from ...core.robustness.robust_aggregation import RobustAggregator
from ..trainers import create_model_trainer
from .fedavg import FedAVGAggregator, FedAVGClientManager, FedAVGServerManager, FedAVGTrainer

ATTACKER_CLIENT_INDEX = 1          # "we always assume the client with client_index=1 as the attacker"


class FedAvgRobustAggregator(FedAVGAggregator):
    def __init__(self, *a, targetted_task_test_loader=None, num_dps_poisoned_dataset=0, **kw):
        super().__init__(*a, **kw)
        args = self.args
        if self.robust is None:
            if not getattr(args, "defense_type", None):
                args.defense_type = "norm_diff_clipping"
            self.robust = RobustAggregator(args)
        self.targetted_task_test_loader = targetted_task_test_loader
        self.num_dps_poisoned_dataset = num_dps_poisoned_dataset
        freq = max(int(getattr(args, "attack_freq", 10) or 10), 1)
        # the reference's list is 1-based while the round counter it is tested against is 0-based; kept as is
        self.adversary_fl_rounds = [i for i in range(1, int(args.comm_round) + 1) if (i - 1) % freq == 0]
        self.attack_history = []

    def client_sampling(self, round_idx, client_num_in_total, client_num_per_round):
        num_clients = min(client_num_per_round, client_num_in_total)
        np.random.seed(round_idx)
        picked = np.random.choice(range(client_num_in_total), num_clients, replace=False)
        if round_idx in self.adversary_fl_rounds:
            # the attacker is guaranteed to participate: it takes the first worker, the sampled clients follow
            picked = np.array([ATTACKER_CLIENT_INDEX] + list(picked))
        logging.info("client_indexes = %s", picked)
        return picked

    def test_on_server_for_all_clients(self, round_idx):
        super().test_on_server_for_all_clients(round_idx)          # raw task (train / test accuracy of the global model)
        freq = int(getattr(self.args, "frequency_of_the_test", 1) or 1)
        if not (round_idx % freq == 0 or round_idx == int(self.args.comm_round) - 1):
            return
        self.test_target_accuracy(round_idx)

    def test_target_accuracy(self, round_idx):
        """Targeted-task accuracy = fraction of triggered samples classified as the attacker's label (backdoor success)."""
        if self.targetted_task_test_loader is None:
            return None
        m = self.trainer.test(self.targetted_task_test_loader, self.device, self.args)
        acc = m["test_correct"] / max(m["test_total"], 1)
        rec = {"round": round_idx, "targeted_acc": acc, "attacker_round": round_idx in self.adversary_fl_rounds}
        self.attack_history.append(rec)
        if self.history and self.history[-1].get("round") == round_idx:
            self.history[-1]["targeted_acc"] = acc
        self.metrics.log({"Targeted/Acc": acc, "round": round_idx})
        logging.info("targetted-task accuracy, round %d: %.4f", round_idx, acc)
        return acc


class FedAvgRobustTrainer(FedAVGTrainer):
    def __init__(self, *a, poisoned_train_loader=None, num_dps_poisoned_dataset=0, **kw):
        super().__init__(*a, **kw)
        self.poisoned_train_loader, self.num_dps_poisoned_dataset = poisoned_train_loader, num_dps_poisoned_dataset

    def update_dataset(self, client_index):
        super().update_dataset(client_index)
        if int(client_index) == ATTACKER_CLIENT_INDEX and self.poisoned_train_loader is not None:
            self.train_local = self.poisoned_train_loader
            self.local_sample_number = self.num_dps_poisoned_dataset


def FedML_FedAvgRobust_distributed(args, process_id, worker_number, comm, device, dataset, model, model_trainer=None,  # noqa: N802
                                   backend=None, poisoned_train_loader=None, targetted_task_test_loader=None,
                                   num_dps_poisoned_dataset=None, **_):
    """Reference entry point (FedAvgRobustAPI.py): rank 0 = defending server, ranks > 0 = clients, client 1 poisoned."""
    (train_num, test_num, train_global, test_global, local_num, train_local, test_local, class_num) = list(dataset)[:8]
    backend = backend or getattr(args, "mp_transport", "INPROC")
    if poisoned_train_loader is None:
        from ...data import load_poisoned_dataset_from_edge_case_examples
        poisoned_train_loader, _clean, targetted_task_test_loader, num_dps_poisoned_dataset = \
            load_poisoned_dataset_from_edge_case_examples(args)
    trainer = model_trainer or create_model_trainer(model, args)
    if process_id == 0:
        agg = FedAvgRobustAggregator(train_global, test_global, train_num, train_local, test_local, local_num,
                                     worker_number - 1, device, args, trainer,
                                     targetted_task_test_loader=targetted_task_test_loader,
                                     num_dps_poisoned_dataset=num_dps_poisoned_dataset)
        mgr = FedAVGServerManager(args, agg, comm, process_id, worker_number, backend)
        mgr.send_init_msg()
        mgr.run()
        return agg
    trainer.set_id(process_id - 1)
    tr = FedAvgRobustTrainer(process_id - 1, train_local, local_num, test_local, train_num, device, args, trainer,
                             poisoned_train_loader=poisoned_train_loader, num_dps_poisoned_dataset=num_dps_poisoned_dataset)
    mgr = FedAVGClientManager(args, tr, comm, process_id, worker_number, backend)
    mgr.run()
    return tr
