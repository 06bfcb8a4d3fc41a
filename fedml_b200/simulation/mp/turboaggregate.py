"""TurboAggregate over the managers: FedAvg whose uploads are pairwise-masked field vectors
(reference: mpi_p2p_mp/turboaggregate/*.py — a FedAvg clone plus the mpc toolbox; the protocol is ours,
see core/security/secagg.py).  Clients agree on seeds through the server-relayed public keys, upload
``mask(q(n_i/N * w_i))``; the server only ever adds masked vectors."""
import torch

from ...core import aggregation as A
from ...core.distributed.communication.message import Message
from ...core.security.mpc_function import DEFAULT_PRIME
from ...core.security.secagg import SecAggClient, SecAggServer
from ..trainers import create_model_trainer
from .fedavg import FedAVGAggregator, FedAVGClientManager, FedAVGServerManager, FedAVGTrainer, MyMessage


class _TAAggregator(FedAVGAggregator):
    def aggregate(self, indices=None):
        idx = sorted(self.model_dict) if indices is None else list(indices)
        masked = {i: self.model_dict[i] for i in idx}
        scale = float(getattr(self.args, "secagg_scale", 2 ** 16))
        summed = SecAggServer(DEFAULT_PRIME).aggregate(masked, scale, list(range(self.worker_num)))
        # the last coordinate carries sum_i n_i / S: dividing by it turns the secure sum into the weighted mean
        mean = summed[:-1] / summed[-1].clamp_min(1e-12)
        w_global = self.get_global_model_params()
        agg = A.cast_like(A.unvectorize_like(mean, w_global, weights_only=False), w_global)
        self.set_global_model_params(agg)
        self.model_dict.clear(); self.sample_num_dict.clear()
        return agg


class _TAClientManager(FedAVGClientManager):
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        n = self.size - 1
        self.sec = SecAggClient(self.rank - 1, DEFAULT_PRIME)
        peers = [SecAggClient(i, DEFAULT_PRIME) for i in range(n)]       # deterministic key material per id
        self.sec.agree({c.cid: c.pk for c in peers})
        self.n_clients = n

    def send_model_to_server(self, receive_id, weights, local_sample_num):
        vec = A.vectorize_weight(weights, weights_only=False)
        S = float(getattr(self.args, "secagg_weight_scale", 1000.0))
        scale = float(getattr(self.args, "secagg_scale", 2 ** 16))
        # upload n_i/S * [w_i ; 1]: the server learns only sum n_i w_i / S and sum n_i / S
        w = float(local_sample_num) / S
        ext = torch.cat([vec, torch.ones(1, dtype=vec.dtype, device=vec.device)])
        masked = self.sec.mask(ext, w, scale, list(range(self.n_clients)))
        m = Message(MyMessage.MSG_TYPE_C2S_SEND_MODEL_TO_SERVER, self.get_sender_id(), receive_id)
        m.add_params(MyMessage.MSG_ARG_KEY_MODEL_PARAMS, masked)
        m.add_params(MyMessage.MSG_ARG_KEY_NUM_SAMPLES, local_sample_num)
        self.send_message(m)


def FedML_TurboAggregate_distributed(args, process_id, worker_number, comm, device, dataset, model, model_trainer=None,  # noqa: N802
                                     backend="INPROC", **_):
    (train_num, test_num, train_global, test_global, local_num, train_local, test_local, class_num) = list(dataset)[:8]
    trainer = model_trainer or create_model_trainer(model, args)
    if process_id == 0:
        agg = _TAAggregator(train_global, test_global, train_num, train_local, test_local, local_num, worker_number - 1,
                            device, args, trainer)
        mgr = FedAVGServerManager(args, agg, comm, 0, worker_number, backend)
        mgr.send_init_msg(); mgr.run()
        return agg
    tr = FedAVGTrainer(process_id - 1, train_local, local_num, test_local, train_num, device, args, trainer)
    mgr = _TAClientManager(args, tr, comm, process_id, worker_number, backend)
    mgr.run()
    return tr
