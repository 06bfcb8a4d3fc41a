"""TurboAggregate over the managers: FedAvg whose uploads are pairwise-masked field vectors
(reference: mpi_p2p_mp/turboaggregate/*.py — a FedAvg clone plus the mpc toolbox; the protocol is ours,
see core/security/secagg.py).  Every client draws its private key from the OS entropy source and publishes only the
public key; the server relays the key table with the first model broadcast; clients derive the pairwise seeds and upload
``mask(q(n_i/N * w_i))`` with the round index in the mask PRG; the server only ever adds masked vectors.  This manager
flow uses pairwise masks without dropout recovery (all workers must report; an honest-but-curious server learns the sum
only); the full double-masking protocol with Shamir shares is ``core.security.secagg.secure_weighted_mean``, used by the
single-process TurboAggregate."""
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


MSG_TYPE_C2S_PUBLIC_KEY = "ta_c2s_public_key"
MSG_ARG_KEY_PUBLIC_KEY = "ta_public_key"
MSG_ARG_KEY_PUBLIC_KEYS = "ta_public_keys"


class _TAServerManager(FedAVGServerManager):
    """Relays the clients' public keys: the first model broadcast waits until every client has published its key and
    carries the {client: public key} table.  Private keys never leave their client; the server sees masked vectors only."""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.public_keys = {}

    def register_message_receive_handlers(self):
        super().register_message_receive_handlers()
        self.register_message_receive_handler(MSG_TYPE_C2S_PUBLIC_KEY, self._on_public_key)

    def _on_public_key(self, msg_params):
        self.public_keys[str(int(msg_params.get(MyMessage.MSG_ARG_KEY_SENDER)) - 1)] = int(msg_params.get(MSG_ARG_KEY_PUBLIC_KEY))
        if len(self.public_keys) == self.size - 1:
            self.send_init_msg()

    def _send_model(self, mtype, receive_id, params, client_index):
        m = Message(mtype, self.get_sender_id(), receive_id)
        m.add_params(MyMessage.MSG_ARG_KEY_MODEL_PARAMS, params)
        m.add_params(MyMessage.MSG_ARG_KEY_CLIENT_INDEX, str(client_index))
        if mtype == MyMessage.MSG_TYPE_S2C_INIT_CONFIG:
            m.add_params(MSG_ARG_KEY_PUBLIC_KEYS, dict(self.public_keys))
        self.send_message(m)


class _TAClientManager(FedAVGClientManager):
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.n_clients = self.size - 1
        self.sec = SecAggClient(self.rank - 1, DEFAULT_PRIME)          # private key from the OS entropy source

    def publish_key(self):
        m = Message(MSG_TYPE_C2S_PUBLIC_KEY, self.get_sender_id(), 0)
        m.add_params(MSG_ARG_KEY_PUBLIC_KEY, int(self.sec.pk))
        self.send_message(m)

    def handle_message_init(self, msg_params):
        pks = msg_params.get(MSG_ARG_KEY_PUBLIC_KEYS) or {}
        self.sec.agree({int(k): int(v) for k, v in pks.items()})
        super().handle_message_init(msg_params)

    def send_model_to_server(self, receive_id, weights, local_sample_num):
        vec = A.vectorize_weight(weights, weights_only=False)
        S = float(getattr(self.args, "secagg_weight_scale", 1000.0))
        scale = float(getattr(self.args, "secagg_scale", 2 ** 16))
        # upload n_i/S * [w_i ; 1]: the server learns only sum n_i w_i / S and sum n_i / S
        w = float(local_sample_num) / S
        ext = torch.cat([vec, torch.ones(1, dtype=vec.dtype, device=vec.device)])
        # the round index is part of the mask PRG: no mask is used twice
        masked = self.sec.mask(ext, w, scale, list(range(self.n_clients)), round_idx=self.round_idx)
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
        mgr = _TAServerManager(args, agg, comm, 0, worker_number, backend)
        mgr.run()                                   # the first broadcast goes out once every public key has arrived
        return agg
    tr = FedAVGTrainer(process_id - 1, train_local, local_num, test_local, train_num, device, args, trainer)
    mgr = _TAClientManager(args, tr, comm, process_id, worker_number, backend)
    mgr.publish_key()
    mgr.run()
    return tr
