"""Vertical FL over the managers (reference: mpi_p2p_mp/classical_vertical_fl/*.py, guest_trainer.py:91-137):
per batch the hosts upload their logit components, the guest (rank 0, owns the labels) sums them, computes
BCE-with-logits and sends the common gradient back down."""
import torch
import torch.nn as nn

from ...core.distributed.communication.message import Message
from ...core.distributed.managers import ClientManager, ServerManager
from ...data.tabular_readers import load_vertical_dataset
from ...models.finance import DenseModel, LocalModel

MSG_H2G_LOGITS, MSG_G2H_GRAD, MSG_FINISH = 1, 2, 6


class _Party:
    def __init__(self, dim, hid, lr, bias):
        self.local, self.dense = LocalModel(dim, hid, lr), DenseModel(hid, 1, lr, bias=bias)
        self.opt = torch.optim.SGD(list(self.local.parameters()) + list(self.dense.parameters()), lr=lr)

    def forward(self, x):
        self.out = self.dense(self.local(x))
        return self.out.detach()

    def backward(self, g):
        self.opt.zero_grad(set_to_none=True); self.out.backward(g); self.opt.step()


def _data(args):
    if not hasattr(args, "vfl_samples"):
        args.vfl_samples = 2000
    dims, tr, te = load_vertical_dataset(args)
    return dims, (tr, te)


class GuestManager(ServerManager):
    def __init__(self, args, comm, rank, size, backend):
        super().__init__(args, comm, rank, size, backend)
        dims, ((self.xs, self.y), (self.xs_te, self.y_te)) = _data(args)
        lr = float(getattr(args, "learning_rate", 0.05))
        self.party = _Party(dims[0], 16, lr, True)
        self.bs, self.n = int(getattr(args, "batch_size", 64)), self.y.shape[0]
        self.n_batches = (self.n + self.bs - 1) // self.bs
        self.epochs, self.step, self.pending = int(getattr(args, "comm_round", 2)), 0, {}
        self.history = []

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_H2G_LOGITS, self.handle_logits)

    def handle_logits(self, p):
        self.pending[p["sender"]] = p["logits"]
        if len(self.pending) < self.size - 1:
            return
        b = self.step % self.n_batches
        sl = slice(b * self.bs, min((b + 1) * self.bs, self.n))
        own = self.party.forward(self.xs[0][sl])
        logit = (own + sum(self.pending.values())).requires_grad_(True)
        loss = nn.functional.binary_cross_entropy_with_logits(logit.squeeze(1), self.y[sl])
        loss.backward()
        self.party.backward(logit.grad)
        self.history.append(float(loss.detach()))
        self.pending = {}
        self.step += 1
        done = self.step == self.epochs * self.n_batches
        for r in range(1, self.size):
            m = Message(MSG_FINISH if done else MSG_G2H_GRAD, 0, r); m.add_params("grad", logit.grad); self.send_message(m)
        if done:
            self.finish()


class HostManager(ClientManager):
    def __init__(self, args, comm, rank, size, backend):
        super().__init__(args, comm, rank, size, backend)
        dims, ((self.xs, _), _) = _data(args)
        self.x = self.xs[rank]
        self.party = _Party(dims[rank], 16, float(getattr(args, "learning_rate", 0.05)), False)
        self.bs, self.n, self.step = int(getattr(args, "batch_size", 64)), self.x.shape[0], 0
        self.n_batches = (self.n + self.bs - 1) // self.bs

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_G2H_GRAD, self.handle_grad)
        self.register_message_receive_handler(MSG_FINISH, self.handle_finish)

    def send_logits(self):
        b = self.step % self.n_batches
        sl = slice(b * self.bs, min((b + 1) * self.bs, self.n))
        m = Message(MSG_H2G_LOGITS, self.rank, 0); m.add_params("logits", self.party.forward(self.x[sl])); self.send_message(m)

    def handle_grad(self, p):
        self.party.backward(p["grad"]); self.step += 1; self.send_logits()

    def handle_finish(self, p):
        self.party.backward(p["grad"]); self.finish()


def FedML_VFL_distributed(args, process_id, worker_number, comm, device=None, dataset=None, model=None,  # noqa: N802
                          model_trainer=None, backend="INPROC", **_):
    if process_id == 0:
        mgr = GuestManager(args, comm, 0, worker_number, backend); mgr.run(); return mgr
    mgr = HostManager(args, comm, process_id, worker_number, backend)
    mgr.register_message_receive_handlers(); mgr.send_logits(); mgr.run()
    return mgr
