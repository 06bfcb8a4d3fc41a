"""Split learning (reference: mpi_p2p_mp/split_nn/*.py, 404 LoC): the model is cut into a client part and a
server part.  A client forwards activations + labels, the server computes the loss, back-propagates through
its part and returns the gradient of the activations; clients take turns in a relay ring (the finished client
passes a semaphore to the next, client_manager.py:41-79).  SGD lr 0.1, momentum 0.9, wd 5e-4 are hard-coded in
the reference (client.py / server.py) and kept as defaults here."""
import torch
import torch.nn as nn

from ...core.distributed.communication.message import Message
from ...core.distributed.managers import ClientManager, ServerManager

MSG_C2S_ACTS, MSG_S2C_GRADS, MSG_C2C_SEMAPHORE, MSG_C2S_VALIDATION_MODE, MSG_C2S_VALIDATION_OVER, MSG_FINISH = 1, 2, 3, 4, 5, 6


def split_model(model: nn.Module, cut: int = 1):
    """Cut a Sequential-like model after ``cut`` top-level children."""
    kids = list(model.children())
    return nn.Sequential(*kids[:cut]), nn.Sequential(*kids[cut:])


class SplitNNServerManager(ServerManager):
    def __init__(self, args, server_model, comm, rank, size, backend, device):
        super().__init__(args, comm, rank, size, backend)
        self.model, self.device = server_model.to(device), device
        self.opt = torch.optim.SGD(self.model.parameters(), lr=float(getattr(args, "learning_rate", 0.1)), momentum=0.9,
                                   weight_decay=5e-4)
        self.criterion = nn.CrossEntropyLoss()
        self.train_mode, self.correct, self.total, self.loss_sum = True, 0, 0, 0.0
        self.history, self.finished_clients = [], 0

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_C2S_ACTS, self.handle_acts)
        self.register_message_receive_handler(MSG_C2S_VALIDATION_MODE, lambda p: setattr(self, "train_mode", False))
        self.register_message_receive_handler(MSG_C2S_VALIDATION_OVER, self.handle_validation_over)
        self.register_message_receive_handler(MSG_FINISH, self.handle_finish)

    def handle_acts(self, p):
        acts, labels = p["acts"].to(self.device), p["labels"].to(self.device)
        if self.train_mode:
            self.model.train()
            acts.requires_grad_(True)
            self.opt.zero_grad(set_to_none=True)
            out = self.model(acts)
            loss = self.criterion(out, labels)
            loss.backward()
            self.opt.step()
            m = Message(MSG_S2C_GRADS, 0, p["sender"]); m.add_params("grads", acts.grad.detach().cpu()); self.send_message(m)
        else:
            self.model.eval()
            with torch.no_grad():
                out = self.model(acts)
            self.correct += int((out.argmax(1) == labels).sum()); self.total += labels.numel()
            self.loss_sum += float(self.criterion(out, labels)) * labels.numel()

    def handle_validation_over(self, p):
        self.history.append({"client": p["sender"], "acc": self.correct / max(self.total, 1), "loss": self.loss_sum / max(self.total, 1)})
        self.correct = self.total = 0; self.loss_sum = 0.0
        self.train_mode = True

    def handle_finish(self, p):
        self.finished_clients += 1
        if self.finished_clients == self.size - 1:
            self.finish()


class SplitNNClientManager(ClientManager):
    def __init__(self, args, client_model, train_batches, test_batches, comm, rank, size, backend, device):
        super().__init__(args, comm, rank, size, backend)
        self.model, self.device = client_model.to(device), device
        self.opt = torch.optim.SGD(self.model.parameters(), lr=float(getattr(args, "learning_rate", 0.1)), momentum=0.9,
                                   weight_decay=5e-4)
        self.train_batches, self.test_batches = list(train_batches), list(test_batches)
        self.epochs, self.epoch, self.batch_idx = int(getattr(args, "epochs", 1)), 0, 0
        self.node_right = 1 if rank == size - 1 else rank + 1

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_C2C_SEMAPHORE, self.handle_semaphore)
        self.register_message_receive_handler(MSG_S2C_GRADS, self.handle_grads)

    def start_if_first(self):
        if self.rank == 1:
            self._forward_next()

    def handle_semaphore(self, p):
        self.batch_idx = 0
        self._forward_next()

    def _forward_next(self):
        x, y = self.train_batches[self.batch_idx]
        self.model.train()
        self.acts = self.model(x.to(self.device))
        m = Message(MSG_C2S_ACTS, self.rank, 0); m.add_params("acts", self.acts.detach().cpu()); m.add_params("labels", y)
        self.send_message(m)

    def handle_grads(self, p):
        self.opt.zero_grad(set_to_none=True)
        self.acts.backward(p["grads"].to(self.device))
        self.opt.step()
        self.batch_idx += 1
        if self.batch_idx < len(self.train_batches):
            self._forward_next()
            return
        # epoch over: validate, then hand the semaphore to the right neighbour (or finish)
        self.send_message(Message(MSG_C2S_VALIDATION_MODE, self.rank, 0))
        self.model.eval()
        with torch.no_grad():
            for x, y in self.test_batches[:4]:
                m = Message(MSG_C2S_ACTS, self.rank, 0); m.add_params("acts", self.model(x.to(self.device)).cpu()); m.add_params("labels", y)
                self.send_message(m)
        self.send_message(Message(MSG_C2S_VALIDATION_OVER, self.rank, 0))
        self.epoch += 1
        last_client = self.rank == self.size - 1
        if self.epoch == self.epochs:
            if not last_client:
                self.send_message(Message(MSG_C2C_SEMAPHORE, self.rank, self.node_right))
            self.send_message(Message(MSG_FINISH, self.rank, 0))
            self.finish()
        else:
            self.send_message(Message(MSG_C2C_SEMAPHORE, self.rank, self.node_right))


def SplitNN_distributed(args, process_id, worker_number, comm, device, dataset, model, model_trainer=None,  # noqa: N802
                        backend="INPROC", **_):
    client_model, server_model = split_model(model, int(getattr(args, "split_cut", 1)))
    if process_id == 0:
        mgr = SplitNNServerManager(args, server_model, comm, 0, worker_number, backend, device)
        mgr.run()
        return mgr
    train_local, test_local = dataset[5], dataset[6]
    cid = process_id - 1
    mgr = SplitNNClientManager(args, client_model, train_local[cid], test_local[cid], comm, process_id, worker_number,
                               backend, device)
    mgr.register_message_receive_handlers()
    mgr.start_if_first()
    mgr.run()
    return mgr
