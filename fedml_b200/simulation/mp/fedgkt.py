"""FedGKT — group knowledge transfer (reference: mpi_p2p_mp/fedgkt/*.py, 1007 LoC).

Clients train a small extractor+head (ResNet-8) with CE + alpha * KL to the server's logits and upload
per-batch feature maps, logits and labels (GKTClientTrainer.py:68-); the server trains the large trunk on
those features with CE + alpha * KL to the client logits and returns per-client logits
(GKTServerTrainer.py:121-367); checkpoints ``last.pth`` / ``best.pth`` with
``{state_dict, optim_dict, epoch, test_accTop1, test_accTop5}`` (:262-290).  Losses ``KL_Loss`` / ``CE_Loss`` with
temperature follow fedgkt/utils.py:78-124."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from ...core.distributed.communication.message import Message
from ...core.distributed.managers import ClientManager, ServerManager
from ...models.cv import ResNet56Server, ResNet8Client
from ...utils.checkpoint import save_checkpoint

MSG_C2S_FEATURES, MSG_S2C_LOGITS, MSG_FINISH = 1, 2, 6


class KL_Loss(nn.Module):  # noqa: N801
    def __init__(self, temperature=1.0):
        super().__init__()
        self.T = temperature

    def forward(self, output_batch, teacher_outputs):
        s = F.log_softmax(output_batch / self.T, dim=1)
        t = F.softmax(teacher_outputs / self.T, dim=1) + 1e-7
        return self.T * self.T * F.kl_div(s, t, reduction="batchmean")


class CE_Loss(nn.Module):  # noqa: N801
    def __init__(self, temperature=1.0):
        super().__init__()
        self.T = temperature

    def forward(self, output_batch, teacher_outputs):
        s = F.log_softmax(output_batch / self.T, dim=1)
        t = F.softmax(teacher_outputs / self.T, dim=1)
        return -self.T * self.T * torch.sum(s * t) / teacher_outputs.size(0)


class GKTServerManager(ServerManager):
    def __init__(self, args, comm, rank, size, backend, device, class_num):
        super().__init__(args, comm, rank, size, backend)
        self.device = device
        self.model = ResNet56Server(class_num, tuple(getattr(args, "gkt_server_layers", (1, 1, 1)))).to(device)
        if getattr(args, "multi_gpu_server", False) and torch.cuda.device_count() > 1:
            self.model = nn.DataParallel(self.model)
        self.opt = torch.optim.SGD(self.model.parameters(), lr=float(getattr(args, "learning_rate", 0.01)), momentum=0.9,
                                   weight_decay=5e-4)
        self.alpha, self.T = float(getattr(args, "alpha", 1.0)), float(getattr(args, "temperature", 3.0))
        self.kl, self.ce = KL_Loss(self.T), nn.CrossEntropyLoss()
        self.round_num, self.round_idx, self.uploads = int(args.comm_round), 0, {}
        self.epochs_server = int(getattr(args, "epochs_server", 1))
        self.best_acc, self.history = 0.0, []

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_C2S_FEATURES, self.handle_features)

    def handle_features(self, p):
        self.uploads[p["sender"]] = p
        if len(self.uploads) < self.size - 1:
            return
        self.model.train()
        distill = bool(getattr(self.args, "whether_distill_on_the_server", True))
        for _ in range(self.epochs_server):
            for cid, up in self.uploads.items():
                for feat, logit_c, y in zip(up["features"], up["logits"], up["labels"]):
                    feat, logit_c, y = feat.to(self.device), logit_c.to(self.device), y.to(self.device)
                    out = self.model(feat)
                    loss = self.ce(out, y) + (self.alpha * self.kl(out, logit_c) if distill else 0.0)
                    self.opt.zero_grad(set_to_none=True); loss.backward(); self.opt.step()
        self.model.eval()
        correct = total = 0
        reply = {}
        with torch.no_grad():
            for cid, up in self.uploads.items():
                outs = []
                for feat, y in zip(up["features"], up["labels"]):
                    o = self.model(feat.to(self.device)).cpu(); outs.append(o)
                    correct += int((o.argmax(1) == y).sum()); total += y.numel()
                reply[cid] = outs
        acc = correct / max(total, 1)
        self.history.append({"round": self.round_idx, "server_acc_on_uploaded_features": acc})
        state = {"state_dict": (self.model.module if isinstance(self.model, nn.DataParallel) else self.model).state_dict(),
                 "optim_dict": self.opt.state_dict(), "epoch": self.round_idx + 1, "test_accTop1": acc, "test_accTop5": acc}
        save_checkpoint(state, acc >= self.best_acc, getattr(self.args, "checkpoint_dir", None) or "./checkpoint")
        self.best_acc = max(self.best_acc, acc)
        self.uploads = {}
        self.round_idx += 1
        done = self.round_idx == self.round_num
        for cid, outs in reply.items():
            m = Message(MSG_FINISH if done else MSG_S2C_LOGITS, 0, cid); m.add_params("logits", outs); self.send_message(m)
        if done:
            self.finish()


class GKTClientManager(ClientManager):
    def __init__(self, args, comm, rank, size, backend, device, class_num, train_batches):
        super().__init__(args, comm, rank, size, backend)
        self.device, self.model = device, ResNet8Client(class_num).to(device)
        self.opt = torch.optim.SGD(self.model.parameters(), lr=float(getattr(args, "learning_rate", 0.01)), momentum=0.9,
                                   weight_decay=5e-4)
        self.alpha, self.T = float(getattr(args, "alpha", 1.0)), float(getattr(args, "temperature", 3.0))
        self.kl, self.ce = KL_Loss(self.T), nn.CrossEntropyLoss()
        self.batches = list(train_batches)
        self.server_logits = None
        self.epochs_client = int(getattr(args, "epochs_client", 1))

    def register_message_receive_handlers(self):
        self.register_message_receive_handler(MSG_S2C_LOGITS, self.handle_logits)
        self.register_message_receive_handler(MSG_FINISH, lambda p: self.finish())

    def handle_logits(self, p):
        self.server_logits = p["logits"]
        self.train_and_upload()

    def train_and_upload(self):
        self.model.train()
        for _ in range(self.epochs_client):
            for i, (x, y) in enumerate(self.batches):
                x, y = x.to(self.device), y.to(self.device)
                out, _ = self.model(x)
                loss = self.ce(out, y)
                if self.server_logits is not None:
                    loss = loss + self.alpha * self.kl(out, self.server_logits[i].to(self.device))
                self.opt.zero_grad(set_to_none=True); loss.backward(); self.opt.step()
        self.model.eval()
        feats, logits, labels = [], [], []
        with torch.no_grad():
            for x, y in self.batches:
                o, f = self.model(x.to(self.device))
                feats.append(f.cpu()); logits.append(o.cpu()); labels.append(y)
        m = Message(MSG_C2S_FEATURES, self.rank, 0)
        m.add_params("features", feats); m.add_params("logits", logits); m.add_params("labels", labels)
        self.send_message(m)


def FedML_FedGKT_distributed(args, process_id, worker_number, comm, device, dataset, model=None, model_trainer=None,  # noqa: N802
                             backend="INPROC", **_):
    class_num = dataset[7]
    if process_id == 0:
        mgr = GKTServerManager(args, comm, 0, worker_number, backend, device, class_num); mgr.run(); return mgr
    mgr = GKTClientManager(args, comm, process_id, worker_number, backend, device, class_num, dataset[5][process_id - 1])
    mgr.register_message_receive_handlers(); mgr.train_and_upload(); mgr.run()
    return mgr
