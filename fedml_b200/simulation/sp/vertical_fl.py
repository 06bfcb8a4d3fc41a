"""Classical vertical FL (feature-partitioned): one guest with labels, several hosts
(reference: single_process/classical_vertical_fl/vfl.py:20-60, party_models.py:12-121, vfl_fixture.py).
Per batch: every party computes its logit component from its own features; the guest sums them, applies
BCE-with-logits, and sends the common gradient d(loss)/d(logit) back; each party back-propagates through its
local extractor + dense scorer with its own learning rate."""
import numpy as np
import torch
import torch.nn as nn

from ...data.tabular_readers import load_vertical_dataset
from ...models.finance import DenseModel, LocalModel


class VFLParty:
    def __init__(self, local_model: LocalModel, dense: DenseModel, lr):
        self.local, self.dense = local_model, dense
        self.opt = torch.optim.SGD(list(local_model.parameters()) + list(dense.parameters()), lr=lr)

    def forward(self, x):
        self._out = self.dense(self.local(x))
        return self._out.detach()

    def backward(self, grad):
        self.opt.zero_grad(set_to_none=True)
        self._out.backward(grad)
        self.opt.step()

    def predict(self, x):
        with torch.no_grad():
            return self.dense(self.local(x))


class VerticalFLAPI:
    def __init__(self, args, device=None, dataset=None, model=None, *_, **__):
        self.args = args
        dims, (self.xs, self.y), (self.xs_te, self.y_te) = load_vertical_dataset(args)      # real files when present
        lr = float(getattr(args, "learning_rate", 0.05))
        hid = int(getattr(args, "vfl_hidden", 16))
        self.parties = [VFLParty(LocalModel(d, hid, lr), DenseModel(hid, 1, lr, bias=(i == 0)), lr) for i, d in enumerate(dims)]
        self.history = []

    def fit_batch(self, idx):
        comps = [p.forward(x[idx]) for p, x in zip(self.parties, self.xs)]
        logit = torch.stack(comps).sum(0).requires_grad_(True)
        loss = nn.functional.binary_cross_entropy_with_logits(logit.squeeze(1), self.y[idx])
        loss.backward()
        for p in self.parties:                       # the guest broadcasts the common gradient
            p.backward(logit.grad)
        return float(loss)

    def evaluate(self):
        logit = torch.stack([p.predict(x) for p, x in zip(self.parties, self.xs_te)]).sum(0).squeeze(1)
        return float(((logit > 0).float() == self.y_te).float().mean())

    def train(self):
        bs, n = int(getattr(self.args, "batch_size", 64)), self.y.shape[0]
        g = torch.Generator().manual_seed(int(getattr(self.args, "random_seed", 0)))
        for ep in range(int(getattr(self.args, "comm_round", getattr(self.args, "epochs", 5)))):
            perm = torch.randperm(n, generator=g)
            losses = [self.fit_batch(perm[i:i + bs]) for i in range(0, n, bs)]
            self.history.append({"round": ep, "train_loss": float(np.mean(losses)), "test_acc": self.evaluate()})
        return self.history
