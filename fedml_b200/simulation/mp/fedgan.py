"""FedGAN (reference: mpi_p2p_mp/fedgan/*.py): every client runs a local GAN loop (BCE, Adam; MyModelTrainer.py:26-91)
and the server averages generator *and* discriminator dicts (FedGANAggregator.py:69-99) — with both nets under one
module (``MNISTGan``) that is exactly the FedAvg protocol with a GAN trainer."""
import torch
import torch.nn as nn

from ...core.alg_frame import ClientTrainer
from ...models.cv import MNISTGan
from .fedavg import FedML_FedAvg_distributed


class GANTrainer(ClientTrainer):
    def get_model_params(self):
        return {k: v.detach().clone() for k, v in self.model.state_dict().items()}

    def set_model_params(self, p):
        self.model.load_state_dict(p)

    def train(self, train_data, device, args):
        m = self.model.to(device); m.train()
        bce = nn.BCELoss()
        og = torch.optim.Adam(m.netg.parameters(), lr=float(getattr(args, "learning_rate", 2e-4)), betas=(0.5, 0.999))
        od = torch.optim.Adam(m.netd.parameters(), lr=float(getattr(args, "learning_rate", 2e-4)), betas=(0.5, 0.999))
        for _ in range(int(getattr(args, "epochs", 1))):
            for x, _y in train_data:
                x = x.to(device).reshape(x.shape[0], -1) * 2 - 1
                b = x.shape[0]
                ones, zeros = torch.ones(b, 1, device=device), torch.zeros(b, 1, device=device)
                z = torch.randn(b, m.netg.latent, device=device)
                fake = m.netg(z)
                od.zero_grad(set_to_none=True)
                ld = bce(m.netd(x), ones) + bce(m.netd(fake.detach()), zeros)
                ld.backward(); od.step()
                og.zero_grad(set_to_none=True)
                lg = bce(m.netd(fake), ones)
                lg.backward(); og.step()
        self.last = (float(ld), float(lg))

    def test(self, test_data, device, args):
        m = self.model.to(device); m.eval()
        with torch.no_grad():
            z = torch.randn(64, m.netg.latent, device=device)
            score = float(m.netd(m.netg(z)).mean())
        return {"test_correct": score, "test_loss": 1.0 - score, "test_total": 1}


def FedML_FedGan_distributed(args, process_id, worker_number, comm, device, dataset, model=None, model_trainer=None,  # noqa: N802
                             backend="INPROC", **kw):
    model = model if isinstance(model, MNISTGan) else MNISTGan()
    return FedML_FedAvg_distributed(args, process_id, worker_number, comm, device, dataset, model,
                                    model_trainer or GANTrainer(model, args), backend=backend, **kw)
