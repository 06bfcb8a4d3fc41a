"""FedNAS — federated DARTS (reference: mpi_p2p_mp/fednas/*.py, 894 LoC): in the search stage clients alternate
architecture-alpha steps (on a held-out half of their data) and weight steps (FedNASTrainer.py:43-165); the
server averages weights *and* alphas (FedNASAggregator.py:66-123) and records the genotype every round.  ``stage: train``
trains the derived network (``NetworkCIFAR`` of ``args.arch``, default the cell the FedNAS paper reports) with the auxiliary
head loss and a drop-path probability that grows with the round (FedNASTrainer.py:167-245)."""
import torch
import torch.nn as nn

from ...core.alg_frame import ClientTrainer
from ...models.darts import Network
from .fedavg import FedML_FedAvg_distributed


class FedNASTrainer(ClientTrainer):
    def get_model_params(self):
        return {k: v.detach().clone() for k, v in self.model.state_dict().items()}

    def set_model_params(self, p):
        self.model.load_state_dict(p)

    def train(self, train_data, device, args):
        m = self.model.to(device)
        m.train()
        batches = list(train_data)
        half = max(1, len(batches) // 2)
        w_batches, a_batches = batches[:half], batches[half:] or batches[:half]
        crit = nn.CrossEntropyLoss()
        search = str(getattr(args, "stage", "search")) == "search" and hasattr(m, "arch_parameters")
        weights = list(m.weight_parameters()) if hasattr(m, "weight_parameters") else list(m.parameters())
        w_opt = torch.optim.SGD(weights, lr=float(getattr(args, "learning_rate", 0.025)), momentum=0.9,
                                weight_decay=float(getattr(args, "weight_decay", 3e-4)))
        from ...models.darts import Architect
        lr = float(getattr(args, "learning_rate", 0.025))
        architect = Architect(m, crit, float(getattr(args, "arch_learning_rate", 3e-4)), float(getattr(args, "arch_weight_decay", 1e-3)),
                              momentum=0.9, weight_decay=float(getattr(args, "weight_decay", 3e-4))) if search else None
        # arch_order 2 = the unrolled (second-order) architect step of the reference (architect.py), 1 = first order
        unrolled = int(getattr(args, "arch_order", 1) or 1) == 2 or bool(getattr(args, "unrolled", False))
        aux_w = float(getattr(args, "auxiliary_weight", 0.4))
        lam_train, lam_valid = float(getattr(args, "lambda_train_regularizer", 1.0)), float(getattr(args, "lambda_valid_regularizer", 1.0))
        if not search and hasattr(m, "drop_path_prob"):     # train stage: drop-path probability ramps up over the rounds
            frac = float(getattr(args, "round_idx", 0) or 0) / max(1, int(getattr(args, "comm_round", 1)))
            m.drop_path_prob = float(getattr(args, "drop_path_prob", 0.2)) * frac
        if hasattr(m, "set_tau"):                       # GDAS: temperature annealed linearly over the rounds
            t_hi, t_lo = float(getattr(args, "tau_max", 10.0)), float(getattr(args, "tau_min", 1.0))
            frac = float(getattr(args, "round_idx", 0) or 0) / max(1, int(getattr(args, "comm_round", 1)) - 1)
            m.set_tau(t_hi - (t_hi - t_lo) * min(1.0, frac))
        for _ in range(int(getattr(args, "epochs", 1))):
            for i, (x, y) in enumerate(w_batches):
                if search:   # alpha step on held-out data (first or second order), then weight step on training data
                    xa, ya = a_batches[i % len(a_batches)]
                    xd, yd, xad, yad = x.to(device), y.to(device), xa.to(device), ya.to(device)
                    if unrolled:
                        architect.step(xd, yd, xad, yad, lr, w_opt, unrolled=True)
                    else:        # the reference's local search: mixed-level first-order rule (FedNASTrainer.py:129-136)
                        architect.step_v2(xd, yd, xad, yad, lam_train, lam_valid)
                w_opt.zero_grad(set_to_none=True)
                out = m(x.to(device))
                if isinstance(out, tuple):                   # derived network: (logits, auxiliary logits or None)
                    loss = crit(out[0], y.to(device))
                    if out[1] is not None:
                        loss = loss + aux_w * crit(out[1], y.to(device))
                else:
                    loss = crit(out, y.to(device))
                loss.backward()
                nn.utils.clip_grad_norm_(weights, float(getattr(args, "grad_clip", 5.0)))
                w_opt.step()

    def test(self, test_data, device, args):
        m = self.model.to(device)
        m.eval()
        c = l = n = 0.0
        with torch.no_grad():
            for x, y in test_data:
                o = m(x.to(device))
                o = o[0] if isinstance(o, tuple) else o
                l += float(nn.functional.cross_entropy(o, y.to(device), reduction="sum"))
                c += float((o.argmax(1).cpu() == y).sum())
                n += len(y)
        return {"test_correct": c, "test_loss": l, "test_total": n, "genotype": str(m.genotype()[0]) if hasattr(m, "genotype") else ""}


def _search_kwargs(args, class_num):
    """Search-network shape from the run arguments (reference defaults: 16 channels, 8 layers, 4 steps; the defaults here are the
    small configuration the CI-sized runs use)."""
    return dict(C=int(getattr(args, "init_channels", 8)), num_classes=class_num, layers=int(getattr(args, "layers", 3)),
                steps=int(getattr(args, "nas_steps", 2)), multiplier=int(getattr(args, "nas_multiplier", 2)))


def FedML_FedNAS_distributed(args, process_id, worker_number, comm, device, dataset, model=None, model_trainer=None,  # noqa: N802
                             backend="INPROC", **kw):
    if str(getattr(args, "stage", "search")) == "train":
        from ...models import darts as D
        arch = getattr(args, "arch", "FedNAS_V1")
        genotype = arch if isinstance(arch, D.Genotype) else getattr(D, str(arch))
        if not isinstance(model, D.NetworkCIFAR):
            model = D.NetworkCIFAR(int(getattr(args, "init_channels", 16)), dataset[7], int(getattr(args, "layers", 8)),
                                   bool(getattr(args, "auxiliary", True)), genotype)
        return FedML_FedAvg_distributed(args, process_id, worker_number, comm, device, dataset, model,
                                        model_trainer or FedNASTrainer(model, args), backend=backend, **kw)
    if not isinstance(model, Network) and str(getattr(args, "nas_space", "darts")).lower() == "gdas":
        from ...models.darts import NetworkGDAS
        model = NetworkGDAS(**_search_kwargs(args, dataset[7]))
    if not isinstance(model, Network):
        model = Network(**_search_kwargs(args, dataset[7]))
    out = FedML_FedAvg_distributed(args, process_id, worker_number, comm, device, dataset, model,
                                   model_trainer or FedNASTrainer(model, args), backend=backend, **kw)
    if process_id == 0:
        out.genotype, out.normal_cnn_count, out.reduce_cnn_count = out.trainer.model.genotype()
    return out
