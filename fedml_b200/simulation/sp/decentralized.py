"""Decentralized online learning: DSGD and PushSum over a gossip topology
(reference: single_process/decentralized/*.py — client_dsgd.py:68-122, client_pushsum.py:72-156,
decentralized_fl_api.py:11-17).  Streaming logistic regression: every iteration each node takes a gradient
step on its next sample, then mixes with its in-neighbours  x_i <- W_ii x_i + sum_j W_ij x_j.  PushSum also
tracks the de-biasing weight ``omega`` (z = x / omega is the estimate) and may re-draw a directed topology
every step (``time_varying``).  Metric: average regret = running mean of the per-sample losses."""
import numpy as np
import torch

from ...core.distributed.topology import AsymmetricTopologyManager, SymmetricTopologyManager


def cal_regret(client_list, client_number, t):
    return sum(c.get_regret() for c in client_list) / (client_number * (t + 1))


class _Node:
    def __init__(self, model_dim, cid, stream_x, stream_y, lr, wd):
        self.id, self.x_stream, self.y_stream, self.lr, self.wd = cid, stream_x, stream_y, lr, wd
        self.w = torch.zeros(model_dim + 1)          # weights + bias, the "model_x" of the reference
        self.omega = 1.0
        self.loss_sum = 0.0

    def z(self):
        return self.w / self.omega

    def train_step(self, t):
        x, y = self.x_stream[t % len(self.x_stream)], self.y_stream[t % len(self.y_stream)]
        z = self.z()
        logit = (z[:-1] * x).sum() + z[-1]
        p = torch.sigmoid(logit)
        self.loss_sum += float(torch.nn.functional.binary_cross_entropy(p.clamp(1e-7, 1 - 1e-7), y))
        g = (p - y)
        grad = torch.cat([g * x, g.view(1)]) + self.wd * z
        self.w = self.w - self.lr * grad

    def get_regret(self):
        return self.loss_sum


class DecentralizedFLAPI:
    def __init__(self, args, device, dataset, model=None, *_, **__):
        self.args = args
        self.n = int(getattr(args, "client_num_in_total", 8))
        self.iters = int(getattr(args, "iteration_number", 200))
        self.mode = str(getattr(args, "mode", "DOL")).upper()          # DOL = DSGD, PUSHSUM
        self.time_varying = bool(getattr(args, "time_varying", False))
        self.b_symmetric = bool(getattr(args, "b_symmetric", self.mode != "PUSHSUM"))
        streams = self._streams(dataset)
        dim = streams[0][0].shape[1]
        lr, wd = float(getattr(args, "learning_rate", 0.1)), float(getattr(args, "weight_decay", 0.0))
        self.nodes = [_Node(dim, i, sx, sy, lr, wd) for i, (sx, sy) in enumerate(streams)]
        self.regret = []
        self._make_topology()

    def _streams(self, dataset):
        local = dataset[5] if isinstance(dataset, (list, tuple)) else None
        out = []
        for i in range(self.n):
            if local is not None and i in local and len(local[i]):
                xs = torch.cat([b[0].reshape(b[0].shape[0], -1).float() for b in local[i]])
                ys = torch.cat([(b[1] % 2).float() for b in local[i]])      # binary task: parity of the label
            else:
                g = torch.Generator().manual_seed(i)
                xs = torch.randn(256, 16, generator=g); ys = (xs[:, 0] > 0).float()
            out.append((xs, ys))
        return out

    def _make_topology(self):
        k_u = int(getattr(self.args, "topology_neighbors_num_undirected", 4))
        k_d = int(getattr(self.args, "topology_neighbors_num_directed", 2))
        tm = SymmetricTopologyManager(self.n, k_u) if self.b_symmetric else AsymmetricTopologyManager(self.n, k_u, k_d)
        tm.generate_topology()
        self.topology = tm

    def train(self):
        W = torch.as_tensor(np.asarray(self.topology.topology), dtype=torch.float32)
        for t in range(self.iters):
            for node in self.nodes:
                node.train_step(t)
            if self.time_varying and not self.b_symmetric:
                self._make_topology()
                W = torch.as_tensor(np.asarray(self.topology.topology), dtype=torch.float32)
            X = torch.stack([n.w for n in self.nodes])
            if self.mode == "PUSHSUM":
                # column-stochastic push: node j sends W[j, i] of its mass to i
                om = torch.tensor([n.omega for n in self.nodes])
                Xn, omn = W.t() @ X, W.t() @ om
                for i, n in enumerate(self.nodes):
                    n.w, n.omega = Xn[i], float(omn[i])
            else:
                Xn = W @ X
                for i, n in enumerate(self.nodes):
                    n.w = Xn[i]
            self.regret.append(cal_regret(self.nodes, self.n, t))
        return self.regret

    def consensus_distance(self):
        Z = torch.stack([n.z() for n in self.nodes])
        return float((Z - Z.mean(0)).norm(dim=1).max())
