"""DARTS search space for FedNAS (reference: model/cv/darts/*.py, 2507 LoC: model_search.py, operations.py,
genotypes.py, architect.py).  A compact but complete differentiable-NAS network: mixed operations over the
standard primitive set, normal/reduction cells with continuous architecture parameters ``alphas_normal`` /
``alphas_reduce`` (exposed through ``arch_parameters()``), ``genotype()`` decoding by the top-2 strongest
non-``none`` edges per node."""
from collections import namedtuple

import torch
import torch.nn as nn
import torch.nn.functional as F

Genotype = namedtuple("Genotype", "normal normal_concat reduce reduce_concat")
PRIMITIVES = ["none", "max_pool_3x3", "avg_pool_3x3", "skip_connect", "sep_conv_3x3", "sep_conv_5x5", "dil_conv_3x3",
              "dil_conv_5x5"]


class _Zero(nn.Module):
    def __init__(self, stride):
        super().__init__()
        self.stride = stride

    def forward(self, x):
        return x.mul(0.0) if self.stride == 1 else x[:, :, ::self.stride, ::self.stride].mul(0.0)


class _FactorizedReduce(nn.Module):
    def __init__(self, ci, co):
        super().__init__()
        self.c1 = nn.Conv2d(ci, co // 2, 1, 2, 0, bias=False)
        self.c2 = nn.Conv2d(ci, co - co // 2, 1, 2, 0, bias=False)
        self.bn = nn.BatchNorm2d(co, affine=False)

    def forward(self, x):
        x = F.relu(x)
        shifted = F.pad(x, (0, 1, 0, 1))[:, :, 1:, 1:]
        return self.bn(torch.cat([self.c1(x), self.c2(shifted)], 1))


def _sep(ci, co, k, s, p):
    return nn.Sequential(nn.ReLU(False), nn.Conv2d(ci, ci, k, s, p, groups=ci, bias=False), nn.Conv2d(ci, ci, 1, bias=False),
                         nn.BatchNorm2d(ci, affine=False), nn.ReLU(False), nn.Conv2d(ci, ci, k, 1, p, groups=ci, bias=False),
                         nn.Conv2d(ci, co, 1, bias=False), nn.BatchNorm2d(co, affine=False))


def _dil(ci, co, k, s, p, d):
    return nn.Sequential(nn.ReLU(False), nn.Conv2d(ci, ci, k, s, p, dilation=d, groups=ci, bias=False),
                         nn.Conv2d(ci, co, 1, bias=False), nn.BatchNorm2d(co, affine=False))


OPS = {
    "none": lambda C, s: _Zero(s),
    "avg_pool_3x3": lambda C, s: nn.Sequential(nn.AvgPool2d(3, s, 1, count_include_pad=False), nn.BatchNorm2d(C, affine=False)),
    "max_pool_3x3": lambda C, s: nn.Sequential(nn.MaxPool2d(3, s, 1), nn.BatchNorm2d(C, affine=False)),
    "skip_connect": lambda C, s: nn.Identity() if s == 1 else _FactorizedReduce(C, C),
    "sep_conv_3x3": lambda C, s: _sep(C, C, 3, s, 1),
    "sep_conv_5x5": lambda C, s: _sep(C, C, 5, s, 2),
    "dil_conv_3x3": lambda C, s: _dil(C, C, 3, s, 2, 2),
    "dil_conv_5x5": lambda C, s: _dil(C, C, 5, s, 4, 2),
}


class MixedOp(nn.Module):
    def __init__(self, C, stride):
        super().__init__()
        self._ops = nn.ModuleList(OPS[p](C, stride) for p in PRIMITIVES)

    def forward(self, x, weights):
        return sum(w * op(x) for w, op in zip(weights, self._ops))


class Cell(nn.Module):
    def __init__(self, steps, multiplier, Cpp, Cp, C, reduction, reduction_prev):
        super().__init__()
        self.reduction, self._steps, self._multiplier = reduction, steps, multiplier
        self.pre0 = _FactorizedReduce(Cpp, C) if reduction_prev else \
            nn.Sequential(nn.ReLU(False), nn.Conv2d(Cpp, C, 1, bias=False), nn.BatchNorm2d(C, affine=False))
        self.pre1 = nn.Sequential(nn.ReLU(False), nn.Conv2d(Cp, C, 1, bias=False), nn.BatchNorm2d(C, affine=False))
        self._ops = nn.ModuleList()
        for i in range(steps):
            for j in range(2 + i):
                self._ops.append(MixedOp(C, 2 if reduction and j < 2 else 1))

    def forward(self, s0, s1, weights):
        states, off = [self.pre0(s0), self.pre1(s1)], 0
        for _ in range(self._steps):
            s = sum(self._ops[off + j](h, weights[off + j]) for j, h in enumerate(states))
            off += len(states)
            states.append(s)
        return torch.cat(states[-self._multiplier:], dim=1)


class Network(nn.Module):
    """DARTS search network (``model_search.Network``)."""

    def __init__(self, C=8, num_classes=10, layers=3, steps=2, multiplier=2, stem_multiplier=3):
        super().__init__()
        self._steps, self._multiplier = steps, multiplier
        Cc = stem_multiplier * C
        self.stem = nn.Sequential(nn.Conv2d(3, Cc, 3, padding=1, bias=False), nn.BatchNorm2d(Cc))
        Cpp, Cp, Cc = Cc, Cc, C
        self.cells, red_prev = nn.ModuleList(), False
        for i in range(layers):
            red = layers >= 3 and i in (layers // 3, 2 * layers // 3)
            if red:
                Cc *= 2
            self.cells.append(Cell(steps, multiplier, Cpp, Cp, Cc, red, red_prev))
            red_prev = red
            Cpp, Cp = Cp, multiplier * Cc
        self.classifier = nn.Linear(Cp, num_classes)
        k = sum(2 + i for i in range(steps))
        self.alphas_normal = nn.Parameter(1e-3 * torch.randn(k, len(PRIMITIVES)))
        self.alphas_reduce = nn.Parameter(1e-3 * torch.randn(k, len(PRIMITIVES)))

    def arch_parameters(self):
        return [self.alphas_normal, self.alphas_reduce]

    def weight_parameters(self):
        ids = {id(p) for p in self.arch_parameters()}
        return [p for p in self.parameters() if id(p) not in ids]

    def forward(self, x):
        s0 = s1 = self.stem(x)
        for cell in self.cells:
            w = F.softmax(self.alphas_reduce if cell.reduction else self.alphas_normal, dim=-1)
            s0, s1 = s1, cell(s0, s1, w)
        return self.classifier(F.adaptive_avg_pool2d(s1, 1).flatten(1))

    def genotype(self):
        none = PRIMITIVES.index("none")

        def parse(weights):
            gene, n, start = [], 2, 0
            for i in range(self._steps):
                end = start + n
                W = weights[start:end]
                edges = sorted(range(i + 2), key=lambda x: -max(W[x][k] for k in range(len(W[x])) if k != none))[:2]
                for j in edges:
                    k_best = max((k for k in range(len(W[j])) if k != none), key=lambda k: W[j][k])
                    gene.append((PRIMITIVES[k_best], j))
                start, n = end, n + 1
            return gene

        concat = list(range(2 + self._steps - self._multiplier, self._steps + 2))
        return Genotype(parse(F.softmax(self.alphas_normal, -1).tolist()), concat,
                        parse(F.softmax(self.alphas_reduce, -1).tolist()), concat)


class NetworkGDAS(Network):
    """GDAS search space (reference: model/cv/darts/model_search_gdas.py): every edge samples ONE operation per forward
    pass with the Gumbel-softmax straight-through estimator (temperature ``tau``, annealed by the trainer via
    ``set_tau``), so a step costs one operation per edge instead of all eight, while the architecture parameters still get
    gradients through the soft probabilities."""

    def __init__(self, *a, tau: float = 10.0, **kw):
        super().__init__(*a, **kw)
        self.tau = float(tau)

    def set_tau(self, tau: float):
        self.tau = float(tau)

    def get_tau(self) -> float:
        return self.tau

    def _edge_weights(self, alphas):
        if not self.training:
            idx = alphas.argmax(dim=-1, keepdim=True)
            return torch.zeros_like(alphas).scatter_(-1, idx, 1.0)
        return F.gumbel_softmax(alphas, tau=self.tau, hard=True, dim=-1)     # one-hot forward, soft backward

    def forward(self, x):
        s0 = s1 = self.stem(x)
        wn, wr = self._edge_weights(self.alphas_normal), self._edge_weights(self.alphas_reduce)
        for cell in self.cells:
            s0, s1 = s1, _gdas_cell(cell, s0, s1, wr if cell.reduction else wn)
        return self.classifier(F.adaptive_avg_pool2d(s1, 1).flatten(1))


def _gdas_cell(cell: Cell, s0, s1, weights):
    """Cell forward that evaluates only the sampled operation of every edge."""
    states, off = [cell.pre0(s0), cell.pre1(s1)], 0
    for _ in range(cell._steps):
        acc = 0
        for j, h in enumerate(states):
            w = weights[off + j]
            k = int(w.argmax())
            acc = acc + w[k] * cell._ops[off + j]._ops[k](h)
        off += len(states)
        states.append(acc)
    return torch.cat(states[-cell._multiplier:], dim=1)


class Architect:
    """Architecture step of DARTS (reference: model/cv/darts/architect.py).  ``step(..., unrolled=False)`` is the
    first-order rule (gradient of the validation loss at the current weights); ``unrolled=True`` is the second-order rule:
    with w' = w - eta * (momentum * buf + grad_w L_train(w, a) + wd * w),

        d a = grad_a L_val(w', a) - eta * [grad_a L_train(w + e v, a) - grad_a L_train(w - e v, a)] / (2 e),
        v = grad_w' L_val(w', a),  e = 0.01 / ||v||

    (the finite-difference Hessian-vector product of the DARTS paper).  Implemented with ``torch.func.functional_call`` on
    a parameter dict, so no second model instance is built."""

    def __init__(self, model: Network, criterion=None, arch_learning_rate=3e-4, arch_weight_decay=1e-3, momentum=0.9,
                 weight_decay=3e-4):
        self.model, self.criterion = model, criterion or nn.CrossEntropyLoss()
        self.momentum, self.weight_decay = float(momentum), float(weight_decay)
        self.optimizer = torch.optim.Adam(model.arch_parameters(), lr=arch_learning_rate, betas=(0.5, 0.999),
                                          weight_decay=arch_weight_decay)

    # -- helpers -----------------------------------------------------------------------------------------------------
    def _names(self):
        arch_ids = {id(p) for p in self.model.arch_parameters()}
        wn = [n for n, p in self.model.named_parameters() if id(p) not in arch_ids]
        an = [n for n, p in self.model.named_parameters() if id(p) in arch_ids]
        return wn, an

    def _loss(self, params: dict, x, y):
        from torch.func import functional_call
        buffers = {n: b for n, b in self.model.named_buffers()}
        return self.criterion(functional_call(self.model, {**params, **buffers}, (x,)), y)

    def arch_gradient(self, x_train, y_train, x_valid, y_valid, eta: float, network_optimizer=None, unrolled: bool = True):
        """-> list of gradients, one per architecture parameter."""
        wn, an = self._names()
        named = dict(self.model.named_parameters())
        w = {n: named[n] for n in wn}
        a = {n: named[n] for n in an}
        if not unrolled:
            gs = torch.autograd.grad(self._loss({**w, **a}, x_valid, y_valid), list(a.values()), allow_unused=True)
            return [torch.zeros_like(p) if gi is None else gi for gi, p in zip(gs, a.values())]
        # virtual step of the weights
        g = torch.autograd.grad(self._loss({**w, **a}, x_train, y_train), list(w.values()), allow_unused=True)
        w_prime = {}
        for (n, p), gi in zip(w.items(), g):
            gi = torch.zeros_like(p) if gi is None else gi
            buf = None
            if network_optimizer is not None:
                buf = network_optimizer.state.get(p, {}).get("momentum_buffer")
            mom = self.momentum * buf if buf is not None else 0.0
            w_prime[n] = (p - eta * (mom + gi + self.weight_decay * p)).detach().requires_grad_(True)
        a_leaf = {n: p.detach().requires_grad_(True) for n, p in a.items()}
        grads = torch.autograd.grad(self._loss({**w_prime, **a_leaf}, x_valid, y_valid),
                                    list(a_leaf.values()) + list(w_prime.values()), allow_unused=True)
        da = [gi if gi is not None else torch.zeros_like(p) for gi, p in zip(grads[:len(a_leaf)], a_leaf.values())]
        v = [gi if gi is not None else torch.zeros_like(p) for gi, p in zip(grads[len(a_leaf):], w_prime.values())]
        vnorm = torch.sqrt(sum((vi * vi).sum() for vi in v))
        eps = 0.01 / float(vnorm.clamp_min(1e-12))
        hv = []
        for sign in (1.0, -1.0):
            w_shift = {n: (p + sign * eps * vi).detach() for (n, p), vi in zip(w.items(), v)}
            a_req = {n: p.detach().requires_grad_(True) for n, p in a.items()}
            hv.append(torch.autograd.grad(self._loss({**w_shift, **a_req}, x_train, y_train), list(a_req.values()), allow_unused=True))
        out = []
        for d, gp, gm, p in zip(da, hv[0], hv[1], a.values()):
            gp = torch.zeros_like(p) if gp is None else gp
            gm = torch.zeros_like(p) if gm is None else gm
            out.append(d - eta * (gp - gm) / (2 * eps))
        return out

    def step(self, x_train, y_train, x_valid, y_valid, eta: float, network_optimizer=None, unrolled: bool = False):
        grads = self.arch_gradient(x_train, y_train, x_valid, y_valid, eta, network_optimizer, unrolled)
        self.optimizer.zero_grad(set_to_none=True)
        for p, gi in zip(self.model.arch_parameters(), grads):
            p.grad = gi.detach().clone()
        self.optimizer.step()


# ============================================================================ evaluation network of a found architecture
# Published cells as (operation, input state) pairs per node: DARTS first / second order (Liu et al. 2019) and the cell the
# FedNAS paper reports (reference model/cv/darts/genotypes.py:74-148).
def _genotype(normal: str, reduce: str) -> Genotype:
    parse = lambda s: [(op, int(j)) for op, j in (tok.split("@") for tok in s.split())]      # noqa: E731
    return Genotype(parse(normal), [2, 3, 4, 5], parse(reduce), [2, 3, 4, 5])


DARTS_V1 = _genotype("sep_conv_3x3@1 sep_conv_3x3@0 skip_connect@0 sep_conv_3x3@1 skip_connect@0 sep_conv_3x3@1 sep_conv_3x3@0 skip_connect@2",
                     "max_pool_3x3@0 max_pool_3x3@1 skip_connect@2 max_pool_3x3@0 max_pool_3x3@0 skip_connect@2 skip_connect@2 avg_pool_3x3@0")
DARTS_V2 = _genotype("sep_conv_3x3@0 sep_conv_3x3@1 sep_conv_3x3@0 sep_conv_3x3@1 sep_conv_3x3@1 skip_connect@0 skip_connect@0 dil_conv_3x3@2",
                     "max_pool_3x3@0 max_pool_3x3@1 skip_connect@2 max_pool_3x3@1 max_pool_3x3@0 skip_connect@2 skip_connect@2 max_pool_3x3@1")
DARTS = DARTS_V2
FedNAS_V1 = _genotype("sep_conv_3x3@1 sep_conv_3x3@0 sep_conv_3x3@2 sep_conv_5x5@0 sep_conv_3x3@1 sep_conv_5x5@3 dil_conv_5x5@3 sep_conv_3x3@4",
                      "max_pool_3x3@0 skip_connect@1 max_pool_3x3@0 max_pool_3x3@2 max_pool_3x3@0 dil_conv_5x5@1 max_pool_3x3@0 dil_conv_5x5@2")


def drop_path(x, drop_prob: float):
    """Zeroes whole samples of a branch with probability ``drop_prob`` and rescales the rest (training only)."""
    if drop_prob <= 0.0:
        return x
    keep = 1.0 - drop_prob
    mask = torch.empty(x.size(0), 1, 1, 1, device=x.device, dtype=x.dtype).bernoulli_(keep)
    return x / keep * mask


def _eval_op(name, C, stride):
    """Operation of a derived cell: affine BatchNorm (the search network uses affine=False so that the architecture weights
    carry the scale)."""
    op = OPS[name](C, stride)
    for m in op.modules():
        if isinstance(m, nn.BatchNorm2d) and not m.affine:
            m.affine = True
            m.weight, m.bias = nn.Parameter(torch.ones(m.num_features)), nn.Parameter(torch.zeros(m.num_features))
    return op


class DerivedCell(nn.Module):
    def __init__(self, genotype: Genotype, Cpp, Cp, C, reduction, reduction_prev):
        super().__init__()
        self.reduction = reduction
        relu_conv_bn = lambda ci: nn.Sequential(nn.ReLU(False), nn.Conv2d(ci, C, 1, bias=False), nn.BatchNorm2d(C))      # noqa: E731
        self.pre0 = _FactorizedReduce(Cpp, C) if reduction_prev else relu_conv_bn(Cpp)
        self.pre1 = relu_conv_bn(Cp)
        genes, self._concat = (genotype.reduce, list(genotype.reduce_concat)) if reduction else (genotype.normal, list(genotype.normal_concat))
        self._indices = [j for _, j in genes]
        self._ops = nn.ModuleList(_eval_op(name, C, 2 if reduction and j < 2 else 1) for name, j in genes)
        self.multiplier = len(self._concat)

    def forward(self, s0, s1, drop_prob=0.0):
        states = [self.pre0(s0), self.pre1(s1)]
        for i in range(len(self._ops) // 2):
            hs = []
            for k in (2 * i, 2 * i + 1):
                h = self._ops[k](states[self._indices[k]])
                if self.training and drop_prob > 0.0 and not isinstance(self._ops[k], nn.Identity):
                    h = drop_path(h, drop_prob)
                hs.append(h)
            states.append(hs[0] + hs[1])
        return torch.cat([states[i] for i in self._concat], dim=1)


class AuxiliaryHeadCIFAR(nn.Module):
    """Auxiliary classifier on the 8x8 feature map at two thirds of the depth (training-time regulariser)."""

    def __init__(self, C, num_classes):
        super().__init__()
        self.features = nn.Sequential(nn.ReLU(True), nn.AvgPool2d(5, stride=3, padding=0, count_include_pad=False),
                                      nn.Conv2d(C, 128, 1, bias=False), nn.BatchNorm2d(128), nn.ReLU(True),
                                      nn.Conv2d(128, 768, 2, bias=False), nn.BatchNorm2d(768), nn.ReLU(True))
        self.classifier = nn.Linear(768, num_classes)

    def forward(self, x):
        return self.classifier(self.features(x).flatten(1))


class NetworkCIFAR(nn.Module):
    """The network that is trained from scratch after the search (reference model/cv/darts/model.py:111-176): ``layers`` derived
    cells of ``genotype`` with reductions at 1/3 and 2/3 of the depth; returns ``(logits, aux_logits)``."""

    def __init__(self, C=36, num_classes=10, layers=20, auxiliary=False, genotype: Genotype = DARTS_V2):
        super().__init__()
        self._layers, self._auxiliary, self.drop_path_prob = layers, auxiliary, 0.0
        Cc = 3 * C
        self.stem = nn.Sequential(nn.Conv2d(3, Cc, 3, padding=1, bias=False), nn.BatchNorm2d(Cc))
        Cpp, Cp, Cc = Cc, Cc, C
        self.cells = nn.ModuleList()
        red_prev, C_aux = False, None
        for i in range(layers):
            red = i in (layers // 3, 2 * layers // 3)
            if red:
                Cc *= 2
            cell = DerivedCell(genotype, Cpp, Cp, Cc, red, red_prev)
            self.cells.append(cell)
            red_prev = red
            Cpp, Cp = Cp, cell.multiplier * Cc
            if i == 2 * layers // 3:
                C_aux = Cp
        self.auxiliary_head = AuxiliaryHeadCIFAR(C_aux, num_classes) if auxiliary else None
        self.global_pooling = nn.AdaptiveAvgPool2d(1)
        self.classifier = nn.Linear(Cp, num_classes)

    def forward(self, x):
        aux = None
        s0 = s1 = self.stem(x)
        for i, cell in enumerate(self.cells):
            s0, s1 = s1, cell(s0, s1, self.drop_path_prob)
            if i == 2 * self._layers // 3 and self.auxiliary_head is not None and self.training:
                aux = self.auxiliary_head(s1)
        return self.classifier(self.global_pooling(s1).flatten(1)), aux


def count_parameters_in_MB(model) -> float:      # noqa: N802  (the reference's name)
    return sum(p.numel() for n, p in model.named_parameters() if "auxiliary" not in n) / 1e6
