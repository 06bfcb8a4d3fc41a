"""DARTS / GDAS search space, derived networks and genotypes for FedNAS, with the reference's parameter tree.

Reference: model/cv/darts/{operations,genotypes,model_search,model_search_gdas,model,architect,utils}.py.  A FedNAS run exchanges
``state_dict``s of the search network (weights and ``alphas_*``) between server and clients, and the derived network is trained
from a ``Genotype``; so module names (``cells.3._ops.5._ops.4.op.1.weight``, ``preprocess0``, ``conv_1`` ...), constructor
signatures and ``genotype()``'s return value follow the reference, and ``tests/test_model_zoo.py`` loads the reference's files
next to these and compares outputs.  What is different: one ``Network`` serves DARTS and GDAS (sub-class), the architecture step
(``Architect``) is functional (no model copy), the factorised reduction also accepts odd feature-map sizes.
"""
from collections import namedtuple

import torch
import torch.nn as nn
import torch.nn.functional as F

Genotype = namedtuple("Genotype", "normal normal_concat reduce reduce_concat")
PRIMITIVES = ["none", "max_pool_3x3", "avg_pool_3x3", "skip_connect", "sep_conv_3x3", "sep_conv_5x5", "dil_conv_3x3",
              "dil_conv_5x5"]


# ------------------------------------------------------------------------------------------------ operations
class ReLUConvBN(nn.Module):
    def __init__(self, C_in, C_out, kernel_size, stride, padding, affine=True):  # noqa: N803
        super().__init__()
        self.op = nn.Sequential(nn.ReLU(inplace=False), nn.Conv2d(C_in, C_out, kernel_size, stride=stride, padding=padding, bias=False),
                                nn.BatchNorm2d(C_out, affine=affine))

    def forward(self, x):
        return self.op(x)


class DilConv(nn.Module):
    """ReLU, dilated depthwise convolution, 1x1 convolution, BN."""

    def __init__(self, C_in, C_out, kernel_size, stride, padding, dilation, affine=True):  # noqa: N803
        super().__init__()
        self.op = nn.Sequential(
            nn.ReLU(inplace=False),
            nn.Conv2d(C_in, C_in, kernel_size, stride=stride, padding=padding, dilation=dilation, groups=C_in, bias=False),
            nn.Conv2d(C_in, C_out, 1, bias=False), nn.BatchNorm2d(C_out, affine=affine))

    def forward(self, x):
        return self.op(x)


class SepConv(nn.Module):
    """Two (ReLU, depthwise, pointwise, BN) rounds; only the first one strides."""

    def __init__(self, C_in, C_out, kernel_size, stride, padding, affine=True):  # noqa: N803
        super().__init__()
        dw = lambda s: nn.Conv2d(C_in, C_in, kernel_size, stride=s, padding=padding, groups=C_in, bias=False)      # noqa: E731
        self.op = nn.Sequential(nn.ReLU(inplace=False), dw(stride), nn.Conv2d(C_in, C_in, 1, bias=False), nn.BatchNorm2d(C_in, affine=affine),
                                nn.ReLU(inplace=False), dw(1), nn.Conv2d(C_in, C_out, 1, bias=False), nn.BatchNorm2d(C_out, affine=affine))

    def forward(self, x):
        return self.op(x)


class Identity(nn.Module):
    def forward(self, x):
        return x


class Zero(nn.Module):
    def __init__(self, stride):
        super().__init__()
        self.stride = stride

    def forward(self, x):
        return x.mul(0.0) if self.stride == 1 else x[:, :, ::self.stride, ::self.stride].mul(0.0)


class FactorizedReduce(nn.Module):
    """Halves the resolution with two 1x1 stride-2 convolutions on the even and the odd pixel grid, concatenated."""

    def __init__(self, C_in, C_out, affine=True):  # noqa: N803
        super().__init__()
        assert C_out % 2 == 0
        self.relu = nn.ReLU(inplace=False)
        self.conv_1 = nn.Conv2d(C_in, C_out // 2, 1, stride=2, padding=0, bias=False)
        self.conv_2 = nn.Conv2d(C_in, C_out // 2, 1, stride=2, padding=0, bias=False)
        self.bn = nn.BatchNorm2d(C_out, affine=affine)

    def forward(self, x):
        x = self.relu(x)
        odd = F.pad(x, (0, 1, 0, 1))[:, :, 1:, 1:] if (x.shape[2] % 2 or x.shape[3] % 2) else x[:, :, 1:, 1:]
        return self.bn(torch.cat([self.conv_1(x), self.conv_2(odd)], dim=1))


OPS = {
    "none": lambda C, stride, affine: Zero(stride),
    "avg_pool_3x3": lambda C, stride, affine: nn.AvgPool2d(3, stride=stride, padding=1, count_include_pad=False),
    "max_pool_3x3": lambda C, stride, affine: nn.MaxPool2d(3, stride=stride, padding=1),
    "skip_connect": lambda C, stride, affine: Identity() if stride == 1 else FactorizedReduce(C, C, affine=affine),
    "sep_conv_3x3": lambda C, stride, affine: SepConv(C, C, 3, stride, 1, affine=affine),
    "sep_conv_5x5": lambda C, stride, affine: SepConv(C, C, 5, stride, 2, affine=affine),
    "sep_conv_7x7": lambda C, stride, affine: SepConv(C, C, 7, stride, 3, affine=affine),
    "dil_conv_3x3": lambda C, stride, affine: DilConv(C, C, 3, stride, 2, 2, affine=affine),
    "dil_conv_5x5": lambda C, stride, affine: DilConv(C, C, 5, stride, 4, 2, affine=affine),
    "conv_7x1_1x7": lambda C, stride, affine: nn.Sequential(
        nn.ReLU(inplace=False), nn.Conv2d(C, C, (1, 7), stride=(1, stride), padding=(0, 3), bias=False),
        nn.Conv2d(C, C, (7, 1), stride=(stride, 1), padding=(3, 0), bias=False), nn.BatchNorm2d(C, affine=affine)),
}


def _search_op(primitive, C, stride):
    """Operation of the search network: no affine BN (the architecture weights carry the scale); pooling gets a BN behind it."""
    op = OPS[primitive](C, stride, False)
    return nn.Sequential(op, nn.BatchNorm2d(C, affine=False)) if "pool" in primitive else op


# ------------------------------------------------------------------------------------------------ search network
class MixedOp(nn.Module):
    def __init__(self, C, stride):
        super().__init__()
        self._ops = nn.ModuleList(_search_op(p, C, stride) for p in PRIMITIVES)

    def forward(self, x, weights):
        return sum(w * op(x) for w, op in zip(weights, self._ops))


class _CellBase(nn.Module):
    def __init__(self, steps, multiplier, C_prev_prev, C_prev, C, reduction, reduction_prev):  # noqa: N803
        super().__init__()
        self.reduction, self._steps, self._multiplier = reduction, steps, multiplier
        self.preprocess0 = FactorizedReduce(C_prev_prev, C, affine=False) if reduction_prev else ReLUConvBN(C_prev_prev, C, 1, 1, 0, affine=False)
        self.preprocess1 = ReLUConvBN(C_prev, C, 1, 1, 0, affine=False)
        self._ops = nn.ModuleList()
        self._bns = nn.ModuleList()

    def _edges(self):
        """(edge index, input state index, stride) of every edge, node by node."""
        k = 0
        for i in range(self._steps):
            for j in range(2 + i):
                yield k, j, 2 if self.reduction and j < 2 else 1
                k += 1


class Cell(_CellBase):
    """Search cell: every edge is a :class:`MixedOp`; ``weights[k]`` are the (softmax or one-hot) operation weights of edge k."""

    def __init__(self, steps, multiplier, C_prev_prev, C_prev, C, reduction, reduction_prev):  # noqa: N803
        super().__init__(steps, multiplier, C_prev_prev, C_prev, C, reduction, reduction_prev)
        for _, _, stride in self._edges():
            self._ops.append(MixedOp(C, stride))

    def forward(self, s0, s1, weights):
        states, off = [self.preprocess0(s0), self.preprocess1(s1)], 0
        for _ in range(self._steps):
            s = sum(self._ops[off + j](h, weights[off + j]) for j, h in enumerate(states))
            off += len(states)
            states.append(s)
        return torch.cat(states[-self._multiplier:], dim=1)


class InnerCell(_CellBase):
    """Cell with the currently strongest operation on every edge (argmax over ``weights``, ``none`` included) - the discrete
    network whose size ``Network.get_current_model_size`` reports."""

    def __init__(self, steps, multiplier, C_prev_prev, C_prev, C, reduction, reduction_prev, weights):  # noqa: N803
        super().__init__(steps, multiplier, C_prev_prev, C_prev, C, reduction, reduction_prev)
        names = list(OPS.keys())
        for k, _, stride in self._edges():
            self._ops.append(_search_op(names[int(weights.data[k].argmax())], C, stride))

    def forward(self, s0, s1):
        states, off = [self.preprocess0(s0), self.preprocess1(s1)], 0
        for _ in range(self._steps):
            s = sum(self._ops[off + j](h) for j, h in enumerate(states))
            off += len(states)
            states.append(s)
        return torch.cat(states[-self._multiplier:], dim=1)


def _stack(make_cell, C, layers, multiplier, stem_multiplier):
    """Stem + ``layers`` cells with reductions (channels doubled) at 1/3 and 2/3 of the depth -> (stem, cells, C_out)."""
    c_curr = stem_multiplier * C
    stem = nn.Sequential(nn.Conv2d(3, c_curr, 3, padding=1, bias=False), nn.BatchNorm2d(c_curr))
    c_pp, c_p, c_curr = c_curr, c_curr, C
    cells, reduction_prev = nn.ModuleList(), False
    for i in range(layers):
        reduction = i in (layers // 3, 2 * layers // 3)
        if reduction:
            c_curr *= 2
        cells.append(make_cell(c_pp, c_p, c_curr, reduction, reduction_prev))
        reduction_prev = reduction
        c_pp, c_p = c_p, multiplier * c_curr
    return stem, cells, c_p


class ModelForModelSizeMeasure(nn.Module):
    def __init__(self, C, num_classes, layers, criterion, alphas_normal, alphas_reduce, steps=4, multiplier=4, stem_multiplier=3):  # noqa: N803
        super().__init__()
        mk = lambda pp, p, c, red, red_prev: InnerCell(steps, multiplier, pp, p, c, red, red_prev, alphas_reduce if red else alphas_normal)  # noqa: E731
        self.stem, self.cells, c_out = _stack(mk, C, layers, multiplier, stem_multiplier)
        self.global_pooling = nn.AdaptiveAvgPool2d(1)
        self.classifier = nn.Linear(c_out, num_classes)

    def forward(self, x):
        s0 = s1 = self.stem(x)
        for cell in self.cells:
            s0, s1 = s1, cell(s0, s1)
        return self.classifier(self.global_pooling(s1).flatten(1))


class Network(nn.Module):
    """DARTS search network: ``Network(C, num_classes, layers, criterion, device, steps=4, multiplier=4, stem_multiplier=3)``."""

    def __init__(self, C=16, num_classes=10, layers=8, criterion=None, device=None, steps=4, multiplier=4, stem_multiplier=3):  # noqa: N803
        super().__init__()
        self._C, self._num_classes, self._layers = C, num_classes, layers
        self._criterion = criterion if criterion is not None else nn.CrossEntropyLoss()
        self._steps, self._multiplier, self._stem_multiplier = steps, multiplier, stem_multiplier
        self.device = device if device is not None else torch.device("cpu")
        mk = lambda pp, p, c, red, red_prev: Cell(steps, multiplier, pp, p, c, red, red_prev)      # noqa: E731
        self.stem, self.cells, c_out = _stack(mk, C, layers, multiplier, stem_multiplier)
        self.global_pooling = nn.AdaptiveAvgPool2d(1)
        self.classifier = nn.Linear(c_out, num_classes)
        self._initialize_alphas()

    # -- architecture parameters
    def _edge_count(self):
        return sum(2 + i for i in range(self._steps))

    def _initialize_alphas(self):
        k, n = self._edge_count(), len(PRIMITIVES)
        self.alphas_normal = nn.Parameter(1e-3 * torch.randn(k, n))
        self.alphas_reduce = nn.Parameter(1e-3 * torch.randn(k, n))
        self._arch_parameters = [self.alphas_normal, self.alphas_reduce]

    def arch_parameters(self):
        return [self.alphas_normal, self.alphas_reduce]

    def new_arch_parameters(self):
        k, n = self._edge_count(), len(PRIMITIVES)
        return [nn.Parameter(1e-3 * torch.randn(k, n, device=self.device)) for _ in range(2)]

    def weight_parameters(self):
        ids = {id(p) for p in self.arch_parameters()}
        return [p for p in self.parameters() if id(p) not in ids]

    def new(self):
        """A fresh network of the same shape carrying this one's architecture parameters."""
        other = type(self)(self._C, self._num_classes, self._layers, self._criterion, self.device, self._steps, self._multiplier,
                           self._stem_multiplier).to(self.device)
        for x, y in zip(other.arch_parameters(), self.arch_parameters()):
            x.data.copy_(y.data)
        return other

    # -- forward
    def _weights(self, alphas):
        return F.softmax(alphas, dim=-1)

    def _run_cell(self, cell, s0, s1, weights):
        return cell(s0, s1, weights)

    def forward(self, x):
        s0 = s1 = self.stem(x)
        wn, wr = self._weights(self.alphas_normal), self._weights(self.alphas_reduce)
        for cell in self.cells:
            s0, s1 = s1, self._run_cell(cell, s0, s1, wr if cell.reduction else wn)
        return self.classifier(self.global_pooling(s1).flatten(1))

    def _loss(self, x, target):
        return self._criterion(self(x), target)

    # -- decoding
    def genotype(self):
        """``(Genotype, parametric ops in the normal cell, parametric ops in the reduction cell)``: per node the two incoming
        edges with the strongest non-``none`` operation, each with that operation."""
        none = PRIMITIVES.index("none")

        def parse(weights):
            gene, conv_count, start = [], 0, 0
            for i in range(self._steps):
                rows = weights[start:start + i + 2]
                strength = lambda j: max(w for k, w in enumerate(rows[j]) if k != none)      # noqa: E731
                for j in sorted(range(i + 2), key=lambda j: -strength(j))[:2]:
                    k_best = max((k for k in range(len(rows[j])) if k != none), key=lambda k: (rows[j][k], -k))
                    conv_count += k_best >= 4
                    gene.append((PRIMITIVES[k_best], j))
                start += i + 2
            return gene, conv_count

        with torch.no_grad():
            normal, n_count = parse(F.softmax(self.alphas_normal, dim=-1).tolist())
            reduce, r_count = parse(F.softmax(self.alphas_reduce, dim=-1).tolist())
        concat = range(2 + self._steps - self._multiplier, self._steps + 2)
        return Genotype(normal=normal, normal_concat=concat, reduce=reduce, reduce_concat=concat), n_count, r_count

    def get_current_model_size(self) -> float:
        """Parameters (in millions, auxiliary heads excluded) of the discrete network the current alphas select."""
        return count_parameters_in_MB(ModelForModelSizeMeasure(self._C, self._num_classes, self._layers, self._criterion, self.alphas_normal,
                                                               self.alphas_reduce, self._steps, self._multiplier, self._stem_multiplier))


class NetworkGDAS(Network):
    """GDAS search space (reference: model/cv/darts/model_search_gdas.py ``Network_GumbelSoftmax``): every edge samples ONE
    operation per forward pass with the Gumbel-softmax straight-through estimator (temperature ``tau``, annealed by the
    trainer via ``set_tau``), so a step costs one operation per edge instead of all eight, while the architecture parameters
    still get gradients through the soft probabilities.  In eval mode the strongest operation is taken (deterministic)."""

    def __init__(self, *a, tau: float = 5.0, **kw):
        super().__init__(*a, **kw)
        self.tau = float(tau)

    def set_tau(self, tau: float):
        self.tau = float(tau)

    def get_tau(self) -> float:
        return self.tau

    def _weights(self, alphas):
        if not self.training:
            idx = alphas.argmax(dim=-1, keepdim=True)
            return torch.zeros_like(alphas).scatter_(-1, idx, 1.0)
        return F.gumbel_softmax(alphas, tau=self.tau, hard=True, dim=-1)     # one-hot forward, soft backward

    def _run_cell(self, cell, s0, s1, weights):
        """Cell forward that evaluates only the sampled operation of every edge."""
        states, off = [cell.preprocess0(s0), cell.preprocess1(s1)], 0
        for _ in range(cell._steps):
            acc = 0
            for j, h in enumerate(states):
                w = weights[off + j]
                k = int(w.argmax())
                acc = acc + w[k] * cell._ops[off + j]._ops[k](h)
            off += len(states)
            states.append(acc)
        return torch.cat(states[-cell._multiplier:], dim=1)


Network_GumbelSoftmax = NetworkGDAS      # noqa: N816  (the reference's class name)


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
        """Also accepts the reference's form ``Architect(model, criterion, args, device)``: ``args`` then supplies
        ``arch_learning_rate``, ``arch_weight_decay``, ``momentum`` and ``weight_decay``."""
        if hasattr(arch_learning_rate, "arch_learning_rate"):
            args = arch_learning_rate
            self.device = arch_weight_decay if isinstance(arch_weight_decay, (torch.device, str)) else None
            arch_learning_rate, arch_weight_decay = float(args.arch_learning_rate), float(getattr(args, "arch_weight_decay", 1e-3))
            momentum, weight_decay = float(getattr(args, "momentum", 0.9)), float(getattr(args, "weight_decay", 3e-4))
        self.model, self.criterion = model, criterion or nn.CrossEntropyLoss()
        self.momentum, self.weight_decay = float(momentum), float(weight_decay)
        self.network_momentum, self.network_weight_decay = self.momentum, self.weight_decay
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

    def _apply(self, grads):
        self.optimizer.zero_grad(set_to_none=True)
        for p, gi in zip(self.model.arch_parameters(), grads):
            p.grad = gi.detach().clone()
        self.optimizer.step()

    def step(self, x_train, y_train, x_valid, y_valid, eta: float, network_optimizer=None, unrolled: bool = False):
        self._apply(self.arch_gradient(x_train, y_train, x_valid, y_valid, eta, network_optimizer, unrolled))

    # -- mixed-level rules (MiLeNAS; reference architect.py:83-225, 364-455): the architecture follows the validation loss AND,
    #    weighted by lambda_train, the training loss, which removes the need for the second-order term in practice
    def _alpha_grad(self, x, y):
        gs = torch.autograd.grad(self.criterion(self.model(x), y), self.model.arch_parameters(), allow_unused=True)
        return [torch.zeros_like(p) if g is None else g for g, p in zip(gs, self.model.arch_parameters())]

    def step_v2(self, input_train, target_train, input_valid, target_valid, lambda_train_regularizer=1.0, lambda_valid_regularizer=1.0):
        """alpha <- Adam(grad_a L_val + lambda_train * grad_a L_train), both at the current weights.  ``lambda_valid`` is part of
        the reference's signature and, like there, does not enter the first-order rule."""
        g_train = self._alpha_grad(input_train, target_train)
        g_val = self._alpha_grad(input_valid, target_valid)
        self._apply([gv + lambda_train_regularizer * gt for gv, gt in zip(g_val, g_train)])

    def step_wa(self, input_train, target_train, input_valid, target_valid, lambda_regularizer=1.0):
        self.step_v2(input_train, target_train, input_valid, target_valid, lambda_regularizer)

    def step_single_level(self, input_train, target_train):
        """Single-level optimisation: the architecture follows the training loss only."""
        self._apply(self._alpha_grad(input_train, target_train))

    def step_AOS(self, input_train, target_train, input_valid, target_valid):  # noqa: N802
        """Alternating optimisation, first order: the architecture follows the validation loss only."""
        self._apply(self._alpha_grad(input_valid, target_valid))

    def step_v2_2ndorder(self, input_train, target_train, input_valid, target_valid, eta, network_optimizer,
                         lambda_train_regularizer=1.0, lambda_valid_regularizer=1.0):
        """Mixed-level rule with the unrolled weights: ``lambda_valid * d/da L_val(w', a) + lambda_train * d/da L_train(w', a)``
        where both terms include the finite-difference second-order correction of :meth:`arch_gradient`."""
        g_val = self.arch_gradient(input_train, target_train, input_valid, target_valid, eta, network_optimizer, unrolled=True)
        g_train = self.arch_gradient(input_train, target_train, input_train, target_train, eta, network_optimizer, unrolled=True)
        self._apply([lambda_valid_regularizer * gv + lambda_train_regularizer * gt for gv, gt in zip(g_val, g_train)])


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


def _genotype_n(normal: str, normal_concat, reduce: str, reduce_concat) -> Genotype:
    parse = lambda s: [(op, int(j)) for op, j in (tok.split("@") for tok in s.split())]      # noqa: E731
    return Genotype(parse(normal), list(normal_concat), parse(reduce), list(reduce_concat))


# five-node cells found by reinforcement learning (Zoph et al. 2018) and by evolution (Real et al. 2018), as DARTS lists them
NASNet = _genotype_n("sep_conv_5x5@1 sep_conv_3x3@0 sep_conv_5x5@0 sep_conv_3x3@0 avg_pool_3x3@1 skip_connect@0 avg_pool_3x3@0 "
                     "avg_pool_3x3@0 sep_conv_3x3@1 skip_connect@1", [2, 3, 4, 5, 6],
                     "sep_conv_5x5@1 sep_conv_7x7@0 max_pool_3x3@1 sep_conv_7x7@0 avg_pool_3x3@1 sep_conv_5x5@0 skip_connect@3 "
                     "avg_pool_3x3@2 sep_conv_3x3@2 max_pool_3x3@1", [4, 5, 6])
AmoebaNet = _genotype_n("avg_pool_3x3@0 max_pool_3x3@1 sep_conv_3x3@0 sep_conv_5x5@2 sep_conv_3x3@0 avg_pool_3x3@3 sep_conv_3x3@1 "
                        "skip_connect@1 skip_connect@0 avg_pool_3x3@1", [4, 5, 6],
                        "avg_pool_3x3@0 sep_conv_3x3@1 max_pool_3x3@0 sep_conv_7x7@2 sep_conv_7x7@0 avg_pool_3x3@1 max_pool_3x3@0 "
                        "max_pool_3x3@1 conv_7x1_1x7@0 sep_conv_3x3@5", [3, 4, 6])
FedNAS_V1 = _genotype("sep_conv_3x3@1 sep_conv_3x3@0 sep_conv_3x3@2 sep_conv_5x5@0 sep_conv_3x3@1 sep_conv_5x5@3 dil_conv_5x5@3 sep_conv_3x3@4",
                      "max_pool_3x3@0 skip_connect@1 max_pool_3x3@0 max_pool_3x3@2 max_pool_3x3@0 dil_conv_5x5@1 max_pool_3x3@0 dil_conv_5x5@2")


def drop_path(x, drop_prob: float):
    """Zeroes whole samples of a branch with probability ``drop_prob`` and rescales the rest (training only)."""
    if drop_prob <= 0.0:
        return x
    keep = 1.0 - drop_prob
    mask = torch.empty(x.size(0), 1, 1, 1, device=x.device, dtype=x.dtype).bernoulli_(keep)
    return x / keep * mask


class DerivedCell(nn.Module):
    def __init__(self, genotype: Genotype, Cpp, Cp, C, reduction, reduction_prev):
        super().__init__()
        self.reduction = reduction
        self.preprocess0 = FactorizedReduce(Cpp, C) if reduction_prev else ReLUConvBN(Cpp, C, 1, 1, 0)
        self.preprocess1 = ReLUConvBN(Cp, C, 1, 1, 0)
        genes, self._concat = (genotype.reduce, list(genotype.reduce_concat)) if reduction else (genotype.normal, list(genotype.normal_concat))
        self._indices = [j for _, j in genes]
        self._ops = nn.ModuleList(OPS[name](C, 2 if reduction and j < 2 else 1, True) for name, j in genes)
        self._steps, self.multiplier = len(genes) // 2, len(self._concat)

    def forward(self, s0, s1, drop_prob=0.0):
        states = [self.preprocess0(s0), self.preprocess1(s1)]
        for i in range(self._steps):
            hs = []
            for k in (2 * i, 2 * i + 1):
                h = self._ops[k](states[self._indices[k]])
                if self.training and drop_prob > 0.0 and not isinstance(self._ops[k], Identity):
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
        if auxiliary:
            self.auxiliary_head = AuxiliaryHeadCIFAR(C_aux, num_classes)
        self.global_pooling = nn.AdaptiveAvgPool2d(1)
        self.classifier = nn.Linear(Cp, num_classes)

    def forward(self, x):
        aux = None
        s0 = s1 = self.stem(x)
        for i, cell in enumerate(self.cells):
            s0, s1 = s1, cell(s0, s1, self.drop_path_prob)
            if i == 2 * self._layers // 3 and self._auxiliary and self.training:
                aux = self.auxiliary_head(s1)
        return self.classifier(self.global_pooling(s1).flatten(1)), aux


class AuxiliaryHeadImageNet(nn.Module):
    """Auxiliary classifier for 14x14 feature maps (the second convolution has no BN, as published)."""

    def __init__(self, C, num_classes):
        super().__init__()
        self.features = nn.Sequential(nn.ReLU(True), nn.AvgPool2d(5, stride=2, padding=0, count_include_pad=False),
                                      nn.Conv2d(C, 128, 1, bias=False), nn.BatchNorm2d(128), nn.ReLU(True),
                                      nn.Conv2d(128, 768, 2, bias=False), nn.ReLU(True))
        self.classifier = nn.Linear(768, num_classes)

    def forward(self, x):
        return self.classifier(self.features(x).flatten(1))


class NetworkImageNet(nn.Module):
    """Derived network for 224x224 inputs (reference model/cv/darts/model.py:160-216): a three-convolution stem that reduces the
    resolution eight-fold (``stem0`` 4x, ``stem1`` 2x), then the cells; 7x7 average pool."""

    def __init__(self, C, num_classes, layers, auxiliary, genotype):
        super().__init__()
        self._layers, self._auxiliary, self.drop_path_prob = layers, auxiliary, 0.5
        self.stem0 = nn.Sequential(nn.Conv2d(3, C // 2, 3, stride=2, padding=1, bias=False), nn.BatchNorm2d(C // 2), nn.ReLU(True),
                                   nn.Conv2d(C // 2, C, 3, stride=2, padding=1, bias=False), nn.BatchNorm2d(C))
        self.stem1 = nn.Sequential(nn.ReLU(True), nn.Conv2d(C, C, 3, stride=2, padding=1, bias=False), nn.BatchNorm2d(C))
        Cpp, Cp, Cc = C, C, C
        self.cells = nn.ModuleList()
        red_prev, C_aux = True, None
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
        if auxiliary:
            self.auxiliary_head = AuxiliaryHeadImageNet(C_aux, num_classes)
        self.global_pooling = nn.AvgPool2d(7)
        self.classifier = nn.Linear(Cp, num_classes)

    def forward(self, x):
        aux = None
        s0 = self.stem0(x)
        s1 = self.stem1(s0)
        for i, cell in enumerate(self.cells):
            s0, s1 = s1, cell(s0, s1, self.drop_path_prob)
            if i == 2 * self._layers // 3 and self._auxiliary and self.training:
                aux = self.auxiliary_head(s1)
        return self.classifier(self.global_pooling(s1).flatten(1)), aux


# ============================================================================ training-script utilities (darts/utils.py)
def count_parameters_in_MB(model) -> float:      # noqa: N802  (the reference's name)
    return sum(p.numel() for n, p in model.named_parameters() if "auxiliary" not in n) / 1e6


class AvgrageMeter:      # (sic) the reference's spelling
    """Running mean of a metric weighted by batch size."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.avg, self.sum, self.cnt = 0, 0, 0

    def update(self, val, n=1):
        self.sum += val * n
        self.cnt += n
        self.avg = self.sum / self.cnt


def accuracy(output, target, topk=(1,)):
    """Top-k accuracies in percent, one 0-dim tensor per k."""
    _, pred = output.topk(max(topk), dim=1, largest=True, sorted=True)
    hit = pred.eq(target.view(-1, 1))
    return [hit[:, :k].any(dim=1).float().sum() * (100.0 / target.size(0)) for k in topk]


def save(model, model_path):
    torch.save(model.state_dict(), model_path)


def load(model, model_path):
    model.load_state_dict(torch.load(model_path, map_location="cpu"))


def save_checkpoint(state, is_best, save):  # noqa: A002
    import os
    import shutil
    filename = os.path.join(save, "checkpoint.pth.tar")
    torch.save(state, filename)
    if is_best:
        shutil.copyfile(filename, os.path.join(save, "model_best.pth.tar"))


def create_exp_dir(path, scripts_to_save=None):
    import os
    import shutil
    os.makedirs(path, exist_ok=True)
    if scripts_to_save:
        os.makedirs(os.path.join(path, "scripts"), exist_ok=True)
        for script in scripts_to_save:
            shutil.copyfile(script, os.path.join(path, "scripts", os.path.basename(script)))


def plot(genotype, filename=None) -> str:
    """Graphviz DOT text of one cell (``genotype.normal`` or ``genotype.reduce``: a list of (operation, input) pairs): the two
    cell inputs ``c_{k-2}``, ``c_{k-1}``, one node per step with its two incoming operation edges, and ``c_{k}`` collecting
    the steps (reference: model/cv/darts/visualize.py).  Written to ``filename`` (+ ``.dot``) when given; rendering is left to
    the ``dot`` tool - the python graphviz binding is not required."""
    steps = len(genotype) // 2
    lines = ["digraph cell {", "  rankdir=LR;", '  node [style=filled, shape=rect, fontname="times"];',
             '  "c_{k-2}" [fillcolor=darkseagreen2];', '  "c_{k-1}" [fillcolor=darkseagreen2];']
    lines += [f'  "{i}" [fillcolor=lightblue];' for i in range(steps)]
    name = lambda j: "c_{k-2}" if j == 0 else "c_{k-1}" if j == 1 else str(j - 2)      # noqa: E731
    for i in range(steps):
        for op, j in genotype[2 * i:2 * i + 2]:
            lines.append(f'  "{name(j)}" -> "{i}" [label="{op}"];')
    lines.append('  "c_{k}" [fillcolor=palegoldenrod];')
    lines += [f'  "{i}" -> "c_{{k}}";' for i in range(steps)]
    lines.append("}")
    text = "\n".join(lines) + "\n"
    if filename:
        with open(filename if filename.endswith(".dot") else filename + ".dot", "w") as fh:
            fh.write(text)
    return text
